from .mnist import DataSet, Datasets, read_data_sets, read_idx, read_idx_python, dense_to_one_hot, synthetic_arrays, find_data_dir  # noqa: F401
