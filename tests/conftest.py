import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

os.environ.setdefault("MASTER_ADDR", "127.0.0.1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run on the GPU box with `-m gpu`)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 GPUs on the box")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        have_gpu = torch.cuda.is_available()
        ngpu = torch.cuda.device_count() if have_gpu else 0
    except Exception:
        have_gpu, ngpu = False, 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device in this container")
    skip_multi = pytest.mark.skip(reason="needs >= 2 GPUs")
    for item in items:
        if "gpu" in item.keywords and not have_gpu:
            item.add_marker(skip_gpu)
        if "multigpu" in item.keywords and ngpu < 2:
            item.add_marker(skip_multi)


@pytest.fixture(scope="session")
def data_dir():
    from distributedtensorflowexample_b200.data import find_data_dir
    return find_data_dir()
