from . import oracle  # noqa: F401
from .mlp import MLPSpec, build_mlp_variables, reference_mlp_spec  # noqa: F401
