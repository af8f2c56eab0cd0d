"""The model family of the reference: sigmoid MLPs declared through the variable store
(reference worker.py:46-57 ``build_net``).

``reference_mlp_spec()`` is 784 -> dense(100, sigmoid) -> dense(10) -> softmax, kernels
N(0,1), biases zero, variable names ``<scope>/dense/{kernel,bias}``, ``<scope>/dense_1/...``.
The "wide" benchmark family (784-8192-8192-10) is the same builder with more layers.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Tuple

from ..variables import VariableDef, VariableStore, random_normal_initializer, zeros_initializer


@dataclass(frozen=True)
class MLPSpec:
    dims: Tuple[int, ...]            # (in, hidden..., out)
    hidden_activation: str = "sigmoid"
    param_dtype: str = "float32"

    @property
    def num_layers(self) -> int:
        return len(self.dims) - 1

    @property
    def num_params(self) -> int:
        return sum(a * b + b for a, b in zip(self.dims[:-1], self.dims[1:]))

    def layer_shapes(self) -> List[Tuple[Tuple[int, int], Tuple[int]]]:
        return [((a, b), (b,)) for a, b in zip(self.dims[:-1], self.dims[1:])]


def reference_mlp_spec() -> MLPSpec:
    return MLPSpec((784, 100, 10))


def wide_mlp_spec(dtype: str = "bfloat16") -> MLPSpec:
    return MLPSpec((784, 8192, 8192, 10), param_dtype=dtype)


def spec_from_flags(hidden: str, in_dim: int = 784, out_dim: int = 10, dtype: str = "float32") -> MLPSpec:
    widths = [int(h) for h in str(hidden).split(",") if h.strip()]
    return MLPSpec((in_dim, *widths, out_dim), param_dtype=dtype)


def build_mlp_variables(store: VariableStore, spec: MLPSpec) -> List[VariableDef]:
    """Declare the layer variables in the current scope/device context, tf.layers.dense naming."""
    out: List[VariableDef] = []
    for (wshape, bshape) in spec.layer_shapes():
        layer = store.unique_layer_name("dense")
        with store.variable_scope(layer):
            out.append(store.get_variable("kernel", wshape, spec.param_dtype, random_normal_initializer()))
            out.append(store.get_variable("bias", bshape, spec.param_dtype, zeros_initializer()))
    return out
