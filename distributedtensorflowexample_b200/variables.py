"""Variable scopes, collections and ``get_vars`` (reference utils.py:3-8, worker.py:27-35;
TF1 tf.variable_scope / tf.get_variable / GraphKeys - SURVEY C18/T6).

A ``VariableStore`` is the graph-less equivalent of TF1's variable collections: variables
are declared by name inside nested scopes and device scopes, recorded in creation order,
and looked up by *regex prefix match on the name* exactly like ``tf.get_collection(key, scope)``.
Declaring a variable does not allocate it; allocation happens when the store is bound to a
parameter back-end (``parallel.ps``), which honours each variable's placement.
"""
from __future__ import annotations

import contextlib
import re
from dataclasses import dataclass
from typing import Callable, Dict, Iterator, List, Optional, Sequence, Tuple, Union

import numpy as np

from .placement import DeviceSpec, OpInfo

TRAINABLE_VARIABLES = "trainable_variables"
GLOBAL_VARIABLES = "variables"

_DTYPE_BYTES = {"float32": 4, "int32": 4, "bfloat16": 2, "float64": 8, "int64": 8}


@dataclass
class Initializer:
    kind: str                      # 'random_normal' | 'zeros' | 'constant'
    mean: float = 0.0
    stddev: float = 1.0
    value: float = 0.0

    def __call__(self, shape: Sequence[int], dtype: str, rng: np.random.Generator) -> np.ndarray:
        npdt = np.float32 if dtype in ("float32", "bfloat16") else np.dtype(dtype)
        if self.kind == "random_normal":
            return rng.normal(self.mean, self.stddev, size=tuple(shape)).astype(npdt)
        if self.kind == "zeros":
            return np.zeros(tuple(shape), dtype=npdt)
        if self.kind == "constant":
            return np.full(tuple(shape), self.value, dtype=npdt)
        raise ValueError(self.kind)


def random_normal_initializer(mean: float = 0.0, stddev: float = 1.0) -> Initializer:
    """TF1 tf.random_normal_initializer(): mean 0, **stddev 1.0** (reference worker.py:51,53)."""
    return Initializer("random_normal", mean=mean, stddev=stddev)


def zeros_initializer() -> Initializer:
    return Initializer("zeros")


def constant_initializer(value: float = 0.0) -> Initializer:
    return Initializer("constant", value=value)


@dataclass
class VariableDef:
    name: str
    shape: Tuple[int, ...]
    dtype: str
    trainable: bool
    initializer: Initializer
    device: DeviceSpec
    index: int                     # creation order
    collections: Tuple[str, ...] = ()

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n

    @property
    def nbytes(self) -> int:
        return self.numel * _DTYPE_BYTES[self.dtype]


DeviceLike = Union[str, DeviceSpec, Callable[[OpInfo], str], None]


class VariableStore:
    def __init__(self) -> None:
        self._vars: Dict[str, VariableDef] = {}
        self._order: List[str] = []
        self._scope: List[str] = []
        self._devices: List[DeviceLike] = []
        self._dense_counters: Dict[str, int] = {}

    # ---- scopes ---------------------------------------------------------------------
    @contextlib.contextmanager
    def variable_scope(self, name: str) -> Iterator[None]:
        self._scope.append(name)
        try:
            yield
        finally:
            self._scope.pop()

    @contextlib.contextmanager
    def device(self, dev: DeviceLike) -> Iterator[None]:
        """``tf.device``: a device string/spec or a device function (e.g. replica_device_setter)."""
        self._devices.append(dev)
        try:
            yield
        finally:
            self._devices.pop()

    @property
    def current_scope(self) -> str:
        return "/".join(self._scope)

    def _resolve_device(self, op: OpInfo) -> DeviceSpec:
        spec = DeviceSpec()
        for dev in self._devices:           # outer -> inner; inner fields override
            if dev is None:
                continue
            if callable(dev):
                s = dev(op)
                inner = DeviceSpec.from_string(s) if isinstance(s, str) else s
            elif isinstance(dev, str):
                inner = DeviceSpec.from_string(dev)
            else:
                inner = dev
            spec = spec.merge_from(inner)
        return spec

    def op_device(self, name: str = "op") -> DeviceSpec:
        """Where a non-variable op created here would be placed."""
        return self._resolve_device(OpInfo(name=name, is_variable=False))

    # ---- declaration ----------------------------------------------------------------
    def get_variable(self, name: str, shape: Sequence[int], dtype: str = "float32",
                     initializer: Optional[Initializer] = None, trainable: bool = True) -> VariableDef:
        full = f"{self.current_scope}/{name}" if self._scope else name
        if full in self._vars:
            raise ValueError(f"variable {full} already exists")
        shape = tuple(int(s) for s in shape)
        n = 1
        for s in shape:
            n *= s
        info = OpInfo(name=full, is_variable=True, nbytes=n * _DTYPE_BYTES[dtype])
        colls = (GLOBAL_VARIABLES, TRAINABLE_VARIABLES) if trainable else (GLOBAL_VARIABLES,)
        v = VariableDef(full, shape, dtype, trainable, initializer or zeros_initializer(),
                        self._resolve_device(info), len(self._order), colls)
        self._vars[full] = v
        self._order.append(full)
        return v

    def unique_layer_name(self, base: str) -> str:
        """tf.layers auto-naming: ``dense``, ``dense_1``, ... per enclosing scope."""
        key = f"{self.current_scope}//{base}"
        k = self._dense_counters.get(key, 0)
        self._dense_counters[key] = k + 1
        return base if k == 0 else f"{base}_{k}"

    # ---- lookup ---------------------------------------------------------------------
    def get_collection(self, key: str, scope: Optional[str] = None) -> List[VariableDef]:
        out = []
        pat = re.compile(scope) if scope else None
        for n in self._order:
            v = self._vars[n]
            if key in v.collections and (pat is None or pat.match(v.name)):
                out.append(v)
        return out

    def __getitem__(self, name: str) -> VariableDef:
        return self._vars[name]

    def __contains__(self, name: str) -> bool:
        return name in self._vars

    def all_variables(self) -> List[VariableDef]:
        return [self._vars[n] for n in self._order]


def get_vars(store: VariableStore, scope: str, trainable: bool = True) -> List[VariableDef]:
    """Reference ``get_vars(scope, trainable=True)`` (utils.py:3-8): creation-ordered variables whose
    name regex-matches ``scope`` at the start, from TRAINABLE_VARIABLES or GLOBAL_VARIABLES."""
    return store.get_collection(TRAINABLE_VARIABLES if trainable else GLOBAL_VARIABLES, scope)
