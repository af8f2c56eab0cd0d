"""numpy restatement of the reference model and its parameter-server update rule - the oracle
every kernel and every end-to-end path is tested against.

Model (reference worker.py:46-68): ``h_0 = x``; ``h_l = sigmoid(h_{l-1} W_l + b_l)`` for hidden
layers; ``logits = h W + b``; loss = mean softmax cross-entropy; SGD ``p -= lr * dL/dp``.
Kernels use TF's ``[in, out]`` layout.  Works in float64 or float32.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import numpy as np


def init_params(dims: Sequence[int], rng: np.random.Generator, dtype=np.float32) -> List[np.ndarray]:
    """N(0,1) kernels, zero biases (reference worker.py:50-53): [W1, b1, W2, b2, ...]."""
    params: List[np.ndarray] = []
    for fan_in, fan_out in zip(dims[:-1], dims[1:]):
        params.append(rng.normal(0.0, 1.0, size=(fan_in, fan_out)).astype(dtype))
        params.append(np.zeros((fan_out,), dtype=dtype))
    return params


def sigmoid(z: np.ndarray) -> np.ndarray:
    return 1.0 / (1.0 + np.exp(-z))


def forward(params: Sequence[np.ndarray], x: np.ndarray) -> Tuple[np.ndarray, List[np.ndarray]]:
    """Returns (logits, activations) with activations[0] = x and one entry per hidden layer."""
    acts = [x]
    n_layers = len(params) // 2
    h = x
    for l in range(n_layers):
        z = h @ params[2 * l] + params[2 * l + 1]
        if l < n_layers - 1:
            h = sigmoid(z)
            acts.append(h)
        else:
            return z, acts
    raise AssertionError


def softmax_xent(logits: np.ndarray, label_ids: np.ndarray) -> Tuple[float, np.ndarray, np.ndarray]:
    """(mean loss, dL/dlogits for the mean loss, per-row loss)."""
    m = logits.max(axis=1, keepdims=True)
    e = np.exp(logits - m)
    se = e.sum(axis=1, keepdims=True)
    lse = (m + np.log(se))[:, 0]
    rows = np.arange(logits.shape[0])
    rowloss = lse - logits[rows, label_ids]
    p = e / se
    dz = p.copy()
    dz[rows, label_ids] -= 1.0
    dz /= logits.shape[0]
    return float(rowloss.mean()), dz, rowloss


def accuracy(logits: np.ndarray, label_ids: np.ndarray) -> float:
    return float((logits.argmax(axis=1) == label_ids).mean())


def gradients(params: Sequence[np.ndarray], x: np.ndarray, label_ids: np.ndarray):
    """(loss, accuracy, grads in the order of params)."""
    logits, acts = forward(params, x)
    loss, dz, _ = softmax_xent(logits, label_ids)
    acc = accuracy(logits, label_ids)
    n_layers = len(params) // 2
    grads: List[np.ndarray] = [None] * len(params)  # type: ignore
    d = dz
    for l in reversed(range(n_layers)):
        h_in = acts[l]
        grads[2 * l] = h_in.T @ d
        grads[2 * l + 1] = d.sum(axis=0)
        if l > 0:
            dh = d @ params[2 * l].T
            d = dh * h_in * (1.0 - h_in)
    return loss, acc, grads


def sgd_step(params: List[np.ndarray], x: np.ndarray, label_ids: np.ndarray, lr: float):
    """One reference training step (pull is implicit: params are the shared copy). In place."""
    loss, acc, grads = gradients(params, x, label_ids)
    for p, g in zip(params, grads):
        p -= (lr * g).astype(p.dtype)
    return loss, acc


class AdamState:
    def __init__(self, params: Sequence[np.ndarray], beta1=0.9, beta2=0.999, eps=1e-8):
        self.m = [np.zeros_like(p) for p in params]
        self.v = [np.zeros_like(p) for p in params]
        self.t = 0
        self.beta1, self.beta2, self.eps = beta1, beta2, eps

    def apply(self, params: List[np.ndarray], grads: Sequence[np.ndarray], lr: float) -> None:
        self.t += 1
        lr_t = lr * np.sqrt(1.0 - self.beta2 ** self.t) / (1.0 - self.beta1 ** self.t)
        for p, g, m, v in zip(params, grads, self.m, self.v):
            m[...] = self.beta1 * m + (1 - self.beta1) * g
            v[...] = self.beta2 * v + (1 - self.beta2) * g * g
            p -= (lr_t * m / (np.sqrt(v) + self.eps)).astype(p.dtype)


def flops_per_step(dims: Sequence[int], batch: int) -> Dict[str, float]:
    fwd = sum(2.0 * batch * a * b for a, b in zip(dims[:-1], dims[1:]))
    # backward: dW for every layer, dX for every layer but the first
    bwd = fwd + sum(2.0 * batch * a * b for a, b in list(zip(dims[:-1], dims[1:]))[1:])
    return {"fwd": fwd, "step": fwd + bwd}
