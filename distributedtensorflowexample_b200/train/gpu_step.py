"""GPU training-step engines: everything between "a batch is on the device" and "global_step
was incremented" (reference worker.py:131-141), with pull and push fused into the
compute kernels.

* ``FusedMlp2Stepper``  - the reference's 784-100-10 model: ONE persistent cooperative kernel per
  launch, any number of steps per launch (``csrc/cuda/mlp2_step.cu``).
* ``LayerwiseStepper``  - any depth/width: one fused kernel per layer and direction; fp32 SIMT
  or (bf16/tf32) tcgen05 GEMMs with bias+sigmoid / sigmoid-grad / push epilogues.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch

from ..models.mlp import MLPSpec
from ..ops import cuda_ops as ops


class DeviceDataset:
    """Training set resident in HBM + on-device epoch sampler: ``next_indices`` yields the
    dataset rows of the next ``nsteps`` batches with TF1 ``DataSet.next_batch`` semantics (reshuffle at
    every epoch boundary, straddling batches take the rest of the old epoch first)."""

    def __init__(self, images, label_ids, device, seed: Optional[int] = None, dtype=torch.float32):
        self.device = torch.device(device)
        self.images = torch.as_tensor(images, dtype=torch.float32).to(self.device).to(dtype).contiguous()
        self.labels = torch.as_tensor(label_ids, dtype=torch.int32).to(self.device).contiguous()
        self.n = self.images.shape[0]
        self._gen = torch.Generator(device=self.device)
        if seed is None:
            self._gen.seed()
        else:
            self._gen.manual_seed(seed)
        self._perm = torch.randperm(self.n, device=self.device, generator=self._gen, dtype=torch.int32)
        self._pos = 0
        self.epochs_completed = 0

    def next_indices(self, batch: int, nsteps: int = 1) -> torch.Tensor:
        need = batch * nsteps
        chunks = []
        while need > 0:
            avail = self.n - self._pos
            if avail == 0:
                self._perm = torch.randperm(self.n, device=self.device, generator=self._gen, dtype=torch.int32)
                self._pos = 0
                self.epochs_completed += 1
                avail = self.n
            take = min(avail, need)
            chunks.append(self._perm[self._pos:self._pos + take])
            self._pos += take
            need -= take
        return chunks[0] if len(chunks) == 1 else torch.cat(chunks)


class FusedMlp2Stepper:
    def __init__(self, spec: MLPSpec, batch: int, params: Sequence[torch.Tensor], global_step: torch.Tensor,
                 lr: float, device, max_steps_per_launch: int = 1, atomic: bool = True, ks: int = 8,
                 remote: bool = True, grad_buffers: Optional[Sequence[torch.Tensor]] = None, tc: bool = False):
        """``tc``: training launches run on the tensor cores (``csrc/cuda/mlp2_step_tc.cu``: tcgen05 kind::tf32, one
        8-CTA cluster, W1 pulled by TMA from the shard); evaluation launches, ``atomic=False`` and synchronous mode
        stay on the fp32 SIMT kernel.
        ``grad_buffers``: synchronous mode - raw gradients go to these 4 buffers (this rank's exchange
        buffer) instead of being pushed; ``parallel.sync.SyncGroup.reduce_scatter_apply`` finishes the step."""
        self.grad_buffers = list(grad_buffers) if grad_buffers is not None else None
        self.cursor = None
        self.sync_exchange = None     # parallel.sync.SyncGroup.exchange_struct(lr): exchange inside the kernel
        if spec.num_layers != 2 or spec.param_dtype != "float32":
            raise ValueError("FusedMlp2Stepper handles 2-layer float32 models")
        D, H, C = spec.dims
        self.spec, self.batch, self.lr, self.atomic = spec, batch, lr, atomic
        self.W1, self.b1, self.W2, self.b2 = params
        self.global_step = global_step
        self.remote = remote      # False only when every shard is in this GPU's own HBM
        self.tc = bool(tc) and self.supports_tc(spec, batch) and atomic and grad_buffers is None
        self.dbg = None
        self.k = ops.Mlp2Step(batch, D, H, C, device, max_steps=max_steps_per_launch, ks=ks)
        self.launches_per_call = 1

    @staticmethod
    def supports(spec: MLPSpec, batch: int) -> bool:
        return (spec.num_layers == 2 and spec.param_dtype == "float32" and batch <= 128 and spec.dims[1] <= 128
                and spec.dims[2] <= 16 and (spec.dims[0] + 7) // 8 <= 148)

    @staticmethod
    def supports_tc(spec: MLPSpec, batch: int) -> bool:
        if spec.num_layers != 2 or spec.param_dtype != "float32" or spec.dims[0] % 4:
            return False
        D, H, C = spec.dims
        return bool(ops.lib().dtfe_mlp2_step_tc_supported(batch, D, H, C))

    def run(self, X: torch.Tensor, labels: torch.Tensor, nsteps: int = 1, batch_idx: Optional[torch.Tensor] = None,
            rows_per_step: int = 0, train: bool = True, out_offset: int = 0, outs=None) -> None:
        """Results land in ``self.k.loss/acc/step[out_offset : out_offset+nsteps]`` (device), or in the three
        tensors ``outs`` when given."""
        self.k.run(self.W1, self.b1, self.W2, self.b2, self.global_step, X, labels, self.lr, nsteps=nsteps,
                   batch_idx=batch_idx, rows_per_step=rows_per_step, atomic=self.atomic, train=train,
                   out_offset=out_offset, remote=self.remote, dbg=self.dbg, grads=self.grad_buffers,
                   cursor=self.cursor, sync_exchange=self.sync_exchange, outs=outs,
                   tc=self.tc and train and self.sync_exchange is None)

    @property
    def loss(self): return self.k.loss
    @property
    def acc(self): return self.k.acc
    @property
    def step(self): return self.k.step


class LayerwiseStepper:
    """fp32 SIMT layer-by-layer engine (any MLPSpec).  Order of operations per step:
    forward (weights read from the shards = pull) -> xent fwd/bwd -> for each layer, last to
    first: dX against the *pre-update* weights, then dW/db with the push fused -> global_step."""

    def __init__(self, spec: MLPSpec, batch: int, params: Sequence[torch.Tensor], global_step: torch.Tensor,
                 lr: float, device, atomic: bool = True, grad_buffers: Optional[Sequence[torch.Tensor]] = None,
                 adam: Optional[dict] = None, tf32: bool = False):
        """``tf32``: run the GEMMs whose operands satisfy TMA's 16-byte pitch rule on the tensor cores
        (tcgen05 kind::tf32, fp32 accumulation, weights pulled by TMA straight from the fp32 shard); the
        rest - here the 10-wide output layer - stays on the exact fp32 SIMT kernels.

        ``adam``: {'slots': [...], 'm': [...], 'v': [...], 't': int32 tensor, 'apply': bool, betas/eps} - gradients
        are pushed (red.add, unscaled) into the owner-side slots; the rank with ``apply=True`` (the chief)
        then runs the owner-side Adam kernel over every variable, consuming and clearing the slots."""
        self.adam = adam
        self._adam_t = None        # read from adam['t'] (the global/adam_t shard variable) at the first apply
        self.tf32 = tf32
        if tf32:
            from ..ops import gemm_tc
            self._tc = gemm_tc
        self.spec, self.batch, self.lr, self.atomic = spec, batch, lr, atomic
        self.params = list(params)
        self.global_step = global_step
        dev = torch.device(device)
        dims = spec.dims
        f32 = dict(device=dev, dtype=torch.float32)
        self.acts = [torch.empty(batch, d, **f32) for d in dims[1:-1]]
        self.logits = torch.empty(batch, dims[-1], **f32)
        self.dz = torch.empty(batch, dims[-1], **f32)
        self.dacts = [torch.empty(batch, d, **f32) for d in dims[1:-1]]
        self.rowloss = torch.empty(batch, **f32)
        self.rowcorrect = torch.empty(batch, device=dev, dtype=torch.int32)
        self.loss = torch.zeros(1, **f32)
        self.acc = torch.zeros(1, **f32)
        self.step = torch.zeros(1, device=dev, dtype=torch.int32)
        # sync mode: gradients go to local buffers instead of being pushed
        self.grad_buffers = list(grad_buffers) if grad_buffers is not None else None
        self.launches_per_call = 2 * spec.num_layers + 2 + (spec.num_layers - 1) + spec.num_layers + 1

    def forward(self, X: torch.Tensor, acts=None, logits=None) -> torch.Tensor:
        acts = self.acts if acts is None else acts
        logits = self.logits if logits is None else logits
        h = X
        L = self.spec.num_layers
        for l in range(L):
            W, b = self.params[2 * l], self.params[2 * l + 1]
            last = l == L - 1
            out = logits if last else acts[l]
            if self.tf32 and self._tc_ok(h, W):
                self._tc.dense_fwd_tc(h, W, b, ops.ACT_NONE if last else ops.ACT_SIGMOID, out)
            else:
                ops.dense_fwd(h, W, b, ops.ACT_NONE if last else ops.ACT_SIGMOID, out=out, shard_reads=True)
            h = out
        return logits

    @staticmethod
    def _tc_ok(*mats) -> bool:
        return all(m.dim() == 2 and m.stride(1) == 1 and m.stride(0) % 4 == 0 and m.data_ptr() % 16 == 0 for m in mats)

    def run(self, X: torch.Tensor, labels: torch.Tensor, train: bool = True) -> None:
        B = X.shape[0]
        self.forward(X)
        ops.softmax_xent(self.logits, labels, 1.0 / B, self.dz if train else None, self.rowloss, self.rowcorrect)
        ops.mean_loss_acc(self.rowloss, self.rowcorrect, B, self.loss, self.acc)
        if not train:
            return
        alpha = -self.lr
        d = self.dz
        L = self.spec.num_layers
        for l in reversed(range(L)):
            W, b = self.params[2 * l], self.params[2 * l + 1]
            h_in = X if l == 0 else self.acts[l - 1]
            d_prev = None
            tc = self.tf32 and self._tc_ok(h_in, d, W)
            if l > 0:
                if tc:
                    d_prev = self._tc.dense_bwd_dx_tc(d, W, h_in, self.dacts[l - 1])
                else:
                    d_prev = ops.dense_bwd_dx(d, W, h_in, ops.ACT_SIGMOID, out=self.dacts[l - 1], shard_reads=True)
            if tc and self.adam is None and self.grad_buffers is None and self.atomic:
                self._tc.dense_bwd_dw_push_tc(h_in, d, W, alpha)
                ops.bias_grad_push(d, b, alpha, atomic=True)
                d = d_prev
                continue
            if self.adam is not None:
                ops.dense_bwd_dw_push(h_in, d, self.adam["slots"][2 * l], 1.0, atomic=True)
                ops.bias_grad_push(d, self.adam["slots"][2 * l + 1], 1.0, atomic=True)
            elif self.grad_buffers is None:
                ops.dense_bwd_dw_push(h_in, d, W, alpha, atomic=self.atomic)
                ops.bias_grad_push(d, b, alpha, atomic=self.atomic)
            else:
                ops.dense_bwd_dw_push(h_in, d, None, 0.0, grad_out=self.grad_buffers[2 * l])
                ops.bias_grad_push(d, None, 0.0, grad_out=self.grad_buffers[2 * l + 1])
            d = d_prev
        if self.adam is not None and self.adam.get("apply", True):
            import math
            t_var = self.adam.get("t")
            if self._adam_t is None:     # after the chief's init-or-restore: continue the restored count
                self._adam_t = int(t_var.item()) if t_var is not None else 0
            self._adam_t += 1
            b1, b2, eps = self.adam.get("beta1", 0.9), self.adam.get("beta2", 0.999), self.adam.get("eps", 1e-8)
            lr_t = self.lr * math.sqrt(1.0 - b2 ** self._adam_t) / (1.0 - b1 ** self._adam_t)
            for i, p in enumerate(self.params):
                ops.adam_apply(p.view(-1), self.adam["m"][i].view(-1), self.adam["v"][i].view(-1),
                               self.adam["slots"][i].view(-1), lr_t, b1, b2, eps)
            if t_var is not None:       # checkpointed next to m and v (TF1 keeps beta powers; the gloo ps keeps this count)
                ops.global_step_add(t_var, 1)
        if self.grad_buffers is None:
            ops.global_step_add(self.global_step, 1, self.step)

    def eval_counts(self, X: torch.Tensor, labels: torch.Tensor, chunk: int = 16384):
        """Test-set accuracy with the current shard weights (reference worker.py:150-154)."""
        dev = X.device
        correct = torch.zeros(1, device=dev, dtype=torch.int32)
        total = 0
        cnt = torch.zeros(1, device=dev, dtype=torch.int32)
        for s in range(0, X.shape[0], chunk):
            xb, yb = X[s:s + chunk], labels[s:s + chunk]
            m = xb.shape[0]
            acts = [torch.empty(m, d, device=dev, dtype=torch.float32) for d in self.spec.dims[1:-1]]
            logits = torch.empty(m, self.spec.dims[-1], device=dev, dtype=torch.float32)
            self.forward(xb, acts, logits)
            rc = torch.empty(m, device=dev, dtype=torch.int32)
            ops.softmax_xent(logits, yb, 1.0, None, None, rc)
            ops.mean_loss_acc(None, rc, m, None, None, cnt)
            correct += cnt
            total += m
        return correct, total


class TcStepper:
    """bf16 tensor-core engine for any MLPSpec with ``param_dtype='bfloat16'`` (the wide benchmark
    family 784-8192-8192-10): every GEMM runs on ``tcgen05.mma`` via ``ops.gemm_tc`` with its
    elementwise neighbour fused - bias+sigmoid into the forward GEMM, the sigmoid gradient into the dX
    GEMM, the gradient push + SGD apply (``red.add`` into the shard, possibly over NVLink) into the dW
    GEMM - and the forward GEMMs read the weights straight from the owning shard through TMA (pull).

    Matrices whose width is not a multiple of 8 (the 10 logits) are kept with a padded row pitch."""

    def __init__(self, spec: MLPSpec, batch: int, params: Sequence[torch.Tensor], global_step: torch.Tensor,
                 lr: float, device, tile_n: int = 0, stage_all: bool = False):
        """``stage_all``: some shard is on a peer GPU - pull EVERY parameter (also biases and small matrices)
        into a local replica once per step.  Peer reads bypass the local L2, and a GEMM epilogue that reads a
        remote bias for each of its tiles was measured at +100 us per layer."""
        from ..ops import gemm_tc
        self.G = gemm_tc
        self.spec, self.batch, self.lr, self.tile_n = spec, batch, lr, tile_n
        self.params = list(params)      # tensors, or lists of row blocks for partitioned variables
        self.global_step = global_step
        dev = torch.device(device)
        dims = spec.dims
        pad8 = lambda n: (n + 7) // 8 * 8
        bf = dict(device=dev, dtype=torch.bfloat16)
        # A partitioned weight is read many times per step by the GEMMs (once per M tile), and peer reads
        # bypass the local L2, so its blocks are pulled ONCE per step into a local replica (the reference's
        # sync_op, worker.py:81-85, as one L1-bypassing vector copy per block) and the GEMMs read that.
        self.stage_all = stage_all
        self.replicas = {}
        self._pull_stream = None
        self._pending = set()
        for i, p in enumerate(self.params):
            if isinstance(p, (list, tuple)):
                rows = sum(b.shape[0] for b in p)
                self.replicas[i] = torch.empty(rows, p[0].stride(0), **bf)[:, :p[0].shape[1]]
            elif stage_all:
                if p.dim() == 2:
                    self.replicas[i] = torch.empty(p.shape[0], p.stride(0), **bf)[:, :p.shape[1]]
                else:
                    self.replicas[i] = torch.empty(pad8(p.numel()), **bf)[:p.numel()]
        self.acts = [torch.empty(batch, pad8(d), **bf)[:, :d] for d in dims[1:-1]]
        C = dims[-1]
        self.logits = torch.empty(batch, pad8(C), device=dev, dtype=torch.float32)[:, :C]
        self.dz = torch.empty(batch, pad8(C), device=dev, dtype=torch.float32)[:, :C]
        self.dz_bf = torch.zeros(batch, pad8(C), **bf)[:, :C]
        self.dacts = [torch.empty(batch, pad8(d), **bf)[:, :d] for d in dims[1:-1]]
        self.rowloss = torch.empty(batch, device=dev, dtype=torch.float32)
        self.rowcorrect = torch.empty(batch, device=dev, dtype=torch.int32)
        self.loss = torch.zeros(1, device=dev, dtype=torch.float32)
        self.acc = torch.zeros(1, device=dev, dtype=torch.float32)
        self.step = torch.zeros(1, device=dev, dtype=torch.int32)
        L = spec.num_layers
        self.launches_per_call = L + 2 + (L - 1) + 2 * L + 1
        import os
        # streaming kernels for an output layer of width <= 16 (ops.thin_*): correct, but at 52 / 74 / 102 us (forward,
        # dX, dW+push; scripts/bench_thin.py) only the forward one beats the 128-wide tensor-core tiles (57 / 70 / 45 us
        # for 4096x8192x10), so they are opt-in
        self.thin = os.environ.get("DTFE_THIN") == "1"
        self._thin_scratch = None

    def pull_partitioned(self) -> None:
        """Issue this step's pulls on side streams, one event per variable: the copy of layer l+1's weights over
        NVLink runs under layer l's GEMM; ``_w`` makes the consumer wait for its own event.  Blocks go round-robin
        over a few streams.  ``DTFE_PULL_CE=1`` issues them as ``cudaMemcpyAsync`` calls (copy engines, no SM taken
        from the GEMMs) instead of the L1-bypassing SM copy kernel; measured at 4 GPUs that is slightly slower
        (2.02 vs 1.96 ms per step), so the kernel stays the default."""
        if not self.replicas:
            return
        import os
        if self._pull_stream is None:
            dev = self.loss.device
            self._pull_streams = [torch.cuda.Stream(device=dev) for _ in range(4)]
            self._pull_stream = self._pull_streams[0]
            self._pull_events = {i: [torch.cuda.Event() for _ in self._pull_streams] for i in self.replicas}
            self._pull_sm = os.environ.get("DTFE_PULL_CE") != "1"
        main = torch.cuda.current_stream()
        for st in self._pull_streams:
            st.wait_stream(main)                   # previous step's readers of the replicas are done
        # small things first (biases, thin matrices), then the big matrices in layer order
        order = sorted(self.replicas, key=lambda i: (self._nbytes(i) > (1 << 20), i))
        rr = 0
        for i in order:
            rep = self.replicas[i]
            blocks = self.params[i] if isinstance(self.params[i], (list, tuple)) else [self.params[i]]
            rp = blocks[0].shape[0]
            used = set()
            for k, blk in enumerate(blocks):
                if blk.dim() == 2:
                    dst = rep[k * rp:(k + 1) * rp]
                    # blocks and replica rows share the pitch, so each block is one contiguous byte range
                    n = rp * blk.stride(0)
                else:
                    dst, n = rep, (blk.numel() + 1) // 2 * 2
                si = rr % len(self._pull_streams); rr += 1
                used.add(si)
                with torch.cuda.stream(self._pull_streams[si]):
                    if self._pull_sm:
                        ops.pull(torch.as_strided(dst, (n,), (1,)).view(torch.float32),
                                 torch.as_strided(blk, (n,), (1,)).view(torch.float32))
                    else:
                        ops.pull_ce(dst, blk, n * 2)
            for si in used:
                self._pull_events[i][si].record(self._pull_streams[si])
            self._pull_used = getattr(self, "_pull_used", {})
            self._pull_used[i] = used
        self._pending = set(self.replicas)

    def _nbytes(self, i) -> int:
        p = self.params[i]
        blocks = p if isinstance(p, (list, tuple)) else [p]
        return sum(b.numel() for b in blocks) * 2

    def _w(self, i):
        if i in self.replicas:
            if i in self._pending:
                for si in self._pull_used[i]:
                    torch.cuda.current_stream().wait_event(self._pull_events[i][si])
                self._pending.discard(i)
            return self.replicas[i]
        return self.params[i]

    def forward(self, X: torch.Tensor, acts=None, logits=None) -> torch.Tensor:
        acts = self.acts if acts is None else acts
        logits = self.logits if logits is None else logits
        h = X
        L = self.spec.num_layers
        self.pull_partitioned()
        for l in range(L):
            W, b = self._w(2 * l), self._w(2 * l + 1)
            last = l == L - 1
            out = logits if last else acts[l]
            if last and self.thin and ops.thin_ok(h, W):
                ops.thin_fwd(h, W, b, out)             # output layer of width <= 16: streaming kernel, not 128-wide tiles
            else:
                self.G.dense_fwd_tc(h, W, b, ops.ACT_NONE if last else ops.ACT_SIGMOID, out, tile_n=self.tile_n)
            h = out
        return logits

    def run(self, X: torch.Tensor, labels: torch.Tensor, train: bool = True) -> None:
        B = X.shape[0]
        self.forward(X)
        ops.softmax_xent_bf16grad(self.logits, labels, 1.0 / B, self.dz, self.dz_bf, self.rowloss, self.rowcorrect)
        ops.mean_loss_acc(self.rowloss, self.rowcorrect, B, self.loss, self.acc)
        if not train:
            return
        alpha = -self.lr
        d = self.dz_bf
        L = self.spec.num_layers
        for l in reversed(range(L)):
            W, b = self.params[2 * l], self.params[2 * l + 1]
            h_in = X if l == 0 else self.acts[l - 1]
            d_prev = None
            thin = l == L - 1 and l > 0 and self.thin and ops.thin_ok(h_in, W)
            if l > 0:   # against the pre-update weights, so before this layer's push
                if thin:
                    d_prev = ops.thin_dx(d, self._w(2 * l), h_in, self.dacts[l - 1])
                else:
                    d_prev = self.G.dense_bwd_dx_tc(d, self._w(2 * l), h_in, self.dacts[l - 1], tile_n=self.tile_n)
            if thin:
                if self._thin_scratch is None:
                    self._thin_scratch = ops.thin_dw_scratch(h_in.shape[0], h_in.shape[1], h_in.device)
                ops.thin_dw_push(h_in, d, W, alpha, self._thin_scratch)
            else:
                self.G.dense_bwd_dw_push_tc(h_in, d, W, alpha, tile_n=self.tile_n, tma_push=self.stage_all)
            ops.bias_grad_push_bf16(d, b, alpha)
            d = d_prev
        ops.global_step_add(self.global_step, 1, self.step)
