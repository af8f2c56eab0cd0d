"""Per-kernel CUDA-event breakdown of one wide-MLP step (run under torchrun with N ranks, or alone)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from distributedtensorflowexample_b200.models.mlp import MLPSpec
from distributedtensorflowexample_b200.parallel.spmd import build_spmd_store
from distributedtensorflowexample_b200.train.gpu_step import TcStepper
from distributedtensorflowexample_b200.ops import cuda_ops as ops

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1: dist.init_process_group("nccl", device_id=dev)
dims = (784, 8192, 8192, 10); B = 4096
spec = MLPSpec(dims, param_dtype="bfloat16")
model, store = build_spmd_store(spec, rank, world, dev, seed=0, ps_strategy="greedy_bytes",
                                partition_min_bytes=(1 << 20) if world > 1 else 0)
names = [v.name for v in model.global_trainable]
st = TcStepper(spec, B, [store.param_handle(n) for n in names], store.tensor("global/global_step"), 0.001, dev,
               stage_all=not store.all_local())
X = torch.rand(B, 784, device=dev).bfloat16(); Y = torch.randint(0, 10, (B,), device=dev, dtype=torch.int32)
G = st.G
events = []
def mark(name):
    e = torch.cuda.Event(enable_timing=True); e.record(); events.append((name, e))
def step(record):
    global events
    events = []
    m = mark if record else (lambda n: None)
    m("start")
    st.pull_partitioned()
    if st._pull_stream is not None:
        for ps in st._pull_streams: torch.cuda.current_stream().wait_stream(ps)
    st._pending = set(); m("pull (not overlapped here)")
    h = X
    L = 3
    for l in range(L):
        W, b = st._w(2 * l), st._w(2 * l + 1)
        out = st.logits if l == L - 1 else st.acts[l]
        G.dense_fwd_tc(h, W, b, 0 if l == L - 1 else 1, out); m(f"fwd{l+1}")
        h = out
    ops.softmax_xent_bf16grad(st.logits, Y, 1.0 / B, st.dz, st.dz_bf, st.rowloss, st.rowcorrect)
    ops.mean_loss_acc(st.rowloss, st.rowcorrect, B, st.loss, st.acc); m("xent+mean")
    d = st.dz_bf
    for l in reversed(range(L)):
        W, b = st.params[2 * l], st.params[2 * l + 1]
        h_in = X if l == 0 else st.acts[l - 1]
        d_prev = None
        if l > 0:
            d_prev = G.dense_bwd_dx_tc(d, st._w(2 * l), h_in, st.dacts[l - 1]); m(f"dX{l+1}")
        G.dense_bwd_dw_push_tc(h_in, d, W, -0.001); m(f"dW{l+1} push")
        ops.bias_grad_push_bf16(d, b, -0.001); m(f"db{l+1} push")
        d = d_prev
    ops.global_step_add(st.global_step, 1, st.step); m("global_step")
for _ in range(3): step(False)
if world > 1: dist.barrier()
torch.cuda.synchronize()
step(True); torch.cuda.synchronize()
rows = {events[i][0]: round(events[i - 1][1].elapsed_time(events[i][1]) * 1e3, 1) for i in range(1, len(events))}
rows["total_us"] = round(events[0][1].elapsed_time(events[-1][1]) * 1e3, 1)
placement = {n.split("global/")[1]: (store.shard_of(n)) for n in names}
out = [None] * world
if world > 1: dist.all_gather_object(out, (rank, rows))
else: out = [(0, rows)]
if rank == 0:
    print(json.dumps({"placement(-1=row-partitioned)": placement}))
    for r, d in out: print(json.dumps({"rank": r, **d}))
if world > 1: dist.barrier(); dist.destroy_process_group()
