#!/bin/bash
# 2-GPU validation batch (run on a box with >= 2 GPUs): multi-GPU tests, launcher topologies, sync/async benches.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_launcher.py "tests/test_gpu_train.py::test_fused_step_allreduce_exchange_single_rank" -x -q > gpurun_out/pytest_gpu2.txt 2>&1
echo "rc_pytest=$?"; tail -15 gpurun_out/pytest_gpu2.txt
timeout 300 $TR --master-port 29601 bench.py --gpus 2 --steps 500 --warmup 20 --mode sync --repeats 9 > gpurun_out/sync2_allreduce.json 2> gpurun_out/sync2_allreduce.err
echo "rc_sync_ar=$?"; tail -3 gpurun_out/sync2_allreduce.err
timeout 300 $TR --master-port 29602 bench.py --gpus 2 --steps 500 --warmup 20 --mode sync --sync-mode reduce_scatter --repeats 9 --no-baseline > gpurun_out/sync2_rs.json 2> gpurun_out/sync2_rs.err
echo "rc_sync_rs=$?"
timeout 120 python bench.py --gpus 1 --steps 500 --warmup 20 --mode sync --repeats 9 --no-baseline > gpurun_out/sync1_allreduce.json 2> gpurun_out/sync1_allreduce.err
echo "rc_sync1=$?"
timeout 400 $TR --master-port 29603 bench.py --gpus 2 --steps 500 --warmup 20 > gpurun_out/async2.json 2> gpurun_out/async2.err
echo "rc_async2=$?"; tail -3 gpurun_out/async2.err
# the reference's multi-GPU launcher, scaled to this box: 4 workers on 2 GPUs, async then --sync (2 workers)
MODE=headless NUM_WORKERS=4 NUM_GPUS=2 GPU_IDS="0 1" timeout 300 bash run_multi_gpu.sh --training_steps 200000 --data synthetic --logdir /tmp/mg_async/log --base_port 23300 > gpurun_out/run_mg_4w2g.log 2>&1
echo "rc_mg_async=$?"; grep -E "exit codes|speed" gpurun_out/run_mg_4w2g.log | tail -4
MODE=headless NUM_WORKERS=2 NUM_GPUS=2 GPU_IDS="0 1" timeout 300 bash run_multi_gpu.sh --sync --num_ps 2 --training_steps 100000 --data synthetic --logdir /tmp/mg_sync/log --base_port 23400 > gpurun_out/run_mg_sync_2w2g.log 2>&1
echo "rc_mg_sync=$?"; grep -E "exit codes|speed" gpurun_out/run_mg_sync_2w2g.log | tail -4
for f in sync2_allreduce sync2_rs sync1_allreduce async2; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(f"gpurun_out/{sys.argv[1]}.json"))
    print(sys.argv[1], "us/step", round(d["ms_per_step"]*1e3,2), "value", round(d["value"]), "vs", d.get("vs_baseline"), "agree", d.get("replicas_bit_identical"), "entry", (d.get("entry") or {}).get("value"), "e2e", (d.get("e2e") or {}).get("value"))
except Exception as e: print(sys.argv[1], "unreadable", e)
PY
done
