#!/bin/bash
# 4-GPU batch: scaling point N=4 (async with the launcher path + the reference's 16 workers / 4 GPUs topology, sync),
# wide config with copy-engine vs SM pulls, bandwidth sweep at 2 and 4 GPUs, time-to-accuracy at 1/2/4 workers.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
j() { grep -h "^{" "$1" | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
e=d.get('entry') or {}
print('$1','us/step',round(d.get('ms_per_step',0)*1e3,2),'value',round(d.get('value',0)),'vs',d.get('vs_baseline'),'entry',e.get('value'),'frac',e.get('fraction_of_device_timed'),'16w4g',(e.get('reference_topology_16w_4gpu') or {}).get('value'),'e2e',(d.get('e2e') or {}).get('value'),'tflops',d.get('tflops_per_gpu'))
" 2>&1 | tail -1; }
timeout 600 $TR --nproc-per-node 4 --master-port 29701 bench.py --gpus 4 --steps 500 --warmup 20 --entry-multi > gpurun_out/async4.json 2> gpurun_out/async4.err; echo "rc_async4=$?"; j gpurun_out/async4.json
timeout 300 $TR --nproc-per-node 4 --master-port 29702 bench.py --gpus 4 --steps 500 --warmup 20 --mode sync --repeats 9 > gpurun_out/sync4.json 2> gpurun_out/sync4.err; echo "rc_sync4=$?"; j gpurun_out/sync4.json
timeout 300 $TR --nproc-per-node 4 --master-port 29703 bench.py --gpus 4 --config wide --steps 10 --warmup 3 --no-e2e > gpurun_out/wide4.json 2> gpurun_out/wide4.err; echo "rc_wide4=$?"; j gpurun_out/wide4.json
DTFE_PULL_SM=1 timeout 300 $TR --nproc-per-node 4 --master-port 29704 bench.py --gpus 4 --config wide --steps 10 --warmup 3 --no-e2e --no-baseline > gpurun_out/wide4_smpull.json 2> gpurun_out/wide4_smpull.err; echo "rc_wide4sm=$?"; j gpurun_out/wide4_smpull.json
timeout 200 $TR --nproc-per-node 4 --master-port 29705 scripts/profile_wide_step.py > gpurun_out/wide4_kernels.txt 2>&1; echo "rc_widek=$?"; grep '"rank": 0' gpurun_out/wide4_kernels.txt | cut -c1-600
timeout 200 python bench.py --gpus 1 --config wide --steps 10 --warmup 3 --no-e2e > gpurun_out/wide1.json 2> gpurun_out/wide1.err; echo "rc_wide1=$?"; j gpurun_out/wide1.json
timeout 300 $TR --nproc-per-node 2 --master-port 29706 bench.py --gpus 2 --config bw > gpurun_out/bw2.json 2> gpurun_out/bw2.err; echo "rc_bw2=$?"
timeout 300 $TR --nproc-per-node 4 --master-port 29707 bench.py --gpus 4 --config bw > gpurun_out/bw4.json 2> gpurun_out/bw4.err; echo "rc_bw4=$?"
timeout 600 python scripts/tta_launcher.py --gpus 1 2 4 --repeats 2 > gpurun_out/tta_launcher_124.jsonl 2> gpurun_out/tta_launcher_124.err; echo "rc_tta=$?"; cat gpurun_out/tta_launcher_124.jsonl | cut -c1-260
for n in 1 2 4; do timeout 200 $TR --nproc-per-node $n --master-port 2971$n scripts/tta_nccl.py --repeats 2 2>/dev/null | grep "^{" >> gpurun_out/tta_nccl_124.jsonl; done; echo "rc_ttanccl=$?"; cat gpurun_out/tta_nccl_124.jsonl | cut -c1-260
