#!/bin/bash
# 8-GPU batch: the scaling end point for every configuration BASELINE.json names.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
j() { grep -h "^{" "$1" | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
e=d.get('entry') or {}
print('$1','us/step',round(d.get('ms_per_step',0)*1e3,2),'value',round(d.get('value',0)),'vs',d.get('vs_baseline'),'entry',e.get('value'),'frac',e.get('fraction_of_device_timed'),'e2e',(d.get('e2e') or {}).get('value'),'tflops',d.get('tflops_per_gpu'),'agree',d.get('replicas_bit_identical'))
" 2>&1 | tail -1; }
timeout 400 $TR --master-port 29801 bench.py --gpus 8 --steps 500 --warmup 20 > gpurun_out/async8.json 2> gpurun_out/async8.err; echo "rc_async8=$?"; j gpurun_out/async8.json
timeout 200 $TR --master-port 29802 bench.py --gpus 8 --steps 500 --warmup 20 --mode sync --repeats 9 > gpurun_out/sync8_ll.json 2> gpurun_out/sync8_ll.err; echo "rc_sync8=$?"; j gpurun_out/sync8_ll.json
timeout 200 $TR --master-port 29803 bench.py --gpus 8 --steps 500 --warmup 20 --mode sync --sync-mode allreduce --repeats 9 --no-baseline > gpurun_out/sync8_flags.json 2> gpurun_out/sync8_flags.err; echo "rc_sync8f=$?"; j gpurun_out/sync8_flags.json
timeout 300 $TR --master-port 29804 bench.py --gpus 8 --config wide --steps 10 --warmup 3 --no-e2e > gpurun_out/wide8.json 2> gpurun_out/wide8.err; echo "rc_wide8=$?"; j gpurun_out/wide8.json
timeout 300 $TR --master-port 29805 bench.py --gpus 8 --config bw > gpurun_out/bw8.json 2> gpurun_out/bw8.err; echo "rc_bw8=$?"
timeout 300 python scripts/tta_launcher.py --gpus 8 --repeats 2 > gpurun_out/tta_launcher_8.jsonl 2> gpurun_out/tta_launcher_8.err; echo "rc_tta=$?"; cat gpurun_out/tta_launcher_8.jsonl | cut -c1-260
timeout 200 $TR --master-port 29806 scripts/tta_nccl.py --repeats 2 2>/dev/null | grep "^{" > gpurun_out/tta_nccl_8.jsonl; echo "rc_ttanccl=$?"; cat gpurun_out/tta_nccl_8.jsonl | cut -c1-260
timeout 200 python -c "
import json
from distributedtensorflowexample_b200.benchlib import measure_entrypoint
print(json.dumps(measure_entrypoint(8, 8, 800000, sync=True)))" > gpurun_out/entry_sync8.json 2> gpurun_out/entry_sync8.err; echo "rc_entry_sync8=$?"; cut -c1-400 gpurun_out/entry_sync8.json
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv | head -9
