#!/bin/bash
# Single-GPU cluster: 1 ps + num_workers workers sharing one GPU (each gets 0.9/num_workers of it).
# Same knobs as the reference launcher; MODE=headless runs without tmux under a supervising parent.
name=${NAME:-sess}
num_workers=${NUM_WORKERS:-2}
num_gpus=1
GPU_ID=(${GPU_IDS:-0})
mode=${MODE:-tmux}

cd "$(dirname "$0")" || exit 1
exec python -m distributedtensorflowexample_b200.launch.launcher \
    --name "$name" --num_workers "$num_workers" --num_gpus "$num_gpus" \
    --gpu_id "${GPU_ID[*]}" --mode "$mode" "$@"
