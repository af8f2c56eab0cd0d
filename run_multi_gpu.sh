#!/bin/bash
# Multi-GPU cluster: 1 ps + num_workers workers spread over the GPUs listed in GPU_ID;
# worker i runs on GPU_ID[i % num_gpus] (16 workers on 4 GPUs = 4 per GPU, as in the reference).
name=${NAME:-sess}
num_workers=${NUM_WORKERS:-16}
num_gpus=${NUM_GPUS:-4}
GPU_ID=(${GPU_IDS:-0 1 2 3})
mode=${MODE:-tmux}

cd "$(dirname "$0")" || exit 1
exec python -m distributedtensorflowexample_b200.launch.launcher \
    --name "$name" --num_workers "$num_workers" --num_gpus "$num_gpus" \
    --gpu_id "${GPU_ID[*]}" --mode "$mode" "$@"
