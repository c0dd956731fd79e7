#!/bin/bash
# 8-GPU timing diagnostics (run under `gpurun --gpus 8`): the default step, then with one bucket update
# skipped at a time (DDPB_ABLATE makes the step WRONG on purpose: timing only), then with the CUDA-IPC
# communicator.  One JSON line per run in gpurun_out/ablate8_*.json.
set -u
mkdir -p gpurun_out
run() { timeout 200 python bench.py --gpus 8 --steps 200 --warmup 20; }
for a in none fl conv2 conv1 all; do
  case $a in none) v="";; all) v="fl,conv2,conv1";; *) v=$a;; esac
  DDPB_ABLATE=$v run > gpurun_out/ablate8_$a.json 2> gpurun_out/ablate8_$a.err
  echo "$a $(tail -1 gpurun_out/ablate8_$a.json | grep -o '"ms_per_step": [0-9.]*' | head -1)"
done
timeout 200 python bench.py --gpus 1 --steps 200 --warmup 20 > gpurun_out/ablate8_n1.json 2>/dev/null
echo "n1 $(tail -1 gpurun_out/ablate8_n1.json | grep -o '"ms_per_step": [0-9.]*' | head -1)"
DDPB_ABLATE=fl,conv2,conv1 timeout 200 python bench.py --gpus 1 --steps 200 --warmup 20 > gpurun_out/ablate8_n1_all.json 2>/dev/null
echo "n1_all $(tail -1 gpurun_out/ablate8_n1_all.json | grep -o '"ms_per_step": [0-9.]*' | head -1)"
