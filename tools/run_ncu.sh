#!/bin/bash
# ONE ncu --set full capture of two eagerly launched training steps (run under gpurun, 1 GPU).
# Usage: bash tools/run_ncu.sh [per-rank batch, default 4096] [tag]
set -u
B=${1:-4096}; TAG=${2:-prof_b$B}
mkdir -p gpurun_out
# the program must exit 0 without ncu first
timeout 300 python bench.py --batch_size $B --steps 2 --warmup 3 --no_graph > gpurun_out/plain_$TAG.log 2>&1
rc=$?; echo "plain eager rc=$rc"; [ $rc -eq 0 ] || exit $rc
# 13 of our kernels per step (+1 weight pack at start-up): skip the three warm-up steps
timeout 900 ncu --set full --clock-control none --import-source on \
  -k 'regex:^(conv|linear|pack|reduce|sgd|softmax|wgrad|zero)' --launch-skip 42 -c 14 -f \
  -o gpurun_out/$TAG python bench.py --batch_size $B --steps 2 --warmup 3 --no_graph > gpurun_out/ncu_$TAG.log 2>&1
echo "ncu rc=$?"; grep -E "PROF|WARN|ERROR" gpurun_out/ncu_$TAG.log | tail -5 | cut -c1-200; ls -la gpurun_out/$TAG.ncu-rep
