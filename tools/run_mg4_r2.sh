#!/bin/bash
# Round-2 4-GPU session (run under `gpurun --gpus 4`): both bench arms at 1/2/4 GPUs with the final code,
# the bucketed scenario, the all-reduce sweep at world sizes 2 and 4 (multicast communicator).
set -u
mkdir -p gpurun_out
for n in 1 2 4; do
  timeout 300 python bench.py --gpus $n --steps 200 --warmup 20 > gpurun_out/ours_b64_n$n.json 2> gpurun_out/ours_b64_n$n.err
  tail -1 gpurun_out/ours_b64_n$n.json | grep -o "\"ms_per_step\": [0-9.]*" | head -1
  timeout 300 python bench.py --impl reference --gpus $n --steps 200 --warmup 20 > gpurun_out/ref_b64_n$n.json 2> gpurun_out/ref_b64_n$n.err
  tail -1 gpurun_out/ref_b64_n$n.json | grep -o "\"ms_per_step\": [0-9.]*" | head -1
done
DDPB_NVLS=0 timeout 300 python bench.py --gpus 4 --steps 200 --warmup 20 > gpurun_out/ours_b64_n4_ipc.json 2>/dev/null
tail -1 gpurun_out/ours_b64_n4_ipc.json | grep -o "\"ms_per_step\": [0-9.]*" | head -1
timeout 300 python bench.py --gpus 1 --steps 200 --warmup 20 --dtype bf16 > gpurun_out/ours_b64_n1_bf16.json 2>/dev/null
timeout 600 python -m pytest tests/test_multi_gpu.py -q -k "bucketed or dp_equiv-4 or dp_equiv-3" 2>&1 | tail -3
for w in 2 4; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $w --master-addr 127.0.0.1 --master-port 2954$w \
    tools/allreduce_sweep.py --multicast --out gpurun_out/sweep_w${w}_mc.json > gpurun_out/sweep_w${w}_mc.log 2>&1
  echo "sweep w=$w rc=$?"
done
