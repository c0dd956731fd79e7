#!/bin/bash
# Round-2 8-GPU session (run under `gpurun --gpus 8`): both bench arms at 8 GPUs, the multi-GPU
# scenarios at the non-power-of-two and the full world size, the all-reduce sweep through the
# multicast communicator.  Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
run() { timeout 300 python bench.py --gpus 8 --steps 200 --warmup 20 "$@"; }
run > gpurun_out/ours_b64_n8.json 2> gpurun_out/ours_b64_n8.err; tail -1 gpurun_out/ours_b64_n8.json | cut -c1-260
run --impl reference > gpurun_out/ref_b64_n8.json 2> gpurun_out/ref_b64_n8.err; tail -1 gpurun_out/ref_b64_n8.json | cut -c1-200
DDPB_NVLS=0 run > gpurun_out/ours_b64_n8_ipc.json 2> gpurun_out/ours_b64_n8_ipc.err; tail -1 gpurun_out/ours_b64_n8_ipc.json | cut -c1-200
run --dtype bf16 > gpurun_out/ours_b64_n8_bf16.json 2> gpurun_out/ours_b64_n8_bf16.err; tail -1 gpurun_out/ours_b64_n8_bf16.json | cut -c1-200
timeout 600 python -m pytest tests/test_multi_gpu.py -q -k "dp_equiv-6 or dp_equiv-8 or fused_sgd-6 or fused_sgd-8 or stress-8 or allreduce-6 or multicast_communicator" 2>&1 | tail -4
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 \
  tools/allreduce_sweep.py --multicast --out gpurun_out/sweep_w8_mc.json > gpurun_out/sweep_w8_mc.log 2>&1
echo "sweep rc=$?"; tail -3 gpurun_out/sweep_w8_mc.log | cut -c1-200
