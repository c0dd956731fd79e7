#!/bin/bash
# Multi-GPU timing diagnostics (run under gpurun --gpus N): where does the scaling overhead come from?
# bench.py at 1 GPU and at all GPUs of the box, then with each bucket update ablated (timing only).
set -u
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
echo "GPUs: $NG"
one() {  # tag n envsetting extra-args...
  local tag=$1 n=$2 setting=$3; shift 3
  if [ $n -eq 1 ]; then
    env $setting timeout 300 python bench.py --gpus 1 "$@" > gpurun_out/abl_$tag.json 2> gpurun_out/abl_$tag.err
  else
    env $setting timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
      --master-port $((29500 + RANDOM % 400)) bench.py --gpus $n "$@" > gpurun_out/abl_$tag.json 2> gpurun_out/abl_$tag.err
  fi
  echo "$tag n=$n [$setting $*] rc=$? $(python - <<PY
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/abl_$tag.json") if l.startswith("{")][-1]
    print(d.get("ms_per_step"), d.get("value"), "e2e", d.get("e2e",{}).get("value"), d.get("clocks",{}).get("reasons"))
except Exception as e:
    print("no json", e)
PY
)"
}
A="--steps 400 --warmup 30"
one n1 1 DDPB_AB=0 $A
one full $NG DDPB_AB=0 $A
one no_fl $NG DDPB_ABLATE=fl $A
one no_conv2 $NG DDPB_ABLATE=conv2 $A
one no_conv1 $NG DDPB_ABLATE=conv1 $A
one nvls $NG DDPB_NVLS=1 $A
