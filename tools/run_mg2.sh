#!/bin/bash
# Two-GPU validation (run under gpurun --gpus 2): every multi-GPU test at world 2, then bench.py
# at N=2 for the tail-update variants and the multicast communicator.  Output in gpurun_out/.
set -u
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
echo "GPUs: $NG"
timeout 1200 python -m pytest tests/test_multi_gpu.py -m gpu -q --no-header -p no:cacheprovider -rs > gpurun_out/pytest_mg2.txt 2>&1
echo "pytest world=2 rc=$?"; grep -E "passed|failed|rank failure|Error|SKIP|assert" gpurun_out/pytest_mg2.txt | head -20 | cut -c1-400
one() {  # tag envsetting extra-args...
  local tag=$1 setting=$2; shift 2
  env $setting timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
    --master-port $((29500 + RANDOM % 400)) bench.py --gpus 2 "$@" > gpurun_out/mg2_$tag.json 2> gpurun_out/mg2_$tag.err
  echo "$tag [$setting $*] rc=$? $(python - <<PY
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/mg2_$tag.json") if l.startswith("{")][-1]
    print(d.get("ms_per_step"), d.get("value"), "e2e", d.get("e2e",{}).get("value"), d.get("clocks",{}).get("reasons"))
except Exception as e:
    print("no json", e)
PY
)"
  grep -E "Warning|Error|error" gpurun_out/mg2_$tag.err | head -3 | cut -c1-300
}
one push DDPB_TAIL_UPDATE=push --steps 1000 --warmup 50
one pull DDPB_TAIL_UPDATE=pull --steps 1000 --warmup 50
one split DDPB_TAIL_UPDATE=split --steps 1000 --warmup 50
one nvls DDPB_NVLS=1 --steps 1000 --warmup 50 --algo 4
one b4096 DDPB_AB=0 --batch_size 4096 --steps 50 --warmup 5
