#!/bin/bash
# Time-boxed 8-GPU check: correctness of the default path at world 8 (fused kernels, DP equivalence),
# then bench.py at N=8 with the IPC and the multicast communicator.  Output in gpurun_out/.
set -u
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
echo "GPUs: $NG"
timeout 240 python -m pytest tests/test_multi_gpu.py -m gpu -q --no-header -p no:cacheprovider -rs \
  -k "(fused_sgd or dp_equiv) and ${NG}]" > gpurun_out/pytest_final$NG.txt 2>&1
echo "pytest world=$NG rc=$?"; grep -E "passed|failed|rank failure|Error|SKIP|assert" gpurun_out/pytest_final$NG.txt | head -12 | cut -c1-400
one() {  # tag envsetting extra-args...
  local tag=$1 setting=$2; shift 2
  env $setting timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 \
    --master-port $((29500 + RANDOM % 400)) bench.py --gpus $NG "$@" > gpurun_out/final_$tag.json 2> gpurun_out/final_$tag.err
  echo "$tag [$setting $*] rc=$? $(python - <<PY
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/final_$tag.json") if l.startswith("{")][-1]
    print(d.get("ms_per_step"), d.get("value"), "e2e", d.get("e2e",{}).get("value"), d.get("clocks",{}).get("reasons"))
except Exception as e:
    print("no json", e)
PY
)"
  grep -E "Warning|Error|error" gpurun_out/final_$tag.err | head -3 | cut -c1-300
}
one ours_b64_n${NG} DDPB_AB=0 --steps 600 --warmup 30
one ours_b64_n${NG}_nvls DDPB_NVLS=1 --steps 600 --warmup 30
