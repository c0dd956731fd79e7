#!/bin/bash
# One single-GPU validation pass (run under gpurun, 1 GPU): full GPU test suite, the two headline
# benches, kernel-level timing, then ONE ncu --set full capture of a per-rank-batch-4096 step.
# Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --no-header -p no:cacheprovider > gpurun_out/pytest1.txt 2>&1
echo "pytest rc=$?"; grep -E "passed|failed|Error" gpurun_out/pytest1.txt | tail -8 | cut -c1-300
bench() {  # tag args...
  local tag=$1; shift
  timeout 600 python bench.py "$@" > gpurun_out/$tag.json 2> gpurun_out/$tag.err
  echo "$tag rc=$? $(python - <<PY
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/$tag.json") if l.startswith("{")][-1]
    print(d.get("ms_per_step"), d.get("value"), "e2e", d.get("e2e",{}).get("value"), d.get("gpu_launches"), d.get("clocks",{}).get("reasons"))
except Exception as e:
    print("no json", e)
PY
)"
}
bench b64 --steps 1000 --warmup 50
DDPB_TAIL_UPDATE=pull bench b64_pull --steps 1000 --warmup 50
DDPB_TAIL_UPDATE=split bench b64_split --steps 1000 --warmup 50
bench b4096 --batch_size 4096 --steps 50 --warmup 5
