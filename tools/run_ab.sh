#!/bin/bash
# Quick single-GPU A/B: kernel + model tests, then bench.py at both batch sizes for each
# "NAME=VALUE" environment setting given on the command line (plus the default).
# Usage: bash tools/run_ab.sh DDPB_C1_CTAS=1 ...
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_zz_no_alloc.py -m gpu -q --no-header -p no:cacheprovider > gpurun_out/pytest_ab.txt 2>&1
echo "pytest rc=$?"; grep -E "passed|failed|Error|mismatch" gpurun_out/pytest_ab.txt | tail -8 | cut -c1-300
one() {  # tag envsetting batch steps warmup
  local tag=$1 setting=$2 b=$3 k=$4 w=$5
  env $setting timeout 600 python bench.py --batch_size $b --steps $k --warmup $w > gpurun_out/ab_$tag.json 2> gpurun_out/ab_$tag.err
  echo "$tag [$setting] rc=$? $(python - <<PY
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/ab_$tag.json") if l.startswith("{")][-1]
    print(d.get("ms_per_step"), d.get("value"), "e2e", d.get("e2e",{}).get("value"), d.get("clocks",{}).get("reasons"))
except Exception as e:
    print("no json", e)
PY
)"
}
i=0
for setting in "DDPB_AB=default" "$@"; do
  one b64_$i "$setting" 64 1000 50
  one b4096_$i "$setting" 4096 50 5
  i=$((i+1))
done
