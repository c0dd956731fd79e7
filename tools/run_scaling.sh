#!/bin/bash
# One multi-GPU validation + scaling pass (run under gpurun --gpus 8).  Everything lands in gpurun_out/.
# Usage: bash tools/run_scaling.sh [sweep]
set -u
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
echo "GPUs: $NG"
run_bench() {  # impl n batch steps warmup tag
  local impl=$1 n=$2 b=$3 k=$4 w=$5 tag=$6
  if [ "$n" = "1" ]; then
    timeout 300 python bench.py --impl $impl --gpus 1 --batch_size $b --steps $k --warmup $w > gpurun_out/$tag.json 2> gpurun_out/$tag.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
      bench.py --impl $impl --gpus $n --batch_size $b --steps $k --warmup $w > gpurun_out/$tag.json 2> gpurun_out/$tag.err
  fi
  echo "$tag rc=$? $(python - <<PY
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/$tag.json") if l.startswith("{")][-1]
    print(d.get("ms_per_step"), d.get("value"), "e2e", d.get("e2e",{}).get("value"), d.get("clocks",{}).get("reasons"), d.get("clocks",{}).get("samples"))
except Exception as e:
    print("no json", e)
PY
)"
}
# 1. correctness at the full world size
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -q --no-header -p no:cacheprovider -k "(test_scenario and ${NG}]) or cli" > gpurun_out/pytest_mg$NG.txt 2>&1
echo "pytest world=$NG rc=$?"; grep -E "passed|failed|rank failure|Error" gpurun_out/pytest_mg$NG.txt | head -12 | cut -c1-400
# 2. scaling, per-rank batch 64 (BASELINE config 2)
for n in 1 2 4 8; do [ $n -le $NG ] && run_bench ours $n 64 1000 50 ours_b64_n$n; done
for n in 1 2 4 8; do [ $n -le $NG ] && run_bench reference $n 64 300 30 ref_b64_n$n; done
# 3. per-rank batch 4096 (config 3)
for n in 1 $NG; do run_bench ours $n 4096 30 5 ours_b4096_n$n; run_bench reference $n 4096 30 5 ref_b4096_n$n; done
# 4. widened CNN (config 4)
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
  tools/bench_wide.py --batch 256 --steps 30 > gpurun_out/wide_w$NG.json 2> gpurun_out/wide_w$NG.err
echo "wide rc=$?"; cat gpurun_out/wide_w$NG.json | cut -c1-600; tail -3 gpurun_out/wide_w$NG.err | cut -c1-300
# 5. all-reduce sweep (config 5), optional
if [ "${1:-}" = "sweep" ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
    tools/allreduce_sweep.py --max-mb 1024 --out gpurun_out/sweep_w$NG.json > gpurun_out/sweep_w$NG.txt 2> gpurun_out/sweep_w$NG.err
  echo "sweep rc=$?"; tail -32 gpurun_out/sweep_w$NG.txt
fi
