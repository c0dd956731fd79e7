#!/bin/bash
# Launch list (duration + DRAM bytes per kernel) and ONE ncu --set full capture of the native kernels of one
# widened-CNN training step (20 launches: conv1, packs, GEMMs, split-K reduce, column sums, bias+ReLU+max-pool, the small-output head).
# Run under gpurun, 1 GPU.   Usage: bash tools/run_ncu_wide.sh [batch=256] [list|full (default: both)]
set -u
B=${1:-256}
MODE=${2:-both}
mkdir -p gpurun_out
timeout 300 python tools/wide_step.py --batch $B --steps 4 > gpurun_out/plain_wide.log 2>&1
rc=$?; echo "plain rc=$rc"; [ $rc -eq 0 ] || exit $rc
K='regex:^(gemm_split|pack_rows|splitk|colsum|relu_maxpool|conv3x3_small_cin|small_linear|reduce_parts32)'
N=20  # native launches per step
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none \
  -k "$K" --launch-skip $((3 * N)) -c $N --csv --log-file gpurun_out/launches_wide.csv \
  python tools/wide_step.py --batch $B --steps 4 > gpurun_out/ncu_lwide.log 2>&1
echo "ncu list rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/launches_wide.csv")) if len(r)>5]
hdr=None; acc={}
for r in rows:
    if r[0]=="ID": hdr=r; continue
    if hdr and r[0].isdigit():
        d=dict(zip(hdr,r)); k=(int(d["ID"]), d["Kernel Name"].split("(")[0][-52:])
        acc.setdefault(k,{})[d["Metric Name"]]=(float(d["Metric Value"].replace(",","")), d["Metric Unit"])
for (i,n),m in sorted(acc.items()):
    t=m.get("gpu__time_duration.sum",(0,""))
    rd=m.get("dram__bytes_read.sum",(0,"")); wr=m.get("dram__bytes_write.sum",(0,"")); l2=m.get("lts__t_bytes.sum",(0,""))
    print(f"{i:3d} {n:52s} {t[0]:9.1f} {t[1]:3s} dram rd {rd[0]:8.1f} {rd[1]:5s} wr {wr[0]:8.1f} {wr[1]:5s} l2 {l2[0]:8.1f} {l2[1]}")
PY
[ "$MODE" = list ] && exit 0
timeout 900 ncu --set full --clock-control none --import-source on -k "$K" --launch-skip $((3 * N)) -c $N -f \
  -o gpurun_out/wide_fc python tools/wide_step.py --batch $B --steps 4 > gpurun_out/ncu_wide.log 2>&1
echo "ncu full rc=$?"; grep -E "WARN|ERROR" gpurun_out/ncu_wide.log | tail -3 | cut -c1-200; ls -la gpurun_out/wide_fc.ncu-rep
