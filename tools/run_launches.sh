#!/bin/bash
# Per-kernel durations of two eagerly launched steps (ONE ncu pass per kernel; run under gpurun, 1 GPU).
# Usage: bash tools/run_launches.sh [batch=4096]
set -u
B=${1:-4096}
mkdir -p gpurun_out
timeout 300 python bench.py --batch_size $B --steps 2 --warmup 3 --no_graph > gpurun_out/plain_l$B.log 2>&1
rc=$?; echo "plain eager rc=$rc"; [ $rc -eq 0 ] || exit $rc
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
  -k 'regex:^(conv|linear|pack|reduce|sgd|softmax|wgrad|zero)' --launch-skip 42 -c 28 --csv \
  --log-file gpurun_out/launches$B.csv python bench.py --batch_size $B --steps 2 --warmup 3 --no_graph > gpurun_out/ncu_l$B.log 2>&1
echo "ncu rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/launches$B.csv")) if len(r)>5]
hdr=None; acc={}
for r in rows:
    if r[0]=="ID": hdr=r; continue
    if hdr and r[0].isdigit():
        d=dict(zip(hdr,r)); k=(int(d["ID"]), d["Kernel Name"].split("(")[0][-44:])
        acc.setdefault(k,{})[d["Metric Name"]]=(float(d["Metric Value"].replace(",","")), d["Metric Unit"])
for (i,n),m in sorted(acc.items())[:14]:
    t=m.get("gpu__time_duration.sum",(0,""))
    rd=m.get("dram__bytes_read.sum",(0,"")); wr=m.get("dram__bytes_write.sum",(0,""))
    print(f"{i:3d} {n:44s} {t[0]:10.1f} {t[1]:4s} rd {rd[0]:8.1f} {rd[1]:6s} wr {wr[0]:8.1f} {wr[1]}")
PY
