"""Summarise an .ncu-rep (read here with `ncu -i ... --page raw --csv`, no GPU needed) into a
markdown table under profiles/.   python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/ncu_b4096.md "title"
"""
import csv
import json
import os
import subprocess
import sys

METRICS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM % of peak"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe % active"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__grid_size", "grid"),
    ("smsp__inst_executed.sum", "warp instrs"),
]


def main():
    rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    peaks = {}
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    with open(out, "w") as f:
        f.write(f"# {title}\n\n")
        f.write(f"Source: `{os.path.basename(rep)}` captured with `ncu --set full --clock-control none "
                f"--import-source on` on one B200 (cold caches, serialised: compare shares, not absolutes). "
                f"Read with `ncu -i ... --page raw --csv`.\n\n")
        if peaks:
            f.write(f"Roofline denominators (MEASURED_PEAKS.json, measured): HBM copy {peaks.get('hbm_gbs')} GB/s, "
                    f"bf16 cuBLAS {peaks.get('bf16_tflops')} TFLOP/s burst / {peaks.get('bf16_tflops_sustained')} sustained.\n\n")
        f.write("| kernel | " + " | ".join(n for _, n in METRICS) + " | achieved GB/s (of measured HBM) |\n")
        f.write("|---|" + "---:|" * (len(METRICS) + 1) + "\n")
        for r in rows[2:]:
            name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")
            vals = []
            for m, _ in METRICS:
                if m in idx:
                    v = r[idx[m]]
                    try:
                        v = f"{float(v):.4g}"
                    except ValueError:
                        pass
                    vals.append(f"{v} {units[idx[m]]}".strip())
                else:
                    vals.append("n/a")
            try:
                t_us = float(r[idx["gpu__time_duration.sum"]])
                scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}
                rd = float(r[idx["dram__bytes_read.sum"]]) * scale[units[idx["dram__bytes_read.sum"]]]
                wr = float(r[idx["dram__bytes_write.sum"]]) * scale[units[idx["dram__bytes_write.sum"]]]
                gbs = (rd + wr) / (t_us * 1e-6) / 1e9
                frac = f"{gbs:.0f} ({gbs / peaks['hbm_gbs']:.2f})" if peaks else f"{gbs:.0f}"
            except Exception:
                frac = "n/a"
            f.write(f"| `{name}` | " + " | ".join(vals) + f" | {frac} |\n")
    print("wrote", out)


if __name__ == "__main__":
    main()
