"""Regenerate the measured-results section of BASELINE.md from profiles/results/*.json
(the JSON lines bench.py printed on the B200 box).   python tools/make_results_table.py"""
import glob
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RES = os.path.join(ROOT, "profiles", "results")


def load(path):
    for line in open(path):
        if line.startswith("{"):
            last = json.loads(line)
    return last


def main():
    rows = {}
    for f in sorted(glob.glob(os.path.join(RES, "*_b*_n*.json"))):
        try:
            d = load(f)
        except Exception:
            continue
        key = (d["config"]["per_rank_batch"], d["n_gpus"])
        rows.setdefault(key, {})[d["impl"]] = d
    out = []
    out.append("| per-rank batch | GPUs | ours ms/step | ours img/s (device-timed) | ours img/s (end to end) | "
               "reference ms/step | reference img/s | reference e2e img/s | speed-up (device / e2e) | scaling eff. ours |")
    out.append("|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|")
    base = {}
    for (b, n), impls in sorted(rows.items()):
        o, r = impls.get("ours"), impls.get("reference")
        if o and n == 1:
            base[b] = o["value"]
        eff = f"{o['value'] / (n * base[b]):.2f}" if o and b in base else ""
        sp = (f"{o['value'] / r['value']:.1f}x / {o['e2e']['value'] / r['e2e']['value']:.1f}x" if o and r else "")
        out.append(
            f"| {b} | {n} | " +
            (f"{o['ms_per_step']:.4f} | {o['value']:,.0f} | {o['e2e']['value']:,.0f} | " if o else " | | | ") +
            (f"{r['ms_per_step']:.4f} | {r['value']:,.0f} | {r['e2e']['value']:,.0f} | " if r else " | | | ") +
            f"{sp} | {eff} |")
    table = "\n".join(out)
    sweep = ""
    sp = os.path.join(RES, "sweep_w8.json")
    if os.path.exists(sp):
        s = json.load(open(sp))
        lines = ["| dtype | bytes | NCCL µs | ours one-shot µs | ours two-shot µs | bus BW ours / NCCL (GB/s) |",
                 "|---|---:|---:|---:|---:|---:|"]
        for r in s["allreduce"]:
            lines.append(f"| {r['dtype']} | {r['bytes']:,} | {r['nccl_us']:.1f} | {r['oneshot_us']:.1f} | "
                         f"{r['twoshot_us']:.1f} | {r['busbw_gbs_ours']:.0f} / {r['busbw_gbs_nccl']:.0f} |")
        lines += ["", "| bucket bytes | NCCL all-reduce + ÷W + foreach SGD µs | fused one-shot µs | fused two-shot µs | "
                  "NVLink roofline µs (770 GB/s/dir) | fraction of roofline |", "|---:|---:|---:|---:|---:|---:|"]
        for r in s["fused"]:
            lines.append(f"| {r['bytes']:,} | {r['nccl_div_sgd_us']:.1f} | {r['fused_oneshot_us']:.1f} | "
                         f"{r['fused_twoshot_us']:.1f} | {r['roofline_us']:.1f} | {r['fraction_of_roofline']:.2f} |")
        sweep = "\n".join(lines)
    wide = ""
    wp = os.path.join(RES, "wide_w8.json")
    if os.path.exists(wp):
        w = load(wp)
        wide = (f"| {w['world']} | {w['per_rank_batch']} | {w['params']:,} | {w['buckets_mb']} | "
                f"{w['ours_ms_per_step']:.3f} | {w['ours_img_s']:,.0f} | {w['torch_ddp_ms_per_step']:.3f} | "
                f"{w['torch_ddp_img_s']:,.0f} |")
        wide = ("| GPUs | per-rank batch | params | bucket sizes (MiB) | ours ms/step | ours img/s | torch DDP ms/step | "
                "torch DDP img/s |\n|---:|---:|---:|---|---:|---:|---:|---:|\n" + wide)
    path = os.path.join(ROOT, "BASELINE.md")
    text = open(path).read()
    block = ("<!-- RESULTS:BEGIN -->\n### 3.1 SimpleCNN training throughput (configs 2 and 3)\n\n" + table +
             "\n\n### 3.2 All-reduce sweep, 8 GPUs (config 5)\n\n" + sweep +
             "\n\n### 3.3 Widened CNN, 8 GPUs (config 4)\n\n" + wide + "\n<!-- RESULTS:END -->")
    if "<!-- RESULTS:BEGIN -->" in text:
        text = re.sub(r"<!-- RESULTS:BEGIN -->.*<!-- RESULTS:END -->", lambda m: block, text, flags=re.S)
    else:
        text = text.replace("## 4. Graphics", block + "\n\n## 4. Graphics")
    open(path, "w").write(text)
    print(table)


if __name__ == "__main__":
    main()
