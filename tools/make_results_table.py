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
        if f.endswith("_nvls.json"):
            continue
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
    # multicast communicator variants and the per-bucket ablation (timing diagnostics)
    extra = ["| GPUs | communicator | ms/step | img/s |", "|---:|---|---:|---:|"]
    for n in (2, 4, 8):
        for tag, name in (("", "CUDA IPC, peer loads/stores (default)"), ("_nvls", "VMM + multicast, multimem (DDPB_NVLS=1)")):
            fp = os.path.join(RES, f"ours_b64_n{n}{tag}.json")
            if os.path.exists(fp):
                d = load(fp)
                extra.append(f"| {n} | {name} | {d['ms_per_step']:.4f} | {d['value']:,.0f} |")
    abl = ["| 4-GPU run | ms/step |", "|---|---:|"]
    for tag, name in (("n1", "1 GPU, same box"), ("no_fl", "4 GPUs, fl.* update skipped"),
                      ("no_conv2", "4 GPUs, net.2.* update skipped"), ("no_conv1", "4 GPUs, net.0.* update skipped")):
        fp = os.path.join(RES, f"ablate4_{tag}.json")
        if os.path.exists(fp):
            abl.append(f"| {name} | {load(fp)['ms_per_step']:.4f} |")
    kt = []
    import csv
    for b in (64, 4096):
        fp = os.path.join(RES, f"kernel_times_b{b}.csv")
        if not os.path.exists(fp):
            continue
        rows_ = [r for r in csv.reader(open(fp)) if len(r) > 5]
        hdr, acc = None, {}
        for r in rows_:
            if r[0] == "ID":
                hdr = r
                continue
            if hdr and r[0].isdigit():
                d = dict(zip(hdr, r))
                k = (int(d["ID"]), d["Kernel Name"].split("(")[0].replace("void ", "").replace("ddpb::", ""))
                acc.setdefault(k, {})[d["Metric Name"]] = float(d["Metric Value"].replace(",", ""))
        kt.append((b, sorted(acc.items())[:13]))
    ktab = []
    if kt:
        ktab = ["| kernel (one step, eager, serialised under ncu) | " + " | ".join(f"B={b}: µs" for b, _ in kt) +
                " | " + " | ".join(f"B={b}: DRAM MB" for b, _ in kt) + " |", "|---|" + "---:|" * (2 * len(kt))]
        for i in range(13):
            name = kt[0][1][i][0][1]
            us = [f"{k[1][i][1].get('gpu__time_duration.sum', 0) / 1e3:.1f}" for k in kt]
            mb = [f"{(k[1][i][1].get('dram__bytes_read.sum', 0) + k[1][i][1].get('dram__bytes_write.sum', 0)) / 1e6:.1f}" for k in kt]
            ktab.append(f"| `{name}` | " + " | ".join(us) + " | " + " | ".join(mb) + " |")
    path = os.path.join(ROOT, "BASELINE.md")
    text = open(path).read()
    notes = ("\n\nRows come from different gpurun boxes and points of the round (GPU time ran out before one last "
             "full sweep): ours N=1/2/8 at per-rank batch 64 and N=1 at 4096 are the final code; ours N=4 predates "
             "the fused conv2-bucket update; ours 4096 x 8 GPUs predates the conv1 / linear-backward kernel changes "
             "(its own N=1 was 1.2408 ms then, i.e. 0.95 efficiency, not the 0.77 the last column derives from the "
             "newer N=1).  The reference rows are unaffected by our changes.  The driver's end-of-round scaling "
             "run is the authoritative set.  ALL rows of ours were measured with the net.2 bucket update ordered only "
             "after the weight gradient (now `DDPB_NET2_ORDER=legacy`); the default since then also waits for the "
             "data gradient (DESIGN.md §5) and has not been measured.")
    block = ("<!-- RESULTS:BEGIN -->\n### 3.1 SimpleCNN training throughput (configs 2 and 3)\n\n" + table + notes +
             "\n\n### 3.2 All-reduce sweep, 8 GPUs (config 5)\n\n" + sweep +
             "\n\n### 3.3 Widened CNN, 8 GPUs (config 4)\n\n" + wide +
             "\n\n### 3.4 Communicator variants and where the scaling overhead sits (per-rank batch 64)\n\n" +
             "\n".join(extra) + "\n\n" + "\n".join(abl) +
             "\n\nSkipping any single bucket's update buys at most 2 µs: the three update paths (streams) are "
             "balanced, so the 4-GPU overhead (+14 µs over one GPU) is the latency of one exchange, paid once, "
             "not one slow bucket.\n\n### 3.5 Per-kernel times of one step\n\n"
             "Captured with `tools/run_launches.sh` before the conv2 split-K reduce was folded into its update "
             "kernel (rows `wgrad_reduce_kernel` + the second `sgd_kernel` are now one "
             "`reduce_allreduce_sgd_ll_kernel<W, 1>`).\n\n" + "\n".join(ktab) + "\n<!-- RESULTS:END -->")
    if "<!-- RESULTS:BEGIN -->" in text:
        text = re.sub(r"<!-- RESULTS:BEGIN -->.*<!-- RESULTS:END -->", lambda m: block, text, flags=re.S)
    else:
        text = text.replace("## 4. Graphics", block + "\n\n## 4. Graphics")
    open(path, "w").write(text)
    print(table)


if __name__ == "__main__":
    main()
