"""Regenerate BASELINE.md §3 (between the RESULTS markers) from the JSON lines bench.py and the sweep tool
printed on the B200 boxes this round: profiles/results/r2/*.json.     python tools/results_r2.py
"""
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RES = os.path.join(ROOT, "profiles", "results", "r2")


def load(name):
    path = os.path.join(RES, name)
    if not os.path.exists(path):
        return None
    last = None
    for line in open(path):
        if line.startswith("{"):
            last = json.loads(line)
    return last


def throughput_table():
    out = ["| per-rank batch | GPUs | ours ms/step | ours img/s (device) | ours img/s (e2e) | reference ms/step | "
           "reference img/s (device) | reference img/s (e2e) | speed-up device / e2e | scaling eff. ours / ref | "
           "replicas bit-identical | DP-equivalence max err | final loss ours / ref |",
           "|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|---|---:|---|"]
    base = {}
    for b in (64, 4096):
        for n in (1, 2, 4, 8):
            o, r = load(f"ours_b{b}_n{n}.json"), load(f"ref_b{b}_n{n}.json")
            if not o:
                continue
            if n == 1:
                base[b] = (o["value"], r["value"] if r else None)
            eff_o = o["value"] / (n * base[b][0]) if b in base else float("nan")
            eff_r = (r["value"] / (n * base[b][1])) if (r and b in base and base[b][1]) else float("nan")
            c = o["config"]
            row = f"| {b} | {n} | {o['ms_per_step']:.4f} | {o['value']:,.0f} | {o['e2e']['value']:,.0f} | "
            if r:
                row += (f"{r['ms_per_step']:.4f} | {r['value']:,.0f} | {r['e2e']['value']:,.0f} | "
                        f"{o['value'] / r['value']:.2f}x / {o['e2e']['value'] / r['e2e']['value']:.2f}x | "
                        f"{eff_o:.2f} / {eff_r:.2f} | ")
            else:
                row += f" | | | | {eff_o:.2f} / - | "
            row += (f"{c.get('replicas_bit_identical')} | {c.get('dp_equiv_max_err', float('nan')):.1e} | "
                    f"{c.get('final_loss')} / {r['config'].get('final_loss') if r else '-'} |")
            out.append(row)
    return "\n".join(out)


def variants_table():
    out = ["| GPUs | variant | ms/step | img/s |", "|---:|---|---:|---:|"]
    for n in (1, 2, 4, 8):
        for tag, name in (("", "bf16x3 (default mode and communicator: multicast from 4 GPUs up, else CUDA IPC)"),
                          ("_ipc", "bf16x3, CUDA-IPC communicator forced (DDPB_NVLS=0)"),
                          ("_nvls", "bf16x3, VMM + multicast communicator forced (DDPB_NVLS=1)"),
                          ("_bf16", "bf16 fast mode (--dtype bf16; NOT the reference's precision)")):
            d = load(f"ours_b64_n{n}{tag}.json")
            if d:
                out.append(f"| {n} | {name} | {d['ms_per_step']:.4f} | {d['value']:,.0f} |")
    return "\n".join(out)


def sweep_tables():
    blocks = []
    for w in (2, 4, 8):
        path = os.path.join(RES, f"sweep_w{w}_mc.json")
        if not os.path.exists(path):
            continue
        s = json.load(open(path))
        lines = [f"World size {w}:", "",
                 "| dtype | bytes | NCCL µs | ours one-shot µs | ours two-shot µs | ours in-switch (multimem) µs | "
                 "best of ours / NCCL |", "|---|---:|---:|---:|---:|---:|---:|"]
        for r in s["allreduce"]:
            ours = [x for x in (r.get("oneshot_us"), r.get("twoshot_us"), r.get("nvls_us")) if x]
            nv_s = f"{r['nvls_us']:.1f}" if r.get("nvls_us") else "-"
            lines.append(f"| {r['dtype']} | {r['bytes']:,} | {r['nccl_us']:.1f} | {r['oneshot_us']:.1f} | "
                         f"{r['twoshot_us']:.1f} | {nv_s} | {min(ours) / r['nccl_us']:.2f} |")
        if s.get("fused"):
            lines += ["", "| bucket bytes | NCCL all-reduce + ÷W + foreach SGD µs | fused one-shot µs | "
                      "fused two-shot µs | fused in-switch µs | NVLink roofline µs (770 GB/s/dir) | "
                      "roofline / best fused |", "|---:|---:|---:|---:|---:|---:|---:|"]
            for r in s["fused"]:
                best = min(x for x in (r.get("fused_oneshot_us"), r.get("fused_twoshot_us"),
                                       r.get("fused_nvls_us")) if x)
                nv_s = f"{r['fused_nvls_us']:.1f}" if r.get("fused_nvls_us") else "-"
                lines.append(f"| {r['bytes']:,} | {r['nccl_div_sgd_us']:.1f} | {r['fused_oneshot_us']:.1f} | "
                             f"{r['fused_twoshot_us']:.1f} | {nv_s} | {r['roofline_us']:.1f} | "
                             f"{r['roofline_us'] / best:.2f} |")
        blocks.append("\n".join(lines))
    return "\n\n".join(blocks) if blocks else "(not measured this round)"


def main():
    block = ("<!-- RESULTS:BEGIN -->\n### 3.1 SimpleCNN training throughput (configs 2 and 3), round 2\n\n"
             "Ours = the DEFAULT precision mode (`bf16x3`: the reference's numerics or better on every op, "
             "`profiles/precision.md`).  Every rank trains on its own data stream (seed 1234 + rank, both arms); "
             "`replicas bit-identical` = all ranks' parameters compared after the run; `DP-equivalence max err` = "
             "max |param| difference between two steps of the N-rank engine and the same engine on one GPU over "
             "the concatenated batch.  SM clocks 1965 MHz with no throttle reason in every run.\n\n"
             + throughput_table() +
             "\n\n### 3.2 Precision modes and communicator variants (per-rank batch 64)\n\n" + variants_table() +
             "\n\n### 3.3 All-reduce sweep (config 5)\n\n" + sweep_tables() +
             "\n<!-- RESULTS:END -->")
    path = os.path.join(ROOT, "BASELINE.md")
    text = open(path).read()
    text = re.sub(r"<!-- RESULTS:BEGIN -->.*<!-- RESULTS:END -->", lambda m: block, text, flags=re.S)
    open(path, "w").write(text)
    print(throughput_table())
    print(variants_table())


if __name__ == "__main__":
    main()
