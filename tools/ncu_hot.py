"""Top stall locations of one kernel in an .ncu-rep (reads the source page; no GPU needed).
    python tools/ncu_hot.py gpurun_out/prof.ncu-rep conv1_fwd_tc_kernel [top=25] [view=sass|cuda] [skip=0]
(`skip` selects the (skip+1)-th captured launch with that base name, e.g. the dgrad instantiation of
conv3x3_tc_kernel is launch 1 when the forward one is launch 0.)
Prints the SASS (or CUDA source, needs -lineinfo + --import-source on) lines with the most warp
stall samples, their share, and instruction counts, plus the stall-reason totals."""
import csv
import subprocess
import sys


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    view = sys.argv[4] if len(sys.argv) > 4 else "sass"
    skip = sys.argv[5] if len(sys.argv) > 5 else "0"
    cmd = ["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kernel, "--launch-skip", skip,
           "--launch-count", "1"]
    if view == "cuda":
        cmd += ["--print-source", "cuda"]
    raw = subprocess.run(cmd, capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hi = next(i for i, r in enumerate(rows) if r and r[0] in ("Address", "#", "Line"))
    hdr = rows[hi]
    idx = {h: i for i, h in enumerate(hdr)}
    body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    samp = "Warp Stall Sampling (All Samples)"
    tot = sum(int(r[idx[samp]] or 0) for r in body) or 1
    print(f"{kernel}: {len(body)} lines, {tot} samples")
    reasons = [h for h in hdr if h.startswith("stall_")]
    if reasons:
        agg = sorted(((sum(int(r[idx[h]] or 0) for r in body), h) for h in reasons), reverse=True)[:8]
        print("stall reasons:", ", ".join(f"{h[6:]} {100 * v / tot:.1f}%" for v, h in agg if v))
    order = sorted(range(len(body)), key=lambda i: -int(body[i][idx[samp]] or 0))[:top]
    for i in sorted(order):
        r = body[i]
        n = int(r[idx[samp]] or 0)
        why = ""
        if reasons:
            rs = sorted(((int(r[idx[h]] or 0), h[6:]) for h in reasons), reverse=True)[:2]
            why = " ".join(f"{h}:{v}" for v, h in rs if v)
        print(f"{i:5d} {100 * n / tot:5.1f}%  x{r[idx['Instructions Executed']]:>9s}  {r[idx['Source']].strip()[:90]:90s} {why}")


if __name__ == "__main__":
    main()
