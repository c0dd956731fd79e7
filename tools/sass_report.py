"""Write profiles/sass_summary.md (+ full listings of the tensor-core and collective kernels under
profiles/sass/): per-kernel register/shared-memory use from the cubin and the SASS mnemonics that
prove the Blackwell path (UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UBLKCP = bulk async copy/TMA, UTMALDG = tensor-map TMA load,
SYNCS = mbarrier; HMMA would be the legacy mma.sync path and must be 0).
    python tools/sass_report.py        (no GPU needed)
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pytorch-distributed-data-parallel-ddp-trainer_b200", "lib", "libddp_b200.so")
OUT = os.path.join(ROOT, "profiles")
KEEP_FULL = ("conv3x3_tc_kernel", "conv3x3_wgrad_tc_kernel", "allreduce_sgd_oneshot_kernelILi8E",
             "allreduce_sgd_twoshot_kernelILi8E", "allreduce_sgd_nvls_kernelILi8E",
             "reduce_allreduce_sgd_ll_kernelILi8ELi0E", "reduce_allreduce_sgd_ll_kernelILi8ELi1E",
             "linear_bwd_tc_kernel", "linear_fwd_tc_kernel", "conv1_fwd_tc_kernel", "conv1_wgrad_tc_kernel",
             "conv1_wgrad_kernelILi1ELi32ELb1E", "linear_fwd_f32_kernel", "linear_bwd_f32_kernelILi10ELb1ELb1ELb0E",
             "linear_bwd_f32_kernelILi10ELb1ELb0ELb1E", "softmax_xent_kernel", "pack_linear_weights_kernel",
             "pack_conv3x3_split_kernel", "allreduce_kernelIfLi8ELi1E",
             # general linear layers (gemm_tc.cu): the three operand-major pairs at N tile 256, tensor-map TMA
             # (template arguments: A_MN, B_MN, BN, STAGES, TM, MT, CL)
             "gemm_split_tc_kernelILb0ELb0ELi256ELi4ELb1ELi1ELi1E", "gemm_split_tc_kernelILb1ELb0ELi256ELi2ELb1ELi1ELi1E",
             "gemm_split_tc_kernelILb1ELb1ELi256ELi2ELb1ELi1ELi1E",
             "gemm_split_tc_kernelILb1ELb1ELi256ELi2ELb1ELi1ELi2E",   # cluster of two, multicast B tile (opt-in)
             "gemm_split_tc_kernelILb1ELb1ELi256ELi3ELb1ELi2ELi1E",   # two M tiles per CTA (opt-in)
             "small_linear_fwd_kernel", "small_linear_dw_kernel", "pack_rows_split_kernel",
             "conv3x3_small_cin_fwd_kernel", "conv3x3_small_cin_wgrad_kernel", "relu_maxpool2_nhwc_fwd_kernel")
KEYS = ["UTCHMMA", "UTCBAR", "LDTM", "UBLKCP", "UTMALDG", "LDGSTS", "SYNCS", "HMMA", "FFMA", "LDGMC", "LDG", "STG",
        "LDS", "STS", "SHFL", "ATOM", "RED", "MEMBAR", "CCTL"]


def demangle(name):
    try:
        return subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip() or name
    except Exception:
        return name


def short_name(fn):
    """Demangled name without the parameter list, template arguments kept."""
    name = demangle(fn)
    if name.endswith(")"):
        depth = 0
        for i in range(len(name) - 1, -1, -1):
            if name[i] == ")":
                depth += 1
            elif name[i] == "(":
                depth -= 1
                if depth == 0:
                    name = name[:i]
                    break
    return name.replace("void ", "").replace("ddpb::", "").replace("(int)", "")


def main():
    os.makedirs(os.path.join(OUT, "sass"), exist_ok=True)
    for old in os.listdir(os.path.join(OUT, "sass")):
        os.unlink(os.path.join(OUT, "sass", old))
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for line in res.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
            continue
        if cur and "REG:" in line:
            usage[cur] = line.strip()
            cur = None
    funcs = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        if cur is not None:
            funcs[cur].append(line)
    rows = []
    for fn, lines in funcs.items():
        cnt = collections.Counter()
        ninstr = 0
        for l in lines:
            m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", l)
            if not m:
                continue
            ninstr += 1
            op = m.group(1)
            for k in KEYS:  # first match wins: LDGSTS / LDGMC are listed before LDG
                if op.startswith(k):
                    cnt[k] += 1
                    break
        rows.append((fn, ninstr, cnt))
        if any(k in fn for k in KEEP_FULL):
            short = re.sub(r"[^A-Za-z0-9]+", "_", short_name(fn)).strip("_")
            with open(os.path.join(OUT, "sass", short + ".sass"), "w") as f:
                f.write(f"// {demangle(fn)}\n// {usage.get(fn, '')}\n")
                # keep the instruction text, drop the raw encoding words
                for l in lines:
                    if re.match(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", l):
                        continue
                    f.write(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l).rstrip() + "\n")
    with open(os.path.join(OUT, "sass_summary.md"), "w") as f:
        f.write("# SASS summary of libddp_b200.so (sm_100a)\n\n")
        f.write("Generated by `python tools/sass_report.py` from `cuobjdump -sass` / `-res-usage`.\n")
        f.write("`UTCHMMA` = tcgen05.mma, `LDTM` = tcgen05.ld, `UBLKCP` = cp.async.bulk (TMA), `SYNCS` = mbarrier, "
                "`UTCBAR` = tcgen05.commit, `LDGSTS` = cp.async, `LDGMC` = multimem.ld_reduce (in-switch reduction).  "
                "`HMMA` (legacy mma.sync) is 0 everywhere.\n\n")
        f.write("| kernel | instrs | " + " | ".join(KEYS) + " | resources |\n")
        f.write("|---|---:|" + "---:|" * len(KEYS) + "---|\n")
        for fn, n, cnt in rows:
            name = short_name(fn)
            f.write(f"| `{name}` | {n} | " + " | ".join(str(cnt.get(k, 0)) for k in KEYS) +
                    f" | {usage.get(fn, '').replace('|', '/')} |\n")
    print("wrote", os.path.join(OUT, "sass_summary.md"), "functions:", len(rows))


if __name__ == "__main__":
    main()
