"""tcgen05.mma issue-cost microbenchmark (one CTA, one issuing thread, 32 MMAs unrolled).
Prints cycles per MMA for the shapes/majors the conv kernels use.   python tools/mma_bench.py"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from ddp_b200.ops import native as nat

lib = nat.lib()
I, U, P = ctypes.c_int, ctypes.c_uint, ctypes.c_void_p
lib.ddpb_mma_microbench.restype = I
lib.ddpb_mma_microbench.argtypes = [I, I, I, I, U, U, U, U, P, P]
out = torch.zeros(2, dtype=torch.int64, device="cuda")
OUTER = 16
PL = 3072  # stage plane bytes (192 rows x 16 B)
# config -> (label, M, N)
CFG = {0: ("K/K", 128, 256), 1: ("K/K", 128, 128), 2: ("K/K", 128, 64), 3: ("K/K", 128, 32),
       4: ("K/K", 128, 16), 5: ("K/K one accumulator", 128, 64), 6: ("MN/MN", 64, 32),
       7: ("MN/MN", 64, 96), 8: ("MN/MN", 64, 64), 9: ("MN/MN", 64, 8), 10: ("MN/MN", 128, 32),
       11: ("MN/MN one accumulator", 64, 32), 12: ("K/K", 64, 32), 13: ("K/K", 64, 256)}


def run(cfg, ash, bsh, la, sa, lb, sb, note=""):
    label, M, N = CFG[cfg]
    for _ in range(2):
        nat.check(lib.ddpb_mma_microbench(cfg, ash, bsh, OUTER, la, sa, lb, sb, out.data_ptr(), None), label)
        torch.cuda.synchronize()
    issue, total = out.tolist()
    n = OUTER * 32
    print(f"{label:24s} M={M:3d} N={N:3d} {note:22s} issue {issue / n:6.1f}  complete {total / n:6.1f} cyc/mma"
          f"  (floor {max(M, 128) * N / 256:5.1f})", flush=True)


for cfg in (0, 1, 2, 3, 4):
    N = CFG[cfg][2]
    run(cfg, 40, 0, PL, 128, N * 16, 128, "A aligned")
run(2, 49, 0, PL, 128, 64 * 16, 128, "A shifted 9 rows")
run(5, 49, 0, PL, 128, 64 * 16, 128, "A shifted 9 rows")
run(3, 49, 0, PL, 128, 32 * 16, 128, "A shifted (dgrad)")
run(12, 40, 0, PL, 128, 32 * 16, 128, "M=64 K-major")
run(13, 40, 0, PL, 128, 256 * 16, 128, "M=64 K-major")
for cfg in (6, 8, 7):
    run(cfg, 0, 40, 128, 2048, 128, PL, "B aligned")
run(6, 0, 49, 128, 2048, 128, PL, "B shifted 9 rows")
run(7, 0, 49, 128, 2048, 128, PL, "B shifted 9 rows")
run(11, 0, 49, 128, 2048, 128, PL, "B shifted 9 rows")
run(10, 0, 49, 128, 2048, 128, PL, "M=128 (16 planes)")
run(9, 0, 0, 128, 2048, 128, 128, "bias block")
