"""tcgen05.mma issue-cost microbenchmark (one CTA, one issuing thread).
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
lib.ddpb_mma_microbench.argtypes = [I, I, I, I, I, I, I, I, U, U, U, U, P, P]
out = torch.zeros(2, dtype=torch.int64, device="cuda")
ITERS = 512


def run(name, M, N, a_mn, b_mn, ash, bsh, nacc, lbo_a, sbo_a, lbo_b, sbo_b):
    for _ in range(2):
        nat.check(lib.ddpb_mma_microbench(M, N, a_mn, b_mn, ash, bsh, nacc, ITERS, lbo_a, sbo_a, lbo_b,
                                          sbo_b, out.data_ptr(), None), name)
        torch.cuda.synchronize()
    issue, total = out.tolist()
    ideal = max(M, 128) * N / 256
    print(f"{name:52s} M={M:3d} N={N:3d} nacc={nacc}  issue {issue / ITERS:7.1f} cyc/mma   complete {total / ITERS:7.1f} cyc/mma"
          f"   (model floor {ideal:5.1f})", flush=True)


PL = 3328  # stage plane bytes (208 rows x 16 B)
# K-major A (planes), K-major B (weights: LBO = N*16, SBO = 128)
for N in (256, 128, 64, 32, 16):
    run(f"K/K  A aligned      ", 128, N, 0, 0, 40, 0, min(4, 512 // N), PL, 128, N * 16, 128)
run("K/K  A shifted +9 rows (tap)", 128, 64, 0, 0, 49, 0, 4, PL, 128, 64 * 16, 128)
run("K/K  A shifted +9, same accumulator", 128, 64, 0, 0, 49, 0, 1, PL, 128, 64 * 16, 128)
run("K/K  dgrad shape, shifted", 128, 32, 0, 0, 49, 0, 8, PL, 128, 32 * 16, 128)
# MN-major A (dout planes 2048 B) and B (in planes 3328 B): LBO = 128 (K groups), SBO = plane
for N in (32, 96, 128, 256):
    run(f"MN/MN wgrad M=64, B aligned", 64, N, 1, 1, 0, 40, min(8, 512 // N), 128, 2048, 128, PL)
run("MN/MN wgrad M=64 N=32, B shifted +9", 64, 32, 1, 1, 0, 49, 9, 128, 2048, 128, PL)
run("MN/MN wgrad M=64 N=32, B shifted, same acc", 64, 32, 1, 1, 0, 49, 1, 128, 2048, 128, PL)
run("MN/MN M=128 N=32 (two 64-ch halves)", 128, 32, 1, 1, 0, 49, 9, 128, 2048, 128, PL)
run("MN/MN M=64 N=8 (bias)", 64, 8, 1, 1, 0, 0, 8, 128, 2048, 128, 128)
run("MN A / K B  M=64 N=32", 64, 32, 1, 0, 0, 0, 8, 128, 2048, 32 * 16, 128)
run("K A / MN B  M=128 N=32", 128, 32, 0, 1, 40, 49, 8, PL, 128, 128, PL)
