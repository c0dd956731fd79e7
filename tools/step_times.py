"""Warm per-kernel latency of the step's pieces: each piece launched back to back REPS times between
two CUDA events on one stream (so launch overhead overlaps execution and the figure is the kernel's
own duration at steady state, L2 warm).   python tools/step_times.py [batch] [precision]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddp_b200.models.simple_cnn import SimpleCNNB200  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
prec = sys.argv[2] if len(sys.argv) > 2 else "bf16x3"
REPS = 200
dev = torch.device("cuda", 0)
m = SimpleCNNB200(batch_cap=B, device=dev, precision=prec)
nat = m.nat
x = torch.rand(B, 1, 28, 28, device=dev)
y = torch.randint(0, 10, (B,), device=dev)
m.forward_backward(x, y)
torch.cuda.synchronize()
st = torch.cuda.current_stream()
from ddp_b200.models.simple_cnn import C1, C2, IMG_H as H, IMG_W as W, NCLS  # noqa: E402
gs = m.grad_sets[0]
cap = B
pieces = []
if m.x3:
    pieces += [
        ("conv1 fwd", lambda: nat.conv1_fwd_tc_split(x, m.w1, m.b1, m.act1, B, cap, H, W)),
        ("conv2 fwd", lambda: nat.conv3x3_fwd_split(m.act1, m.w2_packed, m.b2, m.act2, C1, C2, B, cap, H, W)),
        ("linear fwd", lambda: nat.linear_fwd_f32(m.act2, m.wf, m.lin_partial, NCLS, C2, B, cap, H, W)),
    ]
else:
    pieces += [
        ("conv1 fwd", lambda: nat.conv1_fwd_tc(x, m.w1, m.b1, m.act1, B, cap, H, W)),
        ("conv2 fwd", lambda: nat.conv3x3_fwd(m.act1, m.w2, m.b2, m.act2, B, cap, H, W)),
        ("linear fwd", lambda: nat.linear_fwd_tc(m.act2, m.lin_packed, m.lin_partial, NCLS, C2, B, cap, H, W)),
    ]
S = pieces[2][1]()
pieces += [
    ("softmax+loss", lambda: nat.softmax_xent(m.lin_partial, S, m.bf, y, m.logits, m.dlogits, gs["bf"], m.loss,
                                              NCLS, B, dl_packed=None if m.x3 else m.dl_packed)),
    ("linear bwd", lambda: m.backward_head(B, None, 0, defer_reduce=True) if m.x3 else m.backward_head(B)),
    ("fl dW reduce", lambda: m.backward_head_reduce(None, 0)),
    ("conv2 wgrad partials", lambda: m.backward_conv2_wgrad_partials(B)),
    ("conv2 dgrad", lambda: m.backward_dgrad(B)),
    ("conv1 wgrad partials", lambda: m.backward_conv1_wgrad_partials(x, B)),
    ("sgd (all params)", lambda: m.sgd_step()),
]
print(f"batch {B} precision {prec}")
total = 0.0
for name, fn in pieces:
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(REPS):
        fn()
    e1.record(st)
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / REPS
    total += us
    print(f"  {name:24s} {us:8.2f} us")
print(f"  {'sum':24s} {total:8.2f} us")
