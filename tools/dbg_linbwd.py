import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddp_b200.ops import native as nat, layout as lay
H = W = 28
batch, cap = int(sys.argv[1]), int(sys.argv[2])
C = 64; K = C * H * W
torch.manual_seed(0)
act = torch.relu(torch.randn(batch, C, H, W, device="cuda"))
Wfl = torch.randn(10, K, device="cuda") * 0.01
bfl = torch.zeros(10, device="cuda")
y = torch.randint(0, 10, (batch,), device="cuda")
act_p = lay.nchw_to_planes(act, cap)
packed = torch.empty(nat.linear_pack_bytes(10, C, H, W), dtype=torch.uint8, device="cuda")
def step(name, fn):
    fn(); torch.cuda.synchronize(); print("ok", name, flush=True)
step("pack", lambda: nat.linear_pack(Wfl, packed, 10, C, H, W))
part = nat.scratch(nat.linear_fwd_tc_scratch_bytes(10, C, cap, H, W))
S = [0]
def f(): S[0] = nat.linear_fwd_tc(act_p, packed, part, 10, C, batch, cap, H, W)
step("fwd", f)
logits = torch.zeros(batch, 10, device="cuda"); dl = torch.zeros(batch, 10, device="cuda")
dbias = torch.zeros(10, device="cuda"); loss = torch.zeros(1, device="cuda")
dlp = torch.zeros(nat.dl_packed_bytes(cap), dtype=torch.uint8, device="cuda")
step("softmax plain", lambda: nat.softmax_xent(part, S[0], bfl, y, logits, dl, dbias, loss, 10, batch))
step("softmax packed", lambda: nat.softmax_xent(part, S[0], bfl, y, logits, dl, dbias, loss, 10, batch, dl_packed=dlp))
dact = nat.alloc_planes(C, cap, H, W)
dW = torch.zeros(10, K, device="cuda")
step("bwd", lambda: nat.linear_bwd_tc(act_p, dlp, packed, dact, dW, 10, C, batch, cap, H, W))
print("all ok")
