import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
from ddp_b200.models.simple_cnn import SimpleCNN, SimpleCNNB200
cuda = torch.device("cuda", 0)
cap = int(sys.argv[1]); sched = [int(a) for a in sys.argv[2:]]
torch.manual_seed(1)
oracle = SimpleCNN(device=cuda)
gpu = SimpleCNNB200(batch_cap=cap, device=cuda, init=False)
gpu.load_state_dict(oracle.state_dict())
torch.manual_seed(2)
xs = torch.rand(8, cap, 1, 28, 28, device=cuda); ys = torch.randint(0, 10, (8, cap), device=cuda)
names = ("fl.bias", "fl.weight", "net.2.bias", "net.2.weight", "net.0.bias", "net.0.weight")
for i, b in enumerate(sched):
    x, y = xs[i % 8, :b].contiguous(), ys[i % 8, :b].contiguous()
    # same weights in both models before every step: isolates the per-step gradient error
    gpu.load_state_dict(oracle.state_dict())
    lo = oracle.forward_backward(x, y)
    lg = gpu.forward_backward(x, y)
    torch.cuda.synchronize()
    errs = []
    for n in names:
        g, w = gpu.arena.grad_view(n), oracle.arena.grad_view(n)
        errs.append(f"{n}={((g - w).norm() / w.norm()).item():.3f}")
    print(f"step {i} b={b} loss {lo.item():.4f}/{lg.item():.4f} " + " ".join(errs), flush=True)
    oracle.sgd_step(0.05)
