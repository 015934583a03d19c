import sys, torch
sys.path.insert(0, '/root/repo')
from riggs_b200.optim import FusedAdam
N = 300000
shapes = {"xyz": (N, 3), "f_dc": (N, 1, 3), "f_rest": (N, 15, 3), "opacity": (N, 1), "scaling": (N, 3), "rotation": (N, 4)}
def mk():
    d = {k: torch.randn(*s, device="cuda").requires_grad_(True) for k, s in shapes.items()}
    for v in d.values(): v.grad = torch.randn_like(v)
    return d, [{"params": [d[k]], "lr": 1e-3, "name": k} for k in shapes]
def bench(opt, k=50):
    for _ in range(5): opt.step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k): opt.step()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / k
d, g = mk(); print("riggs FusedAdam ms", bench(FusedAdam(g, lr=0.0, eps=1e-15)))
d, g = mk(); print("torch Adam (default) ms", bench(torch.optim.Adam(g, lr=0.0, eps=1e-15)))
d, g = mk(); print("torch Adam (fused=True) ms", bench(torch.optim.Adam(g, lr=0.0, eps=1e-15, fused=True)))
n = sum(v.numel() for v in d.values()); print("elements", n, "bytes/step", n * 28, "-> at 6.5 TB/s:", n * 28 / 6.5e12 * 1e3, "ms")
