import sys, torch
sys.path.insert(0, '/root/repo')
from riggs_b200.optim import compact_rows
N = 300000
shapes = [(N, 3), (N, 1, 3), (N, 15, 3), (N, 1), (N, 3), (N, 4)]
ts = []
for s in shapes:
    for _ in range(3):
        ts.append(torch.randn(*s, device="cuda"))
ts += [torch.rand(N, 1, device="cuda"), torch.rand(N, 1, device="cuda"), torch.rand(N, device="cuda")]
keep = torch.rand(N, device="cuda") > 0.1
def bench(fn, k=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / k
print("tensors", len(ts), "bytes", sum(t.numel() * 4 for t in ts))
print("riggs compact_rows ms", bench(lambda: compact_rows(keep, ts)))
print("torch t[keep] x21 ms", bench(lambda: [t[keep] for t in ts]))
