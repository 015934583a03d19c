"""Times the tensor-core MLP kernels alone (inference forward, training forward, backward) at a given size."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from riggs_b200 import skeleton_warp as sw
from riggs_b200.mlp import FusedSkipMLP

N = int(sys.argv[1]) if len(sys.argv) > 1 else 300_000
J = 24
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
torch.manual_seed(0)
wm = sw.WeightMLP(3, J - 1).cuda()
fw = FusedSkipMLP(wm)
x = (torch.randn(N, 3) * 0.6).cuda()


def timed(fn):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


with torch.no_grad():
    t_inf = timed(lambda: fw(x))
xg = x.clone()
def train():
    out = fw(xg)
    out.backward(torch.ones_like(out))
t_train = timed(train)
print(json.dumps({"N": N, "forward_inference_ms": round(t_inf, 4), "forward_plus_backward_ms": round(t_train, 4)}))
