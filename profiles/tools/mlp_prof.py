"""Reads the cycle counters a tuning build (build.py --variant prof -DRG_MLP_PROF=1, RG_LIB=...) of the MLP chain kernel
leaves for CTA 0's second tile: per (layer, quarter) what the issuing warp and one epilogue warp waited for."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from riggs_b200 import _lib, skeleton_warp as sw
from riggs_b200.mlp import FusedSkipMLP
torch.manual_seed(0)
wm = sw.WeightMLP(3, 23).cuda(); fw = FusedSkipMLP(wm)
x = (torch.randn(300_000, 3) * 0.6).cuda()
with torch.no_grad():
    fw(x); fw(x)
torch.cuda.synchronize()
buf = (C.c_ulonglong * 2048)()
L = _lib.lib()
L.rg_debug_mlp_prof_read.restype = C.c_int
assert L.rg_debug_mlp_prof_read(buf, 2048) == 0
a = np.frombuffer(buf, dtype=np.uint64).astype(np.int64)
iss = a[:36 * 8].reshape(36, 8); epi = a[1024:1024 + 36 * 8].reshape(36, 8)
t0 = iss[0, 0]
print("layer q | issuer: start  wait_acc_free wait_act wait_full total | epilogue: start wait_acc_full work")
for i in range(33):
    print(f"{i // 4} {i % 4} | {iss[i, 0] - t0:7d} {iss[i, 1]:6d} {iss[i, 2]:6d} {iss[i, 3]:6d} {iss[i, 4]:6d} | {epi[i, 0] - t0:7d} {epi[i, 1]:6d} {epi[i, 2]:6d}")
