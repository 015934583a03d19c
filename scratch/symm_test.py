import os, time, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank=int(os.environ["RANK"]); world=int(os.environ["WORLD_SIZE"]); local=int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev=torch.device("cuda",local)
os.environ.setdefault("NCCL_DEBUG","WARN")
dist.init_process_group("nccl", device_id=dev)
n=17700024
gname=dist.group.WORLD.group_name
try:
    symm_mem.enable_symm_mem_for_group(gname)
except Exception as e:
    print("enable:", e)
t=symm_mem.empty(n, dtype=torch.float32, device=dev)
hdl=symm_mem.rendezvous(t, gname)
if rank==0:
    print("multicast support", hdl.has_multicast_support, "mc ptr", hex(hdl.multicast_ptr), "offset", getattr(hdl,'offset',None), "buf size", hdl.buffer_size, "ptr0", hex(hdl.buffer_ptrs[0]), "tensor ptr", hex(t.data_ptr()))
ref=torch.empty(n, device=dev)
def fill():
    g=torch.Generator(device=dev).manual_seed(rank+1)
    t.copy_(torch.rand(n, generator=g, device=dev)); ref.copy_(t)
fill(); dist.all_reduce(ref); 
torch.ops.symm_mem.multimem_all_reduce_(t, "sum", gname); torch.cuda.synchronize()
print(rank, "multimem max abs diff vs nccl", float((t-ref).abs().max()))
def bench(fn, k=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/k
a=bench(lambda: dist.all_reduce(ref))
b=bench(lambda: torch.ops.symm_mem.multimem_all_reduce_(t, "sum", gname))
try:
    c=bench(lambda: torch.ops.symm_mem.two_shot_all_reduce_(t, "sum", gname))
except Exception as e:
    c=str(e)[:80]
import sys; sys.path.insert(0,'/root/repo')
from riggs_b200 import _lib
L=_lib.lib()
mc=hdl.multicast_ptr + (t.data_ptr()-hdl.buffer_ptrs[rank])
npad=(n+3)//4*4
def ours():
    hdl.barrier(channel=0)
    rc=L.rg_allreduce_multimem(mc, n - n%4, rank, world, torch.cuda.current_stream().cuda_stream)
    assert rc==0, L.rg_last_error()
    hdl.barrier(channel=1)
fill(); dist.all_reduce(ref); ours(); torch.cuda.synchronize()
print(rank, "own multimem kernel max abs diff vs nccl", float((t[:n-n%4]-ref[:n-n%4]).abs().max()))
d=bench(ours)
def ours_nobar():
    L.rg_allreduce_multimem(mc, n - n%4, rank, world, torch.cuda.current_stream().cuda_stream)
e=bench(ours_nobar)
if rank==0: print("own kernel with barriers ms", d, "kernel only", e)
import ctypes as C
arr=(C.c_void_p*world)(*[hdl.buffer_ptrs[r] + (t.data_ptr()-hdl.buffer_ptrs[rank]) for r in range(world)])
def p2p():
    hdl.barrier(channel=0)
    rc=L.rg_allreduce_p2p(arr, n - n%4, rank, world, torch.cuda.current_stream().cuda_stream)
    assert rc==0, L.rg_last_error()
    hdl.barrier(channel=1)
torch.cuda.synchronize(); dist.barrier(); fill(); torch.cuda.synchronize(); dist.barrier(); dist.all_reduce(ref); p2p(); torch.cuda.synchronize()
print(rank, "own p2p kernel max abs diff vs nccl", float((t[:n-n%4]-ref[:n-n%4]).abs().max()))
f=bench(p2p)
g=bench(lambda: L.rg_allreduce_p2p(arr, n - n%4, rank, world, torch.cuda.current_stream().cuda_stream))
h=bench(lambda: (hdl.barrier(channel=0), hdl.barrier(channel=1)))
if rank==0: print("p2p with barriers ms", f, "kernel only", g, "two barriers alone", h)
if rank==0: print("world",world,"nccl ms",a,"multimem ms",b,"two_shot",c)
dist.destroy_process_group()
