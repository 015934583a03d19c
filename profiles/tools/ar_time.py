"""Times the library's own all-reduce of a C0-sized gradient buffer (70.8 MB) under torchrun; env knobs RG_ALLREDUCE,
RG_AR_BLOCKS, RG_AR_UNROLL select the kernel variant."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
from riggs_b200.dist import FlatGrads
P = 300_000
params = [("xyz", torch.zeros(P, 3, device=dev)), ("f_dc", torch.zeros(P, 1, 3, device=dev)), ("f_rest", torch.zeros(P, 15, 3, device=dev)),
          ("opacity", torch.zeros(P, 1, device=dev)), ("scaling", torch.zeros(P, 3, device=dev)), ("rotation", torch.zeros(P, 4, device=dev))]
flat = FlatGrads(params, device=dev)
flat.flat.normal_()
for _ in range(5):
    flat.all_reduce()
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 50
e0.record()
for _ in range(n):
    flat.all_reduce()
e1.record(); torch.cuda.synchronize()
t = torch.tensor([e0.elapsed_time(e1) / n], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"world": world, "mode": os.environ.get("RG_ALLREDUCE", "default"), "blocks": os.environ.get("RG_AR_BLOCKS", ""),
                      "unroll": os.environ.get("RG_AR_UNROLL", ""), "ms_incl_two_barriers": round(float(t[0]), 4), "bytes": flat.total * 4}))
dist.destroy_process_group()
