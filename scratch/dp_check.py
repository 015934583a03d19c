import os, sys, torch, torch.distributed as dist
sys.path.insert(0, '/root/repo')
from riggs_b200 import synth
from riggs_b200.hotpath import HotPath, GraphedStep
rank=int(os.environ["RANK"]); world=int(os.environ["WORLD_SIZE"]); local=int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev=torch.device("cuda",local)
os.environ.setdefault("NCCL_DEBUG","WARN")
dist.init_process_group("nccl", device_id=dev)
scene = synth.make_config("C0", n_views=8)
hp = HotPath(scene, device=dev)
cam = scene.cameras[rank % len(scene.cameras)].to(dev)
g = torch.Generator().manual_seed(1000 + rank)
target = torch.rand(3, cam.image_height, cam.image_width, generator=g).to(dev)
pose = {k: v.to(dev).requires_grad_(True) for k, v in scene.poses[rank % len(scene.poses)].items()}
flat = hp.enable_flat_grads(); flat.overlap = True
print(rank, "peer mapped", flat.hdl is not None, "can overlap", flat.can_overlap(scene.N), flush=True)
def step(overlap):
    hp.zero_grad()
    for v in pose.values(): v.grad = None
    saved = hp.flat
    if not overlap:
        # backward writes into the views but does not reduce; plain all_reduce afterwards
        from riggs_b200.fused import render_skinned
        out = render_skinned(cam, hp.pc, hp.warp, pose, hp.pipe, hp.bg, grad_out=flat.views, reduce=None)
        from riggs_b200.loss_utils import l1_loss_and_grad
        loss, grad = l1_loss_and_grad(out["render"], target)
        out["render"].backward(grad)
    else:
        hp.step(cam, pose, target)
    hp.reduce_grads()
    torch.cuda.synchronize()
    return flat.flat.clone()
a = step(False); b = step(True); c = step(True)
ref = a.abs().max()
print(rank, "overlap vs plain: max abs diff / max", float((a-b).abs().max()/ref), "repeat", float((b-c).abs().max()/ref), flush=True)
gs = GraphedStep(hp, cam, pose, target)
gs.step(); hp.reduce_grads(); torch.cuda.synchronize()
d = flat.flat.clone()
gs.step(); hp.reduce_grads(); torch.cuda.synchronize()
e = flat.flat.clone()
print(rank, "graph vs plain", float((a-d).abs().max()/ref), "graph repeat", float((d-e).abs().max()/ref), "reduces in graph", gs._reduces_in_graph, flush=True)
dist.destroy_process_group()
