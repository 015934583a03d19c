"""Golden vectors for the image loss: outputs of the REFERENCE's own utils/loss_utils.py (imported from
/root/reference, CPU, float64 and float32) on seeded inputs.  Run in the build container only:
    python tests/golden/make_loss_golden.py
"""
import importlib.util
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("ref_loss_utils", "/root/reference/utils/loss_utils.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

out = {}
for tag, shape, dtype, lam in (("a", (3, 37, 53), torch.float64, 0.2), ("b", (3, 64, 48), torch.float32, 0.2),
                               ("c", (1, 16, 16), torch.float64, 1.0), ("d", (3, 9, 70), torch.float64, 0.5)):
    g = torch.Generator().manual_seed(ord(tag) + 7)
    gt = torch.rand(*shape, generator=g, dtype=torch.float64)
    img = (gt + 0.15 * torch.randn(*shape, generator=g, dtype=torch.float64)).clamp(0, 1)
    img[..., :4, :4] = gt[..., :4, :4]  # a patch of exact agreement: sign(0) in the L1 gradient
    img = img.to(dtype).requires_grad_(True)
    gt = gt.to(dtype)
    l1 = ref.l1_loss(img, gt)
    s = ref.ssim(img, gt)
    loss = (1.0 - lam) * l1 + lam * (1.0 - s)
    loss.backward()
    out[f"{tag}_img"] = img.detach().numpy()
    out[f"{tag}_gt"] = gt.numpy()
    out[f"{tag}_lambda"] = np.float64(lam)
    out[f"{tag}_l1"] = l1.detach().numpy()
    out[f"{tag}_ssim"] = s.detach().numpy()
    out[f"{tag}_loss"] = loss.detach().numpy()
    out[f"{tag}_grad"] = img.grad.numpy()
np.savez_compressed(os.path.join(HERE, "loss_ssim.npz"), **out)
print({k: (v.shape, v.dtype) for k, v in out.items() if k.endswith("grad")})
