"""TEST / BENCH INFRASTRUCTURE ONLY — the reference's mapping iteration, re-assembled around the unmodified reference
CUDA rasterizer (oracle/_ref) the way MonoGS runs it, so bench.py can report "reference CUDA mapping on this box"
beside the product.  Not importable from monogs_b200/.

What is mirrored (MonoGS tree):
  * render() per view of the window: activations on every call (sigmoid / exp / normalize, gaussian_model.py:77-97),
    the autograd wrapper with its zero-filled outputs and gradient tensors (diff_gaussian_rasterization/__init__.py),
  * get_loss_mapping_rgb in eager torch ops (utils/slam_utils.py:224-239), summed over the views, plus the isotropic
    regulariser (utils/slam_backend.py:244-246), ONE autograd backward, torch.optim.Adam over the parameter groups
    (gaussian_model.py:285).
"""
import torch

from oracle import ref_gpu


class _RefRasterizeFull(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, shs, opacities, scales, rotations, cam, bg, sh_degree, slot):
        # one state slot of the shim per view: all forwards of the window precede the single backward
        fw = ref_gpu.forward(cam, bg, means3D, opacities, shs=shs, scales=scales, rotations=rotations, sh_degree=sh_degree,
                             slot=slot)
        ctx.cam, ctx.bg, ctx.fw, ctx.deg, ctx.slot = cam, bg, fw, sh_degree, slot
        ctx.save_for_backward(means3D, shs, scales, rotations)
        return fw["color"], fw["depth"], fw["opacity"]

    @staticmethod
    def backward(ctx, g_color, g_depth, g_opacity):
        means3D, shs, scales, rotations = ctx.saved_tensors
        fw = ctx.fw
        if g_depth is None:
            g_depth = torch.zeros_like(fw["depth"])
        g = ref_gpu.backward(ctx.cam, ctx.bg, fw, means3D, g_color.contiguous(), g_depth.contiguous(), shs=shs,
                             scales=scales, rotations=rotations, sh_degree=ctx.deg, slot=ctx.slot)
        return g["means3D"], g["sh"].view_as(shs), g["opacity"], g["scales"], g["rotations"], None, None, None, None


class RefMapper:
    def __init__(self, raw, lrs, cams, gts, bg, sh_degree=0, thr=0.01):
        self.p = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
        self.opt = torch.optim.Adam([{"params": [p], "lr": lrs[k]} for k, p in self.p.items()], lr=0.0, eps=1e-15)
        self.cams, self.gts, self.bg, self.deg, self.thr = cams, gts, bg, sh_degree, thr

    def loss_and_grads(self):
        """One iteration up to the raw-parameter gradients (no optimizer step): (total loss, {name: grad}, per-view
        forward dicts of ref_gpu.forward with radii / n_touched)."""
        total = self._forward_backward()
        grads = {k: v.grad.detach().clone() for k, v in self.p.items()}
        self.opt.zero_grad(set_to_none=True)
        return total.detach(), grads, self.fws

    def step(self):
        total = self._forward_backward()
        self.opt.step()
        self.opt.zero_grad(set_to_none=True)
        return total.detach()

    def _forward_backward(self):
        p = self.p
        self.fws = []
        total = 0.0
        for slot, (cam, gt) in enumerate(zip(self.cams, self.gts)):
            image, depth, opacity = _RefRasterizeFull.apply(p["xyz"], p["f_dc"], torch.sigmoid(p["opacity"]),
                                                            torch.exp(p["scaling"]),
                                                            torch.nn.functional.normalize(p["rotation"]), cam, self.bg, self.deg, slot)
            self.fws.append(image.grad_fn.fw if image.grad_fn is not None else None)
            mask = (gt.sum(dim=0) > self.thr).view(1, *gt.shape[1:])
            total = total + torch.abs(image * mask - gt * mask).mean()
        s = torch.exp(p["scaling"])
        total = total + 10 * torch.abs(s - s.mean(dim=1).view(-1, 1)).mean()
        total.backward()
        return total
