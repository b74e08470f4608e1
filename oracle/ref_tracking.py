"""TEST / BENCH INFRASTRUCTURE ONLY — the reference's tracking iteration, re-assembled around the unmodified
reference CUDA rasterizer (oracle/_ref) the way MonoGS runs it, so bench.py can report "reference CUDA on this
box" beside the product.  Not importable from monogs_b200/.

What is mirrored (MonoGS tree):
  * the autograd wrapper's work per call: 5 zero-filled outputs, 11 zero-filled gradient tensors, the separate
    `dL_dtau.view(-1,6).sum(0)` reduction (diff_gaussian_rasterization/__init__.py:85-147,150-345;
    rasterize_points.cu:71-75,198-208) — ref_gpu.forward/backward allocate exactly those;
  * the fork's first-order tracking objective in eager torch ops: per-pixel residual -> HuberLoss.apply -> L2 norm
    (utils/slam_utils.py:58-75,188-205; utils/slam_frontend.py:506,596-598);
  * torch.optim.Adam over (cam_rot_delta 3e-3, cam_trans_delta 1e-3, exposure_a 2e-2, exposure_b 2e-2)
    (slam_frontend.py:364-392,453) and update_pose with the pure-torch SE3_exp (utils/pose_utils.py:26-98; lietorch
    is not installed here).
"""
import torch

from oracle import ref_gpu


class _RefRasterize(torch.autograd.Function):
    @staticmethod
    def forward(ctx, theta, rho, cam, bg, sc):
        fw = ref_gpu.forward(cam, bg, sc.means3D, sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations,
                             sh_degree=sc.sh_degree)
        ctx.cam, ctx.bg, ctx.sc, ctx.fw = cam, bg, sc, fw
        return fw["color"], fw["depth"], fw["opacity"]

    @staticmethod
    def backward(ctx, g_color, g_depth, g_opacity):
        cam, bg, sc, fw = ctx.cam, ctx.bg, ctx.sc, ctx.fw
        if g_depth is None:
            g_depth = torch.zeros_like(fw["depth"])
        g = ref_gpu.backward(cam, bg, fw, sc.means3D, g_color.contiguous(), g_depth.contiguous(), shs=sc.shs,
                             scales=sc.scales, rotations=sc.rotations, sh_degree=sc.sh_degree)
        grad_tau = torch.sum(g["tau"].view(-1, 6), dim=0)
        return grad_tau[3:].view(1, -1), grad_tau[:3].view(1, -1), None, None, None


class _Huber(torch.autograd.Function):      # utils/slam_utils.py:58-75
    @staticmethod
    def forward(ctx, x, delta):
        ctx.delta = delta
        ctx.save_for_backward(x)
        return torch.where(x.abs() < delta, x, torch.sqrt(2 * delta * x.abs() - delta ** 2) * torch.sign(x))

    @staticmethod
    def backward(ctx, go):
        x, = ctx.saved_tensors
        d = ctx.delta
        return torch.where(x.abs() < d, go, go * d / torch.sqrt(2 * d * x.abs() - d ** 2)), None


def _so3_exp(theta):
    W = torch.zeros(3, 3, device=theta.device, dtype=theta.dtype)
    W[0, 1], W[0, 2], W[1, 0], W[1, 2], W[2, 0], W[2, 1] = -theta[2], theta[1], theta[2], -theta[0], -theta[1], theta[0]
    W2 = W @ W
    a = torch.norm(theta)
    I = torch.eye(3, device=theta.device, dtype=theta.dtype)
    if a < 1e-5:
        return I + W + 0.5 * W2, I + 0.5 * W + W2 / 6.0
    return (I + (torch.sin(a) / a) * W + ((1 - torch.cos(a)) / a ** 2) * W2,
            I + W * ((1.0 - torch.cos(a)) / a ** 2) + W2 * ((a - torch.sin(a)) / a ** 3))


class RefTracker:
    """mode "huber": the fork's first-order objective, ||Huber(opacity * (image_ab*m - gt*m))||_2 with m = the rgb
    boundary mask (slam_utils.py:188-205, slam_frontend.py:506,596-598).  mode "l1": get_loss_tracking (slam_utils.py:83-113)
    — mean(opacity * |image_ab*m - gt*m|) with m = boundary mask x grad_mask, and with a gt_depth the rgbd mix
    alpha * l1_rgb + (1 - alpha) * mean|depth*dm - gt_depth*dm|, dm = (gt_depth > 0.01) * (opacity > 0.95).
    After step(): .image / .depth / .opacity / .loss of the iteration, .grads = (tau[6] = rho|theta, dab[2])."""

    def __init__(self, sc, cam, bg, T_w2c, make_camera, mode="huber", alpha=0.95):
        self.sc, self.bg, self.make_camera = sc, bg, make_camera
        self.mode, self.alpha = mode, alpha
        self.gt_depth = self.grad_mask = None
        self.cam0 = cam
        self.T = T_w2c.clone().to(bg.device)
        self.rot = torch.zeros(3, device=bg.device, requires_grad=True)
        self.trans = torch.zeros(3, device=bg.device, requires_grad=True)
        self.a = torch.ones(1, device=bg.device, requires_grad=True)      # utils/camera_utils.py:56-58
        self.b = torch.zeros(1, device=bg.device, requires_grad=True)
        self.opt = torch.optim.Adam([{"params": [self.rot], "lr": 3e-3}, {"params": [self.trans], "lr": 1e-3},
                                     {"params": [self.a], "lr": 2e-2}, {"params": [self.b], "lr": 2e-2}])
        self.gt = None
        self.mask = None

    def set_frame(self, gt_image, gt_depth=None, grad_mask=None):
        self.gt = gt_image
        self.mask = (gt_image.sum(dim=0) > 0.01).view(1, *gt_image.shape[1:])
        self.gt_depth, self.grad_mask = gt_depth, grad_mask

    def _camera(self):
        view = self.T.transpose(0, 1).contiguous()
        return self.cam0._replace(viewmatrix=view, projmatrix=(view @ self.cam0.projmatrix_raw).contiguous(), campos=view)

    def step(self):
        cam = self._camera()
        image, depth, opacity = _RefRasterize.apply(self.rot, self.trans, cam, self.bg, self.sc)
        image_ab = (torch.abs(self.a) + 1e-8) * image + self.b
        if self.mode == "huber":
            res = opacity * (image_ab * self.mask - self.gt * self.mask)
            loss = torch.norm(_Huber.apply(res, 0.01))
        else:
            m = self.mask * self.grad_mask if self.grad_mask is not None else self.mask
            loss = (opacity * torch.abs(image_ab * m - self.gt * m)).mean()
            if self.gt_depth is not None:
                dm = (self.gt_depth > 0.01).view(*depth.shape) * (opacity > 0.95).view(*depth.shape)
                loss = self.alpha * loss + (1 - self.alpha) * torch.abs(depth * dm - self.gt_depth * dm).mean()
        self.opt.zero_grad()
        loss.backward()
        self.image, self.depth, self.opacity, self.loss = image.detach(), depth.detach(), opacity.detach(), loss.detach()
        self.grads = (torch.cat([self.trans.grad.view(-1), self.rot.grad.view(-1)]).clone(),
                      torch.cat([self.a.grad.view(-1), self.b.grad.view(-1)]).clone())
        with torch.no_grad():
            self.opt.step()
            tau = torch.cat([self.trans, self.rot], dim=0)
            R, V = _so3_exp(tau[3:])
            E = torch.eye(4, device=tau.device)
            E[:3, :3], E[:3, 3] = R, V @ tau[:3]
            self.T = E @ self.T
            self.rot.data.fill_(0)
            self.trans.data.fill_(0)
        return loss


    # ---- second-order (sketched Gauss-Newton / LM) iteration, utils/slam_frontend.py:455-720 with in_second_order ----
    def begin_second_order(self, initial_lambda=1e-3):
        self.lm = dict(lam=float(initial_lambda), prev=None, x=None)

    def second_order_step(self, sketch_indices, rand_weights, sketch_dim, min_lambda=1e-6, max_lambda=1e7, inc=5.0, dec=5.0):
        """One iteration in the reference's order: apply the pending step (TempCamera.step + update_pose, :474-480), render
        with the reference kernels, scalar loss, lambda (:541-548), Sf and the gradient images through
        oracle.golden.sketch_loss_ref (pinned to the reference's slam_utils by tests/test_oracle_cpu.py), the pose columns
        of SJ through `stack_dim` calls of the reference's own sketch-Jacobian kernel (__init__.py:216-253), and
        torch.linalg.lstsq([SJ; sqrt(lambda) I], -[Sf; 0]) (:668-679) in float64.  Returns (SJ [rows,8], Sf [rows], scalar, x)."""
        from oracle import golden as G
        lm = self.lm
        with torch.no_grad():
            if lm["x"] is not None:
                x = lm["x"]
                tau = x[:6].float()
                R, V = _so3_exp(tau[3:])
                E = torch.eye(4, device=tau.device)
                E[:3, :3], E[:3, 3] = R, V @ tau[:3]
                self.T = E @ self.T
                self.a.data += x[6].float(); self.b.data += x[7].float()
            cam = self._camera()
            sc = self.sc
            fw = ref_gpu.forward(cam, self.bg, sc.means3D, sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations,
                                 sh_degree=sc.sh_degree)
            n_stacks = sketch_indices.shape[0]
            res = G.sketch_loss_ref(fw["color"], fw["opacity"], self.gt, self.a.detach(), self.b.detach(), sketch_indices,
                                    rand_weights, sketch_dim)
            scalar = float(res["scalar_l1"])
            if lm["prev"] is not None:
                lm["lam"] = max(lm["lam"] / dec, min_lambda) if scalar < lm["prev"] else min(lm["lam"] * inc, max_lambda)
            lm["prev"] = scalar
            g_img = res["g_image"].contiguous()
            g_dep = torch.zeros_like(fw["depth"])
            SJ = torch.zeros(n_stacks, sketch_dim, 8, device=g_img.device)
            for j in range(n_stacks):
                tau_j = ref_gpu.backward_sketch(cam, self.bg, fw, sc.means3D, g_img, g_dep, sketch_indices[j].contiguous(), sketch_dim,
                                                shs=sc.shs, scales=sc.scales, rotations=sc.rotations, sh_degree=sc.sh_degree)
                SJ[j, :, :6] = tau_j.double().sum(0).float()
                del tau_j
            SJ[:, :, 6:] = res["dexposure"]
            SJ, Sf = SJ.view(-1, 8).double().cpu(), res["Sf"].reshape(-1).double().cpu()
            A = torch.cat((SJ, torch.eye(8, dtype=torch.float64) * lm["lam"] ** 0.5), dim=0)
            rhs = torch.cat((Sf, torch.zeros(8, dtype=torch.float64)), dim=0)
            x = torch.linalg.lstsq(A, -rhs[:, None]).solution[:, 0]
            lm["x"] = x.to(self.T.device)
            self.image, self.opacity = fw["color"], fw["opacity"]
        return SJ, Sf, scalar, x
