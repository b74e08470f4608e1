"""Fused loss, fused Adam and on-device pose update against their torch restatements
(MonoGS utils/slam_utils.py, torch.optim.Adam, utils/pose_utils.py)."""
import pytest
import torch

from tests import util as U

pytestmark = pytest.mark.gpu


def _images(H=97, W=131, seed=0):
    g = torch.Generator().manual_seed(seed)
    image = torch.rand(3, H, W, generator=g)
    gt = torch.rand(3, H, W, generator=g)
    gt[:, :5, :] = 0.0                               # below rgb_boundary_threshold -> masked out
    depth = torch.rand(1, H, W, generator=g) * 3
    gt_depth = torch.rand(1, H, W, generator=g) * 3
    gt_depth[:, :, :7] = 0.0
    opacity = torch.rand(1, H, W, generator=g)
    opacity[:, 40:, :] = opacity[:, 40:, :] * 0.05 + 0.95
    grad_mask = (torch.rand(1, H, W, generator=g) > 0.4).float()
    return image, gt, depth, gt_depth, opacity, grad_mask


@pytest.mark.parametrize("mode,use_depth,use_mask", [(0, False, True), (0, True, True), (1, False, False),
                                                     (2, False, False), (2, True, False)])
def test_fused_loss_matches_torch(mode, use_depth, use_mask):
    from monogs_b200 import ops
    from oracle import golden as G
    image, gt, depth, gt_depth, opacity, grad_mask = _images()
    a0, b0 = torch.tensor([-0.7]), torch.tensor([0.05])
    # torch restatement on CPU in float64
    im, dp, a, b = [t.double().requires_grad_(True) for t in (image, depth, a0, b0)]
    ref = G.loss_ref(mode, im, dp, opacity.double(), gt.double(), a, b, grad_mask=grad_mask.double() if use_mask else None,
                     gt_depth=gt_depth.double() if use_depth else None)
    ref.backward()
    imc, dpc, ac, bc = [t.cuda().requires_grad_(True) for t in (image, depth, a0, b0)]
    if mode == 0:
        out = ops.tracking_loss(imc, dpc, opacity.cuda(), gt.cuda(), ac, bc, grad_mask=grad_mask.cuda() if use_mask else None,
                                gt_depth=gt_depth.cuda() if use_depth else None)
    elif mode == 1:
        out = ops.tracking_loss_huber(imc, opacity.cuda(), gt.cuda(), ac, bc)
    else:
        out = ops.mapping_loss(imc, dpc, gt.cuda(), ac, bc, gt_depth=gt_depth.cuda() if use_depth else None)
    (out * 1.7).backward()
    assert abs(out.item() - ref.item()) <= 1e-5 * abs(ref.item())
    assert U.rel_err(imc.grad, 1.7 * im.grad) <= 1e-5
    if use_depth:
        assert U.rel_err(dpc.grad, 1.7 * dp.grad) <= 1e-5
    assert U.rel_err(ac.grad, 1.7 * a.grad) <= 1e-4 and U.rel_err(bc.grad, 1.7 * b.grad) <= 1e-4


def test_fused_adam_matches_torch():
    from monogs_b200 import ops
    g = torch.Generator().manual_seed(0)
    shapes = [(5003, 3), (5003, 1, 3), (5003, 15, 3), (5003, 1), (5003, 3), (5003, 4), (3,), (1,)]
    lrs = [1.6e-4, 2.5e-3, 1.25e-4, 0.05, 1e-3, 1e-3, 3e-3, 2e-2]
    ps_ref = [torch.randn(*s, generator=g).cuda().requires_grad_(True) for s in shapes]
    ps_mine = [p.detach().clone().requires_grad_(True) for p in ps_ref]
    opt_ref = torch.optim.Adam([{"params": [p], "lr": lr} for p, lr in zip(ps_ref, lrs)], lr=0.0, eps=1e-15)
    opt_mine = ops.FusedAdam([{"params": [p], "lr": lr} for p, lr in zip(ps_mine, lrs)], lr=0.0, eps=1e-15)
    for it in range(5):
        for pr, pm in zip(ps_ref, ps_mine):
            gr = torch.randn(pr.shape, generator=g).cuda() * (10.0 ** (it - 3))
            pr.grad, pm.grad = gr.clone(), gr.clone()
        opt_ref.step(); opt_mine.step()
    for pr, pm in zip(ps_ref, ps_mine):
        assert U.rel_err(pm, pr) <= 1e-6


def test_pose_update_matches_se3_exp():
    from monogs_b200 import ops, scene as S
    T0 = S.se3_exp(torch.tensor([0.3, -0.2, 0.5, 0.2, 0.1, -0.3])).float()
    for tau in ([0.01, -0.005, 0.008, 0.004, -0.003, 0.002], [1e-7, 0, 0, 0, 1e-7, 0], [0.2, 0.1, -0.3, 0.5, -0.4, 0.3]):
        tau = torch.tensor(tau)
        T = T0.clone().cuda()
        rho, theta = tau[:3].clone().cuda(), tau[3:].clone().cuda()
        proj = S.make_camera(64, 48, 60.0, 60.0, 32.0, 24.0).projmatrix_raw.cuda()
        conv, view, full = ops.pose_update(T, rho, theta, proj_raw=proj, want_matrices=True)
        ref = (S.se3_exp(tau) @ T0.double()).float()
        assert U.rel_err(T, ref) <= 1e-6
        assert rho.abs().sum() == 0 and theta.abs().sum() == 0
        assert int(conv.item()) == int(float((tau ** 2).sum()) < 1e-8)
        assert U.rel_err(view, ref.t()) <= 1e-6
        assert U.rel_err(full, ref.t() @ proj.cpu()) <= 1e-5


@pytest.mark.parametrize("mode,use_depth", [(1, False), (0, False), (0, True)])
def test_fused_track_iteration_matches_composed_entry_points(mode, use_depth):
    """mgs_track_iteration (loss fused into the forward blend, Jacobian into preprocess, one update kernel) follows
    the same trajectory as the iteration composed from the individual C-ABI calls, and tracking reduces the loss."""
    from monogs_b200 import scene as S, rasterizer as RZ
    from monogs_b200.tracking import PoseTracker
    sc, cam = U.small_case("dense", T_w2c=torch.eye(4))
    scd = U.to_dev(sc)
    bg = torch.tensor([0.05, 0.1, 0.15]).cuda()
    T_gt = S.se3_exp(torch.tensor(S.GT_TAU)).float()
    cam_gt = U.to_dev(S.make_camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, T_gt))
    gt, _, gt_depth, _, _, _ = RZ.raster_forward(U.settings(cam_gt, bg, sc.sh_degree), scd.means3D, scd.opacities, scd.shs,
                                                 None, scd.scales, scd.rotations, None)
    runs = []
    for fused, graph in ((True, True), (False, False)):
        tr = PoseTracker(scd.means3D, scd.opacities, scd.shs, scd.scales, scd.rotations, sc.sh_degree, cam, bg, torch.eye(4),
                         mode=mode, use_depth=use_depth, use_graph=graph, fused=fused)
        tr.set_frame(gt, gt_depth=gt_depth if use_depth else None)
        losses = []
        for _ in range(12):
            tr.step()
            losses.append(float(tr.loss))
        runs.append((losses, tr.T.clone(), tr.tau.clone(), tr.params.clone(), tr.color.clone()))
    (l0, T0, tau0, p0, c0), (l1, T1, tau1, p1, c1) = runs
    # the first iterations start from identical states: tight.  Later ones inherit the rounding-level differences of the two
    # paths through Adam, which turns a sign flip of a near-zero gradient (float atomics sum in a different order from run
    # to run) into a full learning-rate step: compared at the trajectory tolerance
    assert all(abs(a - b) <= (1e-4 if i < 3 else 1e-3) * abs(b) + 1e-7 for i, (a, b) in enumerate(zip(l0, l1))), \
        [(i, a, b) for i, (a, b) in enumerate(zip(l0, l1))]
    # 12 Adam steps amplify the rounding-level differences of the two paths (ex2.approx vs expf, atomic summation order):
    # the trajectories are compared at 1e-3, the per-call parity gates live in test_parity_gpu.py
    assert U.rel_err(T0, T1) <= 1e-4 and U.rel_err(p0, p1) <= 1e-3 and U.rel_err(c0, c1) <= 1e-3
    assert U.rel_err(tau0, tau1) <= 2e-3
    assert l0[-1] < l0[0]


@pytest.mark.parametrize("mode,use_depth,masked", [(1, False, False), (0, False, False), (0, True, False), (1, False, True), (0, True, True)])
def test_single_pass_blend_matches_two_pass_first_step(mode, use_depth, masked):
    """The single-pass tracking blend (blend_track.cu: forward-order accumulators, tangent contracted in the epilogue)
    against the two-pass path built from the parity-tested entry points (forward, fused loss, pose-only backward), one
    step from the same state, exact exponent in both: images identical, loss and pose tangent to 1e-4."""
    from monogs_b200 import scene as S, rasterizer as RZ
    from monogs_b200.tracking import PoseTracker
    sc, cam = U.small_case("dense", T_w2c=torch.eye(4))
    sc = sc._replace(shs=sc.shs[:, :1].contiguous(), sh_degree=0)     # SH degree 0: the single-pass kernel's domain
    scd = U.to_dev(sc)
    bg = torch.tensor([0.05, 0.1, 0.15]).cuda()
    T_gt = S.se3_exp(torch.tensor(S.GT_TAU)).float()
    cam_gt = U.to_dev(S.make_camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, T_gt))
    gt, _, gt_depth, _, _, _ = RZ.raster_forward(U.settings(cam_gt, bg, 0), scd.means3D, scd.opacities, scd.shs, None,
                                                 scd.scales, scd.rotations, None)
    out = []
    for fused in (True, False):
        tr = PoseTracker(scd.means3D, scd.opacities, scd.shs, scd.scales, scd.rotations, 0, cam, bg, torch.eye(4),
                         mode=mode, use_depth=use_depth, use_graph=False, fused=fused, fast_math=False)
        # masked: a gradient mask over about half of the frame (Camera.compute_grad_mask) and a band of invalid depth; the
        # single-pass kernel skips the tangent accumulation of zero-weight pixels, which must not change anything
        gm = (torch.rand(1, cam.H, cam.W, generator=torch.Generator().manual_seed(5)) > 0.5).float().cuda() if masked else None
        gd = gt_depth.clone() if use_depth else None
        if masked and use_depth: gd[:, : cam.H // 4] = 0.0
        tr.set_frame(gt, gt_depth=gd, grad_mask=gm)
        tr.step()
        torch.cuda.synchronize()
        out.append((tr.loss.clone(), tr.tau.clone(), tr.dab.clone(), tr.color.clone(), tr.depth.clone(), tr.n_touched.clone()))
    (l0, t0, d0, c0, z0, n0), (l1, t1, d1, c1, z1, n1) = out
    assert torch.equal(c0, c1) and torch.equal(z0, z1) and torch.equal(n0, n1)
    assert U.rel_err(l0, l1) <= 1e-5
    assert U.rel_err(t0, t1) <= 1e-4, (t0, t1)
    assert U.rel_err(d0, d1) <= 1e-4


@pytest.mark.parametrize("name", ["ragged", "packed8", "packed16", "crowded"])
def test_single_pass_blend_edge_scenes(name):
    """The fused tracker's staging (packed per-Gaussian rows gathered by id, sorted ids from both sort kernels) on the scenes
    that stress it: an image that is not a multiple of the tile size, tile lists of ~1.5k / ~3k entries (the register paths
    of the large-list sort, many 128-entry batches per tile) and a list beyond 4096 (the in-place bitonic path) — against
    the two-pass path built from the parity-tested entry points, one step from the same state."""
    from monogs_b200 import scene as S, rasterizer as RZ
    from monogs_b200.tracking import PoseTracker
    sc, cam = U.small_case(name, T_w2c=torch.eye(4))
    sc = sc._replace(shs=sc.shs[:, :1].contiguous(), sh_degree=0)
    scd = U.to_dev(sc)
    bg = torch.tensor([0.05, 0.1, 0.15]).cuda()
    T_gt = S.se3_exp(torch.tensor(S.GT_TAU) * 0.3).float()
    cam_gt = U.to_dev(S.make_camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, T_gt))
    gt, _, _, _, _, _ = RZ.raster_forward(U.settings(cam_gt, bg, 0), scd.means3D, scd.opacities, scd.shs, None, scd.scales,
                                          scd.rotations, None)
    out = []
    for fused in (True, False):
        tr = PoseTracker(scd.means3D, scd.opacities, scd.shs, scd.scales, scd.rotations, 0, cam, bg, torch.eye(4), mode=1,
                         use_depth=False, use_graph=False, fused=fused, fast_math=False)
        tr.set_frame(gt)
        for _ in range(2):      # the second step runs on the workspace the first one left behind (zero-maintained counters)
            tr.step()
        torch.cuda.synchronize()
        assert not tr.check_capacity()[1]
        out.append((tr.loss.clone(), tr.tau.clone(), tr.color.clone(), tr.n_touched.clone()))
    (l0, t0, c0, n0), (l1, t1, c1, n1) = out
    assert U.rel_err(c0, c1) <= 1e-5 and U.rel_err(l0, l1) <= 1e-4
    assert U.rel_err(t0, t1) <= 1e-3, (t0, t1)
    assert int((n0 != n1).sum()) <= max(2, n0.numel() // 500)


def test_host_bound_tracking_pipeline_matches_device_steps():
    """PoseTracker.bind_host / step_host (frame from pinned host memory, loss + tangent back to pinned host memory, one
    graph launch per iteration with the next frame's upload beside the kernels) follows the device-resident step()
    sequence exactly: same kernels on the same frame."""
    from monogs_b200 import scene as S, rasterizer as RZ
    from monogs_b200.tracking import PoseTracker
    sc, cam = U.small_case("dense", T_w2c=torch.eye(4))
    sc = sc._replace(shs=sc.shs[:, :1].contiguous(), sh_degree=0)
    scd = U.to_dev(sc)
    bg = torch.tensor([0.05, 0.1, 0.15]).cuda()
    T_gt = S.se3_exp(torch.tensor(S.GT_TAU)).float()
    cam_gt = U.to_dev(S.make_camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, T_gt))
    gt, _, _, _, _, _ = RZ.raster_forward(U.settings(cam_gt, bg, 0), scd.means3D, scd.opacities, scd.shs, None, scd.scales,
                                          scd.rotations, None)
    mk = lambda: PoseTracker(scd.means3D, scd.opacities, scd.shs, scd.scales, scd.rotations, 0, cam, bg, torch.eye(4))
    a = mk()
    a.set_frame(gt)
    dev_losses, dev_taus = [], []
    for _ in range(6):
        a.step()
        dev_losses.append(float(a.loss)); dev_taus.append(a.tau.cpu().clone())
    b = mk()
    gt_host = gt.cpu().pin_memory()
    loss_host, tau_host = torch.zeros(1).pin_memory(), torch.zeros(6).pin_memory()
    b.bind_host(gt_host, loss_host, tau_host)
    b.set_frame(b._stage)
    for i in range(6):
        b.step_host()
        assert abs(float(loss_host) - dev_losses[i]) <= 1e-5 * abs(dev_losses[i]) + 1e-8, i
        assert U.rel_err(tau_host, dev_taus[i]) <= 1e-3, i
    assert U.rel_err(b.T, a.T) <= 1e-5
    # a new frame written into the host buffer is what the iteration AFTER the next replay sees
    gt_host.mul_(0.5)
    b.step_host()
    assert torch.equal(b._stage.cpu(), gt_host) and not torch.equal(b.gt_image.cpu(), gt_host)
    b.step_host()
    assert torch.equal(b.gt_image.cpu(), gt_host)


def test_lagged_host_pipeline_returns_previous_iteration():
    """step_host_lagged(): iteration k is launched before the host waits for k-1; the slot it returns holds exactly the
    results step() produces for iteration k-1."""
    from monogs_b200 import scene as S, rasterizer as RZ
    from monogs_b200.tracking import PoseTracker
    sc, cam = U.small_case("dense", T_w2c=torch.eye(4))
    sc = sc._replace(shs=sc.shs[:, :1].contiguous(), sh_degree=0)
    scd = U.to_dev(sc)
    bg = torch.tensor([0.05, 0.1, 0.15]).cuda()
    cam_gt = U.to_dev(S.make_camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, S.se3_exp(torch.tensor(S.GT_TAU)).float()))
    gt = RZ.raster_forward(U.settings(cam_gt, bg, 0), scd.means3D, scd.opacities, scd.shs, None, scd.scales, scd.rotations, None)[0]
    mk = lambda: PoseTracker(scd.means3D, scd.opacities, scd.shs, scd.scales, scd.rotations, 0, cam, bg, torch.eye(4))
    a = mk()
    a.set_frame(gt)
    ref = []
    for _ in range(7):
        a.step()
        ref.append((float(a.loss), a.tau.cpu().clone()))
    b = mk()
    loss_host, tau_host = torch.zeros(2).pin_memory(), torch.zeros(2, 6).pin_memory()
    b.bind_host(gt.cpu().pin_memory(), loss_host, tau_host)
    b.set_frame(b._stage)
    got = []
    for k in range(8):
        slot = b.step_host_lagged()
        assert (slot is None) == (k == 0)
        if slot is not None:
            got.append((float(loss_host[slot]), tau_host[slot].clone()))
    for (l0, t0), (l1, t1) in zip(ref, got):
        assert abs(l0 - l1) <= 1e-5 * abs(l0) + 1e-8 and U.rel_err(t1, t0) <= 1e-3


def test_block_bitonic_sort_against_std_sort():
    """block_bitonic_sort (csrc/bitonic_sort.cuh: the per-tile depth sort of every binning path) for 1 / 2 / 4 / 8 / 16 words
    per thread on random unique words with ~0 padding, against std::sort (scripts/tools/bitonic_check.cu, built by
    csrc/build.sh).  The sorted tile lists are also compared bit for bit with the reference's in test_parity_gpu.py."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts", "tools", "bitonic_check")
    assert os.path.exists(exe), "scripts/tools/bitonic_check missing: run monogs_b200/csrc/build.sh"
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout + r.stderr
