"""CPU-only: the CPU oracle (oracle/golden.py) against the committed golden vectors, which are outputs of the
unmodified reference CUDA rasterizer (tests/golden/make_golden.py).  This is the pin of the oracle that works
without a GPU."""
import os

import numpy as np
import pytest
import torch

from tests import util as U
from monogs_b200 import scene as S
from oracle import golden as G

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    z = np.load(os.path.join(GOLD, f"{name}.npz"))
    return {k: torch.from_numpy(z[k]) for k in z.files}


@pytest.mark.parametrize("name", ["tiny", "ragged", "sh3"])
def test_golden_renderer_matches_reference_vectors(name):
    ref = _load(name)
    sc, cam = U.small_case(name)
    gc, gd = S.output_grads(cam.W, cam.H, 1)
    bg = torch.tensor([0.1, 0.2, 0.3])
    res = G.render(U.golden_args(sc), cam, bg, grads=(gc, gd))
    # integers: CPU fp32 has no FMA contraction, so allow (and count) borderline Gaussians
    mism = int((res["pre"]["radii"] != ref["radii"]).sum())
    assert mism <= 2, f"{mism} radii differ"
    exact = mism == 0
    if exact:
        assert int(res["keys"].numel()) == int(ref["R"])
        assert torch.equal(res["point_list"].int(), ref["point_list"])
        assert torch.equal(res["keys"] >> 32, ref["keys"] >> 32)
        assert int(((res["keys"] & 0xFFFFFFFF) - (ref["keys"] & 0xFFFFFFFF)).abs().max()) <= 2
        assert torch.equal(res["ranges"].int(), ref["ranges"])
        assert int((res["n_contrib"].int() != ref["n_contrib"]).sum()) <= 2
        assert int((res["n_touched"].int() != ref["n_touched"]).sum()) <= 2
    ti, tg = (1e-4, 1e-3) if exact else (5e-2, 5e-2)
    for k in ("color", "depth", "opacity"):
        assert U.rel_err(res[k], ref[k]) <= ti, k
    assert U.rel_err(res["final_T"], ref["final_T"]) <= ti
    g2d, gg = res["g2d"], res["grads"]
    vis = (ref["radii"] > 0).float()
    assert U.rel_err(g2d["conic"][:, 0], ref["g_conic"][:, 0, 0]) <= tg
    assert U.rel_err(g2d["conic"][:, 1], ref["g_conic"][:, 0, 1]) <= tg
    assert U.rel_err(g2d["conic"][:, 2], ref["g_conic"][:, 1, 1]) <= tg
    assert U.rel_err(g2d["depth"], ref["g_depth"][:, 0]) <= tg
    assert U.rel_err(g2d["color"], ref["g_color"]) <= tg
    for k in ("means3D", "means2D", "opacity", "scales", "rotations", "tau"):
        assert U.rel_err(gg[k], ref["g_" + k]) <= tg, k
    assert U.rel_err(gg["color"], ref["g_sh"][:, :sc.shs.size(1)]) <= tg
    assert float((gg["means3D"].abs().sum(1) * (1 - vis)).sum()) == 0.0   # Q11: culled Gaussians get zeros


def test_knn_golden():
    ref = _load("knn")["dist2"]
    pts = torch.randn(2000, 3, generator=torch.Generator().manual_seed(7))
    assert U.rel_err(G.knn_dist2(pts), ref) <= 1e-6


def test_pose_tangent_against_finite_differences():
    """Secondary check of the analytic pose gradient (valid away from the quirks Q6/Q8: no clamped alpha in the
    scene, no FoV-clamped Gaussians): d/dtau of <render(Exp(tau) T), G> in float64 central differences."""
    # principal point at the image centre: the reference's mean2D->tau Jacobian (backward.cu:497-511) drops the
    # off-centre projection terms P[8], P[9], which the product reproduces but finite differences would not
    P, W, H, fx, fy, cx, cy = 60, 48, 32, 50.0, 50.0, 24.0, 16.0
    sc = S.make_scene(P, W, H, fx, fy, cx, cy, 0, seed=3)
    sc = sc._replace(opacities=sc.opacities * 0.6, means3D=sc.means3D.double(), scales=sc.scales.double() * 2.0,
                     rotations=sc.rotations.double(), shs=sc.shs.double())
    sc = sc._replace(opacities=sc.opacities.double())
    T0 = S.se3_exp(torch.tensor([0.02, -0.01, 0.03, 0.01, 0.02, -0.01]))
    gc, gd = S.output_grads(W, H, 2)
    gc, gd = gc.double(), gd.double()
    bg = torch.tensor([0.2, 0.3, 0.1], dtype=torch.float64)

    def cam_at(tau):
        T = S.se3_exp(tau) @ T0
        c = S.make_camera(W, H, fx, fy, cx, cy, T)
        return c._replace(viewmatrix=T.t().contiguous(), projmatrix=T.t().contiguous() @ c.projmatrix_raw.double(),
                          projmatrix_raw=c.projmatrix_raw.double(), campos=T.t().contiguous())

    def f(tau):
        r = G.render(U.golden_args(sc), cam_at(tau), bg)
        return float((r["color"] * gc).sum() + (r["depth"] * gd).sum())
    res = G.render(U.golden_args(sc), cam_at(torch.zeros(6, dtype=torch.float64)), bg, grads=(gc, gd))
    analytic = res["grads"]["tau"].sum(0)
    # the renderer is piecewise smooth (alpha < 1/255 cut, ceil of the radius): a tiny step keeps the central
    # difference inside one smooth piece; float64 leaves ~1e-8 relative accuracy at this step
    eps = 1e-8
    num = torch.zeros(6, dtype=torch.float64)
    for k in range(6):
        e = torch.zeros(6, dtype=torch.float64); e[k] = eps
        num[k] = (f(e) - f(-e)) / (2 * eps)
    assert U.rel_err(analytic, num) <= 2e-3, (analytic, num)


def test_loss_restatement_huber_derivative():
    """HuberLoss.backward (slam_utils.py:69-75) is the derivative autograd gives the restated forward."""
    x = torch.linspace(-1, 1, 1001, dtype=torch.float64, requires_grad=True)
    G.huber(x, 0.01).sum().backward()
    ax = x.detach().abs()
    expect = torch.where(ax < 0.01, torch.ones_like(ax), 0.01 / torch.sqrt(torch.clamp_min(2 * 0.01 * ax - 1e-4, 1e-30)))
    assert torch.allclose(x.grad, expect, rtol=1e-9, atol=1e-12)


# ---- the Python-side oracles (rows a12 / a16 / f2 / f4) against the reference's own code -----------------------
# tests/golden/loss_ref.npz = outputs of /root/reference/utils/{slam_utils,camera_utils,pose_utils}.py imported
# unchanged on the CPU (tests/golden/make_loss_golden.py).
def _autograd(mode, z, use_depth=False, use_mask=False, init=False):
    im, dp, a, b = [z[k].clone().requires_grad_(True) for k in ("image", "depth", "a", "b")]
    loss = G.loss_ref(mode, im, dp, z["opacity"], z["gt"], None if init else a, None if init else b,
                      grad_mask=z["grad_mask"] if use_mask else None, gt_depth=z["gt_depth"] if use_depth else None)
    loss.backward()
    return loss.detach(), im.grad, dp.grad, a.grad, b.grad


@pytest.mark.parametrize("name,mode,use_depth,use_mask,init", [
    ("track_rgb", 0, False, True, False), ("track_rgbd", 0, True, True, False), ("huber", 1, False, False, False),
    ("map_rgb", 2, False, False, False), ("map_rgbd", 2, True, False, False), ("map_init", 2, False, False, True)])
def test_loss_restatement_matches_reference_code(name, mode, use_depth, use_mask, init):
    z = _load("loss_ref")
    loss, g_im, g_dp, g_a, g_b = _autograd(mode, z, use_depth, use_mask, init)
    assert abs(float(loss) - float(z[name + "_loss"])) <= 1e-6 * abs(float(z[name + "_loss"]))
    assert torch.allclose(g_im, z[name + "_g_image"], rtol=1e-5, atol=1e-9)
    if use_depth:
        assert torch.allclose(g_dp, z[name + "_g_depth"], rtol=1e-5, atol=1e-9)
    if not init:
        assert torch.allclose(g_a, z[name + "_g_a"], rtol=1e-4, atol=1e-8) and torch.allclose(g_b, z[name + "_g_b"], rtol=1e-4, atol=1e-8)


def test_sketch_objective_restatement_matches_reference_code():
    """Sf, the gradient that reaches the rasterizer and ApplyExposure's sketched exposure columns
    (slam_frontend.py:637-665, slam_utils.py:115-185), reference code vs oracle.golden.sketch_loss_ref."""
    z = _load("loss_ref")
    d = int(z["sk_dexposure"].shape[1])
    res = G.sketch_loss_ref(z["image"], z["opacity"], z["gt"], z["a"], z["b"], z["sk_indices"][0], z["sk_weights"][0], d)
    assert torch.allclose(res["Sf"], z["sk_Sf"][0], rtol=1e-5, atol=1e-8)
    assert torch.allclose(res["g_image"], z["sk_g_image"], rtol=1e-5, atol=1e-10)
    assert torch.allclose(res["dexposure"], z["sk_dexposure"], rtol=1e-4, atol=1e-8)
    assert abs(float(res["scalar_l1"]) - float(z["huber_scalar_l1"])) <= 1e-5 * float(z["huber_scalar_l1"])
    # the un-sketched totals: summing the rows over stacks and buckets gives ApplyExposure's plain exposure gradients
    assert abs(float(res["dexposure"][..., 0].sum()) - float(z["sk_g_a"])) <= 1e-4 * abs(float(z["sk_g_a"])) + 1e-7
    assert abs(float(res["dexposure"][..., 1].sum()) - float(z["sk_g_b"])) <= 1e-4 * abs(float(z["sk_g_b"])) + 1e-7


def test_frame_op_restatements_match_reference_code():
    z = _load("loss_ref")
    frame = z["frame"]
    gray = frame.mean(dim=0, keepdim=True)
    gv, gh = G.image_gradient_ref(gray)
    mv, mh = G.image_gradient_mask_ref(gray)
    assert torch.allclose(gv, z["frame_gv"], rtol=1e-6, atol=1e-7) and torch.allclose(gh, z["frame_gh"], rtol=1e-6, atol=1e-7)
    assert torch.equal(mv, z["frame_mv"]) and torch.equal(mh, z["frame_mh"])
    for ds, patch in (("tum", 0), ("replica", 32)):
        gm, rm, rmm = G.grad_masks_ref(frame, 1.1, 0.01, patch)
        assert torch.equal(gm, z[f"frame_grad_mask_{ds}"])
        assert torch.equal(rm, z[f"frame_rgb_pixel_mask_{ds}"]) and torch.equal(rmm, z[f"frame_rgb_pixel_mask_mapping_{ds}"])
    assert float(G.median_depth_ref(z["med_depth"], z["med_opacity"])) == float(z["med_plain"])
    assert float(G.median_depth_ref(z["med_depth"], z["med_opacity"], z["med_mask"])) == float(z["med_masked"])


def test_se3_exp_restatement_matches_reference_code():
    """monogs_b200.scene.se3_exp (the float64 restatement the pose-update tests use) vs utils/pose_utils.py:SE3_exp."""
    z = _load("loss_ref")
    for tau, ref in zip(z["se3_tau"], z["se3_exp"]):
        assert torch.allclose(S.se3_exp(tau).float(), ref, rtol=1e-6, atol=1e-7)
