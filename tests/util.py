"""Shared helpers for the parity tests: small seeded scenes, running the product through its Python surface
(which calls the C-ABI), the GPU oracle (oracle/_ref) and the CPU oracle (oracle/golden.py)."""
import torch

from monogs_b200 import scene as S

# (name, P, W, H, fx, fy, cx, cy, sh_degree) — small cases the CPU oracle finishes in seconds
SMALL = {
    "tiny": (400, 64, 48, 60.0, 61.0, 31.3, 23.6, 0),
    "ragged": (1500, 150, 100, 120.0, 118.0, 74.2, 50.3, 0),     # W,H not multiples of 16
    "sh3": (1200, 128, 96, 110.0, 110.0, 63.5, 47.5, 3),
    "dense": (3000, 96, 64, 80.0, 80.0, 47.5, 31.5, 1),
    "sh2": (2048, 96, 64, 80.0, 80.0, 47.5, 31.5, 2),            # 9 coefficients: 27-float rows, 16-byte chunks span rows
    "sh3_full": (4096, 128, 96, 110.0, 110.0, 63.5, 47.5, 3),    # P a multiple of 32: every warp on the staging fast path
    "packed8": (12000, 48, 32, 60.0, 60.0, 23.5, 15.5, 0),       # ~1.5k instances per tile: 8 words per thread in the tile sort
    "packed16": (24000, 48, 32, 60.0, 60.0, 23.5, 15.5, 0),      # ~3k per tile: 16 words per thread
    "crowded": (60000, 48, 32, 60.0, 60.0, 23.5, 15.5, 0),      # 6 tiles with > 4096 instances each: the in-L2 tile sort
}


def small_case(name, seed=0, T_w2c=None):
    P, W, H, fx, fy, cx, cy, deg = SMALL[name]
    sc = S.make_scene(P, W, H, fx, fy, cx, cy, deg, seed)
    if T_w2c is None:
        # a non-identity pose so the rotation part of the pose Jacobian is exercised
        T_w2c = S.se3_exp(torch.tensor([0.05, -0.03, 0.04, 0.02, -0.015, 0.01])).float()
    cam = S.make_camera(W, H, fx, fy, cx, cy, T_w2c)
    return sc, cam


def to_dev(nt, dev="cuda"):
    return type(nt)(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in nt])


def settings(cam, bg, sh_degree):
    from monogs_b200.rasterizer import GaussianRasterizationSettings
    return GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, bg, 1.0, cam.viewmatrix,
                                         cam.projmatrix, cam.projmatrix_raw, sh_degree, cam.campos, False, False)


def golden_args(sc, colors_precomp=None, cov3D_precomp=None):
    a = dict(means3D=sc.means3D, opacities=sc.opacities, sh_degree=sc.sh_degree, scale_modifier=1.0)
    if colors_precomp is None: a["shs"] = sc.shs
    else: a["colors_precomp"] = colors_precomp
    if cov3D_precomp is None: a["scales"], a["rotations"] = sc.scales, sc.rotations
    else: a["cov3D_precomp"] = cov3D_precomp
    return a


def rel_err(a, b):
    """max |a-b| / max |b| — the 'relative' of BASELINE.json's 1e-4 / 1e-3 gates."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def elem_err(a, b, rtol):
    """Element-wise companion of rel_err: the largest |a-b| / (rtol*|b| + atol) with atol = rtol * max|b| * 1e-2, i.e.
    torch.allclose(a, b, rtol, atol) holds iff the value is <= 1.  rel_err alone (a max-norm ratio) cannot see a wrong
    value on an element that is small next to the largest one; this gate can, down to 1 % of the largest magnitude."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    atol = rtol * float(b.abs().max()) * 1e-2
    if atol == 0.0:
        return 0.0 if float((a - b).abs().max()) == 0.0 else float("inf")
    return float(((a - b).abs() / (rtol * b.abs() + atol)).max())


def check_close(a, b, tol, name="", outlier_frac=0.0, outlier_factor=1.0):
    """Both gates: max-norm relative error <= tol (BASELINE.json's 1e-4 / 1e-3) and element-wise allclose.
    outlier_frac / outlier_factor relax the element-wise gate for sums whose terms cancel heavily (a window of eight views
    accumulated per Gaussian with float atomics: the REFERENCE's own run-to-run spread reaches the gate there): at most
    that fraction of the elements may exceed the element-wise bound, and none by more than that factor.  The max-norm
    gate is never relaxed."""
    r = rel_err(a, b)
    assert r <= tol, f"{name}: max-norm relative error {r:.3e} > {tol}"
    ad, bd = a.detach().double().cpu(), b.detach().double().cpu()
    atol = tol * float(bd.abs().max()) * 1e-2
    if atol == 0.0:
        assert float((ad - bd).abs().max()) == 0.0, name
        return
    ratio = (ad - bd).abs() / (tol * bd.abs() + atol)
    frac, worst = float((ratio > 1.0).double().mean()), float(ratio.max())
    assert frac <= outlier_frac and worst <= max(outlier_factor, 1.0) or worst <= 1.0, \
        (f"{name}: element-wise gate (rtol {tol}, atol = rtol * max|b| * 1e-2) exceeded by {frac:.2e} of the elements, worst factor "
         f"{worst:.3f} (allowed: {outlier_frac:.1e} of the elements up to a factor {outlier_factor})")
