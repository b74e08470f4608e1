"""mgs_raster_backward_pixels (the select_pixels = True branch of the reference's backward; SURVEY.md kernel K8) against
the reference's OWN renderPixelSampleCUDA (backward.cu:592-797), compiled into oracle/_ref and reached through the same
Rasterizer::backward call the reference wrapper makes (rasterize_points.cu:189,233)."""
import pytest
import torch

from tests import util as U

pytestmark = pytest.mark.gpu

GRAD_TOL = 1e-3     # BASELINE.json north_star: gradients within 1e-3 relative of the reference


def _case(name, seed=0):
    from monogs_b200 import scene as S
    sc, cam = U.small_case(name, seed)
    gc, gd = S.output_grads(cam.W, cam.H, seed + 1)
    bg = torch.tensor([0.1, 0.2, 0.3]).cuda()
    return U.to_dev(sc), U.to_dev(cam), bg, gc.cuda(), gd.cuda()


def _both(sc, cam, bg, gc, gd, sel):
    from monogs_b200 import rasterizer as RZ, _lib as L
    from oracle import ref_gpu
    rs = U.settings(cam, bg, sc.sh_degree)
    color, radii, depth, opacity, n_touched, st = RZ.raster_forward(rs, sc.means3D, sc.opacities, sc.shs, None, sc.scales,
                                                                    sc.rotations, None)
    g = RZ.raster_backward(rs, st, L.GRAD_ALL, sc.means3D, sc.shs, None, sc.scales, sc.rotations, None, radii, gc, gd,
                           want_tau_per_gaussian=True, selected_pixel_indices=sel)
    fw = ref_gpu.forward(cam, bg, sc.means3D, sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations,
                         sh_degree=sc.sh_degree)
    gr = ref_gpu.backward(cam, bg, fw, sc.means3D, gc, gd, shs=sc.shs, scales=sc.scales, rotations=sc.rotations,
                          sh_degree=sc.sh_degree, selected_pixel_indices=sel)
    torch.cuda.synchronize()
    return g, gr, (rs, st, radii)


def _compare(g, gr, M):
    U.check_close(g["means3D"], gr["means3D"], GRAD_TOL, "means3D")
    U.check_close(g["means2D"], gr["means2D"], GRAD_TOL, "means2D")
    U.check_close(g["opacity"], gr["opacity"], GRAD_TOL, "opacity")
    U.check_close(g["color"], gr["sh"][:, :M], GRAD_TOL, "sh")
    U.check_close(g["scales"], gr["scales"], GRAD_TOL, "scales")
    U.check_close(g["rotations"], gr["rotations"], GRAD_TOL, "rotations")
    U.check_close(g["tau_pg"], gr["tau"], GRAD_TOL, "tau per Gaussian")
    U.check_close(g["tau"], gr["tau"].double().sum(0).float(), GRAD_TOL, "tau")


def _short_list_pixels(sc, cam, bg, sel):
    """The listed pixels with at most 128 contributors.  renderPixelSampleCUDA stages a pixel's list in batches of
    PIXEL_SAMPLE_BLOCK_SIZE = 128 and lets the other threads overwrite the batch while thread 0 is still reading its last
    entry (no barrier between the j-loop and the next load, backward.cu:680-700): for longer lists the reference's own
    result changes from run to run (3e-2 relative on the `dense` scene), so only race-free pixels can be compared with it."""
    from oracle import ref_gpu
    ref_gpu.forward(cam, bg, sc.means3D, sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations,
                    sh_degree=sc.sh_degree)
    _, _, n_contrib = ref_gpu.image_state(cam)
    return sel[n_contrib.flatten()[sel.long()] <= 128].contiguous()


@pytest.mark.parametrize("name,frac", [("tiny", 0.05), ("ragged", 0.02), ("sh3", 0.1), ("sh2", 0.3)])
def test_pixel_sample_backward_vs_reference_kernel(name, frac):
    sc, cam, bg, gc, gd = _case(name)
    n = max(1, int(cam.W * cam.H * frac))
    sel = torch.randperm(cam.W * cam.H, generator=torch.Generator().manual_seed(3))[:n].int().cuda()
    sel = _short_list_pixels(sc, cam, bg, sel)
    assert sel.numel() >= 1
    g, gr, _ = _both(sc, cam, bg, gc, gd, sel)
    assert float(gr["means3D"].abs().max()) > 0.0
    _compare(g, gr, sc.shs.size(1))


def test_pixel_sample_backward_long_lists_vs_reference_dense_kernel():
    """Pixels with more than 128 contributors (where the reference's pixel kernel races, see _short_list_pixels): the
    operation is, by definition, the dense backward with the gradient images zeroed outside the list — compared with the
    reference's dense renderCUDA on those masked images."""
    from monogs_b200 import rasterizer as RZ, _lib as L
    from oracle import ref_gpu
    sc, cam, bg, gc, gd = _case("dense")
    sel = torch.randperm(cam.W * cam.H, generator=torch.Generator().manual_seed(3))[:cam.W * cam.H // 2].int().cuda()
    mask = torch.zeros(cam.H * cam.W, device="cuda")
    mask[sel.long()] = 1.0
    mask = mask.view(1, cam.H, cam.W)
    rs = U.settings(cam, bg, sc.sh_degree)
    color, radii, depth, opacity, n_touched, st = RZ.raster_forward(rs, sc.means3D, sc.opacities, sc.shs, None, sc.scales,
                                                                    sc.rotations, None)
    g = RZ.raster_backward(rs, st, L.GRAD_ALL, sc.means3D, sc.shs, None, sc.scales, sc.rotations, None, radii, gc, gd,
                           want_tau_per_gaussian=True, selected_pixel_indices=sel)
    fw = ref_gpu.forward(cam, bg, sc.means3D, sc.opacities, shs=sc.shs, scales=sc.scales, rotations=sc.rotations,
                         sh_degree=sc.sh_degree)
    _, _, n_contrib = ref_gpu.image_state(cam)
    assert int(n_contrib.flatten()[sel.long()].max()) > 128
    gr = ref_gpu.backward(cam, bg, fw, sc.means3D, gc * mask, gd * mask, shs=sc.shs, scales=sc.scales,
                          rotations=sc.rotations, sh_degree=sc.sh_degree)
    torch.cuda.synchronize()
    _compare(g, gr, sc.shs.size(1))


def test_pixel_sample_backward_duplicates_and_single_tile():
    """A pixel listed twice sends its gradient twice (the reference launches one block per list entry); a list confined to
    one tile leaves every Gaussian outside that tile with an exactly zero gradient."""
    sc, cam, bg, gc, gd = _case("ragged")
    ys, xs = torch.meshgrid(torch.arange(16, 32), torch.arange(32, 48), indexing="ij")      # tile (2, 1)
    tile_pixels = (ys * cam.W + xs).flatten().int()
    sel = torch.cat([tile_pixels, tile_pixels[::3], tile_pixels[:5]]).cuda()                 # duplicates, some triple
    g, gr, _ = _both(sc, cam, bg, gc, gd, sel)
    _compare(g, gr, sc.shs.size(1))
    untouched = gr["means2D"].abs().sum(1) == 0
    assert int(untouched.sum()) > 0 and float(g["means2D"][untouched].abs().max()) == 0.0


def test_pixel_sample_backward_equals_masked_dense_backward_c1():
    """At the size of config C1 (320x240, 50k Gaussians) with 2 % of the pixels listed: the result equals the dense
    backward fed with the gradient images zeroed outside the list (the definition of the operation), and an empty list
    gives zero gradients."""
    from monogs_b200 import scene as S, rasterizer as RZ, _lib as L
    sc, cam = S.make_config("C1")
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    gc, gd = S.output_grads(cam.W, cam.H, 4)
    gc, gd = gc.cuda(), gd.cuda()
    bg = torch.zeros(3).cuda()
    rs = U.settings(camd, bg, sc.sh_degree)
    color, radii, depth, opacity, n_touched, st = RZ.raster_forward(rs, scd.means3D, scd.opacities, scd.shs, None, scd.scales,
                                                                    scd.rotations, None)
    n = cam.W * cam.H // 50
    sel = torch.randperm(cam.W * cam.H, generator=torch.Generator().manual_seed(9))[:n].int().cuda()
    mask = torch.zeros(cam.H * cam.W, device="cuda")
    mask[sel.long()] = 1.0
    mask = mask.view(1, cam.H, cam.W)
    args = (rs, st, L.GRAD_ALL, scd.means3D, scd.shs, None, scd.scales, scd.rotations, None, radii)
    g = RZ.raster_backward(*args, gc, gd, selected_pixel_indices=sel)
    gm = RZ.raster_backward(*args, gc * mask, gd * mask)
    for k in ("means3D", "means2D", "opacity", "color", "scales", "rotations", "tau"):
        U.check_close(g[k], gm[k], 1e-5, k)
    g0 = RZ.raster_backward(*args, gc, gd, selected_pixel_indices=torch.zeros(0, dtype=torch.int32, device="cuda"))
    assert all(float(g0[k].abs().max()) == 0.0 for k in ("means3D", "means2D", "opacity", "color", "scales", "rotations", "tau"))
    # indices outside the image are ignored (the reference's `inside` test)
    g1 = RZ.raster_backward(*args, gc, gd, selected_pixel_indices=torch.cat([sel, torch.tensor([cam.W * cam.H, -1, 1 << 30],
                                                                                                  dtype=torch.int32).cuda()]))
    for k in ("means3D", "tau"):
        assert torch.equal(g1[k], g[k]) or U.rel_err(g1[k], g[k]) <= 1e-6
