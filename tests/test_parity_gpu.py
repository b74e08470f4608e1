"""Parity of the CUDA product (through the C-ABI) against
  (1) the GPU oracle: the unmodified reference CUDA rasterizer built into oracle/_ref (bit-exact integers,
      1e-4 images, 1e-3 gradients — BASELINE.json), and
  (2) the CPU oracle oracle/golden.py at sizes it finishes in seconds.
The same tests also pin the CPU oracle against the reference build."""
import pytest
import torch

from tests import util as U

pytestmark = pytest.mark.gpu

IMG_TOL = 1e-4    # forward images, relative (BASELINE.json north_star)
GRAD_TOL = 1e-3   # gradients, relative: float atomics make the accumulation order nondeterministic


def _product(sc, cam, bg, gc, gd, colors_precomp=None, cov3D_precomp=None, mask=None):
    from monogs_b200 import rasterizer as RZ, _lib as L
    rs = U.settings(cam, bg, sc.sh_degree)
    shs = None if colors_precomp is not None else sc.shs
    scales = None if cov3D_precomp is not None else sc.scales
    rots = None if cov3D_precomp is not None else sc.rotations
    color, radii, depth, opacity, n_touched, st = RZ.raster_forward(rs, sc.means3D, sc.opacities, shs, colors_precomp,
                                                                    scales, rots, cov3D_precomp)
    g = RZ.raster_backward(rs, st, L.GRAD_ALL if mask is None else mask, sc.means3D, shs, colors_precomp, scales, rots,
                           cov3D_precomp, radii, gc, gd, want_tau_per_gaussian=True)
    tau_pose, tau_img = RZ.raster_backward_pose(rs, st, sc.means3D, shs, colors_precomp, scales, rots, cov3D_precomp,
                                                radii, gc, gd, want_image=True)
    return dict(color=color, radii=radii, depth=depth, opacity=opacity, n_touched=n_touched, st=st, g=g,
                tau_pose=tau_pose, tau_img=tau_img, rs=rs)


def _ws_views(st, P, cam):
    """Sorted keys / point list / ranges / n_contrib out of the product's workspace (test hook)."""
    import ctypes as C
    from monogs_b200 import _lib as L
    lib = L.lib()
    off = (C.c_size_t * 17)()
    lib.mgs_debug_layout.restype = C.c_int
    lib.mgs_debug_layout.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.POINTER(C.c_size_t)]
    assert lib.mgs_debug_layout(P, cam.W, cam.H, st.num_rendered, off) == 0
    names = ["header", "depths", "means2D", "conic_opacity", "rgb4", "tiles_touched", "tile_count", "final_T",
             "n_contrib", "ranges", "records", "tile_cursor", "list", "total"]
    o = dict(zip(names, list(off)))
    ws, R = st.workspace, st.num_rendered
    tiles = ((cam.W + 15) // 16) * ((cam.H + 15) // 16)
    # what the blend kernels consume: the packed 48-B records (x y depth id | conic opacity | rgb extent), tile by tile
    recs = ws[o["records"]:o["records"] + 48 * R].view(torch.int32).view(R, 12)
    plist = recs[:, 3].contiguous()
    depth_bits = recs[:, 2].long() & 0xFFFFFFFF
    rng = ws[o["ranges"]:o["ranges"] + 8 * tiles].view(torch.int32).view(tiles, 2).long()
    tile_keys = torch.repeat_interleave(torch.arange(tiles, device=ws.device), rng[:, 1] - rng[:, 0])
    assert tile_keys.numel() == R
    # the library counting-sorts the tile half and sorts (depth, id) per tile; the reference's 64-bit key is the pair
    return dict(keys=(tile_keys << 32) | depth_bits, plist=plist,
                ranges=ws[o["ranges"]:o["ranges"] + 8 * tiles].view(torch.int32).view(tiles, 2),
                n_contrib=ws[o["n_contrib"]:o["n_contrib"] + 4 * cam.W * cam.H].view(torch.int32).view(cam.H, cam.W),
                final_T=ws[o["final_T"]:o["final_T"] + 4 * cam.W * cam.H].view(torch.float32).view(cam.H, cam.W))


def _check_against_ref(sc, cam, bg, gc, gd, pr, colors_precomp=None, cov3D_precomp=None):
    from oracle import ref_gpu
    shs = None if colors_precomp is not None else sc.shs
    scales = None if cov3D_precomp is not None else sc.scales
    rots = None if cov3D_precomp is not None else sc.rotations
    fw = ref_gpu.forward(cam, bg, sc.means3D, sc.opacities, shs=shs, colors_precomp=colors_precomp, scales=scales,
                         rotations=rots, cov3D_precomp=cov3D_precomp, sh_degree=sc.sh_degree)
    gr = ref_gpu.backward(cam, bg, fw, sc.means3D, gc, gd, shs=shs, colors_precomp=colors_precomp, scales=scales,
                          rotations=rots, cov3D_precomp=cov3D_precomp, sh_degree=sc.sh_degree)
    torch.cuda.synchronize()
    # bit-exact integers
    assert pr["st"].num_rendered == fw["R"]
    assert torch.equal(pr["radii"], fw["radii"])
    keys_ref, list_ref = ref_gpu.binning(fw)
    ranges_ref, final_T_ref, n_contrib_ref = ref_gpu.image_state(cam)
    mine = _ws_views(pr["st"], sc.means3D.size(0), cam)
    assert torch.equal(mine["keys"], keys_ref), "sorted (tile|depth) keys differ"
    assert torch.equal(mine["plist"], list_ref), "sorted Gaussian list differs"
    assert torch.equal(mine["ranges"], ranges_ref), "tile ranges differ"
    assert torch.equal(mine["n_contrib"], n_contrib_ref)
    assert torch.equal(pr["n_touched"], fw["n_touched"])
    # images
    for k in ("color", "depth", "opacity"):
        U.check_close(pr[k], fw[k], IMG_TOL, k)
    U.check_close(mine["final_T"], final_T_ref, IMG_TOL, "final_T")
    # gradients: max-norm relative error AND an element-wise allclose (U.check_close)
    g = pr["g"]
    U.check_close(g["means3D"], gr["means3D"], GRAD_TOL, "means3D")
    U.check_close(g["means2D"], gr["means2D"], GRAD_TOL, "means2D")
    U.check_close(g["opacity"], gr["opacity"], GRAD_TOL, "opacity")
    if shs is not None: U.check_close(g["color"], gr["sh"][:, :shs.size(1)], GRAD_TOL, "sh")
    else: U.check_close(g["color"], gr["color"], GRAD_TOL, "color")
    if scales is not None:
        U.check_close(g["scales"], gr["scales"], GRAD_TOL, "scales")
        U.check_close(g["rotations"], gr["rotations"], GRAD_TOL, "rotations")
    else:
        U.check_close(g["cov3D"], gr["cov3D"], GRAD_TOL, "cov3D")
    U.check_close(g["tau_pg"], gr["tau"], GRAD_TOL, "tau per Gaussian")
    tau_ref = gr["tau"].double().sum(0).float()
    U.check_close(g["tau"], tau_ref, GRAD_TOL, "tau")
    # pose-only backward == sum of the reference's per-Gaussian dL_dtau
    U.check_close(pr["tau_pose"], tau_ref, GRAD_TOL, "tau (pose-only backward)")
    U.check_close(pr["tau_img"].double().sum((1, 2)).float(), tau_ref, GRAD_TOL, "tau image")
    return fw, gr


def _setup(name, seed=0):
    from monogs_b200 import scene as S
    sc, cam = U.small_case(name, seed)
    gc, gd = S.output_grads(cam.W, cam.H, seed + 1)
    bg = torch.tensor([0.1, 0.2, 0.3])
    return sc, cam, bg, gc, gd


@pytest.mark.parametrize("name", ["tiny", "ragged", "sh3", "sh2", "sh3_full", "dense", "packed8", "packed16", "crowded"])
def test_product_vs_reference_small(name):
    sc, cam, bg, gc, gd = _setup(name)
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    pr = _product(scd, camd, bg.cuda(), gc.cuda(), gd.cuda())
    _check_against_ref(scd, camd, bg.cuda(), gc.cuda(), gd.cuda(), pr)


@pytest.mark.parametrize("name", ["tiny", "sh3"])
def test_cpu_oracle_vs_reference_and_product(name):
    """Pins oracle/golden.py against the reference build, and the product against the CPU oracle."""
    from oracle import golden as G, ref_gpu
    sc, cam, bg, gc, gd = _setup(name)
    res = G.render(U.golden_args(sc), cam, bg, grads=(gc, gd))
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    pr = _product(scd, camd, bg.cuda(), gc.cuda(), gd.cuda())
    fw, gr = _check_against_ref(scd, camd, bg.cuda(), gc.cuda(), gd.cuda(), pr)
    # integer outputs: CPU fp32 without FMA may differ on borderline Gaussians; report, allow a handful
    mism = int((res["pre"]["radii"].cuda() != fw["radii"]).sum())
    assert mism <= max(2, sc.means3D.size(0) // 500), f"{mism} radii differ between CPU oracle and reference"
    if mism == 0:
        keys_ref, list_ref = ref_gpu.binning(fw)
        assert torch.equal(res["point_list"].cuda().int(), list_ref)
        # tile half exact; depth half within 2 ulp (the CPU oracle has no FMA contraction)
        kc = res["keys"].cuda()
        assert torch.equal(kc >> 32, keys_ref >> 32)
        assert int(((kc & 0xFFFFFFFF) - (keys_ref & 0xFFFFFFFF)).abs().max()) <= 2
    tol_i, tol_g = (IMG_TOL, GRAD_TOL) if mism == 0 else (5e-2, 5e-2)
    for k in ("color", "depth", "opacity"):
        assert U.rel_err(res[k], fw[k]) <= tol_i, k
        assert U.rel_err(res[k], pr[k]) <= tol_i, k
    gg = res["grads"]
    for k, kr in (("means3D", "means3D"), ("means2D", "means2D"), ("opacity", "opacity"), ("scales", "scales"),
                  ("rotations", "rotations"), ("tau", "tau")):
        assert U.rel_err(gg[k], gr[kr]) <= tol_g, k
    assert U.rel_err(gg["color"], gr["sh"][:, :sc.shs.size(1)]) <= tol_g
    assert U.rel_err(gg["tau"].sum(0), pr["tau_pose"]) <= tol_g


def test_precomputed_color_and_covariance():
    """colors_precomp / cov3D_precomp variants (forward.cu:207,243; null-pointer options, Q14)."""
    from oracle import golden as G
    sc, cam, bg, gc, gd = _setup("ragged")
    cov = G.cov3d(sc.scales, sc.rotations, 1.0).contiguous()
    cols = torch.rand(sc.means3D.size(0), 3, generator=torch.Generator().manual_seed(5))
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    pr = _product(scd, camd, bg.cuda(), gc.cuda(), gd.cuda(), colors_precomp=cols.cuda(), cov3D_precomp=cov.cuda())
    _check_against_ref(scd, camd, bg.cuda(), gc.cuda(), gd.cuda(), pr, colors_precomp=cols.cuda(), cov3D_precomp=cov.cuda())


def test_empty_and_all_culled():
    """P == 0 is legal and returns zero-filled outputs (rasterize_points.cu:90); a scene entirely behind the
    camera renders the background."""
    from monogs_b200 import rasterizer as RZ, scene as S
    sc, cam, bg, gc, gd = _setup("tiny")
    camd = U.to_dev(cam)
    rs = U.settings(camd, bg.cuda(), 0)
    e = lambda *s: torch.zeros(*s, device="cuda")
    color, radii, depth, opacity, n_touched, st = RZ.raster_forward(rs, e(0, 3), e(0, 1), e(0, 1, 3), None, e(0, 3), e(0, 4), None)
    assert color.abs().sum() == 0 and depth.abs().sum() == 0 and radii.numel() == 0 and st.num_rendered == 0
    scd = U.to_dev(sc)
    behind = scd.means3D.clone(); behind[:, 2] = -behind[:, 2].abs() - 1.0
    color, radii, depth, opacity, n_touched, st = RZ.raster_forward(rs, behind, scd.opacities, scd.shs, None, scd.scales, scd.rotations, None)
    assert st.num_rendered == 0 and int(radii.sum()) == 0
    assert torch.allclose(color, bg.cuda()[:, None, None].expand_as(color))
    assert opacity.abs().sum() == 0


def test_autograd_surface_matches_reference_wrapper():
    """The drop-in package: same call as gaussian_renderer/__init__.py:151-168, gradients for every input,
    grad_theta / grad_rho shaped [1,3] and equal to tau[3:], tau[:3] (__init__.py:339-342)."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from oracle import ref_gpu
    sc, cam, bg, gc, gd = _setup("dense")
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    leaf = lambda t: t.clone().requires_grad_(True)
    means3D, opac, shs, scales, rots = leaf(scd.means3D), leaf(scd.opacities), leaf(scd.shs), leaf(scd.scales), leaf(scd.rotations)
    means2D = torch.zeros_like(means3D, requires_grad=True)
    theta, rho = torch.zeros(3, device="cuda", requires_grad=True), torch.zeros(3, device="cuda", requires_grad=True)
    rs = GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, bg.cuda(), 1.0, camd.viewmatrix,
                                       camd.projmatrix, camd.projmatrix_raw, sc.sh_degree, camd.campos, False, False)
    color, radii, depth, opacity, n_touched = GaussianRasterizer(rs)(
        means3D=means3D, means2D=means2D, shs=shs, colors_precomp=None, opacities=opac, scales=scales, rotations=rots,
        cov3D_precomp=None, theta=theta, rho=rho)
    ((color * gc.cuda()).sum() + (depth * gd.cuda()).sum() + opacity.sum() * 0.123).backward()
    fw = ref_gpu.forward(camd, bg.cuda(), scd.means3D, scd.opacities, shs=scd.shs, scales=scd.scales, rotations=scd.rotations, sh_degree=sc.sh_degree)
    gr = ref_gpu.backward(camd, bg.cuda(), fw, scd.means3D, gc.cuda(), gd.cuda(), shs=scd.shs, scales=scd.scales, rotations=scd.rotations, sh_degree=sc.sh_degree)
    tau = gr["tau"].double().sum(0).float()
    assert theta.grad.shape == (3,) and rho.grad.shape == (3,)
    assert U.rel_err(rho.grad, tau[:3]) <= GRAD_TOL and U.rel_err(theta.grad, tau[3:]) <= GRAD_TOL
    assert U.rel_err(means3D.grad, gr["means3D"]) <= GRAD_TOL
    assert U.rel_err(means2D.grad, gr["means2D"]) <= GRAD_TOL
    assert U.rel_err(shs.grad, gr["sh"][:, :scd.shs.size(1)]) <= GRAD_TOL
    assert U.rel_err(opac.grad, gr["opacity"]) <= GRAD_TOL   # the 0.123*opacity term is dropped like the reference (Q7)
    assert U.rel_err(scales.grad, gr["scales"]) <= GRAD_TOL and U.rel_err(rots.grad, gr["rotations"]) <= GRAD_TOL
    # pose-only path: nothing but theta/rho requires grad (tracking, slam_frontend.py:1799-1800)
    theta2, rho2 = torch.zeros(3, device="cuda", requires_grad=True), torch.zeros(3, device="cuda", requires_grad=True)
    c2, _, d2, _, _ = GaussianRasterizer(rs)(means3D=scd.means3D, means2D=torch.zeros_like(scd.means3D), shs=scd.shs,
                                             opacities=scd.opacities, scales=scd.scales, rotations=scd.rotations,
                                             theta=theta2, rho=rho2)
    ((c2 * gc.cuda()).sum() + (d2 * gd.cuda()).sum()).backward()
    assert U.rel_err(rho2.grad, tau[:3]) <= GRAD_TOL and U.rel_err(theta2.grad, tau[3:]) <= GRAD_TOL
    with pytest.raises(Exception, match="excatly one of either SHs"):
        GaussianRasterizer(rs)(means3D=scd.means3D, means2D=means2D, opacities=scd.opacities, scales=scd.scales, rotations=scd.rotations)


def test_sketch_jacobian_matches_reference():
    """Revives the reference's commented-out check (__init__.py:255-305): row i of the sketch Jacobian equals
    the standard backward with the output gradients masked to sketch_indices == i; here against the reference's
    own sketch kernel summed over Gaussians (__init__.py:253)."""
    from monogs_b200 import rasterizer as RZ
    from oracle import ref_gpu
    sc, cam, bg, gc, gd = _setup("ragged")
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    d, stacks = 8, 3
    g = torch.Generator().manual_seed(3)
    idx = torch.randint(-1, d, (stacks, cam.H, cam.W), generator=g, dtype=torch.int32).cuda()
    pr = _product(scd, camd, bg.cuda(), gc.cuda(), gd.cuda())
    mine = RZ.sketch_reduce(pr["tau_img"], idx, d)
    fw = ref_gpu.forward(camd, bg.cuda(), scd.means3D, scd.opacities, shs=scd.shs, scales=scd.scales, rotations=scd.rotations, sh_degree=sc.sh_degree)
    for j in range(stacks):
        ref_tau = ref_gpu.backward_sketch(camd, bg.cuda(), fw, scd.means3D, gc.cuda(), gd.cuda(), idx[j].contiguous(), d,
                                          shs=scd.shs, scales=scd.scales, rotations=scd.rotations, sh_degree=sc.sh_degree)
        assert U.rel_err(mine[j], ref_tau.double().sum(0).float()) <= GRAD_TOL
    # and the masked-standard-backward identity for one row
    i = 2
    m = (idx[0] == i).float()
    gm = ref_gpu.backward(camd, bg.cuda(), fw, scd.means3D, gc.cuda() * m, gd.cuda() * m, shs=scd.shs, scales=scd.scales, rotations=scd.rotations, sh_degree=sc.sh_degree)
    assert U.rel_err(mine[0, i], gm["tau"].double().sum(0).float()) <= GRAD_TOL


def test_mark_visible_and_knn():
    from diff_gaussian_rasterization import GaussianRasterizer
    from simple_knn._C import distCUDA2
    from oracle import ref_gpu, golden as G
    sc, cam, bg, gc, gd = _setup("dense")
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    vis = GaussianRasterizer(U.settings(camd, bg.cuda(), 0)).markVisible(scd.means3D)
    assert torch.equal(vis, ref_gpu.mark_visible(scd.means3D, camd))
    for P in (1, 3, 4, 129, 5000, 40000):
        pts = torch.randn(P, 3, generator=torch.Generator().manual_seed(P)).cuda() * 2 + 0.5
        mine, ref = distCUDA2(pts), ref_gpu.knn_dist2(pts)
        assert torch.equal(mine, ref), f"knn differs at P={P}: {(mine != ref).sum()} entries"
    pts = torch.randn(3000, 3, generator=torch.Generator().manual_seed(7))
    assert U.rel_err(distCUDA2(pts.cuda()), G.knn_dist2(pts)) <= 1e-6


@pytest.mark.parametrize("cfg", ["C1", "C2", "C3", "C5"])
def test_full_size_configs_vs_reference(cfg):
    """BASELINE.json configs 1, 2, 3 and 5 (2M Gaussians, 1920x1080, SH degree 3) at full size against the reference build."""
    from monogs_b200 import scene as S
    sc, cam = S.make_config(cfg)
    gc, gd = S.output_grads(cam.W, cam.H)
    bg = torch.zeros(3)
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    pr = _product(scd, camd, bg.cuda(), gc.cuda(), gd.cuda())
    _check_against_ref(scd, camd, bg.cuda(), gc.cuda(), gd.cuda(), pr)


def test_full_size_properties_c3():
    """Size-independent properties at config 3 size (500k, 1200x680): sorted keys are non-decreasing, ranges tile
    the list, opacity = 1 - final_T, tau image sums to the full backward's tau, linearity in the output grads."""
    from monogs_b200 import scene as S, rasterizer as RZ, _lib as L
    sc, cam = S.make_config("C3")
    gc, gd = S.output_grads(cam.W, cam.H)
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    bg = torch.tensor([0.3, 0.1, 0.2]).cuda()
    pr = _product(scd, camd, bg, gc.cuda(), gd.cuda(), mask=L.GRAD_MEANS3D | L.GRAD_TAU)
    v = _ws_views(pr["st"], sc.means3D.size(0), cam)
    k = v["keys"]
    assert bool((k[1:] >= k[:-1]).all())
    r = v["ranges"].long()
    nz = r[:, 1] > r[:, 0]
    assert int((r[nz, 1] - r[nz, 0]).sum()) == pr["st"].num_rendered
    assert U.rel_err(pr["opacity"][0], 1 - v["final_T"]) <= 1e-6
    assert U.rel_err(pr["tau_pose"], pr["g"]["tau"]) <= GRAD_TOL
    rs = pr["rs"]
    t2, _ = RZ.raster_backward_pose(rs, pr["st"], scd.means3D, scd.shs, None, scd.scales, scd.rotations, None, pr["radii"],
                                    2.0 * gc.cuda(), 2.0 * gd.cuda())
    assert U.rel_err(t2, 2.0 * pr["tau_pose"]) <= 1e-4


@pytest.mark.parametrize("name", ["sh2", "sh3_full"])
def test_sh_staging_with_unaligned_rows(name):
    """The shared-memory SH staging (sh_stage.cuh) moves 16-byte chunks when a warp's block of rows is 16-byte aligned
    and falls back to element-wise moves otherwise: a coefficient array that starts one float into its storage must
    give bit-identical colours and the same SH gradients (up to the order of the float atomics of the backward blend)."""
    sc, cam, bg, gc, gd = _setup(name)
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    ref = _product(scd, camd, bg.cuda(), gc.cuda(), gd.cuda())
    flat = torch.empty(scd.shs.numel() + 1, device="cuda")
    flat[1:] = scd.shs.reshape(-1)
    shifted = flat[1:].view_as(scd.shs)
    assert shifted.is_contiguous() and shifted.data_ptr() % 16 == 4
    got = _product(scd._replace(shs=shifted), camd, bg.cuda(), gc.cuda(), gd.cuda())
    assert torch.equal(got["color"], ref["color"])
    assert U.rel_err(got["g"]["color"], ref["g"]["color"]) <= 1e-5 and U.rel_err(got["g"]["means3D"], ref["g"]["means3D"]) <= 1e-5


def test_sketch_autograd_branch_matches_reference():
    """The branch MonoGS's second-order phase actually calls (utils/slam_frontend.py:655-666 through
    diff_gaussian_rasterization/__init__.py:216-253): GaussianRasterizer(..., sketch_mode=1, sketch_dim, stack_dim,
    sketch_dtau, sketch_indices[repeat, stack, H, W]) followed by one backward(retain_graph=True) per repeat.  Repeat r
    must use sketch_indices[r], return its rows as sketch_dtau.grad [stack, sketch, 6] and no other gradient; rows are
    checked against the reference's own sketch kernel, stack by stack."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from monogs_b200 import scene as S
    from oracle import ref_gpu
    sc, cam, bg, gc, gd = _setup("ragged")
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    d, stacks, repeats = 8, 3, 2
    g = torch.Generator().manual_seed(3)
    idx = torch.randint(-1, d, (repeats, stacks, cam.H, cam.W), generator=g, dtype=torch.int32).cuda()
    sketch_dtau = torch.empty(stacks, d, 6, device="cuda", requires_grad=True)
    theta, rho = torch.zeros(3, device="cuda", requires_grad=True), torch.zeros(3, device="cuda", requires_grad=True)
    rs = GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, bg.cuda(), 1.0, camd.viewmatrix,
                                       camd.projmatrix, camd.projmatrix_raw, sc.sh_degree, camd.campos, False, False)
    color, radii, depth, opacity, n_touched = GaussianRasterizer(rs)(
        means3D=scd.means3D, means2D=torch.zeros_like(scd.means3D), shs=scd.shs, opacities=scd.opacities, scales=scd.scales,
        rotations=scd.rotations, theta=theta, rho=rho, sketch_mode=1, sketch_dim=d, stack_dim=stacks, sketch_dtau=sketch_dtau,
        sketch_indices=idx)
    fw = ref_gpu.forward(camd, bg.cuda(), scd.means3D, scd.opacities, shs=scd.shs, scales=scd.scales, rotations=scd.rotations,
                         sh_degree=sc.sh_degree)
    for r in range(repeats):
        gcr, gdr = S.output_grads(cam.W, cam.H, 10 + r)      # a different upstream gradient per repeat (rand_weights[r])
        gcr, gdr = gcr.cuda(), gdr.cuda()
        sketch_dtau.grad = None
        ((color * gcr).sum() + (depth * gdr).sum()).backward(retain_graph=True)
        got = sketch_dtau.grad
        assert got.shape == (stacks, d, 6)
        assert theta.grad is None and rho.grad is None          # sketch mode returns no pose gradient (__init__.py:317-325)
        for j in range(stacks):
            ref_tau = ref_gpu.backward_sketch(camd, bg.cuda(), fw, scd.means3D, gcr, gdr, idx[r, j].contiguous(), d, shs=scd.shs,
                                              scales=scd.scales, rotations=scd.rotations, sh_degree=sc.sh_degree)
            U.check_close(got[j], ref_tau.double().sum(0).float(), GRAD_TOL, f"repeat {r} stack {j}")


def test_debug_flag_is_honoured(tmp_path, monkeypatch):
    """debug=True (auxiliary.h:166-173, __init__.py:109-116): every launch sequence is followed by a synchronize + error
    check, results are unchanged, and a failing call dumps its inputs to snapshot_fw.dump before re-raising."""
    import os
    from monogs_b200 import rasterizer as RZ
    sc, cam, bg, gc, gd = _setup("tiny")
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    monkeypatch.chdir(tmp_path)
    ref = _product(scd, camd, bg.cuda(), gc.cuda(), gd.cuda())
    rs = U.settings(camd, bg.cuda(), sc.sh_degree)._replace(debug=True)
    color, radii, depth, opacity, n_touched, st = RZ.raster_forward(rs, scd.means3D, scd.opacities, scd.shs, None, scd.scales,
                                                                    scd.rotations, None)
    assert torch.equal(color, ref["color"]) and torch.equal(radii, ref["radii"])
    assert not os.path.exists("snapshot_fw.dump")
    bad = rs._replace(sh_degree=7)       # rejected by the library: the wrapper must leave a snapshot behind
    with pytest.raises(RuntimeError):
        RZ.raster_forward(bad, scd.means3D, scd.opacities, scd.shs, None, scd.scales, scd.rotations, None)
    assert os.path.exists("snapshot_fw.dump")
    snap = torch.load("snapshot_fw.dump")
    assert torch.equal(snap[1], sc.means3D)


def test_speculative_forward_recovers_from_a_wrong_capacity_guess():
    """The drop-in forward sizes its workspace from the previous call of the same (P, W, H); when the next scene has far
    more tile instances than the guess, the binning + blend are repeated at the right size before the call returns."""
    from monogs_b200 import rasterizer as RZ
    from oracle import ref_gpu
    sc, cam, bg, gc, gd = _setup("dense")
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    rs = U.settings(camd, bg.cuda(), sc.sh_degree)
    small = scd.scales * 0.05
    _, _, _, _, _, st0 = RZ.raster_forward(rs, scd.means3D, scd.opacities, scd.shs, None, small, scd.rotations, None)   # primes the guess
    _, _, _, _, _, st1 = RZ.raster_forward(rs, scd.means3D, scd.opacities, scd.shs, None, small, scd.rotations, None)   # speculative, fits
    assert st1.num_rendered == st0.num_rendered
    color, radii, depth, opacity, n_touched, st = RZ.raster_forward(rs, scd.means3D, scd.opacities, scd.shs, None, scd.scales * 3.0,
                                                                    scd.rotations, None)
    assert st.num_rendered > 1.25 * st0.num_rendered + 4096, "the test scene must overflow the guess"
    fw = ref_gpu.forward(camd, bg.cuda(), scd.means3D, scd.opacities, shs=scd.shs, scales=scd.scales * 3.0, rotations=scd.rotations,
                         sh_degree=sc.sh_degree)
    assert st.num_rendered == fw["R"] and torch.equal(radii, fw["radii"]) and torch.equal(n_touched, fw["n_touched"])
    U.check_close(color, fw["color"], IMG_TOL, "color")
    g = RZ.raster_backward(rs, st, 1, scd.means3D, scd.shs, None, scd.scales * 3.0, scd.rotations, None, radii, gc.cuda(), gd.cuda())
    gr = ref_gpu.backward(camd, bg.cuda(), fw, scd.means3D, gc.cuda(), gd.cuda(), shs=scd.shs, scales=scd.scales * 3.0,
                          rotations=scd.rotations, sh_degree=sc.sh_degree)
    U.check_close(g["means3D"], gr["means3D"], GRAD_TOL, "means3D")


@pytest.mark.parametrize("name,use_depth", [("ragged", True), ("dense", False), ("packed8", True)])
def test_dropin_single_pass_tracking_matches_reference(name, use_depth):
    """Tracking through the drop-in surface (only theta / rho require a gradient): from the second call of a size on, the
    forward is mgs_raster_forward_pose and the backward the deferred per-pixel contraction — images bit-identical to the
    standard forward, tangent <= 1e-3 against the reference's backward, with and without a depth gradient, and through the
    sketch branch."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from monogs_b200 import rasterizer as RZ
    from oracle import ref_gpu
    sc, cam, bg, gc, gd = _setup(name)
    if sc.sh_degree != 0:
        sc = sc._replace(shs=sc.shs[:, :1].contiguous(), sh_degree=0)
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    rs = GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, bg.cuda(), 1.0, camd.viewmatrix,
                                       camd.projmatrix, camd.projmatrix_raw, 0, camd.campos, False, False)
    fw = ref_gpu.forward(camd, bg.cuda(), scd.means3D, scd.opacities, shs=scd.shs, scales=scd.scales, rotations=scd.rotations, sh_degree=0)
    gdz = gd.cuda() if use_depth else torch.zeros_like(gd).cuda()
    gr = ref_gpu.backward(camd, bg.cuda(), fw, scd.means3D, gc.cuda(), gdz, shs=scd.shs, scales=scd.scales, rotations=scd.rotations, sh_degree=0)
    tau_ref = gr["tau"].double().sum(0).float()
    for call in range(3):          # call 0 primes the capacity guess (two-stage forward); 1 and 2 take the single pass
        theta, rho = torch.zeros(3, device="cuda", requires_grad=True), torch.zeros(3, device="cuda", requires_grad=True)
        color, radii, depth, opacity, n_touched = GaussianRasterizer(rs)(
            means3D=scd.means3D, means2D=torch.zeros_like(scd.means3D), shs=scd.shs, opacities=scd.opacities, scales=scd.scales,
            rotations=scd.rotations, theta=theta, rho=rho)
        if call > 0:
            assert color.grad_fn.state.pose_state is not None, "the single-pass forward was taken"
        loss = (color * gc.cuda()).sum() + ((depth * gd.cuda()).sum() if use_depth else 0.0)
        loss.backward()
        assert torch.equal(radii, fw["radii"]) and torch.equal(n_touched, fw["n_touched"])
        U.check_close(color, fw["color"], IMG_TOL, "color"); U.check_close(depth, fw["depth"], IMG_TOL, "depth")
        U.check_close(opacity, fw["opacity"], IMG_TOL, "opacity")
        U.check_close(torch.cat([rho.grad, theta.grad]), tau_ref, GRAD_TOL, f"tau, call {call}")
    # sketch branch on the single-pass state
    d, stacks = 8, 2
    idx = torch.randint(-1, d, (1, stacks, cam.H, cam.W), generator=torch.Generator().manual_seed(3), dtype=torch.int32).cuda()
    sketch_dtau = torch.empty(stacks, d, 6, device="cuda", requires_grad=True)
    color, _, depth, _, _ = GaussianRasterizer(rs)(
        means3D=scd.means3D, means2D=torch.zeros_like(scd.means3D), shs=scd.shs, opacities=scd.opacities, scales=scd.scales,
        rotations=scd.rotations, theta=torch.zeros(3, device="cuda", requires_grad=True), rho=torch.zeros(3, device="cuda", requires_grad=True),
        sketch_mode=1, sketch_dim=d, stack_dim=stacks, sketch_dtau=sketch_dtau, sketch_indices=idx)
    assert color.grad_fn.state.pose_state is not None
    ((color * gc.cuda()).sum() + ((depth * gd.cuda()).sum() if use_depth else 0.0)).backward()
    for j in range(stacks):
        ref_tau = ref_gpu.backward_sketch(camd, bg.cuda(), fw, scd.means3D, gc.cuda(), gdz, idx[0, j].contiguous(), d, shs=scd.shs,
                                          scales=scd.scales, rotations=scd.rotations, sh_degree=0)
        U.check_close(sketch_dtau.grad[j], ref_tau.double().sum(0).float(), GRAD_TOL, f"sketch stack {j}")
    # the switch: with SINGLE_PASS_TRACKING off the same call takes the two-pass backward
    RZ.SINGLE_PASS_TRACKING = False
    try:
        theta = torch.zeros(3, device="cuda", requires_grad=True)
        color, _, _, _, _ = GaussianRasterizer(rs)(means3D=scd.means3D, means2D=torch.zeros_like(scd.means3D), shs=scd.shs,
                                                  opacities=scd.opacities, scales=scd.scales, rotations=scd.rotations, theta=theta,
                                                  rho=torch.zeros(3, device="cuda", requires_grad=True))
        assert color.grad_fn.state.pose_state is None
    finally:
        RZ.SINGLE_PASS_TRACKING = True
