"""MapEngine (mgs_map_*: activations once per iteration, per-view fused loss, accumulated gradients, statistics)
against the same mapping iteration written the reference's way: torch activations (gaussian_model.py:77-97) ->
GaussianRasterizer (the drop-in surface, itself parity-tested against the reference CUDA build) -> get_loss_mapping in
plain torch (utils/slam_utils.py:224-253) -> isotropic regulariser (slam_backend.py:244-246) -> autograd."""
import pytest
import torch

from tests import util as U

pytestmark = pytest.mark.gpu
TOL = 1e-3   # gradients (BASELINE.json); observed ~1e-6


def _raw_params(sc, isotropic=False):
    g = torch.Generator().manual_seed(11)
    p = {"xyz": sc.means3D.clone(), "f_dc": sc.shs[:, :1].clone().contiguous(),
         "opacity": torch.log(sc.opacities.clamp(1e-4, 1 - 1e-4) / (1 - sc.opacities.clamp(1e-4, 1 - 1e-4))),
         "scaling": torch.log(sc.scales), "rotation": sc.rotations * (0.5 + torch.rand(sc.rotations.size(0), 1, generator=g) * 2)}
    if sc.shs.size(1) > 1:
        p["f_rest"] = sc.shs[:, 1:].clone().contiguous()
    if isotropic:
        p["scaling"] = p["scaling"].mean(dim=1, keepdim=True)
    return {k: v.cuda().contiguous().requires_grad_(True) for k, v in p.items()}


def _views(sc, cam, n, with_depth, with_exposure):
    from monogs_b200 import scene as S
    from monogs_b200.mapping import MapView
    from monogs_b200.rasterizer import GaussianRasterizationSettings
    g = torch.Generator().manual_seed(5)
    bg = torch.tensor([0.1, 0.2, 0.3]).cuda()
    out = []
    for i in range(n):
        tau = torch.tensor(S.GT_TAU) * (i - 1) * 2.0
        c = U.to_dev(S.make_camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, S.se3_exp(tau).float()))
        rs = GaussianRasterizationSettings(c.H, c.W, c.tanfovx, c.tanfovy, bg, 1.0, c.viewmatrix, c.projmatrix,
                                           c.projmatrix_raw, sc.sh_degree, c.campos, False, False)
        gt = torch.rand(3, cam.H, cam.W, generator=g).cuda()
        gt[:, : cam.H // 5] = 0.0     # a band below the rgb boundary threshold
        gd = (torch.rand(1, cam.H, cam.W, generator=g) * 4).cuda() if with_depth else None
        if gd is not None: gd[:, :, : cam.W // 6] = 0.0
        a = torch.tensor([0.1 * (i + 1)]).cuda().requires_grad_(True) if with_exposure and i != 1 else None
        b = torch.tensor([0.02 * (i - 1)]).cuda().requires_grad_(True) if a is not None else None
        out.append(MapView(rs, gt, gd, a, b, update_stats=(i != 2)))
    return out


def _reference_iteration(p, views, sh_degree, alpha=0.95, thr=0.01):
    """BackEnd.map's loss construction in plain torch on the drop-in rasterizer."""
    from diff_gaussian_rasterization import GaussianRasterizer
    total = 0
    per_view = []
    for v in views:
        means2D = torch.zeros_like(p["xyz"], requires_grad=True)
        theta, rho = torch.zeros(3, device="cuda", requires_grad=True), torch.zeros(3, device="cuda", requires_grad=True)
        scal = torch.exp(p["scaling"])
        if scal.size(-1) == 1: scal = scal.repeat(1, 3)
        shs = torch.cat((p["f_dc"], p["f_rest"]), dim=1) if "f_rest" in p else p["f_dc"]
        image, radii, depth, opacity, n_touched = GaussianRasterizer(v.settings._replace(sh_degree=sh_degree))(
            means3D=p["xyz"], means2D=means2D, shs=shs, opacities=torch.sigmoid(p["opacity"]), scales=scal,
            rotations=torch.nn.functional.normalize(p["rotation"]), theta=theta, rho=rho)
        image_ab = image if v.exposure_a is None else (torch.abs(v.exposure_a) + 1e-8) * image + v.exposure_b
        m = (v.gt_image.sum(dim=0) > thr)[None]
        l1 = torch.abs(image_ab * m - v.gt_image * m)
        if v.gt_depth is None:
            loss = l1.mean()
        else:
            dm = v.gt_depth > 0.01
            loss = alpha * l1.mean() + (1 - alpha) * torch.abs(depth * dm - v.gt_depth * dm).mean()
        total = total + loss
        per_view.append(dict(means2D=means2D, theta=theta, rho=rho, radii=radii, n_touched=n_touched, loss=loss.detach()))
    scaling = torch.exp(p["scaling"])
    if scaling.size(-1) == 1: scaling = scaling.repeat(1, 3)
    total = total + 10 * torch.abs(scaling - scaling.mean(dim=1).view(-1, 1)).mean()
    total.backward()
    return total.detach(), per_view


@pytest.mark.parametrize("name,with_depth,with_exposure,isotropic,fast",
                         [("tiny", False, False, False, False), ("dense", True, True, False, False),
                          ("ragged", True, True, True, False), ("sh3", False, True, False, False),
                          ("dense", True, True, False, True), ("sh3", False, True, False, True),
                          ("sh2", False, False, False, False), ("sh3_full", True, True, False, True)])
def test_map_engine_matches_autograd_path(name, with_depth, with_exposure, isotropic, fast):
    """fast = False: the exact exponent, integer outputs equal to the autograd path's.  fast = True (the MapEngine
    default): ex2.approx in forward and backward blend; a borderline `T > 0.5` decision may move, so n_touched is
    compared as a count of differing entries."""
    from monogs_b200.mapping import MapEngine
    sc, cam = U.small_case(name)
    p_ref = _raw_params(sc, isotropic)
    views = _views(sc, cam, 3, with_depth, with_exposure)
    total_ref, pv = _reference_iteration(p_ref, views, sc.sh_degree)

    p = {k: v.detach().clone() for k, v in p_ref.items()}
    lrs = {k: 0.0 for k in p}
    eng = MapEngine(p, lrs, sh_degree=sc.sh_degree, W=cam.W, H=cam.H, fast_math=fast)
    res = eng.backward_only(views)
    torch.cuda.synchronize()
    n, ov = eng.check_capacity()
    assert not ov
    assert U.rel_err(eng.loss_total, total_ref.view(1)) <= 1e-5
    for k in p:
        if isotropic and k == "rotation":   # an isotropic covariance does not depend on the rotation: both are rounding noise
            noise = 1e-5 * float(p_ref["scaling"].grad.abs().max())
            assert float(eng.grads[k].abs().max()) <= noise and float(p_ref[k].grad.abs().max()) <= noise
            continue
        assert U.rel_err(eng.grads[k], p_ref[k].grad) <= TOL, k
    P = sc.means3D.size(0)
    acc, den, mr = torch.zeros(P, device="cuda"), torch.zeros(P, device="cuda"), torch.zeros(P, device="cuda")
    for i, v in enumerate(views):
        r = res[i]
        assert torch.equal(r.radii, pv[i]["radii"])
        if fast: assert int((r.n_touched != pv[i]["n_touched"]).sum()) <= max(2, P // 500)
        else: assert torch.equal(r.n_touched, pv[i]["n_touched"])
        assert U.rel_err(r.loss, pv[i]["loss"].view(1)) <= 1e-5
        tau_ref = torch.cat((pv[i]["rho"].grad, pv[i]["theta"].grad))
        assert U.rel_err(r.tau, tau_ref) <= TOL
        if v.exposure_a is not None:
            assert U.rel_err(r.dab, torch.cat((v.exposure_a.grad, v.exposure_b.grad))) <= TOL
        if v.update_stats:   # slam_backend.py:292-299, gaussian_model.py:693-697
            vis = pv[i]["radii"] > 0
            acc[vis] += torch.norm(pv[i]["means2D"].grad[vis, :2], dim=-1)
            den[vis] += 1
            mr[vis] = torch.max(mr[vis], pv[i]["radii"][vis].float())
    assert U.rel_err(eng.xyz_gradient_accum, acc) <= TOL
    assert torch.equal(eng.denom, den) and torch.equal(eng.max_radii2D, mr)


def test_map_engine_step_reduces_loss():
    """A few full iterations (Adam included) on a fixed window drive the mapping loss down."""
    from monogs_b200.mapping import MapEngine
    sc, cam = U.small_case("dense")
    p = {k: v.detach().clone() for k, v in _raw_params(sc).items()}
    views = _views(sc, cam, 3, False, False)
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20, "opacity": 0.05, "scaling": 1e-3, "rotation": 1e-3}
    eng = MapEngine(p, lrs, sh_degree=sc.sh_degree, W=cam.W, H=cam.H)
    losses = []
    for _ in range(10):
        total, _ = eng.step(views)
        losses.append(float(total))
    assert losses[-1] < losses[0]


def test_map_engine_densify_between_steps():
    """Statistics gathered by the views drive densify_and_prune on the arena; moments of surviving Gaussians are carried
    over, new ones start at zero, and mapping continues on the resized arena."""
    from monogs_b200.mapping import MapEngine
    sc, cam = U.small_case("dense")
    p = {k: v.detach().clone() for k, v in _raw_params(sc).items()}
    views = _views(sc, cam, 3, False, False)
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20, "opacity": 0.05, "scaling": 1e-3, "rotation": 1e-3}
    eng = MapEngine(p, lrs, sh_degree=sc.sh_degree, W=cam.W, H=cam.H)
    for _ in range(3):
        eng.step(views)
    P0 = eng.P
    assert float(eng.denom.max()) == 6.0      # two of the three views update the statistics, three iterations
    m_before = eng.opt.state[eng.params["xyz"]]["exp_avg"].clone()
    grads = eng.xyz_gradient_accum / eng.denom
    thr = float(grads[~grads.isnan()].median())
    out = eng.densify_and_prune(thr, 0.05, 6.0, 20.0, generator=torch.Generator(device="cuda").manual_seed(1))
    c = out["counts"]
    assert eng.P == c["kept"] + c["cloned"] + c["children"] and eng.P != P0 and c["cloned"] + c["children"] > 0
    st = eng.opt.state[eng.params["xyz"]]
    assert st["step"] == 3 and float(st["exp_avg"][c["kept"]:].abs().max()) == 0.0
    assert float(st["exp_avg"][: c["kept"]].abs().sum()) > 0 and st["exp_avg"].size(0) == eng.P
    assert float(eng.denom.sum()) == 0.0 and eng.max_radii2D.numel() == eng.P
    l0 = float(eng.step(views)[0])
    for _ in range(5):
        l1 = float(eng.step(views)[0])
    assert l1 < l0
    assert not eng.check_capacity()[1]
    del m_before


@pytest.mark.parametrize("name", ["dense", "sh3"])
def test_map_engine_pipeline_matches_single_stream(name):
    """Seven views through the four-stream pipeline (resource sets reused from the fifth view on, atomic gradient and
    statistics accumulation, two forwards in flight) against the one-stream schedule: same gradients up to the order of
    the float additions, same integer outputs and statistics."""
    from monogs_b200.mapping import MapEngine
    sc, cam = U.small_case(name)
    p0 = {k: v.detach().clone() for k, v in _raw_params(sc).items()}
    views = _views(sc, cam, 7, True, False)
    lrs = {k: 0.0 for k in p0}
    res = {}
    for overlap in (False, True):
        eng = MapEngine({k: v.clone() for k, v in p0.items()}, lrs, sh_degree=sc.sh_degree, W=cam.W, H=cam.H, fast_math=False,
                        overlap=overlap)
        for _ in range(2):      # the second iteration reuses every resource set
            out = eng.backward_only(views)
        torch.cuda.synchronize()
        assert not eng.check_capacity()[1]
        res[overlap] = (eng, out)
    (e0, o0), (e1, o1) = res[False], res[True]
    assert U.rel_err(e1.loss_total, e0.loss_total) <= 1e-6
    for k in p0:
        if k == "rotation" and float(e0.grads[k].abs().max()) == 0.0: continue
        assert U.rel_err(e1.grads[k], e0.grads[k]) <= 1e-5, k
    for i in o0:
        assert torch.equal(o1[i].radii, o0[i].radii) and torch.equal(o1[i].n_touched, o0[i].n_touched)
        assert U.rel_err(o1[i].loss, o0[i].loss) <= 1e-6 and U.rel_err(o1[i].tau, o0[i].tau) <= 1e-5
    assert U.rel_err(e1.xyz_gradient_accum, e0.xyz_gradient_accum) <= 1e-5
    assert torch.equal(e1.denom, e0.denom) and torch.equal(e1.max_radii2D, e0.max_radii2D)


def _lrs():
    return {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20, "opacity": 0.05, "scaling": 1e-3, "rotation": 1e-3}


def _window_T(views):
    return [v.settings.viewmatrix.t().contiguous() for v in views]


def test_window_pose_step_matches_torch_adam_and_se3_exp():
    """The window-pose side of a mapping iteration (slam_backend.py:325-332): keyframe_optimizers.step() + update_pose on
    the device for the views the engine rendered, against torch.optim.Adam (default eps) + the float64 SE3 exponential, from
    the tau / dab the same engine reports without pose optimisation.  A frozen view keeps its pose."""
    from monogs_b200.mapping import MapEngine
    from monogs_b200 import scene as S
    sc, cam = U.small_case("dense")
    p = {k: v.detach().clone() for k, v in _raw_params(sc).items()}
    views = _views(sc, cam, 3, False, True)
    lr0 = {k: 0.0 for k in p}             # Gaussians fixed: the two engines see the same scene in both iterations
    plain = MapEngine({k: v.clone() for k, v in p.items()}, lr0, sh_degree=sc.sh_degree, W=cam.W, H=cam.H, fast_math=False)
    ref_out = plain.backward_only(views)
    T0 = _window_T(views)
    pose_lrs = (5e-4, 1.5e-3, 0.01, 0.01)
    eng = MapEngine({k: v.clone() for k, v in p.items()}, lr0, sh_degree=sc.sh_degree, W=cam.W, H=cam.H, fast_math=False)
    eng.set_window(views, T_w2c=T0, frozen=[True, False, False], pose_lrs=pose_lrs)
    eng.step()
    torch.cuda.synchronize()
    T1, ab1 = eng.window_poses()
    assert torch.equal(T1[0], T0[0])                              # the first keyframe is never moved
    for i in (1, 2):
        tau, dab = ref_out[i].tau.cpu().double(), ref_out[i].dab.cpu().double()
        x = torch.zeros(8, dtype=torch.float64, requires_grad=True)
        x.grad = torch.cat([tau, dab])
        groups = [{"params": [x], "lr": 1.0}]
        m, v = 0.1 * x.grad, 0.001 * x.grad ** 2                  # first Adam step in closed form: -lr * g / (|g| + eps)
        upd = -(m / 0.1) / ((v / 0.001).sqrt() + 1e-8)
        lrs8 = torch.tensor([pose_lrs[0]] * 3 + [pose_lrs[1]] * 3 + [pose_lrs[2], pose_lrs[3]], dtype=torch.float64)
        delta = upd * lrs8
        T_ref = (S.se3_exp(delta[:6]) @ T0[i].cpu().double()).float()
        assert U.rel_err(T1[i], T_ref) <= 1e-5, i
        if views[i].exposure_a is not None:
            a0, b0 = float(views[i].exposure_a), float(views[i].exposure_b)
            assert abs(float(ab1[i, 0]) - (a0 + float(delta[6]))) <= 1e-6 and abs(float(ab1[i, 1]) - (b0 + float(delta[7]))) <= 1e-6
    # the engine's own matrices follow the new pose: the next iteration renders from it
    st = eng._win["poses"]
    assert U.rel_err(st["view"][1], T1[1].t()) <= 1e-6 and U.rel_err(st["proj"][1], T1[1].t() @ st["proj_raw"]) <= 1e-5
    l0 = float(eng.step()[0])
    for _ in range(20):
        l1 = float(eng.step()[0])
    assert l1 < l0            # pose + exposure optimisation alone lowers the window loss


def test_map_engine_graph_replay_matches_eager_and_follows_the_lr_schedule():
    """use_graph: the iteration of an installed window is captured once and replayed.  Same parameters as the eager
    engine after several iterations (up to the order of the float atomics), including a learning-rate change between
    replays (the rates live on the device: GaussianModel.update_learning_rate, gaussian_model.py:298-312)."""
    from monogs_b200.mapping import MapEngine, expon_lr
    sc, cam = U.small_case("dense")
    p = {k: v.detach().clone() for k, v in _raw_params(sc).items()}
    views = _views(sc, cam, 5, False, False)
    engs = []
    for use_graph in (False, True):
        eng = MapEngine({k: v.clone() for k, v in p.items()}, _lrs(), sh_degree=sc.sh_degree, W=cam.W, H=cam.H, fast_math=False,
                        use_graph=use_graph)
        eng.set_window(views)
        losses = []
        for it in range(6):
            lr = eng.update_learning_rate(it * 5000, 1.6e-4 * 6, 1.6e-6 * 6, lr_delay_mult=0.01, max_steps=30000)
            assert abs(lr - expon_lr(it * 5000, 1.6e-4 * 6, 1.6e-6 * 6, lr_delay_mult=0.01, max_steps=30000)) < 1e-12
            losses.append(float(eng.step()[0]))
        torch.cuda.synchronize()
        engs.append((eng, losses))
    (e0, l0), (e1, l1) = engs
    assert e1._graph is not None and e1._graph["graph"] is not None, "the window was replayed from a captured graph"
    for a, b in zip(l0, l1):
        assert abs(a - b) <= 1e-5 * abs(a)
    for k in p:
        d0, d1 = e0.params[k] - p[k], e1.params[k] - p[k]
        frac = float(((d0 - d1).abs() > 1e-3 * d0.abs().max()).float().mean())
        assert frac <= 1e-3, (k, frac)
    # and the schedule really reached the kernel: the xyz update of a step shrinks with the rate
    e = MapEngine({k: v.clone() for k, v in p.items()}, _lrs(), sh_degree=sc.sh_degree, W=cam.W, H=cam.H, use_graph=True)
    e.set_window(views)
    for _ in range(3):
        e.step()
    x0 = e.params["xyz"].clone(); e.step(); d_big = float((e.params["xyz"] - x0).abs().max())
    e.set_learning_rates({"xyz": 1.6e-6})
    x0 = e.params["xyz"].clone(); e.step(); d_small = float((e.params["xyz"] - x0).abs().max())
    assert d_small < 0.05 * d_big


def test_reset_opacity_on_the_arena():
    """GaussianModel.reset_opacity / reset_opacity_nonvisible (gaussian_model.py:364-377) + replace_tensor_to_optimizer:
    values and zeroed opacity moments, step counts kept."""
    import math
    from monogs_b200.mapping import MapEngine
    sc, cam = U.small_case("dense")
    p = {k: v.detach().clone() for k, v in _raw_params(sc).items()}
    views = _views(sc, cam, 3, False, False)
    eng = MapEngine(p, _lrs(), sh_degree=sc.sh_degree, W=cam.W, H=cam.H)
    outs = None
    for _ in range(2):
        _, outs = eng.step(views)
    op_before = eng.params["opacity"].clone()
    filters = [outs[i].radii > 0 for i in outs]
    seen = torch.stack(filters).any(dim=0)
    eng.reset_opacity(0.4, visibility_filters=filters)
    inv = math.log(0.4 / 0.6)
    assert torch.equal(eng.params["opacity"][seen], op_before[seen])
    assert torch.allclose(eng.params["opacity"][~seen], torch.full_like(op_before[~seen], inv))
    st = eng.opt.state[eng.params["opacity"]]
    assert float(st["exp_avg"].abs().max()) == 0.0 and float(st["exp_avg_sq"].abs().max()) == 0.0 and st["step"] == 2
    eng.reset_opacity()
    assert torch.allclose(eng.params["opacity"], torch.full_like(op_before, math.log(0.01 / 0.99)))
    eng.step(views)     # mapping continues


def test_assign_views_and_capacity_poll():
    from monogs_b200.mapping import MapEngine, assign_views
    assert assign_views(8, 4) == [[0, 4], [1, 5], [2, 6], [3, 7]]
    assert assign_views(10, 8, 8) == [[0], [1], [2], [3], [4], [5], [6, 9], [7, 8]]      # extras to the least-loaded ranks
    assert sorted(sum(assign_views(10, 4, 8), [])) == list(range(10))
    sc, cam = U.small_case("dense")
    p = {k: v.detach().clone() for k, v in _raw_params(sc).items()}
    views = _views(sc, cam, 3, False, False)
    eng = MapEngine(p, _lrs(), sh_degree=sc.sh_degree, W=cam.W, H=cam.H, capacity_headroom=1.05)
    eng.poll_every = 2
    eng.step(views)
    cap0 = eng.R_cap
    with torch.no_grad():
        eng.params["scaling"] += 1.2          # every splat 3.3x larger: far more tile instances than the capacity
    for _ in range(8):
        eng.step(views)
    torch.cuda.synchronize()
    eng.poll_capacity(); eng.step(views)
    assert eng.overflow_polls >= 1 and eng.R_cap > cap0, "the sticky overflow word was seen and the workspaces regrown"
    assert not eng.check_capacity()[1]
