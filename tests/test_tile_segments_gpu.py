"""Fixed-segment binning of the fused tracker (mgs_track_args.tile_segments: preprocess scatters into per-tile segments,
tile_sort_segments_kernel sorts them; four kernels per iteration) against the general counting-sort binning
(tile_scan + emit + tile_sort_records, the path that is bit-exact against the reference's duplicateWithKeys / SortPairs /
identifyTileRanges, rasterizer_impl.cu:288-321, in tests/test_parity_gpu.py)."""
import pytest
import torch

from tests import util as U

pytestmark = pytest.mark.gpu


def _scene(name, T_scale=1.0):
    from monogs_b200 import scene as S, rasterizer as RZ
    sc, cam = U.small_case(name, T_w2c=torch.eye(4))
    sc = sc._replace(shs=sc.shs[:, :1].contiguous(), sh_degree=0)
    scd = U.to_dev(sc)
    bg = torch.tensor([0.05, 0.1, 0.15]).cuda()
    T_gt = S.se3_exp(torch.tensor(S.GT_TAU) * T_scale).float()
    cam_gt = U.to_dev(S.make_camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, T_gt))
    gt, _, gt_depth, _, _, _ = RZ.raster_forward(U.settings(cam_gt, bg, 0), scd.means3D, scd.opacities, scd.shs, None,
                                                 scd.scales, scd.rotations, None)
    return scd, cam, bg, gt, gt_depth


def _status(tr):
    import ctypes as C
    from monogs_b200 import _lib as L
    from monogs_b200.rasterizer import _stream
    st = torch.zeros(8, dtype=torch.int32).pin_memory()
    L.check(tr.lib.mgs_raster_status_async(L.ptr(tr.ws), C.c_void_p(st.data_ptr()), _stream(tr.dev)), "status")
    torch.cuda.synchronize()
    return st


@pytest.mark.parametrize("name,mode,use_depth,fast", [("dense", 1, False, False), ("dense", 1, False, True), ("dense", 0, True, False),
                                                      ("ragged", 1, False, False), ("tiny", 0, False, True)])
def test_segments_match_general_binning(name, mode, use_depth, fast):
    """Same sorted lists -> the blend walks the same splats in the same order: images, depth and n_touched are bit-identical,
    loss and tangent differ by the order of the float atomics only; the status words agree; the counters are left zero."""
    from monogs_b200.tracking import PoseTracker
    scd, cam, bg, gt, gt_depth = _scene(name)
    out = []
    for seg in (True, False):
        tr = PoseTracker(scd.means3D, scd.opacities, scd.shs, scd.scales, scd.rotations, 0, cam, bg, torch.eye(4), mode=mode,
                         use_depth=use_depth, use_graph=False, fast_math=fast, tile_segments=seg)
        assert tr.tile_segments == seg and tr.kernels_per_iter == (4 if seg else 7)
        tr.set_frame(gt, gt_depth=gt_depth if use_depth else None)
        tr.step()
        torch.cuda.synchronize()
        first = (tr.loss.clone(), tr.tau.clone(), tr.dab.clone(), tr.color.clone(), tr.depth.clone(), tr.n_touched.clone(), tr.radii.clone())
        for _ in range(5):          # later steps run on the counters the previous one left behind
            tr.step()
        st = _status(tr)
        assert st[1] == 0 and st[4] == 0 and 0 < st[7] <= 1024
        out.append((first, tr.T.clone(), float(tr.loss), st.clone()))
    (f0, T0, l0, s0), (f1, T1, l1, s1) = out
    assert torch.equal(f0[3], f1[3]) and torch.equal(f0[4], f1[4]) and torch.equal(f0[6], f1[6])
    if not fast:
        assert torch.equal(f0[5], f1[5])
    assert U.rel_err(f0[0], f1[0]) <= 1e-6
    assert U.rel_err(f0[1], f1[1]) <= 1e-5, (f0[1], f1[1])
    assert U.rel_err(f0[2], f1[2]) <= 1e-5
    assert U.rel_err(T0, T1) <= 1e-5 and abs(l0 - l1) <= 1e-4 * abs(l1)
    assert s0[0] == s1[0] and s0[7] == s1[7]        # num_rendered and the longest tile list of the last iteration


def test_segments_graph_replay_and_second_order():
    """Graph replay of the four-kernel iteration and the second-order iteration on the same workspace follow the general
    binning's trajectory."""
    from monogs_b200.tracking import PoseTracker
    scd, cam, bg, gt, _ = _scene("dense")
    g = torch.Generator(device="cuda").manual_seed(3)
    idx, w = PoseTracker.gen_sketch_args(cam.H, cam.W, stack_dim=4, sketch_dim=16, device="cuda", generator=g)
    res = []
    for seg in (True, False):
        tr = PoseTracker(scd.means3D, scd.opacities, scd.shs, scd.scales, scd.rotations, 0, cam, bg, torch.eye(4), tile_segments=seg)
        tr.set_frame(gt)
        for _ in range(8):
            tr.step()
        T_first = tr.T.clone()
        tr.begin_second_order(initial_lambda=1e-3)
        for _ in range(3):
            tr.set_sketch(idx, w, 16)
            tr.second_order_step()
        torch.cuda.synchronize()
        assert not tr.check_capacity()[1]
        res.append((T_first, tr.T.clone(), tr.sketch_rows.clone(), tr.lm_state.clone()))
    (a0, a1, r0, s0), (b0, b1, r1, s1) = res
    assert U.rel_err(a0, b0) <= 1e-4 and U.rel_err(a1, b1) <= 1e-4
    assert U.rel_err(r0, r1) <= 2e-3 and abs(float(s0[33]) - float(s1[33])) <= 1e-3 * abs(float(s1[33]))


def test_segment_overflow_is_reported_and_tracker_falls_back():
    """A tile list beyond MGS_TILE_SEGMENT: the constructor's probe keeps such a scene on the general binning; forced into
    segment mode the iteration raises the overflow words, and the tracker leaves segment mode at the next poll."""
    from monogs_b200.tracking import PoseTracker, TILE_SEGMENT
    scd, cam, bg, gt, _ = _scene("packed8", 0.3)
    tr = PoseTracker(scd.means3D, scd.opacities, scd.shs, scd.scales, scd.rotations, 0, cam, bg, torch.eye(4), use_graph=False)
    assert not tr.tile_segments and tr._tile_max > TILE_SEGMENT          # ~1.5k instances per tile
    tr.set_frame(gt)
    tr.step()
    ref_color, ref_tau = tr.color.clone(), tr.tau.clone()
    # force segment mode on a workspace that is large enough for it
    tr2 = PoseTracker(scd.means3D, scd.opacities, scd.shs, scd.scales, scd.rotations, 0, cam, bg, torch.eye(4), use_graph=False)
    tr2.tile_segments = True
    tr2._tile_max = 0
    tr2.headroom_saved, tr2.headroom = tr2.headroom, 0.0                 # (the probe would switch it off again)
    tr2._size_workspace()
    tr2.headroom = tr2.headroom_saved
    assert tr2.tile_segments and tr2.R_cap >= tr2.tiles * TILE_SEGMENT
    tr2.set_frame(gt)
    tr2.step()
    st = _status(tr2)
    assert st[1] == 1 and st[4] == 1 and st[7] > TILE_SEGMENT and st[2] < st[0]     # overflow, sticky, longest list, clipped
    n, over = tr2.check_capacity()
    assert over and not tr2.tile_segments
    tr2.T.copy_(torch.eye(4)); tr2.viewmatrix.copy_(tr2.T.t()); tr2.projmatrix.copy_(tr2.viewmatrix @ tr2.proj_raw)
    tr2.set_frame(gt)
    tr2.step()
    torch.cuda.synchronize()
    assert torch.equal(tr2.color, ref_color) and U.rel_err(tr2.tau, ref_tau) <= 1e-5
    assert not tr2.check_capacity()[1]


@pytest.mark.parametrize("name,with_depth,isotropic,overlap", [("dense", True, False, True), ("sh3", False, False, True),
                                                               ("ragged", True, True, False), ("tiny", False, False, False)])
def test_map_engine_segments_match_general_binning(name, with_depth, isotropic, overlap):
    """mgs_map_view.tile_segments (preprocess_kernel<0, ., ., SEGMENTS> + tile_sort_segments_kernel<records>) against the
    general binning of the same engine: identical radii / n_touched / images, gradients to the order of the float atomics;
    two passes over the window, so the second runs on the counters the first left behind."""
    from monogs_b200.mapping import MapEngine
    from tests.test_map_engine_gpu import _raw_params, _views
    sc, cam = U.small_case(name)
    p0 = _raw_params(sc, isotropic)
    views = _views(sc, cam, 3, with_depth, True)
    out = []
    for seg in (True, False):
        p = {k: v.detach().clone() for k, v in p0.items()}
        eng = MapEngine(p, {k: 0.0 for k in p}, sh_degree=sc.sh_degree, W=cam.W, H=cam.H, fast_math=False, overlap=overlap,
                        tile_segments=seg)
        eng.backward_only(views)
        res = eng.backward_only(views)
        torch.cuda.synchronize()
        assert eng.tile_segments == seg
        n, ov = eng.check_capacity()
        assert not ov and 0 < eng._tile_max <= 1024
        out.append(({k: g.clone() for k, g in eng.grads.items()}, [(r.radii.clone(), r.n_touched.clone(), r.tau.clone(), r.loss.clone()) for r in (res[i] for i in range(len(views)))],
                    eng.loss_total.clone(), eng.xyz_gradient_accum.clone(), n))
    (g0, r0, l0, a0, n0), (g1, r1, l1, a1, n1) = out
    assert n0 == n1 and U.rel_err(l0, l1) <= 1e-6
    for (ra, na, ta, la), (rb, nb, tb, lb) in zip(r0, r1):
        assert torch.equal(ra, rb) and torch.equal(na, nb) and U.rel_err(la, lb) <= 1e-6 and U.rel_err(ta, tb) <= 1e-4
    for k in g0:
        if isotropic and k == "rotation":
            continue
        assert U.rel_err(g0[k], g1[k]) <= 1e-5, k
    assert U.rel_err(a0, a1) <= 1e-5
