"""CPU-only: host-side logic — seeded scenes, camera conventions, view sharding, the gradient arena and the
Python surface's argument checks (reference __init__.py:427-431)."""
import math

import pytest
import torch

from monogs_b200 import scene as S
from monogs_b200.mapping import GradArena, shard_views


def test_scene_is_deterministic_and_shaped():
    a, cam = S.make_config("C1", P=1000)
    b, _ = S.make_config("C1", P=1000)
    for x, y in zip(a, b):
        if isinstance(x, torch.Tensor):
            assert torch.equal(x, y)
    assert a.means3D.shape == (1000, 3) and a.shs.shape == (1000, 1, 3) and a.rotations.shape == (1000, 4)
    assert torch.allclose(a.rotations.norm(dim=1), torch.ones(1000), atol=1e-6)
    assert cam.viewmatrix.shape == (4, 4) and math.isclose(cam.tanfovx, 640 / (2 * 535.4))
    behind = (a.means3D[:, 2] <= 0.2).float().mean()
    assert 0.005 < behind < 0.05     # ~2 % exercise the near cull


def test_projection_matches_monogs_convention():
    cam = S.make_camera(640, 480, 535.4, 539.2, 320.1, 247.6)
    P = cam.projmatrix_raw              # already transposed (camera_utils.py:86-88)
    assert math.isclose(float(P[0, 0]), 2 * 535.4 / 640, rel_tol=1e-6)
    assert math.isclose(float(P[1, 1]), 2 * 539.2 / 480, rel_tol=1e-6)
    assert float(P[2, 3]) == 1.0        # z_sign in column-major memory position [11]
    assert torch.allclose(cam.projmatrix, cam.viewmatrix @ P)
    T = S.se3_exp(torch.tensor(S.GT_TAU))
    assert torch.allclose(T[:3, :3] @ T[:3, :3].t(), torch.eye(3, dtype=torch.float64), atol=1e-12)


def test_shard_views_partitions_the_window():
    for n in (8, 10, 3):
        for world in (1, 2, 4, 8):
            got = sorted(i for r in range(world) for i in shard_views(n, world, r))
            assert got == list(range(n))
            sizes = [len(shard_views(n, world, r)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def test_grad_arena_aliases_param_grads():
    ps = [torch.randn(5, 3, requires_grad=True), torch.randn(7, requires_grad=True)]
    ar = GradArena(ps)
    ar.attach()
    (ps[0].sum() * 2 + (ps[1] ** 2).sum()).backward()
    (ps[0].sum() * 3).backward()         # second view of the window accumulates in place
    assert ps[0].grad.data_ptr() == ar.flat.data_ptr()
    assert torch.allclose(ar.flat[:15], torch.full((15,), 5.0))
    # every view starts on a 16-byte boundary (rows are read as float4 by the kernels): 15 -> 16, 7 -> 8 floats
    assert ps[1].grad.data_ptr() == ar.flat.data_ptr() + 16 * 4
    assert torch.allclose(ar.flat[16:23], 2 * ps[1].detach()) and float(ar.flat[15]) == 0.0
    assert ar.nbytes() == 24 * 4


def test_rasterizer_argument_checks():
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    rs = GaussianRasterizationSettings(48, 64, 0.5, 0.5, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), torch.eye(4),
                                       0, torch.eye(4), False, False)
    assert rs._fields == ("image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
                          "projmatrix", "projmatrix_raw", "sh_degree", "campos", "prefiltered", "debug")
    r = GaussianRasterizer(rs)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(means3D=m, means2D=m, opacities=torch.zeros(4, 1), scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m, opacities=torch.zeros(4, 1), shs=torch.zeros(4, 1, 3))


def test_missing_library_fails_loudly(monkeypatch):
    from monogs_b200 import _lib as L
    monkeypatch.setattr(L, "_lib", None)
    monkeypatch.setattr(L, "LIB_PATH", "/nonexistent/libmonogs_b200.so")
    with pytest.raises(RuntimeError, match="no CPU or eager fallback"):
        L.lib()


def test_grad_arena_layout_and_peer_slices():
    """Host logic of the fused exchange (no device): every parameter view starts on a 16-byte boundary, the segment table
    covers the arena with the loss slot last, views over a sibling buffer share the layout, and the per-rank float4 slices
    (the arithmetic mgs_peer_reduce_adam repeats on the device) tile the arena without gaps or overlap for 1..8 ranks."""
    import torch
    from monogs_b200.mapping import GradArena
    from monogs_b200.peer import PeerBlock
    P = 1001
    shapes = [(P, 3), (P, 1, 3), (P, 15, 3), (P, 1), (P, 3), (P, 4)]
    params = [torch.zeros(s) for s in shapes]
    arena = GradArena(params, extra=1)
    pad4 = lambda n: (n + 3) // 4 * 4
    assert arena.numel == sum(pad4(p.numel()) for p in params) + 4 and arena.numel % 4 == 0
    assert arena.segment_ends[-1] == arena.numel and len(arena.segment_ends) == len(params) + 1
    assert all(e % 4 == 0 for e in arena.segment_ends) and arena.segment_ends == sorted(set(arena.segment_ends))
    for v, p in zip(arena.views, params):
        assert v.shape == p.shape and (v.data_ptr() - arena.flat.data_ptr()) % 16 == 0
    other = torch.arange(arena.numel, dtype=torch.float32)
    for v, w, e, p in zip(arena.views, arena.views_over(other), arena.segment_ends, params):
        assert w.shape == v.shape and float(w.reshape(-1)[0]) == e - pad4(p.numel())
    assert arena.extra.data_ptr() - arena.flat.data_ptr() == 4 * (arena.numel - 4)
    external = torch.zeros(arena.numel)
    assert GradArena(params, extra=1, flat=external).flat.data_ptr() == external.data_ptr()
    for world in range(1, 9):
        for numel in (arena.numel, 56, 4, 8 * 4 * 1000 + 4):
            blk = PeerBlock.__new__(PeerBlock)      # slice arithmetic only: no allocation
            blk.numel, blk.world = numel, world
            cuts = [blk.slice_of(r) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == numel
            for (b0, e0), (b1, e1) in zip(cuts, cuts[1:]):
                assert e0 == b1 and b0 <= e0 and b0 % 4 == 0


def test_committed_profile_table_describes_the_committed_kernels():
    """bench.py takes `roofline.traffic` and the issue-roofline inputs of the dominant kernel from the committed ncu table
    (profiles/r02_kernel_metrics.txt) and refuses a row whose kernel sources changed after the capture (the table's
    header records their sha1).  A kernel edit without a fresh capture would silently drop those numbers from the bench
    line; this test makes it loud."""
    import bench
    row, why = bench.profile_row("blend_track_kernel")
    assert row is not None, why
    assert why is None, why
    assert row["us"] > 0 and row["Minst"] > 0 and 0 < row["lanes"] <= 32


def test_algorithmic_bytes_of_the_segment_iteration():
    """The four kernels of the fixed-segment tracking iteration and the seven of the general binning are both accounted
    for; the segment iteration moves fewer algorithmic bytes (no depth / mean2D arrays, no emit pass)."""
    import bench
    P, P_vis, R, R_cap, W, H = 200_000, 145_000, 546_000, 1_228_800, 640, 480
    seg = bench.algorithmic_bytes(P, P_vis, R, R_cap, W, H, segments=True)
    gen = bench.algorithmic_bytes(P, P_vis, R, R_cap, W, H, segments=False)
    assert set(seg) == {"preprocess", "tile_sort_segments", "blend_track", "track_update"}
    assert {"preprocess", "tile_scan", "emit", "tile_sort_records", "blend_track", "track_update"} <= set(gen)
    fused_gen = sum(gen[k] for k in ("preprocess", "tile_scan", "emit", "tile_sort_records", "tile_sort_records_large", "blend_track", "track_update"))
    assert sum(seg.values()) < fused_gen
    assert seg["preprocess"] == gen["preprocess"] - 12 * P_vis + 8 * R      # + scattered words, - depth / mean2D arrays
