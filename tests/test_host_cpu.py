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
