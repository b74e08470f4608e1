"""mgs_loss_fwd_bwd_static — the fused loss with its scratch in the library's own device memory (no allocation on the
call path; include/monogs_b200.h) — against mgs_loss_fwd_bwd with caller-owned scratch, through the C-ABI.

The file name sorts last on purpose: a failure here must not hide other tests.
Every output tensor and every input is created BEFORE the calls under test; the calls themselves get raw pointers.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

LOSS_TOL = 1e-6      # same kernels, same launch geometry: the results are expected to be bit-identical


def _inputs(H, W, seed, dev):
    g = torch.Generator().manual_seed(seed)
    t = dict(image=torch.rand(3, H, W, generator=g), depth=torch.rand(1, H, W, generator=g) * 3,
             opacity=torch.rand(1, H, W, generator=g) * 0.1 + 0.9, gt_image=torch.rand(3, H, W, generator=g),
             gt_depth=torch.rand(1, H, W, generator=g) * 3, grad_mask=(torch.rand(1, H, W, generator=g) > 0.4).float(),
             exposure_a=torch.tensor([0.9 + 0.01 * seed]), exposure_b=torch.tensor([0.02]))
    t["gt_image"][:, :3, :] = 0.0          # below rgb_boundary_threshold: masked out
    t["gt_depth"][:, :, :4] = 0.0
    return {k: v.to(dev).contiguous() for k, v in t.items()}


def _args(L, mode, use_depth, H, W, t):
    a = L.MgsLossArgs()
    a.mode, a.H, a.W, a.use_depth = mode, H, W, int(use_depth)
    a.rgb_boundary_threshold, a.alpha, a.huber_delta, a.loss_scale = 0.01, 0.95, 0.01, 1.0
    for k in ("image", "opacity", "gt_image", "grad_mask", "exposure_a", "exposure_b"):
        setattr(a, k, t[k].data_ptr())
    a.depth = t["depth"].data_ptr() if use_depth else None
    a.gt_depth = t["gt_depth"].data_ptr() if use_depth else None
    return a


def _outputs(H, W, dev):
    return dict(loss=torch.full((1,), -1.0, device=dev), dab=torch.full((2,), -1.0, device=dev),
                g_img=torch.full((3, H, W), -1.0, device=dev), g_dep=torch.full((1, H, W), -1.0, device=dev))


@pytest.mark.parametrize("mode_name,use_depth", [("LOSS_TRACK_L1", True), ("LOSS_TRACK_HUBER", False), ("LOSS_MAP_L1", True)])
def test_static_scratch_loss_matches_caller_scratch_loss(mode_name, use_depth):
    from monogs_b200 import _lib as L
    lib = L.lib()
    dev = torch.device("cuda", 0)
    mode = getattr(L, mode_name)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    # two calls in a row on different frames (the second smaller, so most of the shared array still holds the first
    # call's partial sums when it runs), then the first frame again
    for H, W, seed in ((120, 160, 1), (48, 64, 2), (120, 160, 1)):
        t = _inputs(H, W, seed, dev)
        a = _args(L, mode, use_depth, H, W, t)
        ref, out = _outputs(H, W, dev), _outputs(H, W, dev)
        scratch = torch.empty(lib.mgs_loss_scratch_bytes(H, W), device=dev, dtype=torch.uint8)
        gd = lambda o: L.ptr(o["g_dep"]) if use_depth else None
        torch.cuda.synchronize(dev)
        L.check(lib.mgs_loss_fwd_bwd(C.byref(a), L.ptr(ref["loss"]), L.ptr(ref["dab"]), L.ptr(ref["g_img"]), gd(ref),
                                     L.ptr(scratch), scratch.numel(), stream), "mgs_loss_fwd_bwd")
        L.check(lib.mgs_loss_fwd_bwd_static(C.byref(a), L.ptr(out["loss"]), L.ptr(out["dab"]), L.ptr(out["g_img"]), gd(out),
                                            stream), "mgs_loss_fwd_bwd_static")
        torch.cuda.synchronize(dev)
        assert float(ref["loss"]) > 0.0
        for k in ("loss", "dab", "g_img") + (("g_dep",) if use_depth else ()):
            r, o = ref[k].double().cpu(), out[k].double().cpu()
            err = float((r - o).abs().max() / r.abs().max().clamp_min(1e-30))
            assert err <= LOSS_TOL, f"{mode_name} {H}x{W} {k}: max-norm relative error {err:.3e}"
        # forward only (no gradient images): the Huber path then clears its own second-stage slots
        ref2, out2 = _outputs(H, W, dev), _outputs(H, W, dev)
        L.check(lib.mgs_loss_fwd_bwd(C.byref(a), L.ptr(ref2["loss"]), L.ptr(ref2["dab"]), None, None, L.ptr(scratch),
                                     scratch.numel(), stream), "mgs_loss_fwd_bwd")
        L.check(lib.mgs_loss_fwd_bwd_static(C.byref(a), L.ptr(out2["loss"]), L.ptr(out2["dab"]), None, None, stream),
                "mgs_loss_fwd_bwd_static")
        torch.cuda.synchronize(dev)
        assert abs(float(ref2["loss"]) - float(out2["loss"])) <= LOSS_TOL * abs(float(ref2["loss"]))
        assert abs(float(ref2["loss"]) - float(ref["loss"])) <= LOSS_TOL * abs(float(ref["loss"]))


def test_static_scratch_loss_rejects_frames_beyond_its_limit():
    from monogs_b200 import _lib as L
    lib = L.lib()
    dev = torch.device("cuda", 0)
    t = _inputs(8, 8, 0, dev)
    a = _args(L, L.LOSS_MAP_L1, False, 8, 8, t)
    a.H, a.W = 2048, 2049                      # one column over MGS_LOSS_STATIC_MAX_PIXELS; rejected before any launch
    out = _outputs(8, 8, dev)
    rc = lib.mgs_loss_fwd_bwd_static(C.byref(a), L.ptr(out["loss"]), L.ptr(out["dab"]), None, None, None)
    assert rc == -2                            # MGS_ENOSPC
    msg = lib.mgs_last_error_string().decode()
    assert "MGS_LOSS_STATIC_MAX_PIXELS" in msg and "mgs_loss_fwd_bwd" in msg, msg
    torch.cuda.synchronize(dev)
    assert float(out["loss"]) == -1.0          # untouched
