"""Host side of the drop-in boundary: the `diff_gaussian_rasterization` Python surface on top of the C-ABI.

Mirrors submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py of the reference:
`GaussianRasterizationSettings` (:392-405), `GaussianRasterizer` (:407-469), `rasterize_gaussians` (:22-59)
and the autograd semantics of `_RasterizeGaussians` (:61-390): same argument order, same 5-tuple of outputs,
same 17-tuple of gradients, `grad_rho` / `grad_theta` shaped [1,3].  torch is used for device memory and
the current stream only; all arithmetic happens in libmonogs_b200.so.
"""
import ctypes as C
import threading
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _lib as L


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    projmatrix_raw: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _f32c(t):
    if t is None or t.numel() == 0:
        return None
    if t.dtype != torch.float32 or not t.is_contiguous():
        t = t.to(torch.float32).contiguous()
    return t


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_raw_device = getattr(torch._C, "_cuda_getDevice", None)


def _stream(dev=None):
    """Handle of torch's current stream on `dev` (a torch.device / index; default: the current device).  Callers pass the
    device of the tensors they hand to the library, so the launch goes to a stream of THAT device even when it is not the
    current one (several GPUs in one process).  torch.cuda.current_stream() builds a Stream object and resolves the device
    through several Python layers (~20 us per call, six calls per drop-in tracking step); the raw accessors return the
    same handle in about a microsecond."""
    if isinstance(dev, torch.device):
        dev = dev.index
    if _raw_stream is not None and _raw_device is not None:
        return C.c_void_p(_raw_stream(_raw_device() if dev is None else dev))
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


class _on:
    """`with _on(dev):` — make `dev` the current CUDA device for the library calls inside (the kernels are launched on the
    device that is current, so it must be the one that owns the pointers); a no-op when it already is."""
    __slots__ = ("dev", "prev")

    def __init__(self, dev):
        self.dev = dev.index if isinstance(dev, torch.device) else dev

    def __enter__(self):
        self.prev = None
        if self.dev is not None and _raw_device is not None and _raw_device() != self.dev:
            self.prev = _raw_device()
            torch.cuda.set_device(self.dev)

    def __exit__(self, *exc):
        if self.prev is not None:
            torch.cuda.set_device(self.prev)


def make_view(rs: GaussianRasterizationSettings, sh_coeffs: int, keep: list) -> L.MgsView:
    """Build the C struct; `keep` collects the (possibly converted) tensors so they outlive the call."""
    dev = rs.viewmatrix.device
    tens = {}
    for name in ("bg", "viewmatrix", "projmatrix", "projmatrix_raw", "campos"):
        t = getattr(rs, name)
        if t is not None and not (t.device == dev and t.dtype == torch.float32 and t.is_contiguous()):   # fast path: as MonoGS passes them
            t = _f32c(t.to(dev))
        keep.append(t)
        tens[name] = t
    v = L.MgsView()
    v.image_height, v.image_width = int(rs.image_height), int(rs.image_width)
    v.tanfovx, v.tanfovy = float(rs.tanfovx), float(rs.tanfovy)
    v.scale_modifier = float(rs.scale_modifier)
    v.sh_degree, v.sh_coeffs, v.prefiltered = int(rs.sh_degree), int(sh_coeffs), int(bool(rs.prefiltered))
    for name, t in tens.items():
        setattr(v, name, t.data_ptr() if t is not None else None)
    return v


class RasterState:
    """What forward leaves for backward: the workspace and the sizes it was carved for."""
    __slots__ = ("workspace", "num_rendered", "P", "sh_coeffs", "pose_state")


class _DevCache(threading.local):
    """Per thread and device: a pinned status slot + the event that guards it, and the last num_rendered seen for a
    (P, W, H) problem — the capacity guess of the next forward of that size."""

    def __init__(self):
        self.by_dev = {}

    def get(self, dev):
        c = self.by_dev.get(dev.index)
        if c is None:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(dev))    # materialises the cudaEvent_t the library records into
            c = self.by_dev[dev.index] = dict(status=torch.zeros(8, dtype=torch.int32).pin_memory(), event=ev, guess={})
        return c


_cache = _DevCache()
CAPACITY_HEADROOM = 1.25     # workspace capacity of a speculative forward = previous num_rendered of that size x this
# Tracking through the drop-in surface (only theta / rho require a gradient, SH degree 0): the forward also accumulates the
# loss-independent sums the pose tangent is made of, and backward is one pass over the pixels (mgs_raster_forward_pose /
# mgs_raster_backward_pose_deferred) instead of a second walk over every tile list.
SINGLE_PASS_TRACKING = True


class _NoGuard:
    def __enter__(self): return self
    def __exit__(self, *exc): return False


_NO_GUARD = _NoGuard()


def _debug_guard(rs, s, what, args_fn, dump):
    """debug=True (auxiliary.h:166-173 CHECK_CUDA, __init__.py:109-116,202-209): wait for the launches of `what` and
    raise on any CUDA error; on failure the inputs are written to `dump` before the exception propagates."""
    if not rs.debug:
        return _NO_GUARD

    class _G:
        def __enter__(self_g):
            self_g.snapshot = _cpu_deep_copy(args_fn()) if rs.debug else None   # copied before they can be corrupted
            return self_g

        def __exit__(self_g, et, ev, tb):
            if not rs.debug:
                return False
            if et is None:
                try:
                    L.check(L.lib().mgs_stream_synchronize(s), what + " (debug synchronize)")
                    return False
                except RuntimeError:
                    torch.save(self_g.snapshot, dump)
                    print(f"\nAn error occured in {what}. Please forward {dump} for debugging.")
                    raise
            torch.save(self_g.snapshot, dump)
            print(f"\nAn error occured in {what}. Please forward {dump} for debugging.")
            return False
    return _G()


def _cpu_deep_copy(args):
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


def raster_forward(rs, means3D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp, single_pass=False):
    """Forward of the drop-in surface.  Returns (color, radii, depth, opacity, n_touched, state).
    single_pass: the caller will only ask for the pose tangent (tracking) — state.pose_state then holds the per-pixel
    sums for mgs_raster_backward_pose_deferred (None when the capacity guess missed, or on the first call of a size).

    The reference blocks the host in the middle of every forward to read num_rendered (rasterizer_impl.cu:284) before it
    can size the binning buffers.  Here the first forward of a (P, W, H) problem does the same (two stages, one read);
    every later one launches the whole pipeline at once into a workspace sized from the previous call's count
    (x CAPACITY_HEADROOM) and only then waits for the count, which the device sent out right after the scan — so the
    device is never idle while the host turns around, num_rendered is exact on return, and in the rare case the guess
    was too small the binning + blend are simply repeated at the right size before returning."""
    lib = L.lib()
    dev = means3D.device
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    P, H, W = means3D.size(0), int(rs.image_height), int(rs.image_width)
    means3D, opacities = _f32c(means3D), _f32c(opacities)
    shs, colors_precomp = _f32c(shs), _f32c(colors_precomp)
    scales, rotations, cov3D_precomp = _f32c(scales), _f32c(rotations), _f32c(cov3D_precomp)
    M = shs.size(1) if shs is not None else 0
    keep = []
    view = make_view(rs, M, keep)
    f = dict(device=dev, dtype=torch.float32)
    color, depth, opacity = torch.empty(3, H, W, **f), torch.empty(1, H, W, **f), torch.empty(1, H, W, **f)
    # both fully written by the library (radii by the preprocess, n_touched zeroed before the blend)
    radii = torch.empty(P, device=dev, dtype=torch.int32)
    n_touched = torch.empty(P, device=dev, dtype=torch.int32)
    st = RasterState()
    st.P, st.sh_coeffs, st.pose_state = P, M, None
    if P == 0:  # rasterize_points.cu:90 — legal, zero-filled outputs
        color.zero_(); depth.zero_(); opacity.zero_()
        st.workspace, st.num_rendered = torch.empty(0, device=dev, dtype=torch.uint8), 0
        return color, radii, depth, opacity, n_touched, st
    with _on(dev):
        s = _stream(dev)
        cache = _cache.get(dev)
        status, key = cache["status"], (P, W, H)
        guess = cache["guess"].get(key)
        snap = lambda: (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3D_precomp,
                        rs.viewmatrix, rs.projmatrix, rs.projmatrix_raw, rs.tanfovx, rs.tanfovy, H, W, shs, rs.sh_degree,
                        rs.campos, rs.prefiltered, rs.debug)
        with _debug_guard(rs, s, "forward", snap, "snapshot_fw.dump"):
            base = lib.mgs_raster_workspace_bytes(P, W, H, 0)
            if guess is None:
                ws = torch.empty(base, device=dev, dtype=torch.uint8)
                L.check(lib.mgs_raster_preprocess(C.byref(view), P, L.ptr(means3D), L.ptr(opacities), L.ptr(shs),
                                                  L.ptr(colors_precomp), L.ptr(scales), L.ptr(rotations), L.ptr(cov3D_precomp),
                                                  L.ptr(radii), L.ptr(ws), ws.numel(), C.c_void_p(status.data_ptr()), s),
                        "mgs_raster_preprocess")
                L.check(lib.mgs_stream_synchronize(s), "mgs_stream_synchronize")
                R, rerun = int(status[0]), True
            elif single_pass and int(rs.sh_degree) == 0:
                R_cap = int(guess * CAPACITY_HEADROOM) + 4096
                ws = torch.empty(lib.mgs_raster_workspace_bytes(P, W, H, R_cap), device=dev, dtype=torch.uint8)
                ps = torch.empty(lib.mgs_raster_pose_state_bytes(P, W, H), device=dev, dtype=torch.uint8)
                L.check(lib.mgs_raster_forward_pose(C.byref(view), P, L.ptr(means3D), L.ptr(opacities), L.ptr(shs),
                                                    L.ptr(colors_precomp), L.ptr(scales), L.ptr(rotations), L.ptr(cov3D_precomp),
                                                    L.ptr(radii), L.ptr(ws), ws.numel(), R_cap, L.ptr(color), L.ptr(depth),
                                                    L.ptr(opacity), L.ptr(n_touched), L.ptr(ps), ps.numel(),
                                                    C.c_void_p(status.data_ptr()), C.c_void_p(cache["event"].cuda_event), s),
                        "mgs_raster_forward_pose")
                cache["event"].synchronize()
                R, rerun = int(status[0]), bool(status[1])
                st.pose_state = None if rerun else ps
            else:
                R_cap = int(guess * CAPACITY_HEADROOM) + 4096
                ws = torch.empty(lib.mgs_raster_workspace_bytes(P, W, H, R_cap), device=dev, dtype=torch.uint8)
                L.check(lib.mgs_raster_forward(C.byref(view), P, L.ptr(means3D), L.ptr(opacities), L.ptr(shs),
                                               L.ptr(colors_precomp), L.ptr(scales), L.ptr(rotations), L.ptr(cov3D_precomp),
                                               L.ptr(radii), L.ptr(ws), ws.numel(), R_cap, L.ptr(color), L.ptr(depth),
                                               L.ptr(opacity), L.ptr(n_touched), C.c_void_p(status.data_ptr()),
                                               C.c_void_p(cache["event"].cuda_event), s), "mgs_raster_forward")
                cache["event"].synchronize()      # the status left the device before emit / sort / blend were queued
                R, rerun = int(status[0]), bool(status[1])
            if rerun:
                need = lib.mgs_raster_workspace_bytes(P, W, H, R)
                if need > ws.numel():
                    big = torch.empty(need, device=dev, dtype=torch.uint8)
                    big[:base].copy_(ws[:base])  # per-Gaussian state sits at R-independent offsets
                    ws = big
                L.check(lib.mgs_raster_render(C.byref(view), P, L.ptr(colors_precomp), L.ptr(radii), L.ptr(ws), ws.numel(), R,
                                              L.ptr(color), L.ptr(depth), L.ptr(opacity), L.ptr(n_touched), s),
                        "mgs_raster_render")
            cache["guess"][key] = R
    st.workspace, st.num_rendered = ws, R
    return color, radii, depth, opacity, n_touched, st


def raster_backward(rs, st, grad_mask, means3D, shs, colors_precomp, scales, rotations, cov3D_precomp, radii,
                    grad_color, grad_depth, want_tau_per_gaussian=False, selected_pixel_indices=None):
    """One call of mgs_raster_backward; returns a dict of the requested gradients.
    selected_pixel_indices (int32 [n], linear pixel indices): the select_pixels = True branch of the reference's
    _C.rasterize_gaussians_backward (rasterize_points.cu:174-189) — only those pixels send gradient
    (mgs_raster_backward_pixels).  None = the dense backward."""
    lib = L.lib()
    dev, P = means3D.device, st.P
    means3D, shs, colors_precomp = _f32c(means3D), _f32c(shs), _f32c(colors_precomp)
    scales, rotations, cov3D_precomp = _f32c(scales), _f32c(rotations), _f32c(cov3D_precomp)
    grad_color, grad_depth = _f32c(grad_color), _f32c(grad_depth)
    keep = []
    view = make_view(rs, st.sh_coeffs, keep)
    f = dict(device=dev, dtype=torch.float32)
    out = {}
    M = st.sh_coeffs
    if grad_mask & L.GRAD_MEANS3D: out["means3D"] = torch.empty(P, 3, **f)
    if grad_mask & L.GRAD_MEANS2D: out["means2D"] = torch.empty(P, 3, **f)
    if grad_mask & L.GRAD_OPACITY: out["opacity"] = torch.empty(P, 1, **f)
    if grad_mask & L.GRAD_COLOR:
        out["color"] = torch.empty(P, M, 3, **f) if shs is not None else torch.empty(P, 3, **f)
    if grad_mask & L.GRAD_COV:
        if scales is not None:
            out["scales"], out["rotations"] = torch.empty(P, 3, **f), torch.empty(P, 4, **f)
        else:
            out["cov3D"] = torch.empty(P, 6, **f)
    if grad_mask & L.GRAD_TAU:
        out["tau"] = torch.empty(6, **f)
        if want_tau_per_gaussian:
            out["tau_pg"] = torch.empty(P, 6, **f)
    if P == 0:
        for t in out.values(): t.zero_()
        return out
    sel = None
    if selected_pixel_indices is not None:
        sel = selected_pixel_indices.to(device=dev, dtype=torch.int32).contiguous()
        H, W = int(rs.image_height), int(rs.image_width)
        scratch = torch.empty(lib.mgs_raster_backward_pixels_scratch_bytes(P, W, H), device=dev, dtype=torch.uint8)
    else:
        scratch = torch.empty(lib.mgs_raster_backward_scratch_bytes(P), device=dev, dtype=torch.uint8)
    g = lambda k: L.ptr(out[k]) if k in out else None
    with _on(dev):
        s = _stream(dev)
        snap = lambda: (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3D_precomp,
                        rs.viewmatrix, rs.projmatrix, rs.projmatrix_raw, rs.tanfovx, rs.tanfovy, grad_color, grad_depth, shs,
                        rs.sh_degree, rs.campos, st.workspace, st.num_rendered, rs.debug)
        with _debug_guard(rs, s, "backward", snap, "snapshot_bw.dump"):
            if sel is not None:
                L.check(lib.mgs_raster_backward_pixels(
                    C.byref(view), P, grad_mask, L.ptr(means3D), L.ptr(shs), L.ptr(colors_precomp), L.ptr(scales),
                    L.ptr(rotations), L.ptr(cov3D_precomp), L.ptr(radii), L.ptr(st.workspace), st.workspace.numel(),
                    L.ptr(scratch), scratch.numel(), C.c_void_p(sel.data_ptr()) if sel.numel() else None, sel.numel(),
                    L.ptr(grad_color), L.ptr(grad_depth), g("means3D"), g("means2D"), g("color"), g("opacity"), g("scales"),
                    g("rotations"), g("cov3D"), g("tau"), g("tau_pg"), s), "mgs_raster_backward_pixels")
                return out
            L.check(lib.mgs_raster_backward(C.byref(view), P, grad_mask, L.ptr(means3D), L.ptr(shs), L.ptr(colors_precomp),
                                            L.ptr(scales), L.ptr(rotations), L.ptr(cov3D_precomp), L.ptr(radii),
                                            L.ptr(st.workspace), st.workspace.numel(), L.ptr(scratch), scratch.numel(),
                                            L.ptr(grad_color), L.ptr(grad_depth), g("means3D"), g("means2D"), g("color"),
                                            g("opacity"), g("scales"), g("rotations"), g("cov3D"), g("tau"), g("tau_pg"),
                                            s), "mgs_raster_backward")
    return out


def raster_backward_pose(rs, st, means3D, shs, colors_precomp, scales, rotations, cov3D_precomp, radii,
                         grad_color, grad_depth, want_image=False):
    """Pose-only backward: returns (tau_sum[6], tau_image[6,H,W] or None)."""
    lib = L.lib()
    dev, P = means3D.device, st.P
    H, W = int(rs.image_height), int(rs.image_width)
    means3D, shs, colors_precomp = _f32c(means3D), _f32c(shs), _f32c(colors_precomp)
    scales, rotations, cov3D_precomp = _f32c(scales), _f32c(rotations), _f32c(cov3D_precomp)
    grad_color, grad_depth = _f32c(grad_color), _f32c(grad_depth)
    keep = []
    view = make_view(rs, st.sh_coeffs, keep)
    tau = torch.empty(6, device=dev, dtype=torch.float32)
    tau_image = torch.empty(6, H, W, device=dev, dtype=torch.float32) if want_image else None
    scratch = torch.empty(max(lib.mgs_raster_backward_pose_scratch_bytes(P, W, H), 256), device=dev, dtype=torch.uint8)
    with _on(dev):
        s = _stream(dev)
        snap = lambda: (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3D_precomp,
                        rs.viewmatrix, rs.projmatrix, rs.projmatrix_raw, rs.tanfovx, rs.tanfovy, grad_color, grad_depth, shs,
                        rs.sh_degree, rs.campos, st.workspace, st.num_rendered, rs.debug)
        with _debug_guard(rs, s, "backward", snap, "snapshot_bw.dump"):
            L.check(lib.mgs_raster_backward_pose(C.byref(view), P, L.ptr(means3D), L.ptr(shs), L.ptr(colors_precomp),
                                                 L.ptr(scales), L.ptr(rotations), L.ptr(cov3D_precomp), L.ptr(radii),
                                                 L.ptr(st.workspace), st.workspace.numel(), L.ptr(scratch), scratch.numel(),
                                                 L.ptr(grad_color), L.ptr(grad_depth), L.ptr(tau_image), L.ptr(tau),
                                                 s), "mgs_raster_backward_pose")
    return tau, tau_image


def raster_backward_pose_deferred(rs, st, grad_color, grad_depth, want_image=False):
    """Pose tangent from the per-pixel sums a single-pass forward left in st.pose_state: (tau_sum[6], tau_image or None)."""
    lib = L.lib()
    H, W = int(rs.image_height), int(rs.image_width)
    dev = st.pose_state.device
    grad_color, grad_depth = _f32c(grad_color), _f32c(grad_depth)
    tau = torch.empty(6, device=dev, dtype=torch.float32)
    tau_image = torch.empty(6, H, W, device=dev, dtype=torch.float32) if want_image else None
    with _on(dev):
        L.check(lib.mgs_raster_backward_pose_deferred(W, H, L.ptr(st.pose_state), L.ptr(grad_color), L.ptr(grad_depth),
                                                      L.ptr(tau_image), L.ptr(tau), _stream(dev)),
                "mgs_raster_backward_pose_deferred")
    return tau, tau_image


def sketch_reduce(tau_image, sketch_indices, sketch_dim):
    """sketch_indices i32[stack,H,W] -> f32[stack, sketch_dim, 6]."""
    lib = L.lib()
    if sketch_indices.dim() == 2:
        sketch_indices = sketch_indices.unsqueeze(0)
    idx = sketch_indices.to(torch.int32).contiguous()
    n_stacks, H, W = idx.shape
    out = torch.empty(n_stacks, sketch_dim, 6, device=tau_image.device, dtype=torch.float32)
    with _on(tau_image.device):
        L.check(lib.mgs_sketch_reduce(L.ptr(tau_image), H, W, n_stacks, sketch_dim, L.ptr(idx), L.ptr(out),
                                      _stream(tau_image.device)), "mgs_sketch_reduce")
    return out


def _opt(t):
    return t if (t is not None and t.numel() > 0) else None


class _RasterizeGaussians(torch.autograd.Function):
    """Autograd node with the reference's argument order (__init__.py:61-84) and gradient tuple (:353-371)."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta, rho,
                raster_settings, num_backward_gaussians, sketch_mode, sketch_dim, stack_dim, sketch_dtau,
                sketch_indices):
        need = ctx.needs_input_grad
        # tracking: nothing but the pose tangent (or its sketched rows) will be asked for
        pose_only = (SINGLE_PASS_TRACKING and not any(need[:8]) and (need[8] or need[9] or (sketch_mode != 0 and need[15])))
        color, radii, depth, opacity, n_touched, st = raster_forward(
            raster_settings, means3D, opacities, _opt(sh), _opt(colors_precomp), _opt(scales), _opt(rotations),
            _opt(cov3Ds_precomp), single_pass=pose_only)
        ctx.raster_settings = raster_settings
        ctx.state = st
        ctx.sketch_mode, ctx.sketch_dim, ctx.stack_dim = sketch_mode, sketch_dim, stack_dim
        ctx.repeat_iter = 0
        ctx.num_rendered = st.num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, sketch_dtau,
                              sketch_indices)
        ctx.mark_non_differentiable(radii, n_touched)
        return color, radii, depth, opacity, n_touched

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_radii, grad_out_depth, grad_out_opacity, grad_n_touched):
        # only grad_out_color and grad_out_depth are consumed (__init__.py:150,185-186)
        rs, st = ctx.raster_settings, ctx.state
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, sketch_dtau, sketch_indices = ctx.saved_tensors
        H, W = int(rs.image_height), int(rs.image_width)
        dev = means3D.device
        if grad_out_color is None:
            grad_out_color = torch.zeros(3, H, W, device=dev)
        depth_grad = grad_out_depth          # None (mono tracking): the deferred contraction then skips the depth planes
        if grad_out_depth is None and st.pose_state is None:
            grad_out_depth = torch.zeros(1, H, W, device=dev)
        need = ctx.needs_input_grad
        g = [None] * 17
        if ctx.sketch_mode != 0:
            # sketched Jacobian rows for stack j = 0..stack_dim-1 of repeat `repeat_iter` (__init__.py:216-253)
            assert ctx.sketch_dim <= 64
            if st.pose_state is not None:
                _, tau_image = raster_backward_pose_deferred(rs, st, grad_out_color, depth_grad, want_image=True)
            else:
                _, tau_image = raster_backward_pose(rs, st, means3D, _opt(sh), _opt(colors_precomp), _opt(scales),
                                                    _opt(rotations), _opt(cov3Ds_precomp), radii, grad_out_color,
                                                    grad_out_depth, want_image=True)
            idx = sketch_indices[ctx.repeat_iter][: ctx.stack_dim]
            g[15] = sketch_reduce(tau_image, idx, ctx.sketch_dim).to(sketch_dtau.dtype).view(sketch_dtau.shape)
            ctx.repeat_iter += 1
            return tuple(g)
        mask = 0
        if need[0]: mask |= L.GRAD_MEANS3D
        if need[1]: mask |= L.GRAD_MEANS2D
        if need[2] or need[3]: mask |= L.GRAD_COLOR
        if need[4]: mask |= L.GRAD_OPACITY
        if need[5] or need[6] or need[7]: mask |= L.GRAD_COV
        want_tau = need[8] or need[9]
        if mask == 0 and want_tau and st.pose_state is not None:
            tau, _ = raster_backward_pose_deferred(rs, st, grad_out_color, depth_grad)
        elif mask == 0 and want_tau:
            tau, _ = raster_backward_pose(rs, st, means3D, _opt(sh), _opt(colors_precomp), _opt(scales),
                                          _opt(rotations), _opt(cov3Ds_precomp), radii, grad_out_color,
                                          grad_out_depth)
        elif mask != 0 or want_tau:
            if want_tau: mask |= L.GRAD_TAU
            out = raster_backward(rs, st, mask, means3D, _opt(sh), _opt(colors_precomp), _opt(scales),
                                  _opt(rotations), _opt(cov3Ds_precomp), radii, grad_out_color, grad_out_depth)
            tau = out.get("tau")
            if need[0]: g[0] = out["means3D"]
            if need[1]: g[1] = out["means2D"]
            if need[2]: g[2] = out["color"]
            if need[3]: g[3] = out["color"]
            if need[4]: g[4] = out["opacity"]
            if need[5]: g[5] = out["scales"]
            if need[6]: g[6] = out["rotations"]
            if need[7]: g[7] = out["cov3D"]
        if want_tau:
            if need[8]: g[8] = tau[3:].view(1, -1)   # grad_theta
            if need[9]: g[9] = tau[:3].view(1, -1)   # grad_rho
        ctx.repeat_iter += 1
        return tuple(g)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta, rho,
                        raster_settings, num_backward_gaussians=-1, sketch_mode=0, sketch_dim=0, stack_dim=0,
                        sketch_dtau=None, sketch_indices=None):
    dev = means3D.device
    if sketch_dtau is None: sketch_dtau = torch.empty(0, device=dev)
    if sketch_indices is None: sketch_indices = torch.empty(0, device=dev, dtype=torch.int32)
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                                     theta, rho, raster_settings, num_backward_gaussians, sketch_mode, sketch_dim,
                                     stack_dim, sketch_dtau, sketch_indices)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """bool[P] near-plane test (reference __init__.py:412-421)."""
        with torch.no_grad():
            rs = self.raster_settings
            pos = _f32c(positions)
            P = pos.size(0)
            out = torch.zeros(P, device=pos.device, dtype=torch.uint8)
            vm, pm = _f32c(rs.viewmatrix), _f32c(rs.projmatrix)
            with _on(pos.device):
                L.check(L.lib().mgs_mark_visible(P, L.ptr(pos), L.ptr(vm), L.ptr(pm), L.ptr(out), _stream(pos.device)),
                        "mgs_mark_visible")
        return out.bool()

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, theta=None, rho=None, num_backward_gaussians=None, sketch_mode=0, sketch_dim=0,
                stack_dim=0, sketch_dtau=None, sketch_indices=None):
        raster_settings = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = lambda: torch.Tensor([])
        shs = empty() if shs is None else shs
        colors_precomp = empty() if colors_precomp is None else colors_precomp
        scales = empty() if scales is None else scales
        rotations = empty() if rotations is None else rotations
        cov3D_precomp = empty() if cov3D_precomp is None else cov3D_precomp
        theta = empty() if theta is None else theta
        rho = empty() if rho is None else rho
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   theta, rho, raster_settings, num_backward_gaussians=num_backward_gaussians,
                                   sketch_mode=sketch_mode, sketch_dim=sketch_dim, stack_dim=stack_dim,
                                   sketch_dtau=sketch_dtau, sketch_indices=sketch_indices)
