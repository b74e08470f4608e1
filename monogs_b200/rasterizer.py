"""Host side of the drop-in boundary: the `diff_gaussian_rasterization` Python surface on top of the C-ABI.

Mirrors submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py of the reference:
`GaussianRasterizationSettings` (:392-405), `GaussianRasterizer` (:407-469), `rasterize_gaussians` (:22-59)
and the autograd semantics of `_RasterizeGaussians` (:61-390): same argument order, same 5-tuple of outputs,
same 17-tuple of gradients, `grad_rho` / `grad_theta` shaped [1,3].  torch is used for device memory and
the current stream only; all arithmetic happens in libmonogs_b200.so.
"""
import ctypes as C
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


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def make_view(rs: GaussianRasterizationSettings, sh_coeffs: int, keep: list) -> L.MgsView:
    """Build the C struct; `keep` collects the (possibly converted) tensors so they outlive the call."""
    dev = rs.viewmatrix.device
    tens = {}
    for name in ("bg", "viewmatrix", "projmatrix", "projmatrix_raw", "campos"):
        t = getattr(rs, name)
        t = _f32c(t.to(dev)) if t is not None else None
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
    __slots__ = ("workspace", "num_rendered", "P", "sh_coeffs")


def raster_forward(rs, means3D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp):
    """Two-stage forward with the reference's single host read of num_rendered in between
    (rasterizer_impl.cu:284).  Returns (color, radii, depth, opacity, n_touched, state)."""
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
    color = torch.empty(3, H, W, device=dev, dtype=torch.float32)
    depth = torch.empty(1, H, W, device=dev, dtype=torch.float32)
    opacity = torch.empty(1, H, W, device=dev, dtype=torch.float32)
    radii = torch.zeros(P, device=dev, dtype=torch.int32)
    n_touched = torch.zeros(P, device=dev, dtype=torch.int32)
    st = RasterState()
    st.P, st.sh_coeffs = P, M
    if P == 0:  # rasterize_points.cu:90 — legal, zero-filled outputs
        color.zero_(); depth.zero_(); opacity.zero_()
        st.workspace, st.num_rendered = torch.empty(0, device=dev, dtype=torch.uint8), 0
        return color, radii, depth, opacity, n_touched, st
    s = _stream()
    base = lib.mgs_raster_workspace_bytes(P, W, H, 0)
    ws = torch.empty(base, device=dev, dtype=torch.uint8)
    n_host = torch.empty(1, dtype=torch.int32).pin_memory()
    L.check(lib.mgs_raster_preprocess(C.byref(view), P, L.ptr(means3D), L.ptr(opacities), L.ptr(shs),
                                      L.ptr(colors_precomp), L.ptr(scales), L.ptr(rotations), L.ptr(cov3D_precomp),
                                      L.ptr(radii), L.ptr(ws), ws.numel(), C.c_void_p(n_host.data_ptr()), s),
            "mgs_raster_preprocess")
    torch.cuda.current_stream().synchronize()
    R = int(n_host.item())
    need = lib.mgs_raster_workspace_bytes(P, W, H, R)
    if need > ws.numel():
        big = torch.empty(need, device=dev, dtype=torch.uint8)
        big[:base].copy_(ws)  # per-Gaussian state sits at R-independent offsets
        ws = big
    L.check(lib.mgs_raster_render(C.byref(view), P, L.ptr(colors_precomp), L.ptr(radii), L.ptr(ws), ws.numel(), R,
                                  L.ptr(color), L.ptr(depth), L.ptr(opacity), L.ptr(n_touched), s),
            "mgs_raster_render")
    st.workspace, st.num_rendered = ws, R
    return color, radii, depth, opacity, n_touched, st


def raster_backward(rs, st, grad_mask, means3D, shs, colors_precomp, scales, rotations, cov3D_precomp, radii,
                    grad_color, grad_depth, want_tau_per_gaussian=False):
    """One call of mgs_raster_backward; returns a dict of the requested gradients."""
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
    scratch = torch.empty(lib.mgs_raster_backward_scratch_bytes(P), device=dev, dtype=torch.uint8)
    g = lambda k: L.ptr(out[k]) if k in out else None
    L.check(lib.mgs_raster_backward(C.byref(view), P, grad_mask, L.ptr(means3D), L.ptr(shs), L.ptr(colors_precomp),
                                    L.ptr(scales), L.ptr(rotations), L.ptr(cov3D_precomp), L.ptr(radii),
                                    L.ptr(st.workspace), st.workspace.numel(), L.ptr(scratch), scratch.numel(),
                                    L.ptr(grad_color), L.ptr(grad_depth), g("means3D"), g("means2D"), g("color"),
                                    g("opacity"), g("scales"), g("rotations"), g("cov3D"), g("tau"), g("tau_pg"),
                                    _stream()), "mgs_raster_backward")
    return out
