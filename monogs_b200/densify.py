"""Host side of the arena densify / prune (mgs_densify_plan + mgs_densify_apply).

Mirrors GaussianModel.densify_and_prune and prune_points (gaussian_splatting/scene/gaussian_model.py:470-691) on a dict
of raw parameter tensors {xyz, f_dc, f_rest, opacity, scaling, rotation} and their Adam moments.  torch allocates the new
arrays and (optionally) draws the split noise; every decision, copy and the split arithmetic run in the library."""
import ctypes as C
from typing import Dict, Optional

import torch

from . import _lib as L
from .rasterizer import _stream

NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def _gaussians(params):
    g = L.MgsGaussians()
    g.P = params["xyz"].size(0)
    g.sh_coeffs = 1 + (params["f_rest"].size(1) if params.get("f_rest") is not None and params["f_rest"].numel() > 0 else 0)
    g.sh_degree = 0
    g.isotropic = int(params["scaling"].size(-1) == 1)
    g.xyz, g.f_dc = params["xyz"].data_ptr(), params["f_dc"].data_ptr()
    g.f_rest = params["f_rest"].data_ptr() if g.sh_coeffs > 1 else None
    g.opacity_raw, g.scaling_raw, g.rotation_raw = params["opacity"].data_ptr(), params["scaling"].data_ptr(), params["rotation"].data_ptr()
    return g


def _state(params, m, v, kf, nobs):
    st = L.MgsGaussianState()
    for i, k in enumerate(NAMES):
        t = params.get(k)
        st.param[i] = t.data_ptr() if t is not None and t.numel() > 0 else None
        st.exp_avg[i] = m[k].data_ptr() if m is not None and m.get(k) is not None and m[k].numel() > 0 else None
        st.exp_avg_sq[i] = v[k].data_ptr() if v is not None and v.get(k) is not None and v[k].numel() > 0 else None
    st.unique_kfIDs = kf.data_ptr() if kf is not None else None
    st.n_obs = nobs.data_ptr() if nobs is not None else None
    return st


def _run(params, exp_avg, exp_avg_sq, cfg, grad_accum, denom, prune_mask, unique_kfIDs, n_obs, noise, generator):
    lib = L.lib()
    for k in NAMES:
        if params.get(k) is not None:
            assert params[k].is_contiguous() and params[k].dtype == torch.float32
    dev = params["xyz"].device
    P = params["xyz"].size(0)
    g = _gaussians(params)
    scratch = torch.empty(lib.mgs_densify_scratch_bytes(P), device=dev, dtype=torch.uint8)
    totals = torch.zeros(4, dtype=torch.int32).pin_memory()
    pm = prune_mask.to(torch.uint8).contiguous() if prune_mask is not None else None
    L.check(lib.mgs_densify_plan(C.byref(g), L.ptr(grad_accum), L.ptr(denom), L.ptr(pm), C.byref(cfg), L.ptr(scratch),
                                 scratch.numel(), C.c_void_p(totals.data_ptr()), _stream()), "mgs_densify_plan")
    torch.cuda.current_stream().synchronize()          # the new size is needed on the host to allocate
    n_keep, n_clone, n_child, n_sel = (int(x) for x in totals)
    n_split = cfg.n_split if cfg.mode == 0 else 0
    P_new = n_keep + n_clone + n_split * n_child
    if noise is None and n_sel > 0 and n_split > 0:
        noise = torch.randn(n_split * n_sel, 3, device=dev, generator=generator)
    if noise is not None:
        assert noise.shape == (n_split * n_sel, 3) and noise.is_contiguous()
    new = lambda t: torch.empty((P_new,) + tuple(t.shape[1:]), device=dev, dtype=t.dtype)
    out_p = {k: new(params[k]) for k in NAMES if params.get(k) is not None}
    out_m = {k: new(params[k]) for k in out_p} if exp_avg is not None else None
    out_v = {k: new(params[k]) for k in out_p} if exp_avg_sq is not None else None
    out_kf = new(unique_kfIDs) if unique_kfIDs is not None else None
    out_no = new(n_obs) if n_obs is not None else None
    if P_new > 0:
        src = _state(params, exp_avg, exp_avg_sq, unique_kfIDs, n_obs)
        dst = _state(out_p, out_m, out_v, out_kf, out_no)
        L.check(lib.mgs_densify_apply(C.byref(g), C.byref(src), C.byref(dst), C.byref(cfg), L.ptr(noise), L.ptr(scratch),
                                      scratch.numel(), _stream()), "mgs_densify_apply")
    return dict(params=out_p, exp_avg=out_m, exp_avg_sq=out_v, unique_kfIDs=out_kf, n_obs=out_no,
                counts=dict(kept=n_keep, cloned=n_clone, children=n_split * n_child, split_selected=n_sel))


def densify_and_prune(params: Dict[str, torch.Tensor], exp_avg: Optional[Dict[str, torch.Tensor]],
                      exp_avg_sq: Optional[Dict[str, torch.Tensor]], xyz_gradient_accum: torch.Tensor, denom: torch.Tensor,
                      max_grad: float, min_opacity: float, extent: float, max_screen_size: Optional[float],
                      percent_dense: float = 0.01, n_split: int = 2, unique_kfIDs=None, n_obs=None, noise=None,
                      generator=None):
    """GaussianModel.densify_and_prune (gaussian_model.py:676-691).  All tensors live on the CUDA device; unique_kfIDs and
    n_obs are i32[P].  Returns the new arrays; the caller resets xyz_gradient_accum / denom / max_radii2D to zeros of
    the new size (densification_postfix, :592-594)."""
    cfg = L.MgsDensifyCfg()
    cfg.mode, cfg.n_split = 0, n_split
    cfg.grad_threshold, cfg.min_opacity, cfg.extent, cfg.percent_dense = max_grad, min_opacity, extent, percent_dense
    cfg.max_screen_size = float(max_screen_size) if max_screen_size else 0.0
    return _run(params, exp_avg, exp_avg_sq, cfg, xyz_gradient_accum.contiguous().view(-1), denom.contiguous().view(-1),
                None, unique_kfIDs, n_obs, noise, generator)


def prune_points(params, exp_avg, exp_avg_sq, mask: torch.Tensor, unique_kfIDs=None, n_obs=None):
    """GaussianModel.prune_points(mask) (gaussian_model.py:512-527): drops the Gaussians where mask is True."""
    cfg = L.MgsDensifyCfg()
    cfg.mode, cfg.n_split = 1, 0
    return _run(params, exp_avg, exp_avg_sq, cfg, None, None, mask, unique_kfIDs, n_obs, None, None)
