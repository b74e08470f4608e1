"""Host wrappers of the per-frame image preprocessing and covisibility kernels (csrc/frame_ops.cu).

Reference behaviour mirrored (MonoGS tree): Camera.compute_grad_mask (utils/camera_utils.py:110-142) with image_gradient /
image_gradient_mask (utils/slam_utils.py:7-41), get_median_depth (utils/slam_utils.py:286-297), and the visibility-overlap
counts of is_keyframe / add_to_window / the mapping overlap test (utils/slam_frontend.py:1713-1741, 1918-1932).
torch allocates; every number is computed in libmonogs_b200.so, nothing synchronises the host."""
import ctypes as C

import torch

from . import _lib as L
from .rasterizer import _f32c, _stream


def _scratch(dev):
    return torch.empty(L.lib().mgs_median_scratch_bytes() + 256, device=dev, dtype=torch.uint8)


def grad_intensity(image: torch.Tensor, eps: float = 0.01) -> torch.Tensor:
    """f32[H,W]: masked Scharr gradient magnitude of the channel-mean of image f32[C,H,W]."""
    image = _f32c(image)
    Cc, H, W = image.shape
    out = torch.empty(H, W, device=image.device, dtype=torch.float32)
    L.check(L.lib().mgs_frame_grad_intensity(L.ptr(image), Cc, H, W, eps, L.ptr(out), _stream()), "mgs_frame_grad_intensity")
    return out


def frame_masks(image: torch.Tensor, edge_threshold: float, rgb_boundary_threshold: float = 0.01, patch_median: bool = False):
    """(grad_mask, rgb_pixel_mask, rgb_pixel_mask_mapping), each f32[1,H,W] of 0/1 — Camera.compute_grad_mask.
    patch_median=True is the Replica variant (32x32 patch medians)."""
    image = _f32c(image)
    _, H, W = image.shape
    f = dict(device=image.device, dtype=torch.float32)
    inten = torch.empty(H, W, **f)
    gm, rm, rmm = torch.empty(1, H, W, **f), torch.empty(1, H, W, **f), torch.empty(1, H, W, **f)
    sc = _scratch(image.device)
    L.check(L.lib().mgs_frame_masks(L.ptr(image), H, W, edge_threshold, rgb_boundary_threshold, int(patch_median), L.ptr(inten),
                                    L.ptr(gm), L.ptr(rm), L.ptr(rmm), L.ptr(sc), sc.numel(), _stream()), "mgs_frame_masks")
    return gm, rm, rmm


def median(x: torch.Tensor) -> torch.Tensor:
    """torch.median(x) (lower median) as a device f32[1]; no sort, no host synchronisation."""
    x = _f32c(x)
    out = torch.empty(1, device=x.device, dtype=torch.float32)
    sc = _scratch(x.device)
    L.check(L.lib().mgs_median(L.ptr(x), x.numel(), L.ptr(out), L.ptr(sc), sc.numel(), _stream()), "mgs_median")
    return out


def median_depth(depth: torch.Tensor, opacity: torch.Tensor = None, mask: torch.Tensor = None):
    """get_median_depth: (median f32[1], n_valid i32[1]) over depth > 0 [& opacity > 0.95] [& mask]."""
    depth = _f32c(depth)
    opacity = _f32c(opacity) if opacity is not None else None
    mask = _f32c(mask.float()) if mask is not None else None
    out = torch.empty(1, device=depth.device, dtype=torch.float32)
    nv = torch.empty(1, device=depth.device, dtype=torch.int32)
    sc = _scratch(depth.device)
    L.check(L.lib().mgs_median_depth(L.ptr(depth), L.ptr(opacity), L.ptr(mask), depth.numel(), L.ptr(out), L.ptr(nv), L.ptr(sc),
                                     sc.numel(), _stream()), "mgs_median_depth")
    return out, nv


def pack_visibility(n_touched: torch.Tensor) -> torch.Tensor:
    """i32[(P+31)//32] bit mask of n_touched > 0 (occ_aware_visibility at 1 bit per Gaussian)."""
    nt = n_touched.to(torch.int32).contiguous()
    P = nt.numel()
    bits = torch.zeros((P + 31) // 32, device=nt.device, dtype=torch.int32)
    L.check(L.lib().mgs_visibility_pack(P, L.ptr(nt), L.ptr(bits), _stream()), "mgs_visibility_pack")
    return bits


def covisibility(P: int, cur_bits: torch.Tensor, kf_bits: torch.Tensor) -> torch.Tensor:
    """i32[K+1,3]: rows k < K = (intersection, union, |kf_k|) of the current frame with keyframe k; row K = (|cur|, 0, 0).
    IoU (is_keyframe) = [:,0]/[:,1]; Szymkiewicz-Simpson (add_to_window) = [:,0]/min(|cur|, [:,2])."""
    kf_bits = kf_bits.contiguous().view(-1, cur_bits.numel()) if kf_bits.numel() else kf_bits.view(0, cur_bits.numel())
    K = kf_bits.size(0)
    out = torch.empty(K + 1, 3, device=cur_bits.device, dtype=torch.int32)
    L.check(L.lib().mgs_covisibility(P, L.ptr(cur_bits), K, L.ptr(kf_bits), L.ptr(out), _stream()), "mgs_covisibility")
    return out
