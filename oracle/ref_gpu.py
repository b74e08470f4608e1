"""TEST INFRASTRUCTURE ONLY — ctypes driver for oracle/_ref/libref_rasterizer.so, the unmodified reference
CUDA rasterizer + simple-knn built by oracle/Makefile.  It is the GPU oracle of the parity tests and the
"reference CUDA" timing shown beside the bench line.  Nothing under monogs_b200/ imports this module.

Argument order follows CudaRasterizer::Rasterizer::forward / backward
(submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:20-140); output and gradient buffers are
allocated zero-filled here exactly as rasterize_points.cu:71-75,198-208 does.
"""
import ctypes as C
import os

import torch

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_rasterizer.so")
_lib = None


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_PATH)
        _lib.ref_forward.restype = C.c_int
    return _lib


def _p(t):
    return C.c_void_p(t.data_ptr()) if (t is not None and t.numel() > 0) else None


def forward(cam, bg, means3D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
            cov3D_precomp=None, sh_degree=0, scale_modifier=1.0, slot=0):
    dev = means3D.device
    P, W, H = means3D.size(0), cam.W, cam.H
    M = shs.size(1) if shs is not None else 0
    color = torch.zeros(3, H, W, device=dev)
    depth = torch.zeros(1, H, W, device=dev)
    opacity = torch.zeros(1, H, W, device=dev)
    radii = torch.zeros(P, device=dev, dtype=torch.int32)
    n_touched = torch.zeros(P, device=dev, dtype=torch.int32)
    R = lib().ref_forward(slot, P, sh_degree, M, _p(bg), W, H, _p(means3D), _p(shs), _p(colors_precomp), _p(opacities),
                          _p(scales), C.c_float(scale_modifier), _p(rotations), _p(cov3D_precomp), _p(cam.viewmatrix),
                          _p(cam.projmatrix), _p(cam.campos), C.c_float(cam.tanfovx), C.c_float(cam.tanfovy), 0,
                          _p(color), _p(depth), _p(opacity), _p(radii), _p(n_touched))
    return dict(color=color, depth=depth, opacity=opacity, radii=radii, n_touched=n_touched, R=R, P=P, M=M)


def backward(cam, bg, fw, means3D, grad_color, grad_depth, shs=None, colors_precomp=None, scales=None, rotations=None,
             cov3D_precomp=None, sh_degree=0, scale_modifier=1.0, slot=0, selected_pixel_indices=None):
    """selected_pixel_indices (int32 [n] on the device, n > 0): the reference's select_pixels branch
    (renderPixelSampleCUDA) instead of the dense renderCUDA."""
    dev = means3D.device
    P, M, W, H = fw["P"], fw["M"], cam.W, cam.H
    z = lambda *s: torch.zeros(*s, device=dev)
    g = dict(means2D=z(P, 3), conic=z(P, 2, 2), opacity=z(P, 1), color=z(P, 3), depth=z(P, 1), means3D=z(P, 3),
             cov3D=z(P, 6), sh=z(P, max(M, 1), 3), scales=z(P, 3), rotations=z(P, 4), tau=z(P, 6))
    if selected_pixel_indices is not None:
        sel = selected_pixel_indices.to(device=dev, dtype=torch.int32).contiguous()
        lib().ref_backward_pixels(slot, sel.numel(), _p(sel), P, sh_degree, M, fw["R"], _p(bg), W, H, _p(means3D), _p(shs),
                                  _p(colors_precomp), _p(scales), C.c_float(scale_modifier), _p(rotations),
                                  _p(cov3D_precomp), _p(cam.viewmatrix), _p(cam.projmatrix), _p(cam.projmatrix_raw),
                                  _p(cam.campos), C.c_float(cam.tanfovx), C.c_float(cam.tanfovy), _p(fw["radii"]),
                                  _p(grad_color), _p(grad_depth), _p(g["means2D"]), _p(g["conic"]), _p(g["opacity"]),
                                  _p(g["color"]), _p(g["depth"]), _p(g["means3D"]), _p(g["cov3D"]), _p(g["sh"]),
                                  _p(g["scales"]), _p(g["rotations"]), _p(g["tau"]))
        torch.cuda.synchronize()
        return g
    lib().ref_backward(slot, P, sh_degree, M, fw["R"], _p(bg), W, H, _p(means3D), _p(shs), _p(colors_precomp), _p(scales),
                       C.c_float(scale_modifier), _p(rotations), _p(cov3D_precomp), _p(cam.viewmatrix),
                       _p(cam.projmatrix), _p(cam.projmatrix_raw), _p(cam.campos), C.c_float(cam.tanfovx),
                       C.c_float(cam.tanfovy), _p(fw["radii"]), _p(grad_color), _p(grad_depth), _p(g["means2D"]),
                       _p(g["conic"]), _p(g["opacity"]), _p(g["color"]), _p(g["depth"]), _p(g["means3D"]),
                       _p(g["cov3D"]), _p(g["sh"]), _p(g["scales"]), _p(g["rotations"]), _p(g["tau"]))
    return g


def backward_sketch(cam, bg, fw, means3D, grad_color, grad_depth, sketch_indices, sketch_dim, shs=None,
                    colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, sh_degree=0,
                    scale_modifier=1.0, slot=0):
    """Returns df_dtau [P, d, 6] (the only output the reference wrapper consumes, __init__.py:253)."""
    dev = means3D.device
    P, M, W, H, d = fw["P"], fw["M"], cam.W, cam.H, sketch_dim
    z = lambda *s: torch.zeros(*s, device=dev)
    g = dict(means2D=z(P, d, 3), conic=z(P, d, 2, 2), opacity=z(P, d, 1), color=z(P, d, 3), depth=z(P, d, 1),
             means3D=z(P, d, 3), cov3D=z(P, d, 6), sh=z(P, d, max(M, 1), 3), scales=z(P, d, 3), rotations=z(P, d, 4),
             tau=z(P, d, 6))
    lib().ref_backward_sketch(slot, 1, d, _p(sketch_indices), P, sh_degree, M, fw["R"], _p(bg), W, H, _p(means3D),
                              _p(shs), _p(colors_precomp), _p(scales), C.c_float(scale_modifier), _p(rotations),
                              _p(cov3D_precomp), _p(cam.viewmatrix), _p(cam.projmatrix), _p(cam.projmatrix_raw),
                              _p(cam.campos), C.c_float(cam.tanfovx), C.c_float(cam.tanfovy), _p(fw["radii"]),
                              _p(grad_color), _p(grad_depth), _p(g["means2D"]), _p(g["conic"]), _p(g["opacity"]),
                              _p(g["color"]), _p(g["depth"]), _p(g["means3D"]), _p(g["cov3D"]), _p(g["sh"]),
                              _p(g["scales"]), _p(g["rotations"]), _p(g["tau"]))
    return g["tau"]


def binning(fw, slot=0):
    dev = fw["radii"].device
    R = fw["R"]
    keys = torch.zeros(max(R, 1), device=dev, dtype=torch.int64)
    plist = torch.zeros(max(R, 1), device=dev, dtype=torch.int32)
    if R > 0:
        lib().ref_get_binning(slot, R, _p(keys), _p(plist))
    return keys[:R], plist[:R]


def image_state(cam, slot=0):
    tiles = ((cam.W + 15) // 16) * ((cam.H + 15) // 16)
    ranges = torch.zeros(tiles, 2, device="cuda", dtype=torch.int32)
    final_T = torch.zeros(cam.H, cam.W, device="cuda")
    n_contrib = torch.zeros(cam.H, cam.W, device="cuda", dtype=torch.int32)
    lib().ref_get_image_state(slot, cam.W, cam.H, _p(ranges), tiles, _p(final_T), _p(n_contrib))
    return ranges, final_T, n_contrib


def knn_dist2(points):
    out = torch.zeros(points.size(0), device=points.device)
    lib().ref_knn_dist2(points.size(0), _p(points.contiguous()), _p(out))
    return out


def mark_visible(means3D, cam):
    out = torch.zeros(means3D.size(0), device=means3D.device, dtype=torch.bool)
    lib().ref_mark_visible(means3D.size(0), _p(means3D), _p(cam.viewmatrix), _p(cam.projmatrix), _p(out))
    return out
