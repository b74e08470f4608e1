// Host-side launch functions shared between the translation units of libmonogs_b200.
#pragma once
#include "common.cuh"

namespace mgs {

const char* get_error();

// fused loops: the tile scan runs in the last CTA of the preprocess kernel, no counter memset (see raster_forward.cu)
// in_kernel: run the scan in the last CTA of preprocess_kernel (true) or as its own one-CTA launch behind it (false).
// Measured on the B200: the separate launch wins in the serial tracking iteration (+2 %), the folded tail wins in the mapping
// engine, whose views overlap on several streams (+3 %); MGS_SCAN_FOLD=0|1 overrides both for measurements.
// segment_cap != 0 (fused tracker, packed-row mode): fixed segments of that many instances per tile, scattered by the
// preprocess kernel itself; no scan and no emit — launch_sort_segments() follows instead of launch_binning().
struct FoldScan { int64_t R_cap; bool heavy_first; bool in_kernel; int segment_cap = 0; };
constexpr int TILE_SEGMENT = 1024;      // instances per tile segment (= the longest list the one-CTA-per-tile sort takes)
int launch_sort_segments(const ViewConst& v, void* ws, const WsLayout& L, int64_t R_cap, cudaStream_t s, bool ids_only);
int launch_preprocess(const ViewConst& v, int P, const float* means3D, const float* opacities, const float* shs,
                      const float* colors_precomp, const float* scales, const float* rotations,
                      const float* cov3D_precomp, int32_t* radii, void* ws, const WsLayout& L, cudaStream_t s,
                      float* jac = nullptr, int jac_ncol = 0, const float* sh_rest = nullptr, const FoldScan* fold = nullptr);
int launch_tile_scan(const ViewConst& v, void* ws, const WsLayout& L, int64_t R_cap, cudaStream_t s, bool heavy_first = false);
int launch_binning(const ViewConst& v, int P, const float* colors_precomp, const int32_t* radii, void* ws,
                   const WsLayout& L, int64_t R_cap, cudaStream_t s, bool heavy_first = false, bool scan_done = false,
                   bool clear_counts = false, bool ids_only = false);   // ids_only: sorted Gaussian ids (u32) instead of records
struct LossK;
// optional fused loss epilogue of the forward blend (tracking path): per-pixel loss + un-normalised gradient images
struct FusedLoss { const LossK* k; float grad_scale; float* g_img; float* g_depth; double* partials; bool fast; };
int launch_blend_forward(const ViewConst& v, void* ws, const WsLayout& L, float* out_color, float* out_depth,
                         float* out_opacity, int32_t* n_touched, cudaStream_t s, const FusedLoss* fl = nullptr,
                         bool heavy_first = false, bool clear_counts = false);   // clear_counts: fixed-segment binning (zero each tile's counter)

// second-order epilogue of the single-pass blend: where a pixel's Jacobian-row contribution goes (sketch_indices i32[n_stacks,H,W],
// rows f32[n_stacks * d, 9], zero on entry)
struct SketchRoute { const int32_t* idx; float* rows; int n_stacks, d; };
// single-pass tracking blend (blend_track.cu): forward + loss + pose tangent; SH degree 0 / 6-column Jacobians only
int launch_blend_track(const ViewConst& v, void* ws, const WsLayout& L, const float* jac, float* out_color,
                       float* out_depth, float* out_opacity, int32_t* n_touched, const FusedLoss& fl, bool has_depth,
                       float* tau_sum, cudaStream_t s, const float* sk_weights = nullptr, float sk_scale = 0.f,
                       float* sk_image = nullptr, const struct TrackUpdateK* upd = nullptr, uint32_t* ticket = nullptr,
                       bool deferred = false, bool rows = false,    // rows: ids + packed per-Gaussian rows (fused tracker)
                       bool clear_counts = false,                   // fixed-segment binning: zero each tile's instance counter
                       const SketchRoute* route = nullptr);
int blend_track_warps();      // 8 or 4: CTAs per tile = 8 / warps; loss partial slots = tiles * 8 / warps
// deferred single pass of the drop-in surface: accumulator planes of blend_track<EPI = 2> and their contraction
size_t pose_deferred_bytes(int W, int H);
int launch_pose_contract(const float* ea, int W, int H, const float* g_color, const float* g_depth, float* tau_image,
                         float* tau_sum, cudaStream_t s);

struct BackwardArgs {
  ViewConst v; int P; uint32_t grad_mask;
  const float *means3D, *shs, *colors_precomp, *scales, *rotations, *cov3D_precomp;
  const int32_t* radii; const void* ws; WsLayout L;
  void* scratch; size_t scratch_bytes;
  const float *dL_dcolor_img, *dL_ddepth_img;
  float *dL_dmeans3D, *dL_dmeans2D, *dL_dcolors_or_sh, *dL_dopacity, *dL_dscales, *dL_drotations, *dL_dcov3D;
  float *dL_dtau_sum, *dL_dtau_pg;
  // mapping window (map_fused.cu): accumulate == 1 adds into the gradient outputs (visible Gaussians only) instead of
  // overwriting them; optional densification statistics updated in the same pass (gaussian_model.py:693-697,
  // slam_backend.py:292-299): grad_accum[g] += |dL/dmean2D.xy|, denom[g] += 1, max_radii2D[g] = max(., radii[g])
  int accumulate; float *stat_grad_accum, *stat_denom, *stat_max_radii;
  // split SH layout (mapping engine): shs = f_dc [P,1,3], sh_rest = f_rest [P,M-1,3]; gradients go to dL_dcolors_or_sh
  // ([P,1,3]) and dL_dsh_rest ([P,M-1,3]).  Null = one array [P,M,3].
  const float* sh_rest = nullptr; float* dL_dsh_rest = nullptr;
  bool pre_cleared = false;   // launch_backward_full: the caller has already zeroed the scratch accumulator and dL_dtau_sum on this stream
  bool fast = false;   // blend_backward_full only: exponent variant of the forward that produced ws (blend_exp<FAST>)
  bool heavy_first = false;   // blend_backward_full only: walk the tiles in the scan's heavy-first order (tile_ctr[.. + 3])
  const int32_t* pixel_mult = nullptr;   // blend_backward_full only: i32[H*W] multiplicity of each pixel in the selected-pixel
                                         // list of mgs_raster_backward_pixels (null: every pixel once)
  // sketch mode (sketch_dim > 0): per-pixel row routing, output sketch_dtau[d,6]
  int sketch_dim; const int32_t* sketch_indices; float* sketch_dtau;
};
int launch_backward(const BackwardArgs& a, cudaStream_t s);
int launch_pixel_multiplicity(const int32_t* indices, int n, int npix, int32_t* mult, cudaStream_t s);   // mult[p] = #occurrences of p
int launch_backward_full(const BackwardArgs& a, cudaStream_t s);
int launch_blend_backward_pose(const BackwardArgs& a, const float* jac, bool with_color, cudaStream_t s, bool fast = false);
size_t backward_scratch_bytes(int P, int sketch_dim);
size_t backward_g2d_bytes();      // bytes per Gaussian of the accumulator launch_backward_full zero-fills at the head of its scratch

}  // namespace mgs
