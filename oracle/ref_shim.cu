// TEST INFRASTRUCTURE ONLY — GPU oracle shim.
// Wraps the UNMODIFIED reference rasterizer (compiled from /root/reference where it lies, see
// oracle/Makefile) behind a small C ABI so the parity tests and bench.py can run it on the B200 box through
// ctypes, without torch's extension builder.  This file contains no rasterizer logic of its own: it only
// owns the three arenas the reference asks for (what rasterize_points.cu:27-33,81-87 does with torch
// tensors) and forwards pointers to CudaRasterizer::Rasterizer (rasterizer.h:20-140) and SimpleKNN::knn
// (simple_knn.h).  Nothing in monogs_b200/ may link or load it.
#include <cuda_runtime.h>
#include <functional>
#include <cstdint>
#include <cstdio>
// reference headers, given as absolute paths by the Makefile: the rasterizer directory cannot go on the -I
// path because its own "math.h" would shadow the system header
#include REF_RASTERIZER_IMPL_H   // GeometryState / BinningState / ImageState + rasterizer.h
#include REF_SIMPLE_KNN_H

namespace {
struct Arena {
  char* ptr = nullptr; size_t cap = 0;
  char* get(size_t n) {
    if (n > cap) { if (ptr) cudaFree(ptr); cudaMalloc(&ptr, n); cap = n; }
    return ptr;
  }
};
// slot-indexed arenas so several forward states can stay alive (e.g. a mapping window)
constexpr int kSlots = 16;
Arena g_geom[kSlots], g_bin[kSlots], g_img[kSlots];
}  // namespace

extern "C" {

// returns num_rendered.  Output buffers must be zero-filled by the caller (rasterize_points.cu:71-75).
int ref_forward(int slot, int P, int D, int M, const float* bg, int W, int H, const float* means3D, const float* shs,
                const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                const float* campos, float tan_fovx, float tan_fovy, int prefiltered, float* out_color,
                float* out_depth, float* out_opacity, int* radii, int* n_touched) {
  std::function<char*(size_t)> gf = [slot](size_t n) { return g_geom[slot].get(n); };
  std::function<char*(size_t)> bf = [slot](size_t n) { return g_bin[slot].get(n); };
  std::function<char*(size_t)> imf = [slot](size_t n) { return g_img[slot].get(n); };
  return CudaRasterizer::Rasterizer::forward(gf, bf, imf, P, D, M, bg, W, H, means3D, shs, colors_precomp, opacities,
                                             scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix,
                                             campos, tan_fovx, tan_fovy, prefiltered != 0, out_color, out_depth,
                                             out_opacity, radii, n_touched, false);
}

// Gradient buffers must be zero-filled by the caller (rasterize_points.cu:198-208).
void ref_backward(int slot, int P, int D, int M, int R, const float* bg, int W, int H, const float* means3D,
                  const float* shs, const float* colors_precomp, const float* scales, float scale_modifier,
                  const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                  const float* projmatrix, const float* projmatrix_raw, const float* campos, float tan_fovx,
                  float tan_fovy, const int* radii, const float* dL_dpix, const float* dL_dpix_depth,
                  float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_ddepth,
                  float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                  float* dL_dtau) {
  CudaRasterizer::Rasterizer::backward(P, D, M, R, bg, W, H, means3D, shs, colors_precomp, scales, scale_modifier,
                                       rotations, cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw, campos,
                                       tan_fovx, tan_fovy, radii, g_geom[slot].ptr, g_bin[slot].ptr, g_img[slot].ptr,
                                       0, nullptr, 0, nullptr, nullptr, dL_dpix, dL_dpix_depth, dL_dmean2D, dL_dconic,
                                       dL_dopacity, dL_dcolor, dL_ddepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale,
                                       dL_drot, dL_dtau, false);
}

// The select_pixels = true branch of the reference wrapper (rasterize_points.cu:189,233): same call with the pixel list,
// which makes BACKWARD::render launch renderPixelSampleCUDA (backward.cu:1139-1165) instead of renderCUDA.
void ref_backward_pixels(int slot, int n_selected, int* selected_pixel_indices, int P, int D, int M, int R, const float* bg,
                         int W, int H, const float* means3D, const float* shs, const float* colors_precomp,
                         const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                         const float* viewmatrix, const float* projmatrix, const float* projmatrix_raw,
                         const float* campos, float tan_fovx, float tan_fovy, const int* radii, const float* dL_dpix,
                         const float* dL_dpix_depth, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                         float* dL_dcolor, float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                         float* dL_dscale, float* dL_drot, float* dL_dtau) {
  CudaRasterizer::Rasterizer::backward(P, D, M, R, bg, W, H, means3D, shs, colors_precomp, scales, scale_modifier,
                                       rotations, cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw, campos,
                                       tan_fovx, tan_fovy, radii, g_geom[slot].ptr, g_bin[slot].ptr, g_img[slot].ptr,
                                       n_selected, selected_pixel_indices, 0, nullptr, nullptr, dL_dpix, dL_dpix_depth,
                                       dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_ddepth, dL_dmean3D, dL_dcov3D,
                                       dL_dsh, dL_dscale, dL_drot, dL_dtau, false);
}

// Sketch-Jacobian backward; all df_* buffers are [P, sketch_dim, ...] and zero-filled by the caller.
void ref_backward_sketch(int slot, int sketch_mode, int sketch_dim, const int* sketch_indices, int P, int D, int M,
                         int R, const float* bg, int W, int H, const float* means3D, const float* shs,
                         const float* colors_precomp, const float* scales, float scale_modifier,
                         const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                         const float* projmatrix, const float* projmatrix_raw, const float* campos, float tan_fovx,
                         float tan_fovy, const int* radii, const float* df_dpix, const float* df_dpix_depth,
                         float* df_dmean2D, float* df_dconic, float* df_dopacity, float* df_dcolor, float* df_ddepth,
                         float* df_dmean3D, float* df_dcov3D, float* df_dsh, float* df_dscale, float* df_drot,
                         float* df_dtau) {
  CudaRasterizer::Rasterizer::backwardSketchJacobian(
      sketch_mode, sketch_dim, sketch_indices, P, D, M, R, bg, W, H, means3D, shs, colors_precomp, scales,
      scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw, campos, tan_fovx, tan_fovy,
      radii, g_geom[slot].ptr, g_bin[slot].ptr, g_img[slot].ptr, 0, nullptr, 0, nullptr, nullptr, df_dpix,
      df_dpix_depth, df_dmean2D, df_dconic, df_dopacity, df_dcolor, df_ddepth, df_dmean3D, df_dcov3D, df_dsh,
      df_dscale, df_drot, df_dtau, false);
}

// Copies of the reference's internal binning state for the bit-exact checks (device-to-device).
void ref_get_binning(int slot, int R, uint64_t* keys_sorted, uint32_t* point_list) {
  char* p = g_bin[slot].ptr;
  CudaRasterizer::BinningState b = CudaRasterizer::BinningState::fromChunk(p, R);
  cudaMemcpy(keys_sorted, b.point_list_keys, (size_t)R * 8, cudaMemcpyDeviceToDevice);
  cudaMemcpy(point_list, b.point_list, (size_t)R * 4, cudaMemcpyDeviceToDevice);
}
void ref_get_image_state(int slot, int W, int H, uint32_t* ranges /*uint2[tiles]*/, int tiles, float* final_T,
                         uint32_t* n_contrib) {
  char* p = g_img[slot].ptr;
  CudaRasterizer::ImageState im = CudaRasterizer::ImageState::fromChunk(p, (size_t)W * H);
  cudaMemcpy(ranges, im.ranges, (size_t)tiles * 8, cudaMemcpyDeviceToDevice);
  cudaMemcpy(final_T, im.accum_alpha, (size_t)W * H * 4, cudaMemcpyDeviceToDevice);
  cudaMemcpy(n_contrib, im.n_contrib, (size_t)W * H * 4, cudaMemcpyDeviceToDevice);
}
void ref_get_geometry(int slot, int P, float* depths, float* means2D, float* conic_opacity, float* rgb) {
  char* p = g_geom[slot].ptr;
  CudaRasterizer::GeometryState g = CudaRasterizer::GeometryState::fromChunk(p, P);
  cudaMemcpy(depths, g.depths, (size_t)P * 4, cudaMemcpyDeviceToDevice);
  cudaMemcpy(means2D, g.means2D, (size_t)P * 8, cudaMemcpyDeviceToDevice);
  cudaMemcpy(conic_opacity, g.conic_opacity, (size_t)P * 16, cudaMemcpyDeviceToDevice);
  cudaMemcpy(rgb, g.rgb, (size_t)P * 12, cudaMemcpyDeviceToDevice);
}

void ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present) {
  CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, present);
}

void ref_knn_dist2(int P, float* points, float* mean_dists) { SimpleKNN::knn(P, (float3*)points, mean_dists); }

void ref_release(void) {
  for (int i = 0; i < kSlots; i++) {
    for (Arena* a : {&g_geom[i], &g_bin[i], &g_img[i]}) { if (a->ptr) cudaFree(a->ptr); a->ptr = nullptr; a->cap = 0; }
  }
}
}
