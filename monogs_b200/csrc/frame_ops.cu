// Per-frame image preprocessing and covisibility bookkeeping around the tracking path.
//   * Scharr gradient intensity of the gray image with the 3x3 validity mask  (utils/slam_utils.py:7-41,
//     utils/camera_utils.py:110-118): one kernel instead of two pads, four grouped convolutions and six elementwise ops
//   * exact (lower) median by radix select on the float bits, optionally over a validity predicate
//     (camera_utils.py:131 `img_grad_intensity.median()`, slam_utils.py:286-297 get_median_depth) — no sort, no host sync
//   * Camera.grad_mask / rgb_pixel_mask / rgb_pixel_mask_mapping (camera_utils.py:120-142), global-median and
//     32x32-patch-median (Replica) variants
//   * visibility bit masks and their intersection / union counts for keyframe selection and window pruning
//     (utils/slam_frontend.py:1713-1741, 1918-1932): 1 bit per Gaussian instead of the reference's int64 masks
#include "launchers.h"
#include <cub/block/block_radix_sort.cuh>

namespace mgs {

__device__ __forceinline__ int reflect_idx(int i, int n) { return i < 0 ? -i : (i >= n ? 2 * n - 2 - i : i); }   // pad mode="reflect", 1 px

// gray = mean over channels; Scharr v / h (normalised by 32 = sum of the absolute weights), masked where the 3x3 neighbourhood of |gray| > eps is not
// full; intensity = sqrt(gv^2 + gh^2)
__global__ void __launch_bounds__(256)
grad_intensity_kernel(const float* __restrict__ image, const int C, const int H, const int W, const float eps,
                      float* __restrict__ intensity) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= W || y >= H) return;
  const size_t plane = (size_t)H * W;
  float g[3][3];
  bool full = true;
#pragma unroll
  for (int dy = -1; dy <= 1; dy++)
#pragma unroll
    for (int dx = -1; dx <= 1; dx++) {
      const size_t p = (size_t)reflect_idx(y + dy, H) * W + reflect_idx(x + dx, W);
      float s = 0.f;
      for (int c = 0; c < C; c++) s += image[c * plane + p];
      const float v = s / (float)C;
      g[dy + 1][dx + 1] = v;
      full = full && (fabsf(v) > eps);
    }
  const float nrm = 1.0f / 32.0f;   // 1 / sum|conv_y| (slam_utils.py:16)
  // conv_x = [[3,10,3],[0,0,0],[-3,-10,-3]] -> "v";  conv_y = [[3,0,-3],[10,0,-10],[3,0,-3]] -> "h"  (cross-correlation)
  const float gv = nrm * (3.f * g[0][0] + 10.f * g[0][1] + 3.f * g[0][2] - 3.f * g[2][0] - 10.f * g[2][1] - 3.f * g[2][2]);
  const float gh = nrm * (3.f * g[0][0] - 3.f * g[0][2] + 10.f * g[1][0] - 10.f * g[1][2] + 3.f * g[2][0] - 3.f * g[2][2]);
  intensity[(size_t)y * W + x] = full ? sqrtf(gv * gv + gh * gh) : 0.f;
}

// ---- radix select: k-th smallest of the valid elements, k = (n_valid - 1) / 2 (torch.median = lower median) ----
__device__ __forceinline__ uint32_t float_order_bits(float f) {
  const uint32_t b = __float_as_uint(f);
  return b ^ ((b >> 31) ? 0xFFFFFFFFu : 0x80000000u);
}
__device__ __forceinline__ float order_bits_float(uint32_t u) {
  return __uint_as_float(u ^ ((u >> 31) ? 0x80000000u : 0xFFFFFFFFu));
}
// predicate of get_median_depth (slam_utils.py:286-297): depth > 0 [and opacity > 0.95] [and mask]
struct SelPred { const float* opacity; const float* mask; int depth_rule; };
__device__ __forceinline__ bool sel_valid(const SelPred& p, const float v, const size_t i) {
  if (!p.depth_rule) return v == v;     // plain median: every non-NaN element
  bool ok = v > 0.f;
  if (p.opacity) ok = ok && p.opacity[i] > 0.95f;
  if (p.mask) ok = ok && p.mask[i] != 0.f;
  return ok;
}
struct SelState { uint32_t prefix, k, n_valid, pad; uint32_t hist[256]; };

__global__ void __launch_bounds__(256)
select_hist_kernel(const float* __restrict__ x, const size_t n, const SelPred pred, const int pass, SelState* __restrict__ st) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0u;
  __syncthreads();
  const int shift = 24 - 8 * pass;
  const uint32_t prefix = st->prefix;
  const uint32_t hi_mask = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + 8));
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float v = x[i];
    if (!sel_valid(pred, v, i)) continue;
    const uint32_t u = float_order_bits(v);
    if ((u & hi_mask) == (prefix & hi_mask)) atomicAdd(&h[(u >> shift) & 255u], 1u);
  }
  __syncthreads();
  if (h[threadIdx.x]) atomicAdd(&st->hist[threadIdx.x], h[threadIdx.x]);
}

__global__ void select_pick_kernel(const int pass, SelState* __restrict__ st, float* __restrict__ out, int32_t* __restrict__ n_valid_out) {
  if (pass == 0) {
    uint32_t n = 0;
    for (int b = 0; b < 256; b++) n += st->hist[b];
    st->n_valid = n;
    st->k = n ? (n - 1u) / 2u : 0u;
    if (n_valid_out) n_valid_out[0] = (int32_t)n;
  }
  uint32_t k = st->k, acc = 0;
  int b = 0;
  for (; b < 255; b++) {
    if (acc + st->hist[b] > k) break;
    acc += st->hist[b];
  }
  st->k = k - acc;
  st->prefix |= (uint32_t)b << (24 - 8 * pass);
  for (int i = 0; i < 256; i++) st->hist[i] = 0u;
  if (pass == 3) out[0] = st->n_valid ? order_bits_float(st->prefix) : __int_as_float(0x7fc00000);   // empty -> NaN like torch
}

// grad_mask = intensity > median * edge_threshold;  rgb masks (camera_utils.py:131-142)
__global__ void __launch_bounds__(256)
frame_masks_kernel(const float* __restrict__ image, const float* __restrict__ intensity, const float* __restrict__ median,
                   const int H, const int W, const float edge_threshold, const float rgb_thr,
                   float* __restrict__ grad_mask, float* __restrict__ rgb_pixel_mask, float* __restrict__ rgb_pixel_mask_mapping) {
  const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x, plane = (size_t)H * W;
  if (p >= plane) return;
  const float gm = (median == nullptr) ? grad_mask[p] : (intensity[p] > median[0] * edge_threshold ? 1.f : 0.f);
  const float rm = (image[p] + image[plane + p] + image[2 * plane + p]) > rgb_thr ? 1.f : 0.f;
  if (median != nullptr) grad_mask[p] = gm;
  if (rgb_pixel_mask) rgb_pixel_mask[p] = rm * gm;
  if (rgb_pixel_mask_mapping) rgb_pixel_mask_mapping[p] = rm;
}

// Replica variant: per 32x32 patch, mask = intensity > median(patch) * multiplier; pixels outside whole patches get 0
// (F.unfold / F.fold with stride = size drop the remainder, camera_utils.py:120-129)
__global__ void __launch_bounds__(256)
patch_median_mask_kernel(const float* __restrict__ intensity, const int H, const int W, const float multiplier,
                         float* __restrict__ grad_mask) {
  using Sort = cub::BlockRadixSort<float, 256, 4>;
  __shared__ typename Sort::TempStorage tmp;
  __shared__ float s_med;
  const int px0 = blockIdx.x * 32, py0 = blockIdx.y * 32;
  float v[4], key[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int e = threadIdx.x * 4 + i;             // blocked arrangement over the 1024 pixels of the patch
    v[i] = intensity[(size_t)(py0 + (e >> 5)) * W + px0 + (e & 31)];
    key[i] = v[i];
  }
  Sort(tmp).Sort(key);
  if (threadIdx.x == 127) s_med = key[3];          // element 511 = lower median of 1024
  __syncthreads();
  const float thr = s_med * multiplier;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int e = threadIdx.x * 4 + i;
    grad_mask[(size_t)(py0 + (e >> 5)) * W + px0 + (e & 31)] = v[i] > thr ? 1.f : 0.f;
  }
}

// ---- covisibility ----
__global__ void __launch_bounds__(256)
visibility_pack_kernel(const int P, const int32_t* __restrict__ n_touched, uint32_t* __restrict__ bits) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned b = __ballot_sync(0xffffffffu, idx < P && n_touched[idx] > 0);
  if ((threadIdx.x & 31) == 0 && idx < P) bits[idx >> 5] = b;
}

// out[k] = (|cur & kf_k|, |cur | kf_k|, |kf_k|), out[K] = (|cur|, 0, 0)
__global__ void __launch_bounds__(256)
covisibility_kernel(const int words, const uint32_t* __restrict__ cur, const int K, const uint32_t* __restrict__ kf,
                    int32_t* __restrict__ out) {
  const int k = blockIdx.y;
  int inter = 0, uni = 0, cnt = 0;
  for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < words; w += gridDim.x * blockDim.x) {
    const uint32_t a = cur[w];
    if (k < K) { const uint32_t b = kf[(size_t)k * words + w]; inter += __popc(a & b); uni += __popc(a | b); cnt += __popc(b); }
    else inter += __popc(a);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    inter += __shfl_xor_sync(0xffffffffu, inter, o); uni += __shfl_xor_sync(0xffffffffu, uni, o); cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (inter) atomicAdd(&out[3 * k], inter);
    if (uni) atomicAdd(&out[3 * k + 1], uni);
    if (cnt) atomicAdd(&out[3 * k + 2], cnt);
  }
}

}  // namespace mgs

using namespace mgs;

static int run_select(const float* x, size_t n, const SelPred& pred, void* scratch, size_t scratch_bytes, float* out,
                      int32_t* n_valid_out, cudaStream_t s) {
  if (scratch_bytes < sizeof(SelState)) { set_error("median: scratch too small"); return MGS_ENOSPC; }
  SelState* st = reinterpret_cast<SelState*>(scratch);
  MGS_CUDA_TRY(cudaMemsetAsync(st, 0, sizeof(SelState), s));
  const int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  for (int pass = 0; pass < 4; pass++) {
    select_hist_kernel<<<blocks > 0 ? blocks : 1, 256, 0, s>>>(x, n, pred, pass, st);
    MGS_LAUNCH_CHECK();
    select_pick_kernel<<<1, 1, 0, s>>>(pass, st, out, n_valid_out);
    MGS_LAUNCH_CHECK();
  }
  return MGS_OK;
}

extern "C" {

size_t mgs_median_scratch_bytes(void) { return (sizeof(SelState) + 255) / 256 * 256; }

int mgs_median(const float* x, int64_t n, float* out, void* scratch, size_t scratch_bytes, void* stream) {
  if (!x || !out || !scratch || n < 0) { set_error("median: bad arguments"); return MGS_EINVAL; }
  SelPred p{nullptr, nullptr, 0};
  return run_select(x, (size_t)n, p, scratch, scratch_bytes, out, nullptr, (cudaStream_t)stream);
}

int mgs_median_depth(const float* depth, const float* opacity, const float* mask, int64_t n, float* out, int32_t* n_valid_out,
                     void* scratch, size_t scratch_bytes, void* stream) {
  if (!depth || !out || !scratch || n < 0) { set_error("median_depth: bad arguments"); return MGS_EINVAL; }
  SelPred p{opacity, mask, 1};
  return run_select(depth, (size_t)n, p, scratch, scratch_bytes, out, n_valid_out, (cudaStream_t)stream);
}

int mgs_frame_grad_intensity(const float* image, int32_t C, int32_t H, int32_t W, float eps, float* intensity, void* stream) {
  if (!image || !intensity || C <= 0 || H < 2 || W < 2) { set_error("grad_intensity: bad arguments"); return MGS_EINVAL; }
  dim3 grid((W + 31) / 32, (H + 7) / 8);
  grad_intensity_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(image, C, H, W, eps, intensity);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

int mgs_frame_masks(const float* image, int32_t H, int32_t W, float edge_threshold, float rgb_boundary_threshold,
                    int32_t patch_median, float* intensity_scratch, float* grad_mask, float* rgb_pixel_mask,
                    float* rgb_pixel_mask_mapping, void* scratch, size_t scratch_bytes, void* stream) {
  if (!image || !intensity_scratch || !grad_mask || !scratch || H < 2 || W < 2) { set_error("frame_masks: bad arguments"); return MGS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  int rc = mgs_frame_grad_intensity(image, 3, H, W, 0.01f, intensity_scratch, stream);
  if (rc) return rc;
  const size_t npix = (size_t)H * W;
  const float* median = nullptr;
  if (patch_median) {
    MGS_CUDA_TRY(cudaMemsetAsync(grad_mask, 0, npix * sizeof(float), s));
    if (H >= 32 && W >= 32) {
      patch_median_mask_kernel<<<dim3(W / 32, H / 32), 256, 0, s>>>(intensity_scratch, H, W, edge_threshold, grad_mask);
      MGS_LAUNCH_CHECK();
    }
  } else {
    if (scratch_bytes < mgs_median_scratch_bytes() + 256) { set_error("frame_masks: scratch too small"); return MGS_ENOSPC; }
    float* med = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + mgs_median_scratch_bytes());
    SelPred p{nullptr, nullptr, 0};
    if ((rc = run_select(intensity_scratch, npix, p, scratch, scratch_bytes, med, nullptr, s))) return rc;
    median = med;
  }
  frame_masks_kernel<<<(unsigned)((npix + 255) / 256), 256, 0, s>>>(image, intensity_scratch, median, H, W, edge_threshold,
                                                                     rgb_boundary_threshold, grad_mask, rgb_pixel_mask,
                                                                     rgb_pixel_mask_mapping);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

int mgs_visibility_pack(int32_t P, const int32_t* n_touched, uint32_t* bits, void* stream) {
  if (P < 0 || (P > 0 && (!n_touched || !bits))) { set_error("visibility_pack: bad arguments"); return MGS_EINVAL; }
  if (P == 0) return MGS_OK;
  visibility_pack_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, n_touched, bits);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

int mgs_covisibility(int32_t P, const uint32_t* cur_bits, int32_t K, const uint32_t* kf_bits, int32_t* out, void* stream) {
  if (P < 0 || K < 0 || !out || (P > 0 && (!cur_bits || (K > 0 && !kf_bits)))) { set_error("covisibility: bad arguments"); return MGS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  MGS_CUDA_TRY(cudaMemsetAsync(out, 0, (size_t)(K + 1) * 3 * sizeof(int32_t), s));
  if (P == 0) return MGS_OK;
  const int words = (P + 31) / 32;
  const int bx = (words + 255) / 256 < 64 ? (words + 255) / 256 : 64;
  covisibility_kernel<<<dim3(bx, K + 1), 256, 0, s>>>(words, cur_bits, K, kf_bits, out);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

}  // extern "C"
