// markVisible, simple-knn distCUDA2 and the on-device pose update.
//   mark_visible : checkFrustum (rasterizer_impl.cu:54-66)
//   knn_dist2    : SimpleKNN::knn (submodules/simple-knn/simple_knn.cu:46-222) — exact 3-NN mean squared
//                  distance.  Same result set as the reference (exact search), different machinery: no host
//                  round trips for the bounding box, points gathered into Morton order once so the search
//                  streams contiguous float4s through shared memory, 128-point boxes, block-cooperative
//                  pruning with __syncthreads_or.
//   pose_update  : T <- Exp(tau) T, tau = (rho, theta) (utils/pose_utils.py:26-98, math.h:217-339)
#include "launchers.h"
#include "pose_update.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cfloat>

namespace mgs {

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
  const float3 pv = xform4x3(p, view);
  present[idx] = pv.z > 0.2f ? 1 : 0;
}

// ---- knn ------------------------------------------------------------------------------------------
constexpr int KNN_BOX = 128;

struct KnnLayout { size_t bbox, codes, codes_sorted, idx, idx_sorted, pts_sorted, boxes, sort_tmp, sort_tmp_bytes, total; };
static int knn_layout(int P, KnnLayout* L) {
  const size_t Pp = P > 0 ? P : 1;
  size_t off = 0;
  auto take = [&](size_t b) { size_t o = off; off = (off + b + 255) / 256 * 256; return o; };
  L->bbox = take(6 * sizeof(int));
  L->codes = take(Pp * 4); L->codes_sorted = take(Pp * 4); L->idx = take(Pp * 4); L->idx_sorted = take(Pp * 4);
  L->pts_sorted = take(Pp * 16);
  L->boxes = take(((Pp + KNN_BOX - 1) / KNN_BOX) * 32);
  size_t tb = 0;
  cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, tb, (uint32_t*)nullptr, (uint32_t*)nullptr, (uint32_t*)nullptr,
                                                  (uint32_t*)nullptr, (int)Pp);
  if (e != cudaSuccess) { set_error("knn: CUB size query failed: %s", cudaGetErrorString(e)); return MGS_ECUDA; }
  L->sort_tmp_bytes = tb; L->sort_tmp = take(tb);
  L->total = off;
  return MGS_OK;
}

// order-preserving float <-> int so the bounding box can be reduced with integer atomics
__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void knn_bbox_init_kernel(int* bbox) {  // both seeded with the origin (simple_knn.cu:192, Q16)
  if (threadIdx.x < 6) bbox[threadIdx.x] = f2ord(0.f);
}
__global__ void __launch_bounds__(256) knn_bbox_kernel(int P, const float* __restrict__ pts, int* __restrict__ bbox) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  if (idx < P) {
#pragma unroll
    for (int c = 0; c < 3; c++) lo[c] = hi[c] = pts[3 * idx + c];
  }
#pragma unroll
  for (int c = 0; c < 3; c++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
      hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int c = 0; c < 3; c++) { atomicMin(&bbox[c], f2ord(lo[c])); atomicMax(&bbox[3 + c], f2ord(hi[c])); }
  }
}

__device__ __forceinline__ uint32_t spread10(uint32_t x) {
  x = (x | (x << 16)) & 0x030000FF;
  x = (x | (x << 8)) & 0x0300F00F;
  x = (x | (x << 4)) & 0x030C30C3;
  x = (x | (x << 2)) & 0x09249249;
  return x;
}
__global__ void __launch_bounds__(256)
knn_morton_kernel(int P, const float* __restrict__ pts, const int* __restrict__ bbox, uint32_t* __restrict__ codes,
                  uint32_t* __restrict__ idxs) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  uint32_t code = 0;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const float lo = ord2f(bbox[c]), hi = ord2f(bbox[3 + c]);
    const uint32_t q = (uint32_t)(((pts[3 * idx + c] - lo) / (hi - lo)) * ((1 << 10) - 1));
    code |= spread10(q) << c;
  }
  codes[idx] = code;
  idxs[idx] = idx;
}

// gather into Morton order + per-box AABB (one warp-shuffle reduction per 128-point box)
__global__ void __launch_bounds__(KNN_BOX)
knn_gather_boxes_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ idx_sorted,
                        float4* __restrict__ pts_sorted, float4* __restrict__ boxes /* 2 float4 per box */) {
  const int i = blockIdx.x * KNN_BOX + threadIdx.x;
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  if (i < P) {
    const uint32_t src = idx_sorted[i];
    const float x = pts[3 * src], y = pts[3 * src + 1], z = pts[3 * src + 2];
    pts_sorted[i] = make_float4(x, y, z, __uint_as_float(src));
    lo[0] = hi[0] = x; lo[1] = hi[1] = y; lo[2] = hi[2] = z;
  }
  __shared__ float slo[KNN_BOX / 32][3], shi[KNN_BOX / 32][3];
#pragma unroll
  for (int c = 0; c < 3; c++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
      hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
    }
    if ((threadIdx.x & 31) == 0) { slo[threadIdx.x >> 5][c] = lo[c]; shi[threadIdx.x >> 5][c] = hi[c]; }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < KNN_BOX / 32; w++)
      for (int c = 0; c < 3; c++) { slo[0][c] = fminf(slo[0][c], slo[w][c]); shi[0][c] = fmaxf(shi[0][c], shi[w][c]); }
    boxes[2 * blockIdx.x] = make_float4(slo[0][0], slo[0][1], slo[0][2], 0.f);
    boxes[2 * blockIdx.x + 1] = make_float4(shi[0][0], shi[0][1], shi[0][2], 0.f);
  }
}

__device__ __forceinline__ void keep3(float d, float (&best)[3]) {  // insertion into an ascending triple
#pragma unroll
  for (int j = 0; j < 3; j++)
    if (best[j] > d) { const float t = best[j]; best[j] = d; d = t; }
}
__device__ __forceinline__ float dist2(const float4& q, const float4& o) {
  const float dx = o.x - q.x, dy = o.y - q.y, dz = o.z - q.z;
  return dx * dx + dy * dy + dz * dz;   // same expression as updateKBest (simple_knn.cu:132-134)
}
__device__ __forceinline__ float box_dist2(const float4& lo, const float4& hi, const float4& p) {
  float dx = 0.f, dy = 0.f, dz = 0.f;
  if (p.x < lo.x || p.x > hi.x) dx = fminf(fabsf(p.x - lo.x), fabsf(p.x - hi.x));
  if (p.y < lo.y || p.y > hi.y) dy = fminf(fabsf(p.y - lo.y), fabsf(p.y - hi.y));
  if (p.z < lo.z || p.z > hi.z) dz = fminf(fabsf(p.z - lo.z), fabsf(p.z - hi.z));
  return dx * dx + dy * dy + dz * dz;
}

__global__ void __launch_bounds__(KNN_BOX)
knn_search_kernel(int P, const float4* __restrict__ pts_sorted, const float4* __restrict__ boxes, int nboxes,
                  float* __restrict__ out) {
  __shared__ float4 tile[KNN_BOX];
  const int i = blockIdx.x * KNN_BOX + threadIdx.x;
  const bool live = i < P;
  const float4 q = live ? pts_sorted[i] : make_float4(0.f, 0.f, 0.f, 0.f);
  float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
  // own box first: it almost always holds the neighbours and gives a tight rejection radius
  for (int pass = 0; pass < 2; pass++) {
    for (int b = (pass == 0 ? blockIdx.x : 0); b < (pass == 0 ? (int)blockIdx.x + 1 : nboxes); b++) {
      if (pass == 1 && b == (int)blockIdx.x) continue;
      bool need = false;
      if (live) need = !(box_dist2(boxes[2 * b], boxes[2 * b + 1], q) > best[2]);
      if (!__syncthreads_or(need)) continue;
      const int j0 = b * KNN_BOX;
      if (j0 + (int)threadIdx.x < P) tile[threadIdx.x] = pts_sorted[j0 + threadIdx.x];
      __syncthreads();
      if (need) {
        const int cnt = min(KNN_BOX, P - j0);
        for (int j = 0; j < cnt; j++) {
          if (j0 + j == i) continue;
          keep3(dist2(q, tile[j]), best);
        }
      }
      __syncthreads();
    }
  }
  if (live) out[__float_as_uint(q.w)] = (best[0] + best[1] + best[2]) / 3.0f;
}

// ---- pose update ---------------------------------------------------------------------------------
__global__ void pose_update_kernel(float* __restrict__ T, float* __restrict__ rho, float* __restrict__ theta,
                                   const float* __restrict__ proj_raw, float* __restrict__ view_out,
                                   float* __restrict__ proj_out, int32_t* __restrict__ converged, float thr) {
  if (threadIdx.x == 0) pose_update_dev(T, rho, theta, proj_raw, view_out, proj_out, converged, thr);
}

}  // namespace mgs

using namespace mgs;

extern "C" {

int mgs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present,
                     void* stream) {
  (void)projmatrix;  // the reference computes p_proj and never uses it (auxiliary.h:148-154)
  if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) { set_error("mark_visible: bad arguments"); return MGS_EINVAL; }
  if (P == 0) return MGS_OK;
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, means3D, viewmatrix, present);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

size_t mgs_knn_workspace_bytes(int32_t P) {
  KnnLayout L;
  if (knn_layout(P, &L) != MGS_OK) return 0;
  return L.total;
}

int mgs_knn_dist2(int32_t P, const float* points, float* mean_dist2, void* workspace, size_t workspace_bytes, void* stream) {
  if (P < 0 || (P > 0 && (!points || !mean_dist2 || !workspace))) { set_error("knn: bad arguments"); return MGS_EINVAL; }
  if (P == 0) return MGS_OK;
  KnnLayout L;
  int rc = knn_layout(P, &L);
  if (rc) return rc;
  if (workspace_bytes < L.total) { set_error("knn: workspace too small: %zu < %zu", workspace_bytes, L.total); return MGS_ENOSPC; }
  cudaStream_t s = (cudaStream_t)stream;
  int* bbox = ws_ptr<int>(workspace, L.bbox);
  uint32_t *codes = ws_ptr<uint32_t>(workspace, L.codes), *codes_s = ws_ptr<uint32_t>(workspace, L.codes_sorted);
  uint32_t *idx = ws_ptr<uint32_t>(workspace, L.idx), *idx_s = ws_ptr<uint32_t>(workspace, L.idx_sorted);
  float4* pts_s = ws_ptr<float4>(workspace, L.pts_sorted);
  float4* boxes = ws_ptr<float4>(workspace, L.boxes);
  const int nb256 = (P + 255) / 256, nboxes = (P + KNN_BOX - 1) / KNN_BOX;
  knn_bbox_init_kernel<<<1, 32, 0, s>>>(bbox);
  knn_bbox_kernel<<<nb256, 256, 0, s>>>(P, points, bbox);
  knn_morton_kernel<<<nb256, 256, 0, s>>>(P, points, bbox, codes, idx);
  MGS_LAUNCH_CHECK();
  size_t tb = L.sort_tmp_bytes;
  MGS_CUDA_TRY(cub::DeviceRadixSort::SortPairs(ws_ptr<void>(workspace, L.sort_tmp), tb, codes, codes_s, idx, idx_s, P, 0, 30, s));
  knn_gather_boxes_kernel<<<nboxes, KNN_BOX, 0, s>>>(P, points, idx_s, pts_s, boxes);
  knn_search_kernel<<<nboxes, KNN_BOX, 0, s>>>(P, pts_s, boxes, nboxes, mean_dist2);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

int mgs_pose_update(float* T_w2c, float* rho, float* theta, const float* proj_raw, float* viewmatrix_out,
                    float* projmatrix_out, int32_t* converged_out, float converged_threshold, void* stream) {
  if (!T_w2c || !rho || !theta) { set_error("pose_update: null argument"); return MGS_EINVAL; }
  MGS_PROF("pose_update", (cudaStream_t)stream, pose_update_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(
      T_w2c, rho, theta, proj_raw, viewmatrix_out, projmatrix_out, converged_out, converged_threshold));
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

}  // extern "C"
