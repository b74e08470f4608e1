// Warp-cooperative staging of spherical-harmonics rows through shared memory.
//
// The per-Gaussian kernels give one thread one Gaussian.  With SH degree 3 a Gaussian's coefficients are 48 floats
// (192 B), so thread-private accesses are 4-byte loads / stores at a 192-byte lane stride: every instruction touches 32
// different sectors and each sector is fetched or written eight times (ncu at config C5: 16-20 sectors per request, L2
// traffic 4.6x the DRAM traffic, 22 % of the DRAM peak).  The rows of a warp's 32 consecutive Gaussians are, however,
// one contiguous block of memory.  The warp moves that block with 16-byte coalesced accesses between global memory and
// a padded [32][RS] tile in shared memory (RS odd: conflict-free row access), and each thread then works on its row.
//
// Two source layouts: one array [P, M, 3] (the rasterizer API, `rest == nullptr`), or the model's two parameters
// f_dc [P, 1, 3] and f_rest [P, M-1, 3] (gaussian_model.py:41-42) so that the mapping engine needs no concatenated
// copy.  A row in shared memory is always [dc(3) | rest(3(M-1))].
#pragma once
#include "common.cuh"

namespace mgs {

__host__ __device__ inline int sh_row_stride(const int M) { return (3 * M) | 1; }

// e -> (row, offset) for e < 2^16 and row_len < 2^10 with one multiply: row = floor(e / row_len)
struct FastDiv {
  uint32_t magic, d;
  __host__ __device__ explicit FastDiv(uint32_t d_) : magic((uint32_t)((0x100000000ull + d_ - 1) / d_)), d(d_) {}
  __device__ __forceinline__ uint32_t div(uint32_t e) const { return __umulhi(e, magic); }
};

// Fast path of both directions: 32 full rows and a 16-byte aligned block (32 * row_len floats is always a whole number
// of 16-byte chunks; a chunk may span two rows).  Batches of up to 6 chunks per lane keep 6 x 512 B per warp in flight:
// the per-Gaussian kernels run at ~25 % occupancy, so memory-level parallelism has to come from each warp.
constexpr int SH_BATCH = 6;

__device__ __forceinline__ bool sh_fast_path(const void* p, const int n_rows) {
  return n_rows == 32 && (reinterpret_cast<uintptr_t>(p) & 15u) == 0;
}

__device__ __forceinline__ void sh_stage_load_fast(float* __restrict__ tile, const int RS, const float* __restrict__ src,
                                                   const int row_len, const int dst_off, const int lane) {
  const int chunks = 8 * row_len;            // 32 rows * row_len / 4
  const FastDiv fd((uint32_t)row_len);
  const float4* src4 = reinterpret_cast<const float4*>(src);
  for (int c0 = lane; c0 < chunks; c0 += 32 * SH_BATCH) {
    float4 t[SH_BATCH];
#pragma unroll
    for (int u = 0; u < SH_BATCH; u++) {
      const int c = c0 + 32 * u;
      if (c < chunks) t[u] = __ldg(src4 + c);
    }
#pragma unroll
    for (int u = 0; u < SH_BATCH; u++) {
      const int c = c0 + 32 * u;
      if (c < chunks) {
        const int r = (int)fd.div((uint32_t)(4 * c)), k = 4 * c - r * row_len;
        float* q = tile + r * RS + dst_off + k;          // element i of the chunk: row r (+1 past the row end)
        const int wrap = RS - row_len;                   // jump from the end of row r to the start of row r+1
        const float val[4] = {t[u].x, t[u].y, t[u].z, t[u].w};
#pragma unroll
        for (int i = 0; i < 4; i++) q[i + (k + i >= row_len ? wrap : 0)] = val[i];
      }
    }
  }
}

__device__ __forceinline__ void sh_stage_store_fast(const float* __restrict__ tile, const int RS, float* __restrict__ dst,
                                                    const int row_len, const int src_off, const int lane, const int mode,
                                                    const unsigned row_mask) {
  const int chunks = 8 * row_len;
  const FastDiv fd((uint32_t)row_len);
  float4* dst4 = reinterpret_cast<float4*>(dst);
  const int wrap = RS - row_len;
  for (int c0 = lane; c0 < chunks; c0 += 32 * SH_BATCH) {
    float4 t[SH_BATCH];
    bool on[SH_BATCH];
    const float* q[SH_BATCH];
    int kk[SH_BATCH];
#pragma unroll
    for (int u = 0; u < SH_BATCH; u++) {
      const int c = c0 + 32 * u;
      on[u] = false;
      if (c < chunks) {
        const int r = (int)fd.div((uint32_t)(4 * c)), k = 4 * c - r * row_len;
        const int r_last = k + 3 >= row_len ? min(r + 1, 31) : r;
        on[u] = mode == 0 || (((row_mask >> r) | (row_mask >> r_last)) & 1u) != 0u;
        q[u] = tile + r * RS + src_off + k;
        kk[u] = k;
        if (on[u] && mode == 1) t[u] = dst4[c];          // all read-modify-write loads of the batch first
      }
    }
#pragma unroll
    for (int u = 0; u < SH_BATCH; u++) {
      const int c = c0 + 32 * u;
      if (!on[u]) continue;
      float val[4];
#pragma unroll
      for (int i = 0; i < 4; i++) val[i] = q[u][i + (kk[u] + i >= row_len ? wrap : 0)];
      const float4 v4 = make_float4(val[0], val[1], val[2], val[3]);
      if (mode == 0) dst4[c] = v4;
      else if (mode == 1) dst4[c] = make_float4(t[u].x + v4.x, t[u].y + v4.y, t[u].z + v4.z, t[u].w + v4.w);
      else atomicAdd(dst4 + c, v4);
    }
  }
}

// global block [n_rows][row_len] (contiguous) -> tile rows at column dst_off
__device__ __forceinline__ void sh_stage_load(float* __restrict__ tile, const int RS, const float* __restrict__ src,
                                              const int row_len, const int dst_off, const int n_rows, const int lane) {
  if (sh_fast_path(src, n_rows)) { sh_stage_load_fast(tile, RS, src, row_len, dst_off, lane); return; }
  const int total = n_rows * row_len;
  const FastDiv fd((uint32_t)row_len);
  if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
    for (int e = 4 * lane; e < total; e += 128) {
      float val[4];
      if (e + 3 < total) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(src + e));
        val[0] = t.x; val[1] = t.y; val[2] = t.z; val[3] = t.w;
      } else {
#pragma unroll
        for (int u = 0; u < 4; u++) val[u] = e + u < total ? __ldg(src + e + u) : 0.f;
      }
      int r = (int)fd.div((uint32_t)e), k = e - r * row_len;
#pragma unroll
      for (int u = 0; u < 4; u++) {
        if (e + u < total) tile[r * RS + dst_off + k] = val[u];
        if (++k == row_len) { k = 0; r++; }
      }
    }
  } else {
    for (int e = lane; e < total; e += 32) {
      const int r = (int)fd.div((uint32_t)e), k = e - r * row_len;
      tile[r * RS + dst_off + k] = __ldg(src + e);
    }
  }
}

// tile rows at column src_off -> global block.  mode 0: overwrite, 1: +=, 2: atomicAdd.  In modes 1 / 2 a chunk whose
// rows are all outside `row_mask` is skipped (their tile rows must be zero when a chunk can span two rows).
__device__ __forceinline__ void sh_stage_store(const float* __restrict__ tile, const int RS, float* __restrict__ dst,
                                               const int row_len, const int src_off, const int n_rows, const int lane,
                                               const int mode, const unsigned row_mask) {
  if (sh_fast_path(dst, n_rows)) { sh_stage_store_fast(tile, RS, dst, row_len, src_off, lane, mode, row_mask); return; }
  const int total = n_rows * row_len;
  const FastDiv fd((uint32_t)row_len);
  const bool vec = (reinterpret_cast<uintptr_t>(dst) & 15u) == 0;
  for (int e = 4 * lane; e < total; e += 128) {
    const int r0 = (int)fd.div((uint32_t)e);
    const int last = min(e + 3, total - 1);
    const int r1 = (int)fd.div((uint32_t)last);
    if (mode != 0 && ((row_mask >> r0) & 1u) == 0u && ((row_mask >> r1) & 1u) == 0u) continue;
    float val[4];
    int r = r0, k = e - r0 * row_len;
#pragma unroll
    for (int u = 0; u < 4; u++) {
      val[u] = e + u < total ? tile[r * RS + src_off + k] : 0.f;
      if (++k == row_len) { k = 0; r++; }
    }
    if (vec && e + 3 < total) {
      float4* q = reinterpret_cast<float4*>(dst + e);
      if (mode == 0) *q = make_float4(val[0], val[1], val[2], val[3]);
      else if (mode == 1) { float4 t = *q; t.x += val[0]; t.y += val[1]; t.z += val[2]; t.w += val[3]; *q = t; }
      else atomicAdd(q, make_float4(val[0], val[1], val[2], val[3]));
    } else {
#pragma unroll
      for (int u = 0; u < 4; u++) {
        if (e + u >= total) break;
        if (mode == 0) dst[e + u] = val[u];
        else if (mode == 1) dst[e + u] += val[u];
        else atomicAdd(dst + e + u, val[u]);
      }
    }
  }
}

}  // namespace mgs
