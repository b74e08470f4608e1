// Forward pipeline of the rasterizer: per-Gaussian preprocess, scan, key generation, sort, tile ranges +
// packed records, alpha blend.  Replaces CudaRasterizer::Rasterizer::forward (rasterizer_impl.cu:198-344),
// FORWARD::preprocess / FORWARD::render (forward.cu:157-258, 263-392).
#include "launchers.h"
#include "pose_jac.cuh"
#include "sh_stage.cuh"
#include "bitonic_sort.cuh"
#include <cub/block/block_scan.cuh>
#include <cub/block/block_merge_sort.cuh>
#include <cstdarg>
#include <cstdio>
#include <map>
#include <string>
#include <vector>

namespace mgs {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}
const char* get_error() { return g_err; }

// ---- kernel timer ----
struct ProfRec { const char* name; cudaEvent_t e0, e1; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof;
static ProfRec g_cur;
void prof_begin(const char* name, cudaStream_t s) {
  if (!g_prof_on) return;
  g_cur.name = name;
  cudaEventCreate(&g_cur.e0); cudaEventCreate(&g_cur.e1);
  cudaEventRecord(g_cur.e0, s);
}
void prof_end(cudaStream_t s) {
  if (!g_prof_on) return;
  cudaEventRecord(g_cur.e1, s);
  g_prof.push_back(g_cur);
}
thread_local int g_pdl = 0;
bool pdl_allowed() {
  static const bool env_on = [] { const char* e = getenv("MGS_PDL"); return !(e && e[0] == '0'); }();
  return env_on && !g_prof_on;    // the profiler's event records sit between the kernels: nothing to chain
}
void prof_enable(bool on) {
  for (auto& r : g_prof) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  g_prof.clear();
  g_prof_on = on;
}
int prof_report(char* buf, size_t n) {
  cudaDeviceSynchronize();
  std::map<std::string, std::pair<int, double>> agg;
  for (auto& r : g_prof) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) { auto& a = agg[r.name]; a.first++; a.second += ms; }
  }
  size_t off = 0;
  for (auto& kv : agg) {
    int w = snprintf(buf + off, off < n ? n - off : 0, "%s %d %.6f\n", kv.first.c_str(), kv.second.first, kv.second.second);
    if (w < 0 || off + (size_t)w >= n) break;
    off += (size_t)w;
  }
  if (n) buf[off < n ? off : n - 1] = 0;
  return (int)agg.size();
}

static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

int ws_layout(int32_t P, int32_t W, int32_t H, int64_t R_cap, WsLayout* L) {
  if (P < 0 || W <= 0 || H <= 0 || R_cap < 0) { set_error("ws_layout: bad sizes"); return MGS_EINVAL; }
  const size_t Pp = (size_t)(P > 0 ? P : 1), Rp = (size_t)(R_cap > 0 ? R_cap : 1);
  const size_t npix = (size_t)W * H;
  const size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes); return o; };
  // ---- R_cap-independent part (stage 1 and backward use only this) ----
  L->header = take(sizeof(WsHeader));
  L->depths = take(Pp * 4);
  L->means2D = take(Pp * 8);
  L->conic_opacity = take(Pp * 16);
  L->rgb4 = take(Pp * 16);
  L->tiles_touched = take(Pp * 4);
  L->tile_ctr = take(tiles * TILE_CTR_STRIDE * 4);
  L->final_T = take(npix * 4);
  L->n_contrib = take(npix * 4);
  L->ranges = take(tiles * 8);
  // ---- per-instance part ----
  L->records = take(Rp * sizeof(SplatRec));
  L->list = take(Rp * 8);
  L->total = off;
  return MGS_OK;
}

// ------------------------------------------------------------------------------------------------
// K1: per-Gaussian preprocess.  One thread per Gaussian; identical arithmetic order to forward.cu:157-258
// so that radii, tile rects and depth bits match the reference bit for bit.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cov3d_from_scale_rot(const float* __restrict__ scales, const float* __restrict__ rots,
                                                     int idx, float mod, float* cov6) {
  // forward.cu:120-154 — quaternion is used as given (not normalised), order (r,x,y,z) = q[0..3]
  const float sx = mod * scales[3 * idx + 0], sy = mod * scales[3 * idx + 1], sz = mod * scales[3 * idx + 2];
  const float4 q = reinterpret_cast<const float4*>(rots)[idx];
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  M3 S = m3_cols(sx, 0.f, 0.f, 0.f, sy, 0.f, 0.f, 0.f, sz);
  M3 R = m3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                 2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                 2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
  M3 Mm = m3_mul(S, R);
  M3 Sg = m3_mul(m3_t(Mm), Mm);
  cov6[0] = Sg[0][0]; cov6[1] = Sg[0][1]; cov6[2] = Sg[0][2];
  cov6[3] = Sg[1][1]; cov6[4] = Sg[1][2]; cov6[5] = Sg[2][2];
}

struct Cov2DOut { float a, b, c; };
// forward.cu:76-115 — EWA projection of the 3D covariance, FoV clamp 1.3x, +0.3 low-pass
__device__ __forceinline__ Cov2DOut cov2d_project(const float3& mean, const ViewConst& v, const float* cov6) {
  float3 t = xform4x3(mean, v.view);
  const float limx = 1.3f * v.tanfovx, limy = 1.3f * v.tanfovy;
  const float txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = min(limx, max(-limx, txtz)) * t.z;
  t.y = min(limy, max(-limy, tytz)) * t.z;
  M3 J = m3_cols(v.focal_x / t.z, 0.0f, -(v.focal_x * t.x) / (t.z * t.z),
                 0.0f, v.focal_y / t.z, -(v.focal_y * t.y) / (t.z * t.z),
                 0.f, 0.f, 0.f);
  M3 Wm = m3_cols(v.view[0], v.view[4], v.view[8], v.view[1], v.view[5], v.view[9],
                  v.view[2], v.view[6], v.view[10]);
  M3 T = m3_mul(Wm, J);
  M3 Vrk = m3_cols(cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]);
  M3 cov = m3_mul(m3_mul(m3_t(T), m3_t(Vrk)), T);
  Cov2DOut o;
  o.a = cov[0][0] + 0.3f; o.b = cov[0][1]; o.c = cov[1][1] + 0.3f;
  return o;
}

#ifndef MGS_SH_INLINE
#define MGS_SH_INLINE 1      // 0: the weight-array form (64 B of stack per thread)
#endif
#ifndef MGS_COV6_REGS
#define MGS_COV6_REGS 1
#endif
// SH -> RGB (forward.cu:22-73).  Returns the clamped colour and a 3-bit mask of channels clamped at 0.
__device__ __forceinline__ float3 sh_to_rgb(int deg, const float3& pos, const float* __restrict__ campos,
                                            const float* __restrict__ sh /* this Gaussian's row [M,3] */, uint32_t& clamp_mask) {
#if MGS_SH_INLINE
  // Each basis weight is folded into the colour as soon as it is known — c += w_k * sh[k], k = 0, 1, 2, ... in that order,
  // the same multiply-adds as a loop over a weight array — so that no array of 16 weights exists (indexed dynamically it
  // lived in local memory, fully unrolled it cost 12 registers and a resident CTA per SM).
  float c[3];
#pragma unroll
  for (int ch = 0; ch < 3; ch++) c[ch] = kSH0 * sh[ch];
  auto acc = [&](const int k, const float wk) {
#pragma unroll
    for (int ch = 0; ch < 3; ch++) c[ch] = fmaf(wk, sh[3 * k + ch], c[ch]);
  };
  if (deg > 0) {
    float dx = pos.x - campos[0], dy = pos.y - campos[1], dz = pos.z - campos[2];
    const float len = sqrtf(dx * dx + dy * dy + dz * dz);
    const float x = dx / len, y = dy / len, z = dz / len;
    acc(1, -(kSH1 * y)); acc(2, kSH1 * z); acc(3, -(kSH1 * x));
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      acc(4, kSH2[0] * xy); acc(5, kSH2[1] * yz); acc(6, kSH2[2] * (2.0f * zz - xx - yy));
      acc(7, kSH2[3] * xz); acc(8, kSH2[4] * (xx - yy));
      if (deg > 2) {
        acc(9, kSH3[0] * y * (3.0f * xx - yy));
        acc(10, kSH3[1] * xy * z);
        acc(11, kSH3[2] * y * (4.0f * zz - xx - yy));
        acc(12, kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy));
        acc(13, kSH3[4] * x * (4.0f * zz - xx - yy));
        acc(14, kSH3[5] * z * (xx - yy));
        acc(15, kSH3[6] * x * (xx - 3.0f * yy));
      }
    }
  }
#else
  float w[16];
  w[0] = kSH0;
  if (deg > 0) {
    float dx = pos.x - campos[0], dy = pos.y - campos[1], dz = pos.z - campos[2];
    const float len = sqrtf(dx * dx + dy * dy + dz * dz);
    const float x = dx / len, y = dy / len, z = dz / len;
    w[1] = -(kSH1 * y); w[2] = kSH1 * z; w[3] = -(kSH1 * x);
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      w[4] = kSH2[0] * xy; w[5] = kSH2[1] * yz; w[6] = kSH2[2] * (2.0f * zz - xx - yy);
      w[7] = kSH2[3] * xz; w[8] = kSH2[4] * (xx - yy);
      if (deg > 2) {
        w[9] = kSH3[0] * y * (3.0f * xx - yy);
        w[10] = kSH3[1] * xy * z;
        w[11] = kSH3[2] * y * (4.0f * zz - xx - yy);
        w[12] = kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
        w[13] = kSH3[4] * x * (4.0f * zz - xx - yy);
        w[14] = kSH3[5] * z * (xx - yy);
        w[15] = kSH3[6] * x * (xx - 3.0f * yy);
      }
    }
  }
  const int n = (deg + 1) * (deg + 1);
  float c[3];
#pragma unroll
  for (int ch = 0; ch < 3; ch++) c[ch] = w[0] * sh[ch];
  for (int k = 1; k < n; k++) {
#pragma unroll
    for (int ch = 0; ch < 3; ch++) c[ch] = fmaf(w[k], sh[3 * k + ch], c[ch]);
  }
#endif
  clamp_mask = 0;
#pragma unroll
  for (int ch = 0; ch < 3; ch++) {
    c[ch] += 0.5f;
    if (c[ch] < 0.f) clamp_mask |= (1u << ch);
    c[ch] = fmaxf(c[ch], 0.f);
  }
  return make_float3(c[0], c[1], c[2]);
}

// Visit every tile of a splat's rect, four tiles per callback so that the callback can keep four atomics in flight
// (a chain of dependent atomics is latency-bound: ~0.5 us each).  Rects of up to 32 tiles are walked by the owning
// lane; larger ones are broadcast (with the owner's 64-bit payload) and walked by the whole warp, lane-strided, so one
// screen-filling splat does not serialise a warp.  Must be reached by all 32 lanes.
// fn(const uint32_t (&tile)[4], int count, payload of the owning Gaussian).
template <typename F>
__device__ __forceinline__ void for_each_tile(const uint2 rmin, const uint2 rmax, const int gx,
                                              const unsigned long long payload, F&& fn) {
  const int lane = threadIdx.x & 31;
  const uint32_t w = rmax.x - rmin.x, area = w * (rmax.y - rmin.y);
  const bool big = area > 32u;
  if (area != 0u && !big) {
    uint32_t x = rmin.x, y = rmin.y;
    for (uint32_t t = 0; t < area; t += 4u) {
      uint32_t tl[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        tl[u] = y * gx + x;
        if (++x == rmax.x) { x = rmin.x; y++; }
      }
      fn(tl, (int)min(4u, area - t), payload);
    }
  }
  unsigned m = __ballot_sync(0xffffffffu, big);
  while (m) {
    const int src = __ffs(m) - 1;
    m &= m - 1;
    const uint32_t x0 = __shfl_sync(0xffffffffu, rmin.x, src), y0 = __shfl_sync(0xffffffffu, rmin.y, src);
    const uint32_t ww = __shfl_sync(0xffffffffu, w, src), aa = __shfl_sync(0xffffffffu, area, src);
    const unsigned long long pl = __shfl_sync(0xffffffffu, payload, src);
    for (uint32_t t0 = 0; t0 < aa; t0 += 128u) {
      uint32_t tl[4];
      int cnt = 0;
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const uint32_t t = t0 + 32u * u + lane;
        tl[u] = (y0 + t / ww) * gx + x0 + t % ww;
        cnt += t < aa ? 1 : 0;           // valid slots are a prefix: t grows with u
      }
      fn(tl, cnt, pl);
    }
  }
}

#ifndef MGS_EMIT_COOP
#define MGS_EMIT_COOP 4          // emit_kernel: chunks per step of the warp-cooperative rect walk (0: per-lane walk)
#endif
// The same visit, distributed over the warp: the rects of the warp's 32 splats are laid end to end (prefix sum of their
// areas) and lane l takes instance base + l of every 32-instance chunk — whoever owns it (binary search over the prefix,
// five shuffles).  Every lane is busy whatever the mix of rect sizes (the per-lane walk above ran emit_kernel at 11 of 32
// lanes, and its time was the dependent atomic chain of the lane with the largest rect), and U chunks are handed to the
// callback together so that it can keep U atomics per lane in flight.  fn(tile[U], ok[U], payload[U]).
// Must be reached by all 32 lanes.
template <int U, typename F>
__device__ __forceinline__ void for_each_tile_coop(const uint2 rmin, const uint2 rmax, const int gx,
                                                   const unsigned long long payload, F&& fn) {
  const int lane = threadIdx.x & 31;
  const uint32_t w = rmax.x - rmin.x, area = w * (rmax.y - rmin.y);
  uint32_t incl = area;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  const uint32_t excl = incl - area, total = __shfl_sync(0xffffffffu, incl, 31);
  const float inv_w = w > 0u ? 1.0f / (float)w : 0.f;
  const uint32_t pl_lo = (uint32_t)payload, pl_hi = (uint32_t)(payload >> 32);
  for (uint32_t base = 0; base < total; base += 32u * U) {
    uint32_t tile[U];
    bool ok[U];
    unsigned long long pl[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const uint32_t q = base + 32u * u + lane;
      // owner = the last lane whose first instance is <= q (lanes without a rect share their successor's start and lose)
      int lo = 0;
#pragma unroll
      for (int step = 16; step > 0; step >>= 1) {
        const uint32_t e = __shfl_sync(0xffffffffu, excl, (lo + step) & 31);
        if (e <= q) lo += step;       // lo + step <= 31 by construction
      }
      const uint32_t x0 = __shfl_sync(0xffffffffu, rmin.x, lo), y0 = __shfl_sync(0xffffffffu, rmin.y, lo);
      const uint32_t ww = __shfl_sync(0xffffffffu, w, lo), e0 = __shfl_sync(0xffffffffu, excl, lo);
      const float iw = __shfl_sync(0xffffffffu, inv_w, lo);
      const uint32_t plo = __shfl_sync(0xffffffffu, pl_lo, lo), phi = __shfl_sync(0xffffffffu, pl_hi, lo);
      const uint32_t k = q - e0;
      // k / ww for k < 2^20: (k + 0.5) / ww is never within 0.5 / ww of an integer, far outside the float error
      const uint32_t ty = (uint32_t)__float2int_rz(((float)k + 0.5f) * iw);
      ok[u] = q < total;
      tile[u] = (y0 + ty) * gx + x0 + (k - ty * ww);
      pl[u] = ((unsigned long long)phi << 32) | plo;
    }
    fn(tile, ok, pl);
  }
}

// ------------------------------------------------------------------------------------------------
// Binning.  The reference sorts R 64-bit (tile | depth) keys globally (rasterizer_impl.cu:70-111, 306-311).  Tile ids
// are a small dense range, so the tile half is a COUNTING sort: per-tile counts (preprocess_kernel) -> exclusive scan
// (tile ranges, rasterizer_impl.cu:116-138, fall out of it directly) -> scatter of (depth bits << 32 | Gaussian id) into
// the tile's segment -> per-tile sort of that 64-bit word in shared memory.  The reference's radix sort is stable and its
// instances are emitted in Gaussian-index order, so ties on (tile, depth) are ordered by index there too: the result
// is the same permutation, bit for bit.
// ------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 1024, LARGE_LIST = 1024;
// Whole-CTA scan of the per-tile instance counts: tile ranges (clipped to R_cap), scatter cursors, header, the compact
// list of large tiles and (heavy_first) the launch order of the multi-wave blend kernel.  THREADS = 1024 for the
// stand-alone tile_scan_kernel; the fused paths run it with the CTA size of preprocess_kernel in that kernel's LAST CTA
// (FOLD_SCAN), which saves a launch.  Counts are read with ld.global.cg: they were accumulated by L2 atomics of other CTAs.
template <int THREADS>
__device__ __noinline__ void tile_scan_body(uint32_t* __restrict__ tile_ctr, const int tiles, const uint32_t R_cap,
                                               uint2* __restrict__ ranges, WsHeader* __restrict__ hdr, const int heavy_first) {
  constexpr int IPT = 4;                   // THREADS * 4 tiles per sweep; few registers: the body also runs as the tail of preprocess_kernel
  using Scan = cub::BlockScan<uint32_t, THREADS>;
  __shared__ typename Scan::TempStorage tmp;
  __shared__ uint32_t n_large, max_count;
  if (threadIdx.x == 0) { n_large = 0u; max_count = 0u; }
  __syncthreads();
  uint32_t carry = 0;
  uint32_t ck[IPT] = {0u, 0u, 0u, 0u};     // the counts of the (usually only) sweep, kept for the heavy-first passes below
  for (int base = 0; base < tiles; base += THREADS * IPT) {
    const int t0 = base + threadIdx.x * IPT;
    uint32_t c[IPT], sum = 0u, m = 0u;
#pragma unroll
    for (int i = 0; i < IPT; i++) { c[i] = t0 + i < tiles ? __ldcg(&tile_ctr[(size_t)(t0 + i) * TILE_CTR_STRIDE]) : 0u; ck[i] = c[i]; }
#pragma unroll
    for (int i = 0; i < IPT; i++) { sum += c[i]; m = max(m, c[i]); }
    m = __reduce_max_sync(0xffffffffu, m);
    if ((threadIdx.x & 31) == 0 && m > 0u) atomicMax(&max_count, m);
    uint32_t excl, total;
    Scan(tmp).ExclusiveSum(sum, excl, total);
    uint32_t start = carry + excl;
#pragma unroll
    for (int i = 0; i < IPT; i++) {
      const int t = t0 + i;
      if (t < tiles) {
        const uint32_t lo = min(start, R_cap), hi = min(start + c[i], R_cap);
        ranges[t] = (c[i] == 0u) ? make_uint2(0u, 0u) : make_uint2(lo, hi);            // untouched tiles keep {0,0}
        tile_ctr[(size_t)t * TILE_CTR_STRIDE + 1] = start;
        if (hi - lo > (uint32_t)LARGE_LIST)
          tile_ctr[(size_t)atomicAdd(&n_large, 1u) * TILE_CTR_STRIDE + 2] = (uint32_t)t;
        start += c[i];
      }
    }
    carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    hdr->num_rendered = carry;
    hdr->overflow = carry > R_cap ? 1u : 0u;
    hdr->r_used = carry > R_cap ? R_cap : carry;
    hdr->n_large = n_large;
    if (carry > R_cap) hdr->overflow_sticky = 1u;
    if (carry > hdr->max_rendered) hdr->max_rendered = carry;
    if (max_count > hdr->max_tile_count) hdr->max_tile_count = max_count;
  }
  // Heavy-first launch order for per-tile kernels that run in several waves (blend_track_kernel: 3 CTAs per SM): a
  // counting sort of the tiles into 64 buckets of list length, longest lists first, written to
  // tile_ctr[i * TILE_CTR_STRIDE + 3].  CTAs are dispatched in blockIdx order, so the long tiles start at once and the
  // short ones fill the tail of the grid instead of the other way round.
  if (!heavy_first) return;   // only the single-pass tracking blend reads the order (neutral at C3, +1 % at C1 / C2)
  __shared__ uint32_t bucket[64];
  if (threadIdx.x < 64) bucket[threadIdx.x] = 0u;
  __syncthreads();
  // bucket = 63 - floor(63 c / max) in float: any monotone map of the count will do (a 64-bit integer division per tile
  // and pass cost more than the rest of the pass).  With one sweep (tiles <= THREADS * 4: every frame up to 1920x1080 at
  // 1024 threads) the counts are still in registers; otherwise they are read again.
  const float inv = 63.0f / (float)(max_count > 0u ? max_count : 1u);
  auto bucket_of = [&](const uint32_t c) { return 63u - min(63u, (uint32_t)((float)c * inv)); };
  const bool single = tiles <= THREADS * IPT;
  if (single) {
#pragma unroll
    for (int i = 0; i < IPT; i++)
      if ((int)threadIdx.x * IPT + i < tiles) atomicAdd(&bucket[bucket_of(ck[i])], 1u);
  } else {
    for (int t = threadIdx.x; t < tiles; t += THREADS) atomicAdd(&bucket[bucket_of(__ldcg(&tile_ctr[(size_t)t * TILE_CTR_STRIDE]))], 1u);
  }
  __syncthreads();
  if (threadIdx.x < 32) {   // exclusive scan of the 64 bucket sizes
    const uint32_t a = bucket[2 * threadIdx.x], b = bucket[2 * threadIdx.x + 1];
    uint32_t incl = a + b;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)threadIdx.x >= o) incl += up;
    }
    bucket[2 * threadIdx.x] = incl - a - b;
    bucket[2 * threadIdx.x + 1] = incl - b;
  }
  __syncthreads();
  if (single) {
#pragma unroll
    for (int i = 0; i < IPT; i++) {
      const int t = (int)threadIdx.x * IPT + i;
      if (t < tiles) tile_ctr[(size_t)atomicAdd(&bucket[bucket_of(ck[i])], 1u) * TILE_CTR_STRIDE + 3] = (uint32_t)t;
    }
  } else {
    for (int t = threadIdx.x; t < tiles; t += THREADS) {
      const uint32_t pos = atomicAdd(&bucket[bucket_of(__ldcg(&tile_ctr[(size_t)t * TILE_CTR_STRIDE]))], 1u);
      tile_ctr[(size_t)pos * TILE_CTR_STRIDE + 3] = (uint32_t)t;
    }
  }
}

__global__ void __launch_bounds__(SCAN_THREADS)
tile_scan_kernel(uint32_t* __restrict__ tile_ctr, const int tiles, const uint32_t R_cap,
                 uint2* __restrict__ ranges, WsHeader* __restrict__ hdr, const int heavy_first) {
  pdl_enter<PDL_SCAN>();
  tile_scan_body<SCAN_THREADS>(tile_ctr, tiles, R_cap, ranges, hdr, heavy_first);
}

// JAC_NCOL 0: plain; 6 / 9: also build the pose Jacobian of visible Gaussians (two-pass tracking path); 32: the folded
// 32-float basis of the single-pass tracking blend (pose_jac.cuh).
// FOLD_SCAN: the last CTA to finish (ticket in the workspace header, which must have been zero-filled once) runs the tile
// scan in place of a separate tile_scan_kernel launch.
// STAGE: SH rows of the warp's 32 Gaussians go through shared memory with coalesced 16-byte loads (sh_stage.cuh);
// sh_rest != nullptr selects the split layout (shs = f_dc [P,1,3], sh_rest = f_rest [P,M-1,3]).
#ifndef MGS_PRE_BASIS_MIN_CTAS
#define MGS_PRE_BASIS_MIN_CTAS 3
#endif
// SEGMENTS (fused tracker, packed-row mode): the counting sort's scatter happens HERE — every tile owns a fixed segment of
// `seg_cap` words of `list`, and an instance takes slot = atomicAdd(tile counter) of its tile at once, so there is no scan
// between counting and scattering and no emit pass at all (tile_sort_segments_kernel sorts the segments in place of
// tile_scan + emit + tile_sort_records x 2).  Instances beyond seg_cap are dropped; the count still says so (overflow).
#ifndef MGS_SEG_COOP
#define MGS_SEG_COOP 4           // chunks (returning atomics per lane in flight) of the segment scatter's warp-cooperative walk
#endif
template <int JAC_NCOL, bool STAGE, bool FOLD_SCAN, bool SEGMENTS = false>
__global__ void __launch_bounds__(256, (JAC_NCOL == 0 && !STAGE) ? 5 : (((JAC_NCOL == 32 || JAC_NCOL == TRACK_ROW_FLOATS) && !STAGE) ? MGS_PRE_BASIS_MIN_CTAS : 0))   // plain: 48 registers, 5 CTAs per SM; pose basis: 80, 3 CTAs
preprocess_kernel(const ViewConst v, const int P,
                  const float* __restrict__ means3D, const float* __restrict__ opacities,
                  const float* __restrict__ shs, const float* __restrict__ sh_rest, const float* __restrict__ colors_precomp,
                  const float* __restrict__ scales, const float* __restrict__ rotations,
                  const float* __restrict__ cov3D_precomp,
                  int32_t* __restrict__ radii, float* __restrict__ depths, float2* __restrict__ means2D,
                  float4* __restrict__ conic_opacity, float4* __restrict__ rgb4,
                  uint32_t* __restrict__ tiles_touched, uint32_t* __restrict__ tile_ctr,
                  float* __restrict__ jac, const uint32_t R_cap, uint2* __restrict__ ranges, WsHeader* __restrict__ hdr,
                  const int heavy_first, unsigned long long* __restrict__ list, const uint32_t seg_cap) {
  extern __shared__ __align__(16) float sh_tile[];
  pdl_enter<PDL_PRE>();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  int my_radius_i = 0;
  uint32_t touched = 0;
  uint2 rmin = {0u, 0u}, rmax = {0u, 0u};
  // phase 1: geometry and culling
  bool alive = false;
  float3 p = {0.f, 0.f, 0.f}, conic = {0.f, 0.f, 0.f};
  float2 pix = {0.f, 0.f};
  float view_z = 0.f, my_radius = 0.f;
  uint2 r0 = {0u, 0u}, r1 = {0u, 0u};
  uint32_t area = 0;
  if (idx < P) do {
    p = make_float3(means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
    // near cull only (auxiliary.h:139-164)
    const float4 ph = xform4x4(p, v.proj);
    const float pw = 1.0f / (ph.w + 0.0000001f);
    const float3 pproj = {ph.x * pw, ph.y * pw, ph.z * pw};
    const float3 pview = xform4x3(p, v.view);
    if (pview.z <= 0.2f) {
      if (v.prefiltered) { printf("Point is filtered although prefiltered is set. This shouldn't happen!"); __trap(); }
      break;
    }
#if MGS_COV6_REGS
    // the six covariance entries live in registers either way (a pointer that selects between the global array and a
    // local one forces the local one into local memory: 6250 local stores per launch in the round-1 profile)
    float cov6[6];
    if (cov3D_precomp != nullptr) {
#pragma unroll
      for (int k = 0; k < 6; k++) cov6[k] = cov3D_precomp[6 * (size_t)idx + k];
    } else cov3d_from_scale_rot(scales, rotations, idx, v.scale_modifier, cov6);
#else
    float cov6_local[6];
    const float* cov6;
    if (cov3D_precomp != nullptr) cov6 = cov3D_precomp + 6 * (size_t)idx;
    else { cov3d_from_scale_rot(scales, rotations, idx, v.scale_modifier, cov6_local); cov6 = cov6_local; }
#endif
    const Cov2DOut cv = cov2d_project(p, v, cov6);
    const float det = (cv.a * cv.c - cv.b * cv.b);
    if (det == 0.0f) break;
    const float det_inv = 1.f / det;
    conic = make_float3(cv.c * det_inv, -cv.b * det_inv, cv.a * det_inv);
    const float mid = 0.5f * (cv.a + cv.c);
    const float lambda1 = mid + sqrtf(max(0.1f, mid * mid - det));
    const float lambda2 = mid - sqrtf(max(0.1f, mid * mid - det));
    my_radius = ceilf(3.f * sqrtf(max(lambda1, lambda2)));
    pix = make_float2(ndc_to_pix(pproj.x, v.W), ndc_to_pix(pproj.y, v.H));
    tile_rect(pix, (int)my_radius, v.gx, v.gy, r0, r1);
    area = (r1.x - r0.x) * (r1.y - r0.y);
    if (area == 0) break;
    view_z = pview.z;
    alive = true;
  } while (0);
  // phase 2: colour, pose Jacobian, per-Gaussian outputs
  const float* sh_row = nullptr;
  if (STAGE) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int RS = sh_row_stride(v.M);
    float* tile = sh_tile + warp * 32 * RS;
    const int first = blockIdx.x * blockDim.x + warp * 32;
    const int n_rows = min(32, P - first);
    if (__ballot_sync(0xffffffffu, alive) != 0u) {   // warp-uniform
      if (sh_rest == nullptr) sh_stage_load(tile, RS, shs + (size_t)first * 3 * v.M, 3 * v.M, 0, n_rows, lane);
      else {
        sh_stage_load(tile, RS, shs + (size_t)first * 3, 3, 0, n_rows, lane);
        sh_stage_load(tile, RS, sh_rest + (size_t)first * 3 * (v.M - 1), 3 * (v.M - 1), 3, n_rows, lane);
      }
      __syncwarp();
    }
    sh_row = tile + lane * RS;
  } else if (shs != nullptr) sh_row = shs + (size_t)idx * v.M * 3;
  if (alive) {
    uint32_t cm = 0;
    float3 col = {0.f, 0.f, 0.f};
    if (colors_precomp == nullptr) {
      col = sh_to_rgb(v.D, p, v.campos, sh_row, cm);
      // (packed-row mode: colour, conic and opacity live in the row; nothing reads the separate arrays, and the kernel
      // is bound by its strided stores)
      if (JAC_NCOL != TRACK_ROW_FLOATS) rgb4[idx] = make_float4(col.x, col.y, col.z, __uint_as_float(cm));
    }
    if (JAC_NCOL == 32) build_pose_basis(v, idx, p, conic, scales, rotations, cov3D_precomp, jac);
    else if (JAC_NCOL == TRACK_ROW_FLOATS) {
      // the packed row of the fused tracker: record part (what write_record assembles per instance) + basis
      // ... assembled in the CTA's shared-memory tile [256][44] and written out below by the whole CTA: the rows of a CTA's
      // 256 consecutive Gaussians are one contiguous 45 KB block, so the copy is coalesced 16-byte stores, whereas a thread
      // writing its own row issues eleven 16-byte stores at a 176-byte lane stride (the kernel was bound by those)
      build_pose_basis(v, idx, p, conic, scales, rotations, cov3D_precomp, sh_tile, TRACK_ROW_FLOATS, 12, (int)threadIdx.x);
      const float op = opacities[idx];
      float4* row = reinterpret_cast<float4*>(sh_tile + (size_t)threadIdx.x * TRACK_ROW_FLOATS);
      row[0] = make_float4(pix.x, pix.y, view_z, __uint_as_float((uint32_t)idx));
      row[1] = make_float4(conic.x, conic.y, conic.z, op);
      row[2] = make_float4(col.x, col.y, col.z, __uint_as_float(splat_extent(conic.x, conic.y, conic.z, op)));
    }
    else if (JAC_NCOL > 0) build_pose_jac<(JAC_NCOL == 9 ? 9 : 6)>(v, idx, p, shs, scales, rotations, cov3D_precomp, cm, jac);
    if (!(SEGMENTS && JAC_NCOL == TRACK_ROW_FLOATS)) {   // (packed-row segment mode: emit_kernel, their only reader there, does not run)
      depths[idx] = view_z;
      means2D[idx] = pix;
    }
    if (JAC_NCOL != TRACK_ROW_FLOATS) conic_opacity[idx] = make_float4(conic.x, conic.y, conic.z, opacities[idx]);
    my_radius_i = (int)my_radius;
    touched = area;
    rmin = r0; rmax = r1;
  }
  if (idx < P) {
    radii[idx] = my_radius_i;
    if (!SEGMENTS) tiles_touched[idx] = touched;
  }
  if constexpr (JAC_NCOL == TRACK_ROW_FLOATS) {
    __shared__ uint32_t s_alive[8];
    const unsigned am = __ballot_sync(0xffffffffu, alive);
    if ((threadIdx.x & 31) == 0) s_alive[threadIdx.x >> 5] = am;
    __syncthreads();
    constexpr int RC = TRACK_ROW_FLOATS / 4;
    const float4* src4 = reinterpret_cast<const float4*>(sh_tile);
    float4* dst4 = reinterpret_cast<float4*>(jac + (size_t)blockIdx.x * 256 * TRACK_ROW_FLOATS);
#pragma unroll
    for (int q = 0; q < RC; q++) {
      const int e = q * 256 + (int)threadIdx.x, r = e / RC;
      if ((s_alive[r >> 5] >> (r & 31)) & 1u) dst4[e] = src4[e];
    }
  }
  if constexpr (SEGMENTS) {
    // count AND scatter: (depth bits << 32 | id) straight into the tile's fixed segment, warp-cooperative walk of the 32
    // rects with four returning atomics per lane in flight (the walk of emit_kernel)
    const unsigned long long word = alive ? (((unsigned long long)__float_as_uint(view_z) << 32) | (uint32_t)idx) : 0ull;
    for_each_tile_coop<MGS_SEG_COOP>(rmin, rmax, v.gx, word, [&](const uint32_t (&tile)[MGS_SEG_COOP], const bool (&ok)[MGS_SEG_COOP],
                                                                 const unsigned long long (&w)[MGS_SEG_COOP]) {
      uint32_t slot[MGS_SEG_COOP];
#pragma unroll
      for (int u = 0; u < MGS_SEG_COOP; u++)
        if (ok[u]) slot[u] = atomicAdd(&tile_ctr[tile[u] * TILE_CTR_STRIDE], 1u);
#pragma unroll
      for (int u = 0; u < MGS_SEG_COOP; u++)
        if (ok[u] && slot[u] < seg_cap) list[(size_t)tile[u] * seg_cap + slot[u]] = w[u];
    });
    return;
  }
  // instances per tile (the histogram of the counting sort that replaces the reference's global key sort)
  for_each_tile(rmin, rmax, v.gx, 0ull, [&](const uint32_t (&tile)[4], int cnt, unsigned long long) {
#pragma unroll
    for (int u = 0; u < 4; u++)
      if (u < cnt) atomicAdd(&tile_ctr[tile[u] * TILE_CTR_STRIDE], 1u);
  });
  if constexpr (FOLD_SCAN) {
    __shared__ int s_last;
    // the CTA barrier orders every thread's counter atomics before thread 0's fence (fences are cumulative), which orders
    // them before the ticket: one MEMBAR per CTA instead of one per thread
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      s_last = (atomicAdd(&hdr->ticket, 1u) == gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (s_last) {
      __threadfence();
      if constexpr (STAGE) tile_scan_body<128>(tile_ctr, v.gx * v.gy, R_cap, ranges, hdr, heavy_first);
      else tile_scan_body<256>(tile_ctr, v.gx * v.gy, R_cap, ranges, hdr, heavy_first);
      if (threadIdx.x == 0) hdr->ticket = 0u;
    }
  }
}

// (tile_scan_body / tile_scan_kernel are defined above preprocess_kernel, whose last CTA may run the scan itself)
// scatter: one (depth bits << 32 | id) word per (Gaussian, tile) into the tile's segment, slot taken atomically
__global__ void __launch_bounds__(256)
emit_kernel(const int P, const uint32_t R_cap, const int gx, const int gy,
            const float2* __restrict__ means2D, const float* __restrict__ depths, const int32_t* __restrict__ radii,
            uint32_t* __restrict__ tile_ctr, unsigned long long* __restrict__ list) {
  pdl_enter<PDL_EMIT>();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  uint2 rmin = {0u, 0u}, rmax = {0u, 0u};
  unsigned long long word = 0ull;
  if (idx < P) {
    const int rad = radii[idx];
    if (rad > 0) {
      tile_rect(means2D[idx], rad, gx, gy, rmin, rmax);
      // view depth > 0.2, so its bit pattern orders like the float (rasterizer_impl.cu:99-101)
      word = ((unsigned long long)__float_as_uint(depths[idx]) << 32) | (uint32_t)idx;
    }
  }
#if MGS_EMIT_COOP
  for_each_tile_coop<MGS_EMIT_COOP>(rmin, rmax, gx, word, [&](const uint32_t (&tile)[MGS_EMIT_COOP], const bool (&ok)[MGS_EMIT_COOP],
                                                              const unsigned long long (&w)[MGS_EMIT_COOP]) {
    uint32_t slot[MGS_EMIT_COOP];
#pragma unroll
    for (int u = 0; u < MGS_EMIT_COOP; u++)
      if (ok[u]) slot[u] = atomicAdd(&tile_ctr[tile[u] * TILE_CTR_STRIDE + 1], 1u);
#pragma unroll
    for (int u = 0; u < MGS_EMIT_COOP; u++)
      if (ok[u] && slot[u] < R_cap) list[slot[u]] = w[u];
  });
#else
  for_each_tile(rmin, rmax, gx, word, [&](const uint32_t (&tile)[4], int cnt, unsigned long long w) {
    uint32_t slot[4];
#pragma unroll
    for (int u = 0; u < 4; u++)
      if (u < cnt) slot[u] = atomicAdd(&tile_ctr[tile[u] * TILE_CTR_STRIDE + 1], 1u);
#pragma unroll
    for (int u = 0; u < 4; u++)
      if (u < cnt && slot[u] < R_cap) list[slot[u]] = w;
  });
#endif
}

// Per-tile sort of the 64-bit words + record packing.  Lists of up to SORT_CAP words are merge-sorted from registers
// (cub::BlockMergeSort, 2/4/8/16 words per thread by list length; every word is unique, slots >= n are padded with
// ~0 which no real word equals).  Longer lists are sorted in place in global memory (L2) with a bitonic network in
// its "mirror" form: every compare-exchange puts the smaller word at the lower index, so slots >= n act as +inf
// without being stored and n need not be a power of two.
constexpr int SORT_CAP = 4096;
struct U64Less { __device__ __forceinline__ bool operator()(unsigned long long a, unsigned long long b) const { return a < b; } };

// means2D == nullptr: ids-only mode of the fused tracker — the sorted Gaussian ids (u32) are all that is written, into the
// same workspace region; blend_track_kernel<ROWS> gathers the packed per-Gaussian rows by id (pose_jac.cuh).
__device__ __forceinline__ void write_record(const unsigned long long w, void* __restrict__ dst_base, const int e,
                                             const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                                             const float4* __restrict__ rgb4, const float* __restrict__ colors_precomp) {
  if (means2D == nullptr) { reinterpret_cast<uint32_t*>(dst_base)[e] = (uint32_t)w; return; }
  SplatRec* dst = reinterpret_cast<SplatRec*>(dst_base) + e;
  const uint32_t id = (uint32_t)w;
  const float2 xy = means2D[id];
  const float4 co = conic_opacity[id];
  float4 c;
  if (colors_precomp != nullptr) c = make_float4(colors_precomp[3 * id], colors_precomp[3 * id + 1], colors_precomp[3 * id + 2], 0.f);
  else c = rgb4[id];
  float4* out = reinterpret_cast<float4*>(dst);
  out[0] = make_float4(xy.x, xy.y, __uint_as_float((uint32_t)(w >> 32)), __uint_as_float(id));
  out[1] = co;
  out[2] = make_float4(c.x, c.y, c.z, __uint_as_float(splat_extent(co.x, co.y, co.z, co.w)));
}

#ifndef MGS_SORT_768_AS_1024
#define MGS_SORT_768_AS_1024 0    // 1: lists of 513..768 words take the 1024-word bitonic network instead of the 3-words-per-thread merge sort
#endif
#ifndef MGS_SORT_THREADS
#define MGS_SORT_THREADS 256     // CTA size of the short-list sort kernel; measured at C2: 256 -> 33 us, 128 -> 44 us, 64 -> 48 us
#endif
template <int IPT, int THREADS, typename Temp>
__device__ __noinline__ void sort_tile_regs(Temp& temp, const unsigned long long* __restrict__ seg, const int n,
                                            void* __restrict__ dst, const float2* __restrict__ means2D,
                                            const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb4,
                                            const float* __restrict__ colors_precomp) {
  using Sort = cub::BlockMergeSort<unsigned long long, THREADS, IPT>;
  unsigned long long k[IPT];
#pragma unroll
  for (int i = 0; i < IPT; i++) {
    const int e = threadIdx.x * IPT + i;
    k[i] = e < n ? seg[e] : ~0ull;
  }
  // (the exchange buffer lives in the sort's shared storage; an instantiation whose storage is too small is the dead
  // LARGE branch of the short-list kernel and keeps the merge sort)
  if constexpr (MGS_SORT_BITONIC && (IPT & (IPT - 1)) == 0 && sizeof(Temp) >= (size_t)THREADS * IPT * sizeof(unsigned long long)) {
    block_bitonic_sort<IPT, THREADS>(k, reinterpret_cast<unsigned long long*>(&temp));
  } else Sort(reinterpret_cast<typename Sort::TempStorage&>(temp)).Sort(k, U64Less());
#pragma unroll
  for (int i = 0; i < IPT; i++) {
    const int e = threadIdx.x * IPT + i;
    if (e < n) write_record(k[i], dst, e, means2D, conic_opacity, rgb4, colors_precomp);
  }
}

// Two launches cover the tiles.  LARGE = false: one CTA per tile, lists of up to 1024 words (few registers, 8 KB of
// shared memory, full occupancy — the common case).  LARGE = true: a fixed grid walks the list of large tiles the scan
// kernel compacted (tile_ctr[i * TILE_CTR_STRIDE + 2], i < header.n_large); with no large tile it costs an empty wave
// of a few hundred CTAs instead of one CTA per tile.
template <bool LARGE>
__global__ void __launch_bounds__(LARGE ? 256 : MGS_SORT_THREADS)
tile_sort_records_kernel(const uint2* __restrict__ ranges, unsigned long long* __restrict__ list,
                         const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                         const float4* __restrict__ rgb4, const float* __restrict__ colors_precomp,
                         SplatRec* __restrict__ recs, uint32_t* __restrict__ tile_ctr,
                         const WsHeader* __restrict__ hdr, const int clear_counts) {
  constexpr int ST = LARGE ? 256 : MGS_SORT_THREADS, SI = 1024 / MGS_SORT_THREADS;   // SI words per thread cover LARGE_LIST
  using Temp = typename cub::BlockMergeSort<unsigned long long, ST, LARGE ? 16 : SI>::TempStorage;
  __shared__ Temp temp;
  pdl_enter<LARGE ? PDL_SORTL : PDL_SORT>();
  if (!LARGE) {
    // fused loops (tracker, mapping engine) render into the same workspace again and again: the tile's instance counter
    // is dead once the scan has run, so this CTA leaves it zero for the next preprocess instead of a memset per render
    if (clear_counts && threadIdx.x == 0) tile_ctr[(size_t)blockIdx.x * TILE_CTR_STRIDE] = 0u;
    const uint2 range = ranges[blockIdx.x];
    const int n = (int)(range.y - range.x);
    if (n == 0 || n > LARGE_LIST) return;
    unsigned long long* seg = list + range.x;
    void* dst = means2D == nullptr ? (void*)(reinterpret_cast<uint32_t*>(recs) + range.x) : (void*)(recs + range.x);
    if (n <= 256) sort_tile_regs<SI / 4, ST>(temp, seg, n, dst, means2D, conic_opacity, rgb4, colors_precomp);
    else if (n <= 512) sort_tile_regs<SI / 2, ST>(temp, seg, n, dst, means2D, conic_opacity, rgb4, colors_precomp);
    else if (n <= 768 && !MGS_SORT_768_AS_1024) sort_tile_regs<SI * 3 / 4, ST>(temp, seg, n, dst, means2D, conic_opacity, rgb4, colors_precomp);   // lists just over 512 are common at 640x480 / 200k
    else sort_tile_regs<SI, ST>(temp, seg, n, dst, means2D, conic_opacity, rgb4, colors_precomp);
    return;
  }
  const uint32_t n_large = hdr->n_large;
  for (uint32_t li = blockIdx.x; li < n_large; li += gridDim.x) {
    const uint2 range = ranges[tile_ctr[(size_t)li * TILE_CTR_STRIDE + 2]];
    const int n = (int)(range.y - range.x);
    unsigned long long* seg = list + range.x;
    void* dst = means2D == nullptr ? (void*)(reinterpret_cast<uint32_t*>(recs) + range.x) : (void*)(recs + range.x);
    if (n <= SORT_CAP) {
      if (n <= 2048) sort_tile_regs<8, 256>(temp, seg, n, dst, means2D, conic_opacity, rgb4, colors_precomp);
      else sort_tile_regs<16, 256>(temp, seg, n, dst, means2D, conic_opacity, rgb4, colors_precomp);
      __syncthreads();          // temp is reused by the next tile
      continue;
    }
    int N = 1;
    while (N < n) N <<= 1;
    for (int k = 2; k <= N; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        const bool mirror = (j == (k >> 1));
        for (int q = threadIdx.x; q < (N >> 1); q += blockDim.x) {
          const int lo = q & (j - 1), blk = (q - lo) << 1;   // block of 2j slots, offset lo inside its lower half
          const int i = blk + lo;
          const int p = mirror ? blk + 2 * j - 1 - lo : i + j;
          if (p < n) {
            const unsigned long long a = __ldcg(seg + i), b = __ldcg(seg + p);
            if (a > b) { __stcg(seg + i, b); __stcg(seg + p, a); }
          }
        }
        __syncthreads();
      }
    }
    for (int e = threadIdx.x; e < n; e += blockDim.x)
      write_record(__ldcg(seg + e), dst, e, means2D, conic_opacity, rgb4, colors_precomp);
  }
}

// Fixed-segment binning of the fused tracker (preprocess_kernel<.., SEGMENTS>): tile t's instances sit unsorted in
// list[t * cap .. t * cap + min(count, cap)).  CTA b >= 1 sorts the segment of tile b - 1 and writes the sorted Gaussian ids
// (u32) at the same offsets of the record region; CTA 0 meanwhile does what is left of the scan: ranges, the status header
// and the heavy-first launch order of the blend kernel (tile_ctr[i * TILE_CTR_STRIDE + 3]) — no prefix sum, the segment
// starts are a multiplication.  The tile counters are NOT cleared here (CTA 0 reads all of them while the others run); the
// blend kernel, which visits every tile exactly once, leaves them zero for the next iteration.
// IDS: sorted Gaussian ids (fused tracker, packed rows); otherwise 48-byte records gathered by id (mapping engine).
#ifndef MGS_SEG_SORT_MIN_CTAS
#define MGS_SEG_SORT_MIN_CTAS 8   // record variant: 32 registers instead of 40 -> 8 CTAs per SM (mapping C4, 8 views: 457.3 -> 459.1 it/s)
#endif
template <bool IDS>
__global__ void __launch_bounds__(MGS_SORT_THREADS, IDS ? 0 : MGS_SEG_SORT_MIN_CTAS)
tile_sort_segments_kernel(const int tiles, const uint32_t cap, const unsigned long long* __restrict__ list,
                          void* __restrict__ out, const float2* __restrict__ means2D_in, const float4* __restrict__ conic_opacity,
                          const float4* __restrict__ rgb4, uint32_t* __restrict__ tile_ctr, uint2* __restrict__ ranges,
                          WsHeader* __restrict__ hdr) {
  constexpr int ST = MGS_SORT_THREADS, SI = 1024 / MGS_SORT_THREADS;
  using Temp = typename cub::BlockMergeSort<unsigned long long, ST, SI>::TempStorage;
  __shared__ Temp temp;
  pdl_enter<PDL_SORT>();
  if (blockIdx.x != 0) {
    const int tile = (int)blockIdx.x - 1;
    const int n = (int)min(__ldcg(&tile_ctr[(size_t)tile * TILE_CTR_STRIDE]), cap);
    if (n == 0) return;
    const unsigned long long* seg = list + (size_t)tile * cap;
    void* dst = IDS ? (void*)(reinterpret_cast<uint32_t*>(out) + (size_t)tile * cap) : (void*)(reinterpret_cast<SplatRec*>(out) + (size_t)tile * cap);
    const float2* means2D = IDS ? nullptr : means2D_in;      // write_record: null = ids only
    if (n <= 256) sort_tile_regs<SI / 4, ST>(temp, seg, n, dst, means2D, conic_opacity, rgb4, nullptr);
    else if (n <= 512) sort_tile_regs<SI / 2, ST>(temp, seg, n, dst, means2D, conic_opacity, rgb4, nullptr);
    else if (n <= 768 && !MGS_SORT_768_AS_1024) sort_tile_regs<SI * 3 / 4, ST>(temp, seg, n, dst, means2D, conic_opacity, rgb4, nullptr);
    else sort_tile_regs<SI, ST>(temp, seg, n, dst, means2D, conic_opacity, rgb4, nullptr);
    return;
  }
  __shared__ uint32_t s_sum, s_used, s_max, bucket[64];
  if (threadIdx.x == 0) { s_sum = 0u; s_used = 0u; s_max = 0u; }
  if (threadIdx.x < 64) bucket[threadIdx.x] = 0u;
  __syncthreads();
  uint32_t sum = 0u, used = 0u, m = 0u;
  for (int t = threadIdx.x; t < tiles; t += ST) {
    const uint32_t c = __ldcg(&tile_ctr[(size_t)t * TILE_CTR_STRIDE]), n = min(c, cap);
    ranges[t] = (c == 0u) ? make_uint2(0u, 0u) : make_uint2((uint32_t)t * cap, (uint32_t)t * cap + n);
    sum += c; used += n; m = max(m, c);
  }
  sum = __reduce_add_sync(0xffffffffu, sum); used = __reduce_add_sync(0xffffffffu, used); m = __reduce_max_sync(0xffffffffu, m);
  if ((threadIdx.x & 31) == 0) { atomicAdd(&s_sum, sum); atomicAdd(&s_used, used); atomicMax(&s_max, m); }
  __syncthreads();
  if (threadIdx.x == 0) {
    const bool over = s_max > cap;
    hdr->num_rendered = s_sum;
    hdr->overflow = over ? 1u : 0u;
    hdr->r_used = s_used;
    hdr->n_large = 0u;
    if (over) hdr->overflow_sticky = 1u;
    if (s_sum > hdr->max_rendered) hdr->max_rendered = s_sum;
    if (s_max > hdr->max_tile_count) hdr->max_tile_count = s_max;
  }
  // heavy-first order: counting sort of the tiles into 64 buckets of list length, longest first (as tile_scan_body)
  const float inv = 63.0f / (float)(s_max > 0u ? s_max : 1u);
  auto bucket_of = [&](const uint32_t c) { return 63u - min(63u, (uint32_t)((float)c * inv)); };
  for (int t = threadIdx.x; t < tiles; t += ST) atomicAdd(&bucket[bucket_of(__ldcg(&tile_ctr[(size_t)t * TILE_CTR_STRIDE]))], 1u);
  __syncthreads();
  if (threadIdx.x < 32) {   // exclusive scan of the 64 bucket sizes
    const uint32_t a = bucket[2 * threadIdx.x], b = bucket[2 * threadIdx.x + 1];
    uint32_t incl = a + b;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)threadIdx.x >= o) incl += up;
    }
    bucket[2 * threadIdx.x] = incl - a - b;
    bucket[2 * threadIdx.x + 1] = incl - b;
  }
  __syncthreads();
  for (int t = threadIdx.x; t < tiles; t += ST) {
    const uint32_t pos = atomicAdd(&bucket[bucket_of(__ldcg(&tile_ctr[(size_t)t * TILE_CTR_STRIDE]))], 1u);
    tile_ctr[(size_t)pos * TILE_CTR_STRIDE + 3] = (uint32_t)t;
  }
}

int launch_sort_segments(const ViewConst& v, void* ws, const WsLayout& L, int64_t R_cap, cudaStream_t s, bool ids_only) {
  const int tiles = v.gx * v.gy;
  if ((int64_t)tiles * TILE_SEGMENT > R_cap) {
    set_error("tile segments: R_cap %lld < tiles x %d = %lld", (long long)R_cap, TILE_SEGMENT, (long long)tiles * TILE_SEGMENT);
    return MGS_ENOSPC;
  }
  cudaError_t e;
#define MGS_SEG_ARGS tiles + 1, MGS_SORT_THREADS, 0, s, tiles, (uint32_t)TILE_SEGMENT, ws_ptr<unsigned long long>(ws, L.list),          \
      ws_ptr<void>(ws, L.records), ws_ptr<float2>(ws, L.means2D), ws_ptr<float4>(ws, L.conic_opacity), ws_ptr<float4>(ws, L.rgb4),  \
      ws_ptr<uint32_t>(ws, L.tile_ctr), ws_ptr<uint2>(ws, L.ranges), ws_ptr<WsHeader>(ws, L.header)
  MGS_PROF("tile_sort_segments", s, e = ids_only ? launch_chain(tile_sort_segments_kernel<true>, MGS_SEG_ARGS)
                                                 : launch_chain(tile_sort_segments_kernel<false>, MGS_SEG_ARGS));
#undef MGS_SEG_ARGS
  MGS_CUDA_TRY(e);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

int launch_tile_scan(const ViewConst& v, void* ws, const WsLayout& L, int64_t R_cap, cudaStream_t s, bool heavy_first) {
  cudaError_t e;
  MGS_PROF("tile_scan", s, e = launch_chain(tile_scan_kernel, 1, SCAN_THREADS, 0, s,
      ws_ptr<uint32_t>(ws, L.tile_ctr), v.gx * v.gy, (uint32_t)R_cap, ws_ptr<uint2>(ws, L.ranges),
      ws_ptr<WsHeader>(ws, L.header), heavy_first ? 1 : 0));
  MGS_CUDA_TRY(e);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

int launch_preprocess(const ViewConst& v, int P, const float* means3D, const float* opacities, const float* shs,
                      const float* colors_precomp, const float* scales, const float* rotations,
                      const float* cov3D_precomp, int32_t* radii, void* ws, const WsLayout& L, cudaStream_t s,
                      float* jac, int jac_ncol, const float* sh_rest, const FoldScan* fold) {
  if (P == 0) return MGS_OK;
  // fold != nullptr: the workspace is one a fused loop renders into repeatedly — its tile counters were left zero by the
  // previous render's sort kernel (clear_counts) and its header ticket is zero, so no memset and no separate scan launch
  if (fold == nullptr)
    MGS_CUDA_TRY(memset_async(ws_ptr<void>(ws, L.tile_ctr), 0, (size_t)v.gx * v.gy * TILE_CTR_STRIDE * sizeof(uint32_t), s));
  const uint32_t R_cap = fold ? (uint32_t)fold->R_cap : 0u;
  const int heavy = fold ? (fold->heavy_first ? 1 : 0) : 0;
  const uint32_t seg_cap = fold ? (uint32_t)fold->segment_cap : 0u;
  if (seg_cap != 0u && !((jac == nullptr || jac_ncol == TRACK_ROW_FLOATS) && !fold->in_kernel)) {
    set_error("preprocess: fixed tile segments serve the fused tracker's packed-row mode and the mapping engine"); return MGS_EINVAL;
  }
  // fold->in_kernel picks the folded tail or a separate tile_scan launch (launchers.h); either way the workspace is the
  // zero-maintained one of a fused loop: no memset.
  static const int scan_env = [] { const char* e = getenv("MGS_SCAN_FOLD"); return e ? (e[0] == '1' ? 1 : 0) : -1; }();
  const bool scan_separate = fold != nullptr && (seg_cap != 0u || (scan_env >= 0 ? scan_env == 0 : !fold->in_kernel));
  const FoldScan* const fold_req = fold;
  if (scan_separate) fold = nullptr;
#define MGS_PRE_ARGS v, P, means3D, opacities, shs, sh_rest, colors_precomp, scales, rotations, cov3D_precomp, radii,  \
      ws_ptr<float>(ws, L.depths), ws_ptr<float2>(ws, L.means2D), ws_ptr<float4>(ws, L.conic_opacity),               \
      ws_ptr<float4>(ws, L.rgb4), ws_ptr<uint32_t>(ws, L.tiles_touched), ws_ptr<uint32_t>(ws, L.tile_ctr), jac, R_cap, \
      ws_ptr<uint2>(ws, L.ranges), ws_ptr<WsHeader>(ws, L.header), heavy, ws_ptr<unsigned long long>(ws, L.list), seg_cap
  // SH rows of more than one coefficient are staged through shared memory (128 threads: 4 tiles of [32][3M | 1] floats)
  const bool stage = shs != nullptr && colors_precomp == nullptr && v.M > 1 && !(jac != nullptr && (jac_ncol == 32 || jac_ncol == TRACK_ROW_FLOATS));
  if (sh_rest != nullptr && !stage) { set_error("preprocess: split SH layout needs sh_coeffs > 1 and no colors_precomp"); return MGS_EINVAL; }
  if (jac != nullptr && (jac_ncol == 32 || jac_ncol == TRACK_ROW_FLOATS) && fold_req == nullptr) { set_error("preprocess: the folded pose basis is built by the fused SH-degree-0 tracker only"); return MGS_EINVAL; }
  cudaError_t pe = cudaSuccess;
  prof_begin(fold ? "preprocess+scan" : "preprocess", s);
  if (stage) {
    const int T = 128;
    const size_t smem = (size_t)T * sh_row_stride(v.M) * sizeof(float);
    if (fold) {
      if (jac == nullptr) pe = launch_chain(preprocess_kernel<0, true, true>, (P + T - 1) / T, T, smem, s, MGS_PRE_ARGS);
      else if (jac_ncol == 9) pe = launch_chain(preprocess_kernel<9, true, true>, (P + T - 1) / T, T, smem, s, MGS_PRE_ARGS);
      else { set_error("preprocess: unsupported fused variant"); return MGS_EINVAL; }
    } else if (jac == nullptr && seg_cap != 0u) pe = launch_chain(preprocess_kernel<0, true, false, true>, (P + T - 1) / T, T, smem, s, MGS_PRE_ARGS);
    else if (jac == nullptr) pe = launch_chain(preprocess_kernel<0, true, false>, (P + T - 1) / T, T, smem, s, MGS_PRE_ARGS);
    else if (jac_ncol == 9) pe = launch_chain(preprocess_kernel<9, true, false>, (P + T - 1) / T, T, smem, s, MGS_PRE_ARGS);
    else pe = launch_chain(preprocess_kernel<6, true, false>, (P + T - 1) / T, T, smem, s, MGS_PRE_ARGS);
  } else {
    if (fold) {
      if (jac == nullptr) pe = launch_chain(preprocess_kernel<0, false, true>, (P + 255) / 256, 256, 0, s, MGS_PRE_ARGS);
      else if (jac_ncol == 32) pe = launch_chain(preprocess_kernel<32, false, true>, (P + 255) / 256, 256, 0, s, MGS_PRE_ARGS);
      else if (jac_ncol == TRACK_ROW_FLOATS) pe = launch_chain(preprocess_kernel<TRACK_ROW_FLOATS, false, true>, (P + 255) / 256, 256, 256 * TRACK_ROW_FLOATS * sizeof(float), s, MGS_PRE_ARGS);
      else { set_error("preprocess: unsupported fused variant"); return MGS_EINVAL; }
    } else if (jac == nullptr && seg_cap != 0u) pe = launch_chain(preprocess_kernel<0, false, false, true>, (P + 255) / 256, 256, 0, s, MGS_PRE_ARGS);
    else if (jac == nullptr) pe = launch_chain(preprocess_kernel<0, false, false>, (P + 255) / 256, 256, 0, s, MGS_PRE_ARGS);
    else if (jac_ncol == 9) pe = launch_chain(preprocess_kernel<9, false, false>, (P + 255) / 256, 256, 0, s, MGS_PRE_ARGS);
    else if (jac_ncol == 32) pe = launch_chain(preprocess_kernel<32, false, false>, (P + 255) / 256, 256, 0, s, MGS_PRE_ARGS);
    else if (jac_ncol == TRACK_ROW_FLOATS && seg_cap != 0u) pe = launch_chain(preprocess_kernel<TRACK_ROW_FLOATS, false, false, true>, (P + 255) / 256, 256, 256 * TRACK_ROW_FLOATS * sizeof(float), s, MGS_PRE_ARGS);
    else if (jac_ncol == TRACK_ROW_FLOATS) pe = launch_chain(preprocess_kernel<TRACK_ROW_FLOATS, false, false>, (P + 255) / 256, 256, 256 * TRACK_ROW_FLOATS * sizeof(float), s, MGS_PRE_ARGS);
    else pe = launch_chain(preprocess_kernel<6, false, false>, (P + 255) / 256, 256, 0, s, MGS_PRE_ARGS);
  }
  prof_end(s);
#undef MGS_PRE_ARGS
  MGS_CUDA_TRY(pe);
  MGS_LAUNCH_CHECK();
  if (seg_cap != 0u) return MGS_OK;     // no scan: launch_sort_segments() follows
  if (fold_req != nullptr && fold == nullptr) return launch_tile_scan(v, ws, L, fold_req->R_cap, s, fold_req->heavy_first);
  return MGS_OK;
}

int launch_binning(const ViewConst& v, int P, const float* colors_precomp, const int32_t* radii, void* ws,
                   const WsLayout& L, int64_t R_cap, cudaStream_t s, bool heavy_first, bool scan_done, bool clear_counts,
                   bool ids_only) {
  const int tiles = v.gx * v.gy;
  if (!scan_done) {
    int rc = launch_tile_scan(v, ws, L, R_cap, s, heavy_first);   // ranges (clipped to R_cap), scatter cursors, header
    if (rc) return rc;
  }
  if (R_cap == 0 || P == 0) return MGS_OK;
  cudaError_t e;
  MGS_PROF("emit", s, e = launch_chain(emit_kernel, (P + 255) / 256, 256, 0, s,
      P, (uint32_t)R_cap, v.gx, v.gy, ws_ptr<float2>(ws, L.means2D), ws_ptr<float>(ws, L.depths), radii,
      ws_ptr<uint32_t>(ws, L.tile_ctr), ws_ptr<unsigned long long>(ws, L.list)));
  MGS_CUDA_TRY(e);
  MGS_LAUNCH_CHECK();
#define MGS_SORT_ARGS ws_ptr<uint2>(ws, L.ranges), ws_ptr<unsigned long long>(ws, L.list),                                   \
      (ids_only ? (const float2*)nullptr : ws_ptr<float2>(ws, L.means2D)),                                                       \
      ws_ptr<float4>(ws, L.conic_opacity), ws_ptr<float4>(ws, L.rgb4), colors_precomp, ws_ptr<SplatRec>(ws, L.records),     \
      ws_ptr<uint32_t>(ws, L.tile_ctr), ws_ptr<WsHeader>(ws, L.header), clear_counts ? 1 : 0
  MGS_PROF("tile_sort_records", s, e = launch_chain(tile_sort_records_kernel<false>, tiles, MGS_SORT_THREADS, 0, s, MGS_SORT_ARGS));
  MGS_CUDA_TRY(e);
  MGS_LAUNCH_CHECK();
  MGS_PROF("tile_sort_records_large", s,
           e = launch_chain(tile_sort_records_kernel<true>, tiles < 592 ? tiles : 592, 256, 0, s, MGS_SORT_ARGS));
  MGS_CUDA_TRY(e);
  MGS_LAUNCH_CHECK();
#undef MGS_SORT_ARGS
  return MGS_OK;
}

}  // namespace mgs
