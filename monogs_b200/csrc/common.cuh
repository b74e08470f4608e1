// Shared device helpers, the workspace layout and error plumbing for libmonogs_b200.
// Reference paths are relative to /root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/monogs_b200.h"

namespace mgs {

constexpr int TILE_X = 16;   // config.h:16-17 — the tile size is part of the parity contract (keys, ranges)
constexpr int TILE_Y = 16;
constexpr int TILE_PIX = TILE_X * TILE_Y;
// per-tile atomic counters sit one per 128-B line: L2 serialises atomics per line, so densely packed counters
// (38 lines for 1200 tiles) cap the whole chip at a few atomics per clock
constexpr int TILE_CTR_STRIDE = 32;

// ---- error plumbing -------------------------------------------------------------------------
void set_error(const char* fmt, ...);
#define MGS_CUDA_TRY(expr)                                                              \
  do {                                                                                  \
    cudaError_t e__ = (expr);                                                           \
    if (e__ != cudaSuccess) {                                                           \
      mgs::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,               \
                     cudaGetErrorString(e__));                                          \
      return MGS_ECUDA;                                                                 \
    }                                                                                   \
  } while (0)
#define MGS_LAUNCH_CHECK() MGS_CUDA_TRY(cudaGetLastError())

// ---- optional per-kernel CUDA-event timing (mgs_profile_enable / mgs_profile_report); off by default and
//      never used under graph capture.  bench.py uses it for the live roofline numbers. ----
void prof_begin(const char* name, cudaStream_t s);
void prof_end(cudaStream_t s);
#define MGS_PROF(name, s, ...) do { mgs::prof_begin(name, s); __VA_ARGS__; mgs::prof_end(s); } while (0)

// ---- per-instance record: what a tile's blend loop consumes, contiguous per tile after the sort ----
// 48 B, 16-B aligned so a tile's list [start,end) is one legal cp.async.bulk source.
struct __align__(16) SplatRec {
  float x, y, depth; uint32_t id;       // pixel-space mean, view depth (forward.cu:254), Gaussian index
  float cx, cy, cz, opacity;            // conic (inverse 2D covariance) + opacity (forward.cu:257)
  float r, g, b; uint32_t extent;       // colour; half2 (hx, hy): half-extents of the alpha >= 1/255 ellipse
};
static_assert(sizeof(SplatRec) == 48, "SplatRec must stay 48 bytes");

struct WsHeader {
  uint32_t num_rendered;   // last element of the inclusive scan (rasterizer_impl.cu:284)
  uint32_t overflow;       // 1 when num_rendered > R_cap
  uint32_t r_used;         // min(num_rendered, R_cap): instances actually binned
  uint32_t n_large;        // tiles whose list exceeds 1024 words; their ids are tile_ctr[i * TILE_CTR_STRIDE + 2], i < n_large
  // sticky words: only ever raised by the kernels, cleared by mgs_raster_status_reset.  A loop that reuses one
  // workspace for many renders (PoseTracker, MapEngine: several views per workspace and iteration) reads them once in a
  // while instead of after every render, and still learns about ANY render that did not fit.
  uint32_t overflow_sticky;   // 1 once any binning into this workspace saw num_rendered > R_cap
  uint32_t max_rendered;      // largest num_rendered any binning into this workspace saw
  uint32_t ticket;            // CTA ticket of the folded scan (preprocess_kernel<.., FOLD_SCAN>); zero between launches
  uint32_t max_tile_count;    // sticky: longest tile list (instances of one tile) any binning into this workspace saw
  uint32_t pad[56];
};
static_assert(sizeof(WsHeader) == 256, "header is one 256-B slot");

// Host-side view of the workspace.  Everything backward needs lives at offsets that do not depend on
// R_cap (header, per-Gaussian state, image state, then the records), so backward never needs R_cap.
struct WsLayout {
  size_t header, depths, means2D, conic_opacity, rgb4, tiles_touched;
  size_t tile_ctr;                                // u32[tiles][TILE_CTR_STRIDE]: [0] instances per tile, [1] scatter cursor,
                                                  // [2] compact list of large tiles
  size_t final_T, n_contrib, ranges, records;
  size_t list;                                    // u64[R_cap]: (depth bits << 32 | Gaussian id), grouped by tile
  size_t total;
};
// pure host arithmetic (no CUDA call)
int ws_layout(int32_t P, int32_t W, int32_t H, int64_t R_cap, WsLayout* out);

template <typename T> __host__ __device__ inline T* ws_ptr(void* base, size_t off) {
  return reinterpret_cast<T*>(reinterpret_cast<char*>(base) + off);
}
template <typename T> __host__ __device__ inline const T* ws_ptr(const void* base, size_t off) {
  return reinterpret_cast<const T*>(reinterpret_cast<const char*>(base) + off);
}

// ---- small column-major 3x3 (m[c][r]); the product and transpose follow the exact expression order of
//      the matrix type the reference kernels use, so fp32 contraction yields the same bits ----
struct M3 {
  float m[3][3];
  __device__ __forceinline__ float* operator[](int c) { return m[c]; }
  __device__ __forceinline__ const float* operator[](int c) const { return m[c]; }
};
__device__ __forceinline__ M3 m3_cols(float a0, float a1, float a2, float b0, float b1, float b2,
                                      float c0, float c1, float c2) {
  M3 r;
  r.m[0][0] = a0; r.m[0][1] = a1; r.m[0][2] = a2;
  r.m[1][0] = b0; r.m[1][1] = b1; r.m[1][2] = b2;
  r.m[2][0] = c0; r.m[2][1] = c1; r.m[2][2] = c2;
  return r;
}
__device__ __forceinline__ M3 m3_mul(const M3& A, const M3& B) {
  M3 R;
#pragma unroll
  for (int c = 0; c < 3; c++)
#pragma unroll
    for (int r = 0; r < 3; r++)
      R.m[c][r] = A.m[0][r] * B.m[c][0] + A.m[1][r] * B.m[c][1] + A.m[2][r] * B.m[c][2];
  return R;
}
__device__ __forceinline__ M3 m3_t(const M3& A) {
  M3 R;
#pragma unroll
  for (int c = 0; c < 3; c++)
#pragma unroll
    for (int r = 0; r < 3; r++) R.m[c][r] = A.m[r][c];
  return R;
}

// ---- point transforms with the reference's accumulation order (auxiliary.h:58-87) ----
__device__ __forceinline__ float3 xform4x3(const float3& p, const float* __restrict__ mat) {
  float3 t = {
      mat[0] * p.x + mat[4] * p.y + mat[8] * p.z + mat[12],
      mat[1] * p.x + mat[5] * p.y + mat[9] * p.z + mat[13],
      mat[2] * p.x + mat[6] * p.y + mat[10] * p.z + mat[14],
  };
  return t;
}
__device__ __forceinline__ float4 xform4x4(const float3& p, const float* __restrict__ mat) {
  float4 t = {
      mat[0] * p.x + mat[4] * p.y + mat[8] * p.z + mat[12],
      mat[1] * p.x + mat[5] * p.y + mat[9] * p.z + mat[13],
      mat[2] * p.x + mat[6] * p.y + mat[10] * p.z + mat[14],
      mat[3] * p.x + mat[7] * p.y + mat[11] * p.z + mat[15]};
  return t;
}

// NDC -> pixel; evaluated in double like the reference (auxiliary.h:41-44).
__device__ __forceinline__ float ndc_to_pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// Tile rectangle of a splat (auxiliary.h:46-56): float division, truncation, clamp to the grid.
__device__ __forceinline__ void tile_rect(float2 p, int max_radius, int gx, int gy, uint2& rmin, uint2& rmax) {
  rmin.x = min(gx, max(0, (int)((p.x - max_radius) / TILE_X)));
  rmin.y = min(gy, max(0, (int)((p.y - max_radius) / TILE_Y)));
  rmax.x = min(gx, max(0, (int)((p.x + max_radius + TILE_X - 1) / TILE_X)));
  rmax.y = min(gy, max(0, (int)((p.y + max_radius + TILE_Y - 1) / TILE_Y)));
}

// 1 / x for a normal positive x far from the ends of the exponent range (1 - alpha in [0.01, 1]; a positive-definite
// conic entry): the bare MUFU.RCP.  __fdividef(1, x) wraps the same
// instruction in a range fix-up for |x| > 2^126 (compare, select, conditional rescale: 4 more issue slots per call)
// that can never trigger here; the value is the same.
__device__ __forceinline__ float rcp_unit(const float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// log2 of a normal positive x (255 * opacity > 1): the bare MUFU.LG2, without __logf's denormal rescue
__device__ __forceinline__ float lg2_normal(const float x) {
  float r;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// Half-extents (in pixels, rounded UP to fp16) of the axis-aligned box around the region where this splat can
// pass the blend's `alpha >= 1/255` test (forward.cu:352-355): alpha = op*exp(power) >= 1/255  <=>
// d^T conic d <= 2 ln(255 op), an ellipse whose box is sqrt(2 ln(255 op) * cov_xx/yy), cov = conic^-1.
// A warp patch that misses the box holds only pixels the reference would skip, so culling it is exact; the
// 0.2 % + 0.01 px margin dwarfs fp32 rounding of power/exp.  Non-PD conics get an infinite box (never culled),
// op <= 1/255 gets a negative one (always culled: alpha <= op < 1/255 everywhere).
__device__ __forceinline__ uint32_t splat_extent(float cx, float cy, float cz, float op) {
  float hx, hy;
  const float det = cx * cz - cy * cy;
  if (!(op * 255.f > 1.f)) { hx = hy = -1.f; }
  else if (!(det > 0.f) || !(cx > 0.f) || !(cz > 0.f)) { hx = hy = __int_as_float(0x7f800000); }
  else {
    const float two_tau = 2.f * logf(255.f * op);
    hx = sqrtf(two_tau * cz / det) * 1.002f + 0.01f;
    hy = sqrtf(two_tau * cx / det) * 1.002f + 0.01f;
    if (!(hx == hx)) hx = __int_as_float(0x7f800000);
    if (!(hy == hy)) hy = __int_as_float(0x7f800000);
  }
  const __half2 h = __halves2half2(__float2half_ru(hx), __float2half_ru(hy));
  return *reinterpret_cast<const uint32_t*>(&h);
}
// lane-parallel cull test of record `rec` (12 floats) against the pixel box [x0,x1] x [y0,y1]: the half-extent box
// first, then exactly — the minimum of f(d) = d^T conic d over the box against 2 ln(255 op) (alpha >= 1/255 <=>
// f <= 2 ln(255 op), forward.cu:352-355).  f is a positive-definite quadratic centred on the splat, so when the centre
// lies outside the box the minimum sits on the edge(s) facing it: at most two 1-D clamped minimisations.  The box
// holds every pixel centre of the patch, so a patch culled here holds only pixels the reference would skip.
__device__ __forceinline__ bool splat_hits(const float* rec, float x0, float x1, float y0, float y1) {
  const float x = rec[0], y = rec[1];
  const uint32_t e = __float_as_uint(rec[11]);
  const float2 h = __half22float2(*reinterpret_cast<const __half2*>(&e));
  if (!((x + h.x >= x0) && (x - h.x <= x1) && (y + h.y >= y0) && (y - h.y <= y1))) return false;
  const float dxe = fminf(fmaxf(x, x0), x1) - x, dye = fminf(fmaxf(y, y0), y1) - y;   // centre -> nearest box point
  if (dxe == 0.f && dye == 0.f) return true;
  const float a = rec[4], b = rec[5], c = rec[6], op = rec[7];
  if (!(a > 0.f) || !(c > 0.f) || !(a * c - b * b > 0.f)) return true;   // not positive definite: never culled
  float fmin = __int_as_float(0x7f800000);
  if (dxe != 0.f) {
    const float t = fminf(fmaxf(y - b * dxe * rcp_unit(c), y0), y1) - y;
    fmin = a * dxe * dxe + 2.f * b * dxe * t + c * t * t;
  }
  if (dye != 0.f) {
    const float s = fminf(fmaxf(x - b * dye * rcp_unit(a), x0), x1) - x;
    fmin = fminf(fmin, a * s * s + 2.f * b * s * dye + c * dye * dye);
  }
  const float thr = 1.3862943611198906f * lg2_normal(255.f * op);   // 2 ln(255 op); op > 1/255 here (the half-extent box of op <= 1/255 is empty)
  return !(fmin > thr * 1.001f + 1e-3f);         // NaN -> not culled
}

// exp() of the blend exponent.  FAST = false: expf, bit-identical to the reference's images.  FAST = true (fused
// tracking iteration only): one FMUL + MUFU.EX2 (ex2.approx.ftz), relative error ~4e-7 — far inside the 1e-4 image
// tolerance; forward and backward of one iteration always use the same variant, so T is recovered consistently.
template <bool FAST> __device__ __forceinline__ float blend_exp(float power) {
  if (FAST) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(power * 1.4426950408889634f));
    return r;
  }
  return expf(power);
}

// SH basis constants (auxiliary.h:21-39)
__device__ constexpr float kSH0 = 0.28209479177387814f;
__device__ constexpr float kSH1 = 0.4886025119029199f;
__device__ constexpr float kSH2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                      -1.0925484305920792f, 0.5462742152960396f};
__device__ constexpr float kSH3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                      0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                      -0.5900435899266435f};

struct ViewConst {  // by-value kernel parameter: scalars + the pointers of mgs_view
  int W, H, gx, gy;
  float tanfovx, tanfovy, focal_x, focal_y, scale_modifier;
  int D, M, prefiltered;
  const float* bg; const float* view; const float* proj; const float* proj_raw; const float* campos;
};
inline ViewConst make_view_const(const mgs_view* v) {
  ViewConst c;
  c.W = v->image_width; c.H = v->image_height;
  c.gx = (c.W + TILE_X - 1) / TILE_X; c.gy = (c.H + TILE_Y - 1) / TILE_Y;
  c.tanfovx = v->tanfovx; c.tanfovy = v->tanfovy;
  c.focal_y = c.H / (2.0f * v->tanfovy);   // rasterizer_impl.cu:225-226
  c.focal_x = c.W / (2.0f * v->tanfovx);
  c.scale_modifier = v->scale_modifier;
  c.D = v->sh_degree; c.M = v->sh_coeffs; c.prefiltered = v->prefiltered;
  c.bg = v->bg; c.view = v->viewmatrix; c.proj = v->projmatrix; c.proj_raw = v->projmatrix_raw;
  c.campos = v->campos;
  return c;
}

// ---- programmatic dependent launch (PDL) along the fixed kernel chains of the fused entry points -----------------------
// A tracking iteration is seven dependent kernels of 7-200 us, a mapping view eleven: the drain + launch latency between
// two of them is a measurable share.  Every kernel of those chains starts with pdl_enter(): it lets the NEXT kernel of
// the stream be scheduled as soon as all CTAs of this one have started (griddepcontrol.launch_dependents), then blocks
// until the PREVIOUS kernel has completed and its writes are visible (griddepcontrol.wait) — the dependency itself is the
// full stream-order one, only the launch latency of the successor is hidden behind the predecessor.  Both instructions
// are no-ops in a kernel that was launched without the attribute.  Completion is transitive: kernel k+1 completes only
// after its CTAs passed their wait, i.e. after kernel k completed, so a wait on k+1 also orders everything before it.
// Host side: a fused entry point opens a PdlScope; launch_chain() adds the attribute when the previous operation the
// scope put on the stream was one of its own kernels (state 2) and launches normally otherwise (first kernel, or after
// a memset / copy, which call pdl_break() through the wrappers below).  launch_chain() must only ever launch kernels
// that begin with pdl_enter().  MGS_PDL=0 switches the attribute off (A/B measurements); the event profiler does too.
// MGS_PDL_MASK: which kernels release their successor at their START (bits below); the others release it when they exit.
// Measured on the B200 (profiles/r02_notes.txt item 14): releasing early pays only where the successor is small or the
// stream is alone on the GPU — CTAs of a large successor that sit in their wait hold registers and shared memory that the
// mapping engine's other streams need (all bits set: mapping window of 8 views 441 -> 371 it/s).
#ifndef MGS_PDL_MASK
#define MGS_PDL_MASK (PDL_PRE | PDL_SCAN | PDL_EMIT | PDL_TRK)
#endif
enum : int { PDL_PRE = 1, PDL_SCAN = 2, PDL_EMIT = 4, PDL_SORT = 8, PDL_SORTL = 16, PDL_TRK = 32, PDL_UPD = 64, PDL_FWD = 128,
             PDL_LOSSF = 256, PDL_BWD = 512, PDL_PBWD = 1024 };
template <int BIT> __device__ __forceinline__ void pdl_enter() {
  if (MGS_PDL_MASK & BIT) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
extern thread_local int g_pdl;      // 0: no scope; 1: in a scope, the last operation was not a chained kernel; 2: it was
bool pdl_allowed();
struct PdlScope {
  bool outer;
  PdlScope() : outer(g_pdl == 0) { if (outer) g_pdl = pdl_allowed() ? 1 : 0; }
  ~PdlScope() { if (outer) g_pdl = 0; }
};
inline void pdl_break() { if (g_pdl == 2) g_pdl = 1; }
inline cudaError_t memset_async(void* p, int value, size_t bytes, cudaStream_t s) { pdl_break(); return cudaMemsetAsync(p, value, bytes, s); }
inline cudaError_t memcpy_async(void* dst, const void* src, size_t bytes, cudaMemcpyKind kind, cudaStream_t s) {
  pdl_break();
  return cudaMemcpyAsync(dst, src, bytes, kind, s);
}
template <typename... P, typename... A>
inline cudaError_t launch_chain(void (*kern)(P...), const dim3 grid, const dim3 block, const size_t smem, cudaStream_t s, A&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at = {};
  at.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at.val.programmaticStreamSerializationAllowed = 1;
  if (g_pdl == 2) { cfg.attrs = &at; cfg.numAttrs = 1; }
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, static_cast<P>(args)...);
  if (g_pdl == 1) g_pdl = 2;
  return e;
}

// number of significant bits of the tile id (rasterizer_impl.cu:35-50)
// — the reference's bisection returns the bit length of n (and 1 for n == 0).
inline uint32_t tile_id_bits(uint32_t n) { return n ? 32u - (uint32_t)__builtin_clz(n) : 1u; }

}  // namespace mgs
