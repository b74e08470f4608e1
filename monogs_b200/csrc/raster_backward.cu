// Full backward of the rasterizer (mapping path: every Gaussian gradient + pose tangent).
// Replaces BACKWARD::render (backward.cu:800-1027) and BACKWARD::preprocess (backward.cu:154-361,433-567).
//
// blend_backward_full: one CTA per tile, warp = 8x4 pixel patch, back-to-front over the tile's packed
// records (streamed with cp.async.bulk).  The reference pays >=11 block barriers per Gaussian per tile for
// a 256-thread tree reduction; here each warp that has at least one contributing lane reduces its 12
// partials with a 17-shuffle transposing butterfly, lanes add the warp totals into a per-batch shared
// accumulator, and after the batch every thread flushes one Gaussian with three 16-byte vector atomics.
// Two block barriers per 256 Gaussians.
#include "launchers.h"
#include "gaussian_backward.cuh"
#include "tma.cuh"
#include "blend_pair.cuh"
#include "sh_stage.cuh"

namespace mgs {

// per-Gaussian screen-space gradient accumulator (scratch, zeroed per call): 3 x float4
struct __align__(16) G2DGrad {
  float mx, my, cx, cy;     // dL/dmean2D (ndc-scaled, backward.cu:877-878), dL/dconic xx, xy
  float cw, op, dd, pad0;   // dL/dconic yy, dL/dopacity, dL/ddepth
  float r, g, b, pad1;      // dL/dcolor
};
static_assert(sizeof(G2DGrad) == 48, "G2DGrad is 48 bytes");

constexpr int BWD_BATCH = 128;     // records per batch = threads per CTA (one flush thread per record)
constexpr int BWD_THREADS = 128;

// Sum 10 per-lane values across the warp with a transposing butterfly: every exchange step halves the number of live
// values per lane (5+3+2+1+1 = 12 shuffles instead of 10*5).  Afterwards the lane whose reduce10_slot() is k holds the
// warp total of v[k]; other lanes hold duplicates.
__device__ __forceinline__ float warp_reduce10_transpose(const float (&v)[10], const int lane) {
  float b[5], c[3], d[2];
  {
    const bool hi = lane & 16;
#pragma unroll
    for (int i = 0; i < 5; i++) {
      const float keep = hi ? v[i + 5] : v[i], send = hi ? v[i] : v[i + 5];
      b[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
  }
  {
    const bool hi = lane & 8;
#pragma unroll
    for (int i = 0; i < 2; i++) {
      const float keep = hi ? b[i + 3] : b[i], send = hi ? b[i] : b[i + 3];
      c[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    c[2] = b[2] + __shfl_xor_sync(0xffffffffu, b[2], 8);
  }
  {
    const bool hi = lane & 4;
    const float keep = hi ? c[1] : c[0], send = hi ? c[0] : c[1];
    d[0] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    d[1] = c[2] + __shfl_xor_sync(0xffffffffu, c[2], 4);
  }
  const bool hi = lane & 2;
  const float keep = hi ? d[1] : d[0], send = hi ? d[0] : d[1];
  const float e = keep + __shfl_xor_sync(0xffffffffu, send, 2);
  return e + __shfl_xor_sync(0xffffffffu, e, 1);
}
// index of the value whose total this lane owns after warp_reduce10_transpose, or -1
__device__ __forceinline__ int reduce10_slot(const int lane) {
  if (lane & 1) return -1;
  const int half = 5 * (lane >> 4);
  if (lane & 2) return (lane & 12) ? -1 : 2 + half;
  return ((lane & 4) ? 1 : 0) + ((lane & 8) ? 3 : 0) + half;
}

// Two pixels per thread (same column, rows y and y+4 of the warp's 8x8 patch) with packed fp32, see blend_pair.cuh.
// A pixel that skips a splat (outside, behind its last contributor, power > 0 or alpha < 1/255) takes part with
// alpha = 0, which leaves its recurrences untouched exactly:  T / (1 - 0) = T,  and the pending fold
// accum = last_alpha * last_colour + (1 - last_alpha) * accum  is simply applied one splat early and followed by a
// fold with last_alpha = 0.  So the per-pixel updates run unconditionally for both halves, and the "last colour" is
// the colour of the previous record the warp processed: warp-uniform, not per-pixel state.
template <bool FAST>   // FAST: the ex2.approx exponent of a FAST forward (T is recovered with the same alpha)
#ifndef MGS_BWD_MIN_CTAS
#define MGS_BWD_MIN_CTAS 9     // 56 registers, no spills: 9 CTAs per SM instead of 8 (+1 % on the mapping iteration; 10 spills, 12 is 10 % slower)
#endif
__global__ void __launch_bounds__(BWD_THREADS, MGS_BWD_MIN_CTAS)
blend_backward_full_kernel(const int W, const int H, const int gx, const uint32_t* __restrict__ tile_order,
                           const uint2* __restrict__ ranges, const SplatRec* __restrict__ recs,
                           const float* __restrict__ bg, const float* __restrict__ final_T,
                           const uint32_t* __restrict__ n_contrib,
                           const float* __restrict__ dL_dpix, const float* __restrict__ dL_dpix_depth,
                           const int32_t* __restrict__ pixel_mult, G2DGrad* __restrict__ g2d) {
  __shared__ __align__(128) SplatRec stage[2][BWD_BATCH];
  __shared__ __align__(16) float acc[BWD_BATCH * 12];
  __shared__ __align__(8) uint64_t full_bar[2];
  __shared__ int s_max_contrib;

  pdl_enter<PDL_BWD>();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile = tile_order != nullptr ? (int)tile_order[(size_t)blockIdx.x * TILE_CTR_STRIDE + 3] : (int)blockIdx.x;   // heavy first
  const int tx = tile % gx, ty = tile / gx;
  const int px = tx * TILE_X + (warp & 1) * 8 + (lane & 7);
  const int py0 = ty * TILE_Y + (warp >> 1) * 8 + (lane >> 3), py1 = py0 + 4;
  const bool in0 = px < W && py0 < H, in1 = px < W && py1 < H;
  const size_t pix0 = (size_t)py0 * W + px, pix1 = (size_t)py1 * W + px, plane = (size_t)W * H;
  const float2 npx = make_float2(-(float)px, -(float)px), npy = make_float2(-(float)py0, -(float)py1);
  const float patch_x0 = (float)(tx * TILE_X + (warp & 1) * 8), patch_y0 = (float)(ty * TILE_Y + (warp >> 1) * 8);
  const uint2 range = ranges[tile];

  const float2 T_final = make_float2(in0 ? final_T[pix0] : 0.f, in1 ? final_T[pix1] : 0.f);
  int last0 = in0 ? (int)n_contrib[pix0] : 0, last1 = in1 ? (int)n_contrib[pix1] : 0;
  float2 dp0 = make_float2(0.f, 0.f), dp1 = dp0, dp2 = dp0, dpd = dp0;
  if (in0) { dp0.x = dL_dpix[pix0]; dp1.x = dL_dpix[plane + pix0]; dp2.x = dL_dpix[2 * plane + pix0]; dpd.x = dL_dpix_depth[pix0]; }
  if (in1) { dp0.y = dL_dpix[pix1]; dp1.y = dL_dpix[plane + pix1]; dp2.y = dL_dpix[2 * plane + pix1]; dpd.y = dL_dpix_depth[pix1]; }
  if (pixel_mult != nullptr) {
    // pixel-sample backward (the reference's renderPixelSampleCUDA, backward.cu:592-797): only the listed pixels send
    // gradient, a pixel listed k times sends it k times.  A pixel that is not listed gets "no contributor" — every splat
    // is then skipped for it by the test below, the tile's walk starts at the deepest contributor of its LISTED pixels,
    // and a tile without any returns at once.
    const int m0 = in0 ? pixel_mult[pix0] : 0, m1 = in1 ? pixel_mult[pix1] : 0;
    last0 = m0 > 0 ? last0 : 0; last1 = m1 > 0 ? last1 : 0;
    const float2 mf = make_float2((float)m0, (float)m1);
    dp0 = __fmul2_rn(dp0, mf); dp1 = __fmul2_rn(dp1, mf); dp2 = __fmul2_rn(dp2, mf); dpd = __fmul2_rn(dpd, mf);
  }
  // a warp none of whose 64 pixels has a contributor (empty patch, or no listed pixel) has nothing to walk
  const bool warp_live = __ballot_sync(0xffffffffu, last0 > 0 || last1 > 0) != 0u;
  if (tid == 0) {
    s_max_contrib = 0;
    mbar_init(&full_bar[0], 1); mbar_init(&full_bar[1], 1);
    mbar_init_fence();
  }
  __syncthreads();
  {
    int m = max(last0, last1);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0 && m > 0) atomicMax(&s_max_contrib, m);
  }
  __syncthreads();
  // nothing behind the deepest contributor of this tile can receive gradient (it would be skipped by
  // `contributor >= last_contributor`, backward.cu:919), so the walk starts there
  const int n = min(s_max_contrib, (int)(range.y - range.x));
  const int rounds = (n + BWD_BATCH - 1) / BWD_BATCH;
  const SplatRec* src = recs + range.x;

  auto issue = [&](int i) {  // batch i covers list positions [lo, hi), walked from hi-1 down to lo
    const int hi = n - i * BWD_BATCH, lo = max(0, hi - BWD_BATCH);
    const uint32_t bytes = (uint32_t)(hi - lo) * sizeof(SplatRec);
    mbar_expect_tx(&full_bar[i & 1], bytes);
    bulk_g2s(&stage[i & 1][0], src + lo, bytes, &full_bar[i & 1]);
  };
  if (tid == 0 && rounds > 0) issue(0);

  float2 T = T_final;
  const float2 zero2 = make_float2(0.f, 0.f);
  float2 ar0 = zero2, ar1 = zero2, ar2 = zero2, ard = zero2, last_alpha = zero2;   // accum_rec colour / depth
  float lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, ld = 0.f;                                    // last processed record (warp-uniform)
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
  const float b0 = bg[0], b1 = bg[1], b2 = bg[2];
  const float2 bg_dot_dpixel = make_float2(b0 * dp0.x + b1 * dp1.x + b2 * dp2.x, b0 * dp0.y + b1 * dp1.y + b2 * dp2.y);
  const float2 neg_T_final = neg2(T_final);
  const int slot = reduce10_slot(lane);

  for (int i = 0; i < rounds; i++) {
    // zero this thread's accumulator slot; the barrier also retires all reads of the other stage
#pragma unroll
    for (int k = 0; k < 3; k++) reinterpret_cast<float4*>(acc)[tid * 3 + k] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    if (tid == 0 && i + 1 < rounds) issue(i + 1);
    mbar_wait(&full_bar[i & 1], (i >> 1) & 1);
    const int hi = n - i * BWD_BATCH, lo = max(0, hi - BWD_BATCH);
    const int cnt = hi - lo;
    const float4* rec4 = reinterpret_cast<const float4*>(&stage[i & 1][0]);
    const float* recf = reinterpret_cast<const float*>(&stage[i & 1][0]);
    for (int c0 = warp_live ? ((cnt - 1) >> 5) << 5 : -1; c0 >= 0; c0 -= 32) {
     // lane-parallel exact cull of 32 splats against this warp's 8x8 patch (see splat_extent)
     const int jj = c0 + lane;
     unsigned hits = __ballot_sync(0xffffffffu, jj < cnt && splat_hits(recf + 12 * jj, patch_x0, patch_x0 + 7.f, patch_y0, patch_y0 + 7.f));
     while (hits) {
      const int j = c0 + 31 - __clz(hits);
      hits &= ~(1u << (j - c0));
      const int contributor = lo + j;  // list position of this Gaussian
      const float4 g0 = rec4[3 * j + 0];
      const float4 g1 = rec4[3 * j + 1];
      float2 dx, dy, power, G, alpha;
      pair_alpha<FAST>(g0, g1, npx, npy, dx, dy, power, G, alpha);
      const bool skip0 = !in0 || contributor >= last0 || power.x > 0.0f || alpha.x < 1.0f / 255.0f;
      const bool skip1 = !in1 || contributor >= last1 || power.y > 0.0f || alpha.y < 1.0f / 255.0f;
      if (__ballot_sync(0xffffffffu, !(skip0 && skip1)) == 0u) continue;
      const float4 g2 = rec4[3 * j + 2];
      const float2 ae = make_float2(skip0 ? 0.f : alpha.x, skip1 ? 0.f : alpha.y);
      const float2 oma = one_minus2(ae);
      // 1-alpha in [0.01,1]: MUFU.RCP, 1 ulp vs the IEEE divide; exactly 1 for a skipped pixel
      const float2 inv_1ma = make_float2(rcp_unit(oma.x), rcp_unit(oma.y));
      T = __fmul2_rn(T, inv_1ma);
      const float2 w = __fmul2_rn(ae, T);                        // dchannel_dcolor
      const float2 oml = one_minus2(last_alpha);
      ar0 = __ffma2_rn(last_alpha, bcast2(lc0), __fmul2_rn(oml, ar0));
      ar1 = __ffma2_rn(last_alpha, bcast2(lc1), __fmul2_rn(oml, ar1));
      ar2 = __ffma2_rn(last_alpha, bcast2(lc2), __fmul2_rn(oml, ar2));
      ard = __ffma2_rn(last_alpha, bcast2(ld), __fmul2_rn(oml, ard));
      float2 dL_dalpha = __fmul2_rn(__fadd2_rn(bcast2(g2.x), neg2(ar0)), dp0);
      dL_dalpha = __ffma2_rn(__fadd2_rn(bcast2(g2.y), neg2(ar1)), dp1, dL_dalpha);
      dL_dalpha = __ffma2_rn(__fadd2_rn(bcast2(g2.z), neg2(ar2)), dp2, dL_dalpha);
      dL_dalpha = __ffma2_rn(__fadd2_rn(bcast2(g0.z), neg2(ard)), dpd, dL_dalpha);
      lc0 = g2.x; lc1 = g2.y; lc2 = g2.z; ld = g0.z;
      last_alpha = ae;
      dL_dalpha = __fmul2_rn(dL_dalpha, T);
      dL_dalpha = __ffma2_rn(__fmul2_rn(neg_T_final, inv_1ma), bg_dot_dpixel, dL_dalpha);
      // per-lane moments of u = G * dL/dalpha about the Gaussian centre and the colour / depth weights; everything that
      // is constant per Gaussian (opacity, conic, the 0.5 W / 0.5 H factors) is applied once, after the sums, by the flush
      float2 u = __fmul2_rn(G, dL_dalpha);
      u.x = skip0 ? 0.f : u.x; u.y = skip1 ? 0.f : u.y;
      const float2 ux = __fmul2_rn(u, dx), uy = __fmul2_rn(u, dy);
      const float2 uxx = __fmul2_rn(ux, dx), uxy = __fmul2_rn(ux, dy), uyy = __fmul2_rn(uy, dy);
      const float2 wd = __fmul2_rn(w, dpd), w0 = __fmul2_rn(w, dp0), w1 = __fmul2_rn(w, dp1), w2 = __fmul2_rn(w, dp2);
      const float v[10] = {u.x + u.y, ux.x + ux.y, uy.x + uy.y, uxx.x + uxx.y, uxy.x + uxy.y, uyy.x + uyy.y,
                           wd.x + wd.y, w0.x + w0.y, w1.x + w1.y, w2.x + w2.y};
      const float tot = warp_reduce10_transpose(v, lane);
      if (slot >= 0) atomicAdd(&acc[j * 12 + slot], tot);
     }
    }
    __syncthreads();
    if (tid < cnt) {
      const float4 m0 = reinterpret_cast<const float4*>(acc)[tid * 3 + 0];   // sum u, u dx, u dy, u dx dx
      const float4 m1 = reinterpret_cast<const float4*>(acc)[tid * 3 + 1];   // u dx dy, u dy dy, depth, colour r
      const float4 m2 = reinterpret_cast<const float4*>(acc)[tid * 3 + 2];   // colour g, b
      const float4 co = rec4[3 * tid + 1];                                    // conic xyz, opacity
      const uint32_t id = __float_as_uint(rec4[3 * tid + 0].w);
      // dL/dG = opacity * dL/dalpha (clamp at 0.99 deliberately ignored, Q6); dG/d(delta) and dG/d(conic) as in
      // backward.cu:986-1010 with the per-Gaussian factors taken out of the pixel sums
      const float hop = -0.5f * co.w;
      const float4 a0 = make_float4(-co.w * (co.x * m0.y + co.y * m0.z) * ddelx_dx,
                                    -co.w * (co.z * m0.z + co.y * m0.y) * ddely_dy, hop * m0.w, hop * m1.x);
      const float4 a1 = make_float4(hop * m1.y, m0.x, m1.z, 0.f);
      const float4 a2 = make_float4(m1.w, m2.x, m2.y, 0.f);
      float4* dst = reinterpret_cast<float4*>(g2d + id);
      if (a0.x != 0.f || a0.y != 0.f || a0.z != 0.f || a0.w != 0.f) atomicAdd(dst + 0, a0);
      if (a1.x != 0.f || a1.y != 0.f || a1.z != 0.f) atomicAdd(dst + 1, a1);
      if (a2.x != 0.f || a2.y != 0.f || a2.z != 0.f) atomicAdd(dst + 2, a2);
    }
    // the next iteration's zeroing touches only this thread's own slot, and its barrier orders the flush
    // reads of rec4 before stage reuse two iterations later
  }
}


// Block-wide sum of a 6-vector, one atomicAdd per block per component.
__device__ __forceinline__ void block_add6(float (&t)[6], float* __restrict__ out6) {
  __shared__ float part[8][6];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 6; k++) {
    float x = t[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) part[warp][k] = x;
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    float x = 0.f;
    const int nw = (blockDim.x + 31) >> 5;
    for (int w = 0; w < nw; w++) x += part[w][threadIdx.x];
    if (x != 0.f) atomicAdd(&out6[threadIdx.x], x);
  }
}

// K9 + K10 fused: one thread per Gaussian, screen-space gradients -> all parameter gradients + pose tangent.
// STAGE: SH rows in, SH gradient rows out through shared memory with coalesced 16-byte accesses (sh_stage.cuh).
template <bool STAGE>
__global__ void __launch_bounds__(256, 3)   // 80 registers: the kernel is latency-bound, occupancy pays (C5: 474 -> 376 us)
preprocess_backward_kernel(const ViewConst v, const int P, const uint32_t mask,
                           const float* __restrict__ means3D, const float* __restrict__ shs,
                           const float* __restrict__ sh_rest, float* __restrict__ dL_dsh_rest,
                           const float* __restrict__ colors_precomp, const float* __restrict__ scales,
                           const float* __restrict__ rotations, const float* __restrict__ cov3D_precomp,
                           const int32_t* __restrict__ radii, const float4* __restrict__ rgb4,
                           const G2DGrad* __restrict__ g2d,
                           float* __restrict__ dL_dmeans3D, float* __restrict__ dL_dmeans2D,
                           float* __restrict__ dL_dcolors_or_sh, float* __restrict__ dL_dopacity,
                           float* __restrict__ dL_dscales, float* __restrict__ dL_drotations,
                           float* __restrict__ dL_dcov3D, float* __restrict__ tau_sum, float* __restrict__ tau_pg,
                           const int accumulate, float* __restrict__ stat_grad_accum, float* __restrict__ stat_denom,
                           float* __restrict__ stat_max_radii) {
  extern __shared__ __align__(16) float sh_tile[];
  pdl_enter<PDL_PBWD>();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const int RS = sh_row_stride(v.M);
  float* tile = sh_tile + (threadIdx.x >> 5) * 32 * RS;
  const int first = blockIdx.x * blockDim.x + (threadIdx.x & ~31);
  const int n_rows = min(32, P - first);
  const bool want_dsh = shs != nullptr && (mask & MGS_GRAD_COLOR);
  bool vis = false;
  if (idx < P) vis = radii[idx] > 0;
  const unsigned vis_mask = __ballot_sync(0xffffffffu, vis);
  if (STAGE && vis_mask != 0u) {   // warp-uniform
    if (sh_rest == nullptr) sh_stage_load(tile, RS, shs + (size_t)first * 3 * v.M, 3 * v.M, 0, n_rows, lane);
    else {
      sh_stage_load(tile, RS, shs + (size_t)first * 3, 3, 0, n_rows, lane);
      sh_stage_load(tile, RS, sh_rest + (size_t)first * 3 * (v.M - 1), 3 * (v.M - 1), 3, n_rows, lane);
    }
    __syncwarp();
  }
  float tau[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (idx < P) {
    const int rad = radii[idx];
    float3 dmean = {0.f, 0.f, 0.f}, dscale = {0.f, 0.f, 0.f};
    float4 drot = {0.f, 0.f, 0.f, 0.f};
    float dcov6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    G2DGrad gg;
    gg.mx = gg.my = gg.cx = gg.cy = gg.cw = gg.op = gg.dd = gg.r = gg.g = gg.b = 0.f;
    const int Msh = v.M;
    float* dsh = !want_dsh ? nullptr : (STAGE ? tile + lane * RS : dL_dcolors_or_sh + (size_t)idx * Msh * 3);
    const float* sh_row = STAGE ? tile + lane * RS : (shs != nullptr ? shs + (size_t)idx * Msh * 3 : nullptr);
    if (vis) {
      const float4* src = reinterpret_cast<const float4*>(g2d + idx);
      const float4 a0 = src[0], a1 = src[1], a2 = src[2];
      gg.mx = a0.x; gg.my = a0.y; gg.cx = a0.z; gg.cy = a0.w; gg.cw = a1.x; gg.op = a1.y; gg.dd = a1.z;
      gg.r = a2.x; gg.g = a2.y; gg.b = a2.z;
      const float3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
      GaussCtx g;
      if (cov3D_precomp != nullptr) {
#pragma unroll
        for (int k = 0; k < 6; k++) g.cov6[k] = cov3D_precomp[6 * (size_t)idx + k];
      } else cov3d_from_scale_rot_b(scales, rotations, idx, v.scale_modifier, g.cov6);
      gauss_ctx_init(g, v, p);
      ConicChainOut cc;
      conic_chain(g, v, gg.cx, gg.cy, gg.cw, cc);
#pragma unroll
      for (int k = 0; k < 6; k++) { tau[k] = cc.tau[k]; dcov6[k] = cc.dcov6[k]; }
      dmean = cc.dmean;
      ProjCtx q;
      proj_ctx_init(q, v, p);
      const float3 dm2 = mean2d_to_mean3d(q, v.proj, gg.mx, gg.my);
      dmean.x += dm2.x; dmean.y += dm2.y; dmean.z += dm2.z;
      float jx[6], jy[6];
      mean2d_tau_jac(q, v.proj_raw, jx, jy);
#pragma unroll
      for (int k = 0; k < 6; k++) tau[k] += gg.mx * jx[k] + gg.my * jy[k];
      // depth gradient (backward.cu:546-556)
      dmean.x += gg.dd * v.view[2]; dmean.y += gg.dd * v.view[6]; dmean.z += gg.dd * v.view[10];
      tau[2] += gg.dd;
      tau[3] += gg.dd * q.pC.y;
      tau[4] += gg.dd * -q.pC.x;
      if (shs != nullptr) {
        const uint32_t cm = __float_as_uint(rgb4[idx].w);
        const float3 dcol = {(cm & 1u) ? 0.f : gg.r, (cm & 2u) ? 0.f : gg.g, (cm & 4u) ? 0.f : gg.b};
        const float3 dms = sh_backward(v.D, Msh, p, v.campos, sh_row, dcol, dsh, STAGE ? 0 : accumulate);
        dmean.x += dms.x; dmean.y += dms.y; dmean.z += dms.z;
        tau[0] += -dms.x; tau[1] += -dms.y; tau[2] += -dms.z;
      }
      if (scales != nullptr && (mask & MGS_GRAD_COV)) cov3d_backward(scales, rotations, idx, v.scale_modifier, dcov6, dscale, drot);
    } else if (dsh != nullptr && (STAGE || !accumulate)) {   // staged rows of invisible Gaussians must read as zero
      for (int k = 0; k < 3 * Msh; k++) dsh[k] = 0.f;
    }
    if (accumulate) {
      if (vis) {   // sum over the views of a mapping window; Gaussians outside this view contribute nothing
        // accumulate == 2: backward passes of different views may run concurrently -> atomics
        const bool at = accumulate == 2;
        auto add = [at](float* q, const float x) { if (at) atomicAdd(q, x); else *q += x; };
        if (mask & MGS_GRAD_MEANS3D) { add(dL_dmeans3D + 3 * idx, dmean.x); add(dL_dmeans3D + 3 * idx + 1, dmean.y); add(dL_dmeans3D + 3 * idx + 2, dmean.z); }
        if (mask & MGS_GRAD_MEANS2D) { add(dL_dmeans2D + 3 * idx, gg.mx); add(dL_dmeans2D + 3 * idx + 1, gg.my); }
        if (mask & MGS_GRAD_OPACITY) add(dL_dopacity + idx, gg.op);
        if ((mask & MGS_GRAD_COLOR) && shs == nullptr) {
          add(dL_dcolors_or_sh + 3 * idx, gg.r); add(dL_dcolors_or_sh + 3 * idx + 1, gg.g); add(dL_dcolors_or_sh + 3 * idx + 2, gg.b);
        }
        if ((mask & MGS_GRAD_COV) && scales != nullptr) {
          add(dL_dscales + 3 * idx, dscale.x); add(dL_dscales + 3 * idx + 1, dscale.y); add(dL_dscales + 3 * idx + 2, dscale.z);
          if (at) atomicAdd(reinterpret_cast<float4*>(dL_drotations) + idx, drot);
          else {
            float4 d = reinterpret_cast<float4*>(dL_drotations)[idx];
            d.x += drot.x; d.y += drot.y; d.z += drot.z; d.w += drot.w;
            reinterpret_cast<float4*>(dL_drotations)[idx] = d;
          }
        }
        if (stat_grad_accum != nullptr) add(stat_grad_accum + idx, sqrtf(gg.mx * gg.mx + gg.my * gg.my));
        if (stat_denom != nullptr) add(stat_denom + idx, 1.f);
        if (stat_max_radii != nullptr) {   // radii >= 0: the float order is the order of the bit patterns
          if (at) atomicMax(reinterpret_cast<int*>(stat_max_radii) + idx, __float_as_int((float)rad));
          else stat_max_radii[idx] = fmaxf(stat_max_radii[idx], (float)rad);
        }
      }
    } else {
    if (mask & MGS_GRAD_MEANS3D) { dL_dmeans3D[3 * idx] = dmean.x; dL_dmeans3D[3 * idx + 1] = dmean.y; dL_dmeans3D[3 * idx + 2] = dmean.z; }
    if (mask & MGS_GRAD_MEANS2D) { dL_dmeans2D[3 * idx] = gg.mx; dL_dmeans2D[3 * idx + 1] = gg.my; dL_dmeans2D[3 * idx + 2] = 0.f; }
    if (mask & MGS_GRAD_OPACITY) dL_dopacity[idx] = gg.op;
    if ((mask & MGS_GRAD_COLOR) && shs == nullptr) {
      dL_dcolors_or_sh[3 * idx] = gg.r; dL_dcolors_or_sh[3 * idx + 1] = gg.g; dL_dcolors_or_sh[3 * idx + 2] = gg.b;
    }
    if (mask & MGS_GRAD_COV) {
      if (scales != nullptr) {
        dL_dscales[3 * idx] = dscale.x; dL_dscales[3 * idx + 1] = dscale.y; dL_dscales[3 * idx + 2] = dscale.z;
        reinterpret_cast<float4*>(dL_drotations)[idx] = drot;
      }
      if (dL_dcov3D != nullptr) {
#pragma unroll
        for (int k = 0; k < 6; k++) dL_dcov3D[6 * (size_t)idx + k] = dcov6[k];
      }
    }
    }
    if (tau_pg != nullptr) {
#pragma unroll
      for (int k = 0; k < 6; k++) tau_pg[6 * (size_t)idx + k] = tau[k];
    }
  }
  if (STAGE && want_dsh && n_rows > 0 && (accumulate == 0 || vis_mask != 0u)) {
    __syncwarp();
    if (sh_rest == nullptr)
      sh_stage_store(tile, RS, dL_dcolors_or_sh + (size_t)first * 3 * v.M, 3 * v.M, 0, n_rows, lane, accumulate, vis_mask);
    else {
      sh_stage_store(tile, RS, dL_dcolors_or_sh + (size_t)first * 3, 3, 0, n_rows, lane, accumulate, vis_mask);
      sh_stage_store(tile, RS, dL_dsh_rest + (size_t)first * 3 * (v.M - 1), 3 * (v.M - 1), 3, n_rows, lane, accumulate, vis_mask);
    }
  }
  if (tau_sum != nullptr) block_add6(tau, tau_sum);
}

size_t backward_scratch_bytes(int P, int sketch_dim) {
  (void)sketch_dim;
  return ((size_t)(P > 0 ? P : 1) * sizeof(G2DGrad) + 255) / 256 * 256 + 256;
}

// selected-pixel list -> multiplicity plane (indices outside [0, npix) are ignored: the reference's `inside` test)
__global__ void __launch_bounds__(256) pixel_multiplicity_kernel(const int32_t* __restrict__ idx, const int n, const int npix,
                                                                 int32_t* __restrict__ mult) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int p = idx[i];
  if (p >= 0 && p < npix) atomicAdd(&mult[p], 1);
}

int launch_pixel_multiplicity(const int32_t* indices, int n, int npix, int32_t* mult, cudaStream_t s) {
  MGS_CUDA_TRY(memset_async(mult, 0, (size_t)npix * sizeof(int32_t), s));
  if (n > 0) {
    MGS_PROF("pixel_multiplicity", s, pixel_multiplicity_kernel<<<(n + 255) / 256, 256, 0, s>>>(indices, n, npix, mult));
    MGS_LAUNCH_CHECK();
  }
  return MGS_OK;
}

size_t backward_g2d_bytes() { return sizeof(G2DGrad); }

int launch_backward_full(const BackwardArgs& a, cudaStream_t s) {
  const ViewConst& v = a.v;
  const int P = a.P;
  const size_t need = backward_scratch_bytes(P, 0);
  if (a.scratch == nullptr || a.scratch_bytes < need) { set_error("backward scratch too small: %zu < %zu", a.scratch_bytes, need); return MGS_ENOSPC; }
  G2DGrad* g2d = reinterpret_cast<G2DGrad*>(a.scratch);
  if (!a.pre_cleared) {   // (mgs_map_view_fwd_bwd clears both ahead of the view's forward, so that no memset sits between the kernels)
    MGS_CUDA_TRY(memset_async(g2d, 0, (size_t)P * sizeof(G2DGrad), s));
    if (a.dL_dtau_sum) MGS_CUDA_TRY(memset_async(a.dL_dtau_sum, 0, 6 * sizeof(float), s));
  }
  const int tiles = v.gx * v.gy;
  cudaError_t e;
  prof_begin("blend_backward_full", s);
  const uint32_t* order = a.heavy_first ? ws_ptr<uint32_t>(a.ws, a.L.tile_ctr) : nullptr;
#define MGS_BWD_ARGS v.W, v.H, v.gx, order, ws_ptr<uint2>(a.ws, a.L.ranges), ws_ptr<SplatRec>(a.ws, a.L.records), v.bg,   \
      ws_ptr<float>(a.ws, a.L.final_T), ws_ptr<uint32_t>(a.ws, a.L.n_contrib), a.dL_dcolor_img, a.dL_ddepth_img,        \
      a.pixel_mult, g2d
  if (a.fast) e = launch_chain(blend_backward_full_kernel<true>, tiles, BWD_THREADS, 0, s, MGS_BWD_ARGS);
  else e = launch_chain(blend_backward_full_kernel<false>, tiles, BWD_THREADS, 0, s, MGS_BWD_ARGS);
#undef MGS_BWD_ARGS
  prof_end(s);
  MGS_CUDA_TRY(e);
  MGS_LAUNCH_CHECK();
  prof_begin("preprocess_backward", s);
#define MGS_PB_ARGS v, P, a.grad_mask, a.means3D, a.shs, a.sh_rest, a.dL_dsh_rest, a.colors_precomp, a.scales, a.rotations,        \
      a.cov3D_precomp, a.radii, ws_ptr<float4>(a.ws, a.L.rgb4), g2d, a.dL_dmeans3D, a.dL_dmeans2D, a.dL_dcolors_or_sh,      \
      a.dL_dopacity, a.dL_dscales, a.dL_drotations, a.dL_dcov3D, a.dL_dtau_sum, a.dL_dtau_pg, a.accumulate,                \
      a.stat_grad_accum, a.stat_denom, a.stat_max_radii
  const bool stage = a.shs != nullptr && v.M > 1;
  if (a.sh_rest != nullptr && (!stage || ((a.grad_mask & MGS_GRAD_COLOR) && a.dL_dsh_rest == nullptr))) {
    set_error("backward: split SH layout needs sh_coeffs > 1 and dL_dsh_rest"); return MGS_EINVAL;
  }
  if (stage) {
    const int T = 128;
    e = launch_chain(preprocess_backward_kernel<true>, (P + T - 1) / T, T, (size_t)T * sh_row_stride(v.M) * sizeof(float), s, MGS_PB_ARGS);
  } else e = launch_chain(preprocess_backward_kernel<false>, (P + 255) / 256, 256, 0, s, MGS_PB_ARGS);
#undef MGS_PB_ARGS
  prof_end(s);
  MGS_CUDA_TRY(e);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

}  // namespace mgs
