// Single-pass tracking blend: forward alpha blend, tracking loss and the pose tangent of that loss in ONE walk over
// each tile's list (replaces renderCUDA forward.cu:263-392 + the loss of utils/slam_utils.py + renderCUDA backward
// backward.cu:800-1027 + the per-Gaussian chain, for the tracking case where only grad_rho / grad_theta are consumed).
//
// The reference's backward walks the list back to front because dC/dalpha_i needs the colour composited BEHIND splat i.
// With r_i = 1/(1-alpha_i), T_i the transmittance in front of i and Cpre_i the colour accumulated before i,
//     dL/dalpha_i = sum_ch g_ch * r_i * (T_i c_i,ch + Cpre_i,ch)  -  r_i * sum_ch g_ch * out_ch        (g = dL/d out)
// (same value as backward.cu:927-975 including the background term; out_ch = final pixel colour).  Everything that
// depends on the still unknown g and out factors out of the sum over splats, so a pixel accumulates, front to back,
//     E      = sum_i  vp_i                         vp_i = r_i * opacity_i * B_i^T geom_i   (6-vector, B_i = pose_jac.cuh)
//     A_ch   = sum_i (T_i c_i,ch + Cpre_i,ch) vp_i
// and, once its colour and the loss gradient g are known in the epilogue, contracts
//     tau_pixel = sum_ch g_ch * (A_ch - out_ch * E)        (+ the same for depth, + g_d * sum_i alpha_i T_i dz_i/dtau).
// The local Jacobians are exactly the reference's (clamp ignored, Q6; FoV quirk inside B_i, Q8), only the order of
// the sum differs.  One pass instead of two: no second exp per pair, no g / final_T / n_contrib round trip.
//
// vp_i is evaluated from the folded per-Gaussian basis of pose_jac.cuh (32 floats, one 128-byte line):
//     vp = s * (dx * (U + dx * A + dy * Bm) + dy * (V + dy * Cm)),   s = opacity * G / (1 - alpha)
// 15 packed multiply-adds; s is folded into the four weights of the accumulation (1, w_r, w_g, w_b).
#include "launchers.h"
#include "tma.cuh"
#include "loss_pixel.cuh"
#include "pose_jac.cuh"
#include "track_update.cuh"
#include <cstdlib>
#include <cstring>


namespace mgs {

constexpr int TRK_NF = POSE_BASIS_FLOATS;   // 32
constexpr int TRK_STAGES = 3;
// One CTA (8 warps) per 16x16 tile; records and bases are staged per batch of 128 splats by the whole CTA (50 KB of
// shared memory), one CTA barrier per batch.  MGS_TRK_MIN_CTAS = resident CTAs per SM the kernel is compiled for: 3 (80
// registers, no spill) or 4 (64 registers, a few spilled words, 32 warps per SM instead of 24).
// Variants that were built, measured on the B200 and dropped (profiles/r02_notes.txt): half-warp cursors over 4x4 blocks,
// two pixels per thread, half-tile CTAs, and per-warp record rings without any CTA barrier.
#ifndef MGS_TRK_MIN_CTAS
#define MGS_TRK_MIN_CTAS 3
#endif
#ifndef MGS_TRK_MIN_CTAS_DEPTH
#define MGS_TRK_MIN_CTAS_DEPTH 3   // the RGB-D variants carry 9 more accumulators
#endif
template <int NW> struct TrkCfg {
  static constexpr int kThreads = NW * 32;
  static constexpr int kBatch = 128;
  static constexpr size_t kSmem = (size_t)TRK_STAGES * kBatch * sizeof(SplatRec) + 2 * (size_t)kBatch * TRK_NF * sizeof(float);
  static constexpr size_t kSmemRows = 2 * (size_t)kBatch * sizeof(uint32_t) + 2 * (size_t)kBatch * TRACK_ROW_FLOATS * sizeof(float);
};

__device__ __forceinline__ void cp_async4(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

struct TrackTail {            // the folded update (FOLD): last CTA of the grid runs track_update_body
  TrackUpdateK k;
  uint32_t* ticket;           // device u32, zero before the first launch; reset by the last CTA
  SketchRoute route;          // EPI = 1: route.idx != null sends the pixel's row contribution straight to the sketch rows
};

// Pipeline per tile: records of batch i+2 in flight by bulk copy (TMA), pose bases of batch i+1 gathered by cp.async
// (their ids come from the records of i+1, already landed), batch i being blended: one CTA barrier (the early-out
// vote) per batch and no load latency on the blend path.
//
// A warp owns an 8x4 pixel patch and walks the splats of the batch that can reach it (lane-parallel exact cull).
// EPI = 0: the epilogue evaluates the first-order tracking loss and contracts the pose tangent.
// EPI = 1 (sketch): it evaluates the second-order (sketched Gauss-Newton) objective instead and writes this pixel's
//   Jacobian row contribution (tau[6], d/d exposure_a, d/d exposure_b, f) to sk_image[9][H][W]; the rows are formed
//   afterwards by sketch_reduce_kernel, whatever the pixel -> row routing of the stacks (pose_backward.cu).
// EPI = 2 (deferred): no loss in the kernel — the accumulators themselves go to sk_image[EA_PLANES][H][W] (E, A_r, A_g,
//   A_b, A_depth, the depth-column sums F and the final colour / depth), and pose_contract_kernel forms
//   tau = sum_ch g_ch (A_ch - out_ch E) once the caller's loss has produced dL/d out.  This is how the DROP-IN autograd
//   surface gets the single pass: E and A do not depend on the loss, so forward(theta, rho) can build them before any
//   loss exists and backward is one cheap pass over the pixels.
// FOLD: the last CTA to finish (ticket) runs the iteration's serial tail (track_update_body) instead of a separate launch.
// ROWS: `recs` holds the sorted Gaussian ids (u32) and `jac` the packed per-Gaussian rows (TRACK_ROW_FLOATS: record + basis);
//   a batch is then ONE gather by id (cp.async) instead of a bulk copy of per-instance records plus the basis gather.
template <bool HAS_DEPTH, bool FAST, int EPI, int NW, bool FOLD, bool ROWS>   // FAST: ex2.approx exponent, n_touched not produced
__global__ void __launch_bounds__(NW * 32, HAS_DEPTH ? MGS_TRK_MIN_CTAS_DEPTH : MGS_TRK_MIN_CTAS)
blend_track_kernel(const int W, const int H, const int gx, const uint32_t* __restrict__ tile_order, const uint2* __restrict__ ranges,
                   const SplatRec* __restrict__ recs, const float* __restrict__ jac, const float* __restrict__ bg,
                   float* __restrict__ out_color, float* __restrict__ out_depth, float* __restrict__ out_opacity,
                   int32_t* __restrict__ n_touched, const LossK lk, const float loss_grad_scale,
                   double* __restrict__ loss_partials, float* __restrict__ tau_sum,
                   const float* __restrict__ sk_weights, const float sk_scale, float* __restrict__ sk_image,
                   const TrackTail tail, uint32_t* __restrict__ clear_ctr) {
  constexpr bool SKETCH = EPI == 1, DEFER = EPI == 2;
  constexpr int TRK_BATCH = TrkCfg<NW>::kBatch, THREADS = NW * 32;
  extern __shared__ __align__(128) unsigned char trk_smem[];
  SplatRec (*stage)[TRK_BATCH] = reinterpret_cast<SplatRec (*)[TRK_BATCH]>(trk_smem);
  float (*jstage)[TRK_BATCH * TRK_NF] =
      reinterpret_cast<float (*)[TRK_BATCH * TRK_NF]>(trk_smem + (size_t)TRK_STAGES * TRK_BATCH * sizeof(SplatRec));
  __shared__ __align__(8) uint64_t full_bar[TRK_STAGES];
  __shared__ float s_part[NW][6];
  __shared__ int s_last;

  pdl_enter<PDL_TRK>();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile = (int)tile_order[(size_t)blockIdx.x * TILE_CTR_STRIDE + 3];   // heavy tiles first (tile_scan_body)
  const int tx = tile % gx, ty = tile / gx;
  const int wrow = warp >> 1;   // which 4-row band of the tile
  const int px = tx * TILE_X + (warp & 1) * 8 + (lane & 7);
  const int py = ty * TILE_Y + wrow * 4 + (lane >> 3);
  const bool inside = px < W && py < H;
  const float pxf = (float)px, pyf = (float)py;
  const float patch_x0 = (float)(tx * TILE_X + (warp & 1) * 8), patch_y0 = (float)(ty * TILE_Y + wrow * 4);

  const uint2 range = ranges[tile];
  const int n = (int)(range.y - range.x);
  const int rounds = (n + TRK_BATCH - 1) / TRK_BATCH;
  const SplatRec* src = recs + range.x;
  // fixed-segment binning: the tile's instance counter is dead once its range has been read, and this grid visits every
  // tile exactly once, so it leaves the counters zero for the next iteration's preprocess (tile_sort_segments_kernel)
  if (clear_ctr != nullptr && tid == 0) clear_ctr[(size_t)tile * TILE_CTR_STRIDE] = 0u;

  bool done = !inside;
  float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
  // The 6-vector accumulators live as three f32x2 register pairs each: every multiply-add that builds vp and folds it
  // into E / A_ch is a packed FMUL2 / FFMA2 (sm_100 packed fp32; same IEEE operation per half).
  float2 E2[3], A02[3], A12[3], A22[3], Ad2[3];
  float F0 = 0.f, F1 = 0.f, F2 = 0.f;
#pragma unroll
  for (int k = 0; k < 3; k++) { E2[k] = A02[k] = A12[k] = A22[k] = Ad2[k] = make_float2(0.f, 0.f); }

  // one (warp, splat) step: rec4 = the staged records, j = the splat's slot in them, jrow = its 32-float basis
  auto hit = [&](const float4* __restrict__ r4 /* the splat's record: 3 quads */, const float* __restrict__ jrow) {
    const float4 g0 = r4[0], g1 = r4[1];   // x, y, depth, id | conic xyz, opacity
    const float dx = g0.x - pxf, dy = g0.y - pyf;
    const float power = -0.5f * (g1.x * dx * dx + g1.z * dy * dy) - g1.y * dx * dy;
    bool valid = !done && !(power > 0.0f);
    const float G = blend_exp<FAST>(power);
    const float alpha = min(0.99f, g1.w * G);
    valid = valid && !(alpha < 1.0f / 255.0f);
    const float test_T = T * (1 - alpha);
    if (valid && test_T < 0.0001f) { done = true; valid = false; }
    if (!FAST) {
      const unsigned touch = __ballot_sync(0xffffffffu, valid && test_T > 0.5f);
      if (touch != 0u && lane == 0) atomicAdd(&n_touched[__float_as_uint(g0.w)], __popc(touch));
    }
    if (__ballot_sync(0xffffffffu, valid) == 0u) return;
    if (valid) {
      const float4* J4 = reinterpret_cast<const float4*>(jrow);
      const float4 b0 = J4[0], b1 = J4[1], b2 = J4[2], b3 = J4[3], b4 = J4[4], b5 = J4[5], b6 = J4[6], b7 = J4[7];
      const float4 g2 = r4[2];
      // folded basis rows: U = f[0..5], V = f[6..11], A = f[12..17], Bm = f[18..23], Cm = f[24..29], yC, -xC
      const float2 dx2 = make_float2(dx, dx), dy2 = make_float2(dy, dy);
      float2 t1[3], t2[3], vp2[3];
      // t1 = U + dx * A + dy * Bm ; t2 = V + dy * Cm ; vp = dx * t1 + dy * t2
      t1[0] = __ffma2_rn(dx2, make_float2(b3.x, b3.y), make_float2(b0.x, b0.y));
      t1[1] = __ffma2_rn(dx2, make_float2(b3.z, b3.w), make_float2(b0.z, b0.w));
      t1[2] = __ffma2_rn(dx2, make_float2(b4.x, b4.y), make_float2(b1.x, b1.y));
      t1[0] = __ffma2_rn(dy2, make_float2(b4.z, b4.w), t1[0]);
      t1[1] = __ffma2_rn(dy2, make_float2(b5.x, b5.y), t1[1]);
      t1[2] = __ffma2_rn(dy2, make_float2(b5.z, b5.w), t1[2]);
      t2[0] = __ffma2_rn(dy2, make_float2(b6.x, b6.y), make_float2(b1.z, b1.w));
      t2[1] = __ffma2_rn(dy2, make_float2(b6.z, b6.w), make_float2(b2.x, b2.y));
      t2[2] = __ffma2_rn(dy2, make_float2(b7.x, b7.y), make_float2(b2.z, b2.w));
#pragma unroll
      for (int m = 0; m < 3; m++) vp2[m] = __ffma2_rn(dy2, t2[m], __fmul2_rn(dx2, t1[m]));
      // s = opacity * G / (1 - alpha): d alpha / d(geometry) up to the polynomial, times r_i (1-alpha in [0.01,1])
      const float sE = g1.w * G * rcp_unit(1.f - alpha);
      const float s0 = sE * fmaf(T, g2.x, C0), s1 = sE * fmaf(T, g2.y, C1), s2 = sE * fmaf(T, g2.z, C2);
      const float2 sE2 = make_float2(sE, sE), s02 = make_float2(s0, s0), s12 = make_float2(s1, s1), s22 = make_float2(s2, s2);
#pragma unroll
      for (int m = 0; m < 3; m++) {
        E2[m] = __ffma2_rn(sE2, vp2[m], E2[m]);
        A02[m] = __ffma2_rn(s02, vp2[m], A02[m]);
        A12[m] = __ffma2_rn(s12, vp2[m], A12[m]);
        A22[m] = __ffma2_rn(s22, vp2[m], A22[m]);
      }
      if (HAS_DEPTH) {
        const float aT = alpha * T;
        const float sd = sE * fmaf(T, g0.z, D);
        const float2 sd2 = make_float2(sd, sd);
#pragma unroll
        for (int m = 0; m < 3; m++) Ad2[m] = __ffma2_rn(sd2, vp2[m], Ad2[m]);
        // depth column of B: dz/dtau = (0, 0, 1, y_C, -x_C, 0)
        F0 += aT; F1 = fmaf(aT, b7.z, F1); F2 = fmaf(aT, b7.w, F2);
      }
      C0 += g2.x * alpha * T;   // same expression as forward.cu:364-367 / blend_forward.cu: identical image bits
      C1 += g2.y * alpha * T;
      C2 += g2.z * alpha * T;
      D += g0.z * alpha * T;
      T = test_T;
    }
  };

  if constexpr (ROWS) {
    // ---- ids + packed rows: ids of batch i+2 and rows of batch i+1 in flight (cp.async) while batch i is blended ----
    constexpr int RF = TRACK_ROW_FLOATS, RC = RF / 4;
    uint32_t (*idstage)[TRK_BATCH] = reinterpret_cast<uint32_t (*)[TRK_BATCH]>(trk_smem);
    float* rowstage = reinterpret_cast<float*>(trk_smem + 2 * TRK_BATCH * sizeof(uint32_t));
    const uint32_t* ids = reinterpret_cast<const uint32_t*>(recs) + range.x;
    auto load_ids = [&](int i) {
      const int cnt = min(TRK_BATCH, n - i * TRK_BATCH);
      if (tid < cnt) cp_async4(&idstage[i & 1][tid], ids + (size_t)i * TRK_BATCH + tid);
    };
    auto gather = [&](int i) {   // rows of batch i (L2-resident per-Gaussian array) -> rowstage[i & 1], 16 B per cp.async
      const int cnt = min(TRK_BATCH, n - i * TRK_BATCH);
      float4* dst = reinterpret_cast<float4*>(rowstage + (size_t)(i & 1) * TRK_BATCH * RF);
      for (int e = tid; e < cnt * RC; e += THREADS) {
        const int slot = e / RC, k = e - slot * RC;
        cp_async16(dst + e, reinterpret_cast<const float4*>(jac + (size_t)idstage[i & 1][slot] * RF) + k);
      }
    };
    if (rounds > 0) {
      load_ids(0); cp_async_commit(); cp_async_wait_all();
      __syncthreads();
      gather(0);
      if (rounds > 1) load_ids(1);
      cp_async_commit();
    }
    for (int i = 0; i < rounds; i++) {
      cp_async_wait_all();   // this thread's share of the rows of batch i (and of the ids of batch i+1) has landed
      if (__syncthreads_count(done) == THREADS) break;   // CTA vote + publication; nothing is in flight here
      if (i + 1 < rounds) {
        gather(i + 1);
        if (i + 2 < rounds) load_ids(i + 2);
        cp_async_commit();
      }
      const int cnt = min(TRK_BATCH, n - i * TRK_BATCH);
      const float* recf = rowstage + (size_t)(i & 1) * TRK_BATCH * RF;
      const float4* row4 = reinterpret_cast<const float4*>(recf);
      for (int c0 = 0; c0 < cnt; c0 += 32) {
        const int jj = c0 + lane;
        unsigned hits = __ballot_sync(0xffffffffu, jj < cnt && splat_hits(recf + RF * jj, patch_x0, patch_x0 + 7.f, patch_y0, patch_y0 + 3.f));
        if (__ballot_sync(0xffffffffu, !done) == 0u) break;
        while (hits != 0u) {
          const int j = c0 + __ffs(hits) - 1;
          hits &= hits - 1;
          hit(row4 + RC * j, recf + RF * j + 12);
        }
      }
    }
  } else {
    // ---- one CTA per tile: records and bases of a batch are staged by the whole CTA, one barrier per batch ----
    if (tid == 0) {
#pragma unroll
      for (int s = 0; s < TRK_STAGES; s++) mbar_init(&full_bar[s], 1);
      mbar_init_fence();
    }
    __syncthreads();
    auto issue = [&](int i) {   // records of batch i -> stage[i % 3]
      const uint32_t bytes = (uint32_t)min(n - i * TRK_BATCH, TRK_BATCH) * sizeof(SplatRec);
      mbar_expect_tx(&full_bar[i % TRK_STAGES], bytes);
      bulk_g2s(&stage[i % TRK_STAGES][0], src + (size_t)i * TRK_BATCH, bytes, &full_bar[i % TRK_STAGES]);
    };
    auto wait_records = [&](int i) { mbar_wait(&full_bar[i % TRK_STAGES], (i / TRK_STAGES) & 1); };
    auto gather = [&](int i) {  // pose bases of batch i (L2-resident per-Gaussian array) -> jstage[i & 1], 16 B per cp.async
      const int cnt = min(TRK_BATCH, n - i * TRK_BATCH);
      const float4* rec4 = reinterpret_cast<const float4*>(&stage[i % TRK_STAGES][0]);
      float4* dst = reinterpret_cast<float4*>(&jstage[i & 1][0]);
      for (int e = tid; e < cnt * (TRK_NF / 4); e += THREADS) {
        const int slot = e / (TRK_NF / 4), k = e - slot * (TRK_NF / 4);
        const uint32_t id = __float_as_uint(rec4[3 * slot + 0].w);
        cp_async16(dst + e, reinterpret_cast<const float4*>(jac + (size_t)id * TRK_NF) + k);
      }
      cp_async_commit();
    };
    if (tid == 0 && rounds > 0) { issue(0); if (rounds > 1) issue(1); }
    if (rounds > 0) { wait_records(0); gather(0); }
    for (int i = 0; i < rounds; i++) {
      cp_async_wait_all();   // this thread's share of the bases of batch i has landed
      // CTA vote; publishes the bases of batch i and guarantees every warp left the buffers reused below
      if (__syncthreads_count(done) == THREADS) {
        if (tid == 0 && i + 1 < rounds) wait_records(i + 1);   // never exit with a bulk copy in flight
        break;
      }
      if (tid == 0 && i + 2 < rounds) issue(i + 2);
      if (i + 1 < rounds) { wait_records(i + 1); gather(i + 1); }
      const int cnt = min(TRK_BATCH, n - i * TRK_BATCH);
      const float4* rec4 = reinterpret_cast<const float4*>(&stage[i % TRK_STAGES][0]);
      const float* recf = reinterpret_cast<const float*>(&stage[i % TRK_STAGES][0]);
      const float* jst = &jstage[i & 1][0];
      for (int c0 = 0; c0 < cnt; c0 += 32) {
        // lane-parallel exact cull of 32 splats against this warp's 8x4 patch (see splat_extent)
        const int jj = c0 + lane;
        unsigned hits = __ballot_sync(0xffffffffu, jj < cnt && splat_hits(recf + 12 * jj, patch_x0, patch_x0 + 7.f, patch_y0, patch_y0 + 3.f));
        if (__ballot_sync(0xffffffffu, !done) == 0u) break;
        if (hits == 0u) continue;
        while (hits != 0u) {
          const int j = c0 + __ffs(hits) - 1;
          hits &= hits - 1;
          hit(rec4 + 3 * j, jst + j * TRK_NF);
        }
      }
    }
  }

  // epilogue: outputs, loss and its gradient from the registers that hold the blended pixel, tangent contraction
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  float tau[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (inside) {
    const size_t pix = (size_t)py * W + px;
    const size_t plane = (size_t)W * H;
    const float img[3] = {C0 + T * bg[0], C1 + T * bg[1], C2 + T * bg[2]};
    out_color[pix] = img[0];
    out_color[plane + pix] = img[1];
    out_color[2 * plane + pix] = img[2];
    out_depth[pix] = D;
    out_opacity[pix] = 1 - T;
    if (DEFER) {
      // planes: E 0..5 | A_r 6..11 | A_g 12..17 | A_b 18..23 | A_d 24..29 | F 30..32 | out r g b 33..35 | D 36
#pragma unroll
      for (int m = 0; m < 3; m++) {
        sk_image[(2 * m) * plane + pix] = E2[m].x; sk_image[(2 * m + 1) * plane + pix] = E2[m].y;
        sk_image[(6 + 2 * m) * plane + pix] = A02[m].x; sk_image[(7 + 2 * m) * plane + pix] = A02[m].y;
        sk_image[(12 + 2 * m) * plane + pix] = A12[m].x; sk_image[(13 + 2 * m) * plane + pix] = A12[m].y;
        sk_image[(18 + 2 * m) * plane + pix] = A22[m].x; sk_image[(19 + 2 * m) * plane + pix] = A22[m].y;
        sk_image[(24 + 2 * m) * plane + pix] = Ad2[m].x; sk_image[(25 + 2 * m) * plane + pix] = Ad2[m].y;
      }
      sk_image[30 * plane + pix] = F0; sk_image[31 * plane + pix] = F1; sk_image[32 * plane + pix] = F2;
      sk_image[33 * plane + pix] = img[0]; sk_image[34 * plane + pix] = img[1]; sk_image[35 * plane + pix] = img[2];
      sk_image[36 * plane + pix] = D;
    }
    float gi[3] = {0.f, 0.f, 0.f}, gd = 0.f, sk_da = 0.f, sk_db = 0.f, sk_f = 0.f;
    if (DEFER) {
    } else if (SKETCH) {
      float l1;
      sk_f = sketch_pixel(lk, (int)pix, img, 1 - T, sk_weights[pix] * sk_scale, gi, sk_da, sk_db, l1);
      acc[0] = l1;
    } else loss_pixel<true, true>(lk, (int)pix, img, D, 1 - T, loss_grad_scale, acc, gi, gd);
    float E[6], A0[6], A1[6], A2[6], Ad[6];
#pragma unroll
    for (int m = 0; m < 3; m++) {
      E[2 * m] = E2[m].x; E[2 * m + 1] = E2[m].y; A0[2 * m] = A02[m].x; A0[2 * m + 1] = A02[m].y;
      A1[2 * m] = A12[m].x; A1[2 * m + 1] = A12[m].y; A2[2 * m] = A22[m].x; A2[2 * m + 1] = A22[m].y;
      Ad[2 * m] = Ad2[m].x; Ad[2 * m + 1] = Ad2[m].y;
    }
#pragma unroll
    for (int k = 0; k < 6; k++) {
      float t = gi[0] * (A0[k] - img[0] * E[k]);
      t = fmaf(gi[1], A1[k] - img[1] * E[k], t);
      t = fmaf(gi[2], A2[k] - img[2] * E[k], t);
      if (HAS_DEPTH) t = fmaf(gd, Ad[k] - D * E[k], t);
      tau[k] = t;
    }
    if (HAS_DEPTH) { tau[2] = fmaf(gd, F0, tau[2]); tau[3] = fmaf(gd, F1, tau[3]); tau[4] = fmaf(gd, F2, tau[4]); }
    if (SKETCH) {
      if (tail.route.idx != nullptr) {
        // Routed here: the pixel's nine values go to row sketch_indices[j][pix] of every stack j that holds the pixel, as
        // fire-and-forget float reductions in L2 (RED.ADD.F32) spread over the whole run of the kernel, behind the other
        // CTAs' blending.  This replaces the [9][H][W] planes and the reduction kernel that re-read them (52 us at C2, bound
        // by its shared-memory float atomics, which are compare-and-swap loops; profiles/r02_notes.txt item 16).
        const float val[9] = {tau[0], tau[1], tau[2], tau[3], tau[4], tau[5], sk_da, sk_db, sk_f};
        for (int j0 = 0; j0 < tail.route.n_stacks; j0 += 8) {
          int row[8];
#pragma unroll
          for (int u = 0; u < 8; u++) row[u] = j0 + u < tail.route.n_stacks ? __ldg(tail.route.idx + (size_t)(j0 + u) * plane + pix) : -1;
#pragma unroll
          for (int u = 0; u < 8; u++) {
            if (row[u] < 0 || row[u] >= tail.route.d) continue;
            float* o = tail.route.rows + ((size_t)(j0 + u) * tail.route.d + row[u]) * 9;
#pragma unroll
            for (int k = 0; k < 9; k++)
              if (val[k] != 0.f) atomicAdd(o + k, val[k]);
          }
        }
      } else {
#pragma unroll
        for (int k = 0; k < 6; k++) sk_image[k * plane + pix] = tau[k];
        sk_image[6 * plane + pix] = sk_da;
        sk_image[7 * plane + pix] = sk_db;
        sk_image[8 * plane + pix] = sk_f;
      }
    }
  }
  if (DEFER) return;
  loss_block_reduce(acc, loss_partials + (size_t)blockIdx.x * 4);      // one slot per CTA (two per tile when NW = 4)
  if (!SKETCH) {
#pragma unroll
    for (int k = 0; k < 6; k++) {
      float x = tau[k];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      if (lane == 0) s_part[warp][k] = x;
    }
    __syncthreads();
    if (tid < 6) {
      float x = 0.f;
#pragma unroll
      for (int w = 0; w < NW; w++) x += s_part[w][tid];
      if (x != 0.f) atomicAdd(&tau_sum[tid], x);
    }
    if (FOLD) {
      // ticket: this CTA's partial sums and tangent atomics are ordered before its increment, so whoever draws the last
      // ticket sees the sums of every tile and runs the iteration's serial tail in place of a separate launch
      __threadfence();
      __syncthreads();
      if (tid == 0) {
        const uint32_t t = atomicAdd(tail.ticket, 1u);
        s_last = (t == gridDim.x - 1) ? 1 : 0;
      }
      __syncthreads();
      if (s_last) {
        __threadfence();
        track_update_body(tail.k);
        if (tid == 0) *tail.ticket = 0u;
      }
    }
  }
}

template <bool HAS_DEPTH, bool FAST, int EPI, int NW, bool FOLD, bool ROWS = false>
static int launch_trk_t(int tiles, cudaStream_t s, const int W, const int H, const int gx, const uint32_t* tile_order,
                        const uint2* ranges, const SplatRec* recs, const float* jac, const float* bg, float* out_color,
                        float* out_depth, float* out_opacity, int32_t* n_touched, const LossK& lk, float grad_scale,
                        double* partials, float* tau_sum, const float* sk_weights, float sk_scale, float* sk_image,
                        const TrackTail& tail, uint32_t* clear_ctr) {
  // the opt-in to > 48 KB of dynamic shared memory is a per-device function attribute
  static bool attr_set[64] = {};
  int dev = 0;
  MGS_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { set_error("blend_track: device ordinal %d out of range", dev); return MGS_EINVAL; }
  auto kern = blend_track_kernel<HAS_DEPTH, FAST, EPI, NW, FOLD, ROWS>;
  if (!attr_set[dev]) {
    MGS_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(ROWS ? TrkCfg<NW>::kSmemRows : TrkCfg<NW>::kSmem)));
    attr_set[dev] = true;
  }
  MGS_CUDA_TRY(launch_chain(kern, tiles, NW * 32, ROWS ? TrkCfg<NW>::kSmemRows : TrkCfg<NW>::kSmem, s, W, H, gx, tile_order, ranges, recs, jac, bg,
                            out_color, out_depth, out_opacity, n_touched, lk, grad_scale, partials, tau_sum, sk_weights, sk_scale,
                            sk_image, tail, clear_ctr));
  return MGS_OK;
}

// Deferred contraction (EPI = 2): tau_pixel = sum_ch g_ch (A_ch - out_ch E) + g_d (A_d - D E) + g_d F, per pixel, from the
// accumulator planes of the forward; summed over the image (tau_sum) and / or kept per pixel (tau_image, the sketch path).
constexpr int EA_PLANES = 37;
__global__ void __launch_bounds__(256)
pose_contract_kernel(const float* __restrict__ ea, const int npix, const float* __restrict__ g_color,
                     const float* __restrict__ g_depth, float* __restrict__ tau_image, float* __restrict__ tau_sum) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  float tau[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (p < npix) {
    const float g0 = g_color[p], g1 = g_color[npix + p], g2 = g_color[2 * (size_t)npix + p];
    const float gd = g_depth != nullptr ? g_depth[p] : 0.f;
    const float o0 = ea[33 * (size_t)npix + p], o1 = ea[34 * (size_t)npix + p], o2 = ea[35 * (size_t)npix + p], D = ea[36 * (size_t)npix + p];
#pragma unroll
    for (int k = 0; k < 6; k++) {
      const float E = ea[k * (size_t)npix + p];
      float t = g0 * (ea[(6 + k) * (size_t)npix + p] - o0 * E);
      t = fmaf(g1, ea[(12 + k) * (size_t)npix + p] - o1 * E, t);
      t = fmaf(g2, ea[(18 + k) * (size_t)npix + p] - o2 * E, t);
      if (g_depth != nullptr) t = fmaf(gd, ea[(24 + k) * (size_t)npix + p] - D * E, t);
      tau[k] = t;
    }
    if (g_depth != nullptr) {
      tau[2] = fmaf(gd, ea[30 * (size_t)npix + p], tau[2]); tau[3] = fmaf(gd, ea[31 * (size_t)npix + p], tau[3]);
      tau[4] = fmaf(gd, ea[32 * (size_t)npix + p], tau[4]);
    }
    if (tau_image != nullptr) {
#pragma unroll
      for (int k = 0; k < 6; k++) tau_image[k * (size_t)npix + p] = tau[k];
    }
  }
  if (tau_sum == nullptr) return;
  __shared__ float s_part[8][6];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 6; k++) {
    float x = tau[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) s_part[warp][k] = x;
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    float x = 0.f;
#pragma unroll
    for (int w = 0; w < 8; w++) x += s_part[w][threadIdx.x];
    if (x != 0.f) atomicAdd(&tau_sum[threadIdx.x], x);
  }
}

size_t pose_deferred_bytes(int W, int H) { return (size_t)EA_PLANES * W * H * sizeof(float); }

int launch_pose_contract(const float* ea, int W, int H, const float* g_color, const float* g_depth, float* tau_image,
                         float* tau_sum, cudaStream_t s) {
  const int npix = W * H;
  if (tau_sum) MGS_CUDA_TRY(memset_async(tau_sum, 0, 6 * sizeof(float), s));
  MGS_PROF("pose_contract", s, pose_contract_kernel<<<(npix + 255) / 256, 256, 0, s>>>(ea, npix, g_color, g_depth, tau_image, tau_sum));
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

int blend_track_warps() { return 8; }

// tau_sum must be zero on entry (the folded tail re-zeroes it); jac holds the folded per-Gaussian bases
// (POSE_BASIS_FLOATS).  sk_image != nullptr selects the second-order epilogue (mono only): sk_weights f32[H,W] (+-1),
// sk_scale = stack_dim * sketch_dim / npix.  upd != nullptr folds the iteration's update tail into the kernel's last CTA.
int launch_blend_track(const ViewConst& v, void* ws, const WsLayout& L, const float* jac, float* out_color,
                       float* out_depth, float* out_opacity, int32_t* n_touched, const FusedLoss& fl, bool has_depth,
                       float* tau_sum, cudaStream_t s, const float* sk_weights, float sk_scale, float* sk_image,
                       const TrackUpdateK* upd, uint32_t* ticket, bool deferred, bool rows, bool clear_counts, const SketchRoute* route) {
  const int tiles = v.gx * v.gy;
  if (clear_counts && !rows) { set_error("blend_track: clear_counts belongs to the fixed-segment (packed-row) mode"); return MGS_EINVAL; }
  if (rows && (deferred || upd != nullptr)) { set_error("blend_track: the packed-row staging serves the fused tracker's un-folded kernels only"); return MGS_EINVAL; }
  if (deferred) {
    // drop-in forward: exact exponent (n_touched produced), depth accumulators always (the loss is not known yet)
    if (sk_image == nullptr || n_touched == nullptr) { set_error("blend_track: deferred mode needs the accumulator planes and n_touched"); return MGS_EINVAL; }
    TrackTail none;
    memset(&none, 0, sizeof(none));
    LossK lk0{};
    int rc0;
    prof_begin("blend_track_deferred", s);
      rc0 = launch_trk_t<true, false, 2, 8, false>(tiles, s, v.W, v.H, v.gx, ws_ptr<uint32_t>(ws, L.tile_ctr), ws_ptr<uint2>(ws, L.ranges),
                                                  ws_ptr<SplatRec>(ws, L.records), jac, v.bg, out_color, out_depth, out_opacity, n_touched,
                                                  lk0, 0.f, nullptr, nullptr, nullptr, 0.f, sk_image, none, nullptr);
    prof_end(s);
    if (rc0) return rc0;
    MGS_LAUNCH_CHECK();
    return MGS_OK;
  }
  if (sk_image != nullptr && (has_depth || sk_weights == nullptr)) { set_error("blend_track: the sketched epilogue is mono only and needs rand_weights"); return MGS_EINVAL; }
  if (upd != nullptr && (ticket == nullptr || sk_image != nullptr)) { set_error("blend_track: folded tail needs a ticket and the first-order epilogue"); return MGS_EINVAL; }
  TrackTail tail;
  memset(&tail, 0, sizeof(tail));
  if (upd) { tail.k = *upd; tail.ticket = ticket; }
  if (route != nullptr) {
    if (sk_image == nullptr) { set_error("blend_track: a sketch route belongs to the second-order epilogue"); return MGS_EINVAL; }
    tail.route = *route;
  }
#define MGS_TRK_ARGS tiles, s, v.W, v.H, v.gx, ws_ptr<uint32_t>(ws, L.tile_ctr), ws_ptr<uint2>(ws, L.ranges), ws_ptr<SplatRec>(ws, L.records), \
      jac, v.bg, out_color, out_depth, out_opacity, n_touched, *fl.k, fl.grad_scale, fl.partials, tau_sum, sk_weights, sk_scale, sk_image, tail, \
      (clear_counts ? ws_ptr<uint32_t>(ws, L.tile_ctr) : (uint32_t*)nullptr)
#define MGS_TRK_DISPATCH(HD, SK, FO)                                                                                                   \
  do {                                                                                                                                 \
    rc = fl.fast ? launch_trk_t<HD, true, SK, 8, FO>(MGS_TRK_ARGS) : launch_trk_t<HD, false, SK, 8, FO>(MGS_TRK_ARGS);            \
  } while (0)
  int rc = MGS_OK;
  prof_begin(sk_image ? "blend_track_sketch" : "blend_track", s);
#define MGS_TRK_DISPATCH_ROWS(HD, SK)                                                                                                  \
  do {                                                                                                                                 \
    rc = fl.fast ? launch_trk_t<HD, true, SK, 8, false, true>(MGS_TRK_ARGS) : launch_trk_t<HD, false, SK, 8, false, true>(MGS_TRK_ARGS); \
  } while (0)
  if (rows) {
    if (sk_image != nullptr) MGS_TRK_DISPATCH_ROWS(false, 1);
    else if (has_depth) MGS_TRK_DISPATCH_ROWS(true, 0);
    else MGS_TRK_DISPATCH_ROWS(false, 0);
  }
  else if (sk_image != nullptr) MGS_TRK_DISPATCH(false, 1, false);
  else if (has_depth) { if (upd) MGS_TRK_DISPATCH(true, 0, true); else MGS_TRK_DISPATCH(true, 0, false); }
  else { if (upd) MGS_TRK_DISPATCH(false, 0, true); else MGS_TRK_DISPATCH(false, 0, false); }
#undef MGS_TRK_DISPATCH_ROWS
  prof_end(s);
#undef MGS_TRK_DISPATCH
#undef MGS_TRK_ARGS
  if (rc) return rc;
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

}  // namespace mgs
