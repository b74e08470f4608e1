// Per-tile alpha blend, forward (replaces renderCUDA, forward.cu:263-392).
//
// One CTA per 16x16 tile, 128 threads; each warp owns an 8x8 pixel patch and each thread two pixels of one column
// (rows y and y+4), evaluated together with packed fp32 (blend_pair.cuh).  The tile's packed 48-B records are
// contiguous after the sort, so batches of 128 records (6 KB) are streamed global->shared with cp.async.bulk
// (TMA 1-D, UBLKCP) into a two-stage ring tracked by mbarriers; one elected thread issues the copies.  A warp walks
// only the records that can reach its patch (lane-parallel exact cull, splat_hits).  Early termination is decided
// per warp with a ballot and per CTA with __syncthreads_count, like the reference's block vote.
#include "launchers.h"
#include "tma.cuh"
#include "loss_pixel.cuh"
#include "blend_pair.cuh"

namespace mgs {

constexpr int FWD_BATCH = 128;
constexpr int FWD_STAGES = 2;
constexpr int FWD_THREADS = 128;

template <bool FUSED_LOSS, bool FAST, bool TOUCH>   // FAST: ex2.approx exponent; TOUCH: n_touched produced
#ifndef MGS_FWD_MIN_CTAS
#define MGS_FWD_MIN_CTAS 10    // 48 registers (16 B of spills): 10 CTAs per SM instead of 9, +0.7 % on the mapping iteration
#endif
__global__ void __launch_bounds__(FWD_THREADS, MGS_FWD_MIN_CTAS)
blend_forward_kernel(const int W, const int H, const int gx, const uint32_t* __restrict__ tile_order,
                     const uint2* __restrict__ ranges, const SplatRec* __restrict__ recs,
                     const float* __restrict__ bg,
                     float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                     float* __restrict__ out_color, float* __restrict__ out_depth, float* __restrict__ out_opacity,
                     int32_t* __restrict__ n_touched,
                     const LossK lk, const float loss_grad_scale, float* __restrict__ g_img, float* __restrict__ g_depth,
                     double* __restrict__ loss_partials, uint32_t* __restrict__ clear_ctr) {
  __shared__ __align__(128) SplatRec stage[FWD_STAGES][FWD_BATCH];
  __shared__ __align__(8) uint64_t full_bar[FWD_STAGES];

  pdl_enter<PDL_FWD>();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // tile_order != null: longest tile lists first (tile_scan_body, heavy_first) — the grid of a single view is ~1.4 waves,
  // and the partial last wave then holds the SHORTEST lists instead of whatever the raster order left for it
  const int tile = tile_order != nullptr ? (int)tile_order[(size_t)blockIdx.x * TILE_CTR_STRIDE + 3] : (int)blockIdx.x;
  const int tx = tile % gx, ty = tile / gx;
  const int px = tx * TILE_X + (warp & 1) * 8 + (lane & 7);
  const int py0 = ty * TILE_Y + (warp >> 1) * 8 + (lane >> 3), py1 = py0 + 4;
  const bool in0 = px < W && py0 < H, in1 = px < W && py1 < H;
  const float2 npx = make_float2(-(float)px, -(float)px), npy = make_float2(-(float)py0, -(float)py1);
  const float patch_x0 = (float)(tx * TILE_X + (warp & 1) * 8), patch_y0 = (float)(ty * TILE_Y + (warp >> 1) * 8);

  const uint2 range = ranges[tile];
  const int n = (int)(range.y - range.x);
  const int rounds = (n + FWD_BATCH - 1) / FWD_BATCH;
  const SplatRec* src = recs + range.x;
  // fixed-segment binning: this grid visits every tile once and leaves the instance counters zero for the next view
  if (clear_ctr != nullptr && tid == 0) clear_ctr[(size_t)tile * TILE_CTR_STRIDE] = 0u;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < FWD_STAGES; s++) mbar_init(&full_bar[s], 1);
    mbar_init_fence();
  }
  __syncthreads();
  if (tid == 0 && rounds > 0) {
    const uint32_t bytes = (uint32_t)min(n, FWD_BATCH) * sizeof(SplatRec);
    mbar_expect_tx(&full_bar[0], bytes);
    bulk_g2s(&stage[0][0], src, bytes, &full_bar[0]);
  }

  bool done0 = !in0, done1 = !in1;
  float2 T = make_float2(1.f, 1.f), C0 = make_float2(0.f, 0.f), C1 = C0, C2 = C0, D = C0;
  uint32_t last0 = 0, last1 = 0;

  for (int i = 0; i < rounds; i++) {
    // CTA vote; also guarantees every warp left the stage that the next copy overwrites
    if (__syncthreads_count(done0 && done1) == FWD_THREADS) {
      if (tid == 0) mbar_wait(&full_bar[i & 1], (i >> 1) & 1);  // never exit with a bulk copy in flight
      break;
    }
    if (tid == 0 && i + 1 < rounds) {
      const int s = (i + 1) & 1;
      const uint32_t bytes = (uint32_t)min(n - (i + 1) * FWD_BATCH, FWD_BATCH) * sizeof(SplatRec);
      mbar_expect_tx(&full_bar[s], bytes);
      bulk_g2s(&stage[s][0], src + (size_t)(i + 1) * FWD_BATCH, bytes, &full_bar[s]);
    }
    mbar_wait(&full_bar[i & 1], (i >> 1) & 1);
    const float4* rec4 = reinterpret_cast<const float4*>(&stage[i & 1][0]);
    const float* recf = reinterpret_cast<const float*>(&stage[i & 1][0]);
    const int cnt = min(FWD_BATCH, n - i * FWD_BATCH);
    const uint32_t base = (uint32_t)i * FWD_BATCH;
    for (int c0 = 0; c0 < cnt; c0 += 32) {
      // lane-parallel exact cull: which of these 32 splats can reach this warp's 8x8 patch at all?
      const int jj = c0 + lane;
      unsigned hits = __ballot_sync(0xffffffffu, jj < cnt && splat_hits(recf + 12 * jj, patch_x0, patch_x0 + 7.f, patch_y0, patch_y0 + 7.f));
      if (__ballot_sync(0xffffffffu, !(done0 && done1)) == 0u) break;
      while (hits) {
        const int j = c0 + __ffs(hits) - 1;
        hits &= hits - 1;
        const float4 g0 = rec4[3 * j + 0];  // x, y, depth, id
        const float4 g1 = rec4[3 * j + 1];  // conic xyz, opacity
        float2 dx, dy, power, G, alpha;
        pair_alpha<FAST>(g0, g1, npx, npy, dx, dy, power, G, alpha);
        bool v0 = !done0 && !(power.x > 0.0f) && !(alpha.x < 1.0f / 255.0f);
        bool v1 = !done1 && !(power.y > 0.0f) && !(alpha.y < 1.0f / 255.0f);
        const float2 test_T = __fmul2_rn(T, one_minus2(alpha));
        if (v0 && test_T.x < 0.0001f) { done0 = true; v0 = false; }
        if (v1 && test_T.y < 0.0001f) { done1 = true; v1 = false; }
        if (TOUCH) {
          const unsigned t0 = __ballot_sync(0xffffffffu, v0 && test_T.x > 0.5f);
          const unsigned t1 = __ballot_sync(0xffffffffu, v1 && test_T.y > 0.5f);
          if ((t0 | t1) != 0u && lane == 0) atomicAdd(&n_touched[__float_as_uint(g0.w)], __popc(t0) + __popc(t1));
        }
        if (v0 || v1) {
          const float4 g2 = rec4[3 * j + 2];
          // colour += c * alpha * T as (c * alpha) * T + colour (forward.cu:364-367); a masked pixel adds c * 0
          const float2 am = make_float2(v0 ? alpha.x : 0.f, v1 ? alpha.y : 0.f);
          C0 = __ffma2_rn(T, __fmul2_rn(bcast2(g2.x), am), C0);
          C1 = __ffma2_rn(T, __fmul2_rn(bcast2(g2.y), am), C1);
          C2 = __ffma2_rn(T, __fmul2_rn(bcast2(g2.z), am), C2);
          D = __ffma2_rn(T, __fmul2_rn(bcast2(g0.z), am), D);
          if (v0) { T.x = test_T.x; last0 = base + j + 1; }
          if (v1) { T.y = test_T.y; last1 = base + j + 1; }
        }
      }
    }
  }

  const size_t plane = (size_t)W * H;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int k = 0; k < 2; k++) {
    const bool inside = k ? in1 : in0;
    if (!inside) continue;
    const size_t pix = (size_t)(k ? py1 : py0) * W + px;
    const float Tk = k ? T.y : T.x, Dk = k ? D.y : D.x;
    const float img[3] = {(k ? C0.y : C0.x) + Tk * bg[0], (k ? C1.y : C1.x) + Tk * bg[1], (k ? C2.y : C2.x) + Tk * bg[2]};
    final_T[pix] = Tk;
    n_contrib[pix] = k ? last1 : last0;
    out_color[pix] = img[0];
    out_color[plane + pix] = img[1];
    out_color[2 * plane + pix] = img[2];
    out_depth[pix] = Dk;
    out_opacity[pix] = 1 - Tk;
    if (FUSED_LOSS) {
      // loss + its (un-normalised) gradient straight from the registers that hold the blended pixel
      float gi[3], gd;
      loss_pixel<true, true>(lk, (int)pix, img, Dk, 1 - Tk, loss_grad_scale, acc, gi, gd);
      g_img[pix] = gi[0]; g_img[plane + pix] = gi[1]; g_img[2 * plane + pix] = gi[2];
      if (g_depth != nullptr) g_depth[pix] = gd;
    }
  }
  if (FUSED_LOSS) loss_block_reduce(acc, loss_partials + (size_t)tile * 4);
}

int launch_blend_forward(const ViewConst& v, void* ws, const WsLayout& L, float* out_color, float* out_depth,
                         float* out_opacity, int32_t* n_touched, cudaStream_t s, const FusedLoss* fl, bool heavy_first, bool clear_counts) {
  const int tiles = v.gx * v.gy;
  uint32_t* clear_ctr = clear_counts ? ws_ptr<uint32_t>(ws, L.tile_ctr) : nullptr;
  const uint32_t* order = heavy_first ? ws_ptr<uint32_t>(ws, L.tile_ctr) : nullptr;
  cudaError_t e;
  prof_begin("blend_forward", s);
  if (fl == nullptr) {
    LossK none{};
    e = launch_chain(blend_forward_kernel<false, false, true>, tiles, FWD_THREADS, 0, s,
        v.W, v.H, v.gx, order, ws_ptr<uint2>(ws, L.ranges), ws_ptr<SplatRec>(ws, L.records), v.bg, ws_ptr<float>(ws, L.final_T),
        ws_ptr<uint32_t>(ws, L.n_contrib), out_color, out_depth, out_opacity, n_touched, none, 0.f, nullptr, nullptr, nullptr, clear_ctr);
  } else {
#define MGS_FWD_ARGS v.W, v.H, v.gx, order, ws_ptr<uint2>(ws, L.ranges), ws_ptr<SplatRec>(ws, L.records), v.bg,           \
        ws_ptr<float>(ws, L.final_T), ws_ptr<uint32_t>(ws, L.n_contrib), out_color, out_depth, out_opacity, n_touched, \
        *fl->k, fl->grad_scale, fl->g_img, fl->g_depth, fl->partials, clear_ctr
    if (fl->fast && n_touched == nullptr) e = launch_chain(blend_forward_kernel<true, true, false>, tiles, FWD_THREADS, 0, s, MGS_FWD_ARGS);
    else if (fl->fast) e = launch_chain(blend_forward_kernel<true, true, true>, tiles, FWD_THREADS, 0, s, MGS_FWD_ARGS);
    else e = launch_chain(blend_forward_kernel<true, false, true>, tiles, FWD_THREADS, 0, s, MGS_FWD_ARGS);
#undef MGS_FWD_ARGS
  }
  prof_end(s);
  MGS_CUDA_TRY(e);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

}  // namespace mgs
