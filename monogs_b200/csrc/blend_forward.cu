// Per-tile alpha blend, forward (replaces renderCUDA, forward.cu:263-392).
//
// One CTA per 16x16 tile, 256 threads, each warp owns an 8x4 pixel patch so that a small splat touches
// few warps.  The tile's packed 48-B records are contiguous after the sort, so batches of 256 records
// (12 KB) are streamed global->shared with cp.async.bulk (TMA 1-D, UBLKCP) into a two-stage ring tracked
// by mbarriers; one elected thread issues the copies.  Early termination is decided per warp with a
// ballot and per CTA with __syncthreads_count, like the reference's block vote.
#include "launchers.h"
#include "tma.cuh"
#include "loss_pixel.cuh"

namespace mgs {

constexpr int FWD_BATCH = 256;
constexpr int FWD_STAGES = 2;

template <bool FUSED_LOSS, bool FAST>   // FAST: approximate exp2 and no n_touched (tracking does not use it)
__global__ void __launch_bounds__(256)
blend_forward_kernel(const int W, const int H, const int gx,
                     const uint2* __restrict__ ranges, const SplatRec* __restrict__ recs,
                     const float* __restrict__ bg,
                     float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                     float* __restrict__ out_color, float* __restrict__ out_depth, float* __restrict__ out_opacity,
                     int32_t* __restrict__ n_touched,
                     const LossK lk, const float loss_grad_scale, float* __restrict__ g_img, float* __restrict__ g_depth,
                     double* __restrict__ loss_partials) {
  __shared__ __align__(128) SplatRec stage[FWD_STAGES][FWD_BATCH];
  __shared__ __align__(8) uint64_t full_bar[FWD_STAGES];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile = blockIdx.x;
  const int tx = tile % gx, ty = tile / gx;
  const int px = tx * TILE_X + (warp & 1) * 8 + (lane & 7);
  const int py = ty * TILE_Y + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = px < W && py < H;
  const float pxf = (float)px, pyf = (float)py;
  const float patch_x0 = (float)(tx * TILE_X + (warp & 1) * 8), patch_y0 = (float)(ty * TILE_Y + (warp >> 1) * 4);

  const uint2 range = ranges[tile];
  const int n = (int)(range.y - range.x);
  const int rounds = (n + FWD_BATCH - 1) / FWD_BATCH;
  const SplatRec* src = recs + range.x;

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

  bool done = !inside;
  float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
  uint32_t last_contributor = 0;

  for (int i = 0; i < rounds; i++) {
    // CTA vote; also guarantees every warp left the stage that the next copy overwrites
    if (__syncthreads_count(done) == TILE_PIX) {
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
      // lane-parallel exact cull: which of these 32 splats can reach this warp's 8x4 patch at all?
      const int jj = c0 + lane;
      unsigned hits = __ballot_sync(0xffffffffu, jj < cnt && splat_hits(recf + 12 * jj, patch_x0, patch_x0 + 7.f, patch_y0, patch_y0 + 3.f));
      if (__ballot_sync(0xffffffffu, !done) == 0u) break;
      while (hits) {
        const int j = c0 + __ffs(hits) - 1;
        hits &= hits - 1;
        const float4 g0 = rec4[3 * j + 0];  // x, y, depth, id
        const float4 g1 = rec4[3 * j + 1];  // conic xyz, opacity
        const float dx = g0.x - pxf, dy = g0.y - pyf;
        const float power = -0.5f * (g1.x * dx * dx + g1.z * dy * dy) - g1.y * dx * dy;
        bool valid = !done && !(power > 0.0f);
        const float alpha = min(0.99f, g1.w * blend_exp<FAST>(power));
        valid = valid && !(alpha < 1.0f / 255.0f);
        const float test_T = T * (1 - alpha);
        if (valid && test_T < 0.0001f) { done = true; valid = false; }
        unsigned touch = 0u;
        if (!FAST) touch = __ballot_sync(0xffffffffu, valid && test_T > 0.5f);
        if (valid) {
          const float4 g2 = rec4[3 * j + 2];
          C0 += g2.x * alpha * T;
          C1 += g2.y * alpha * T;
          C2 += g2.z * alpha * T;
          D += g0.z * alpha * T;
          T = test_T;
          last_contributor = base + j + 1;
        }
        if (!FAST && touch != 0u && lane == 0) atomicAdd(&n_touched[__float_as_uint(g0.w)], __popc(touch));
      }
    }
  }

  if (inside) {
    const size_t pix = (size_t)py * W + px;
    const size_t plane = (size_t)W * H;
    final_T[pix] = T;
    n_contrib[pix] = last_contributor;
    out_color[pix] = C0 + T * bg[0];
    out_color[plane + pix] = C1 + T * bg[1];
    out_color[2 * plane + pix] = C2 + T * bg[2];
    out_depth[pix] = D;
    out_opacity[pix] = 1 - T;
  }
  if (FUSED_LOSS) {
    // loss + its (un-normalised) gradient straight from the registers that hold the blended pixel
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    if (inside) {
      const size_t pix = (size_t)py * W + px;
      const size_t plane = (size_t)W * H;
      const float img[3] = {C0 + T * bg[0], C1 + T * bg[1], C2 + T * bg[2]};
      float gi[3], gd;
      loss_pixel<true, true>(lk, (int)pix, img, D, 1 - T, loss_grad_scale, acc, gi, gd);
      g_img[pix] = gi[0]; g_img[plane + pix] = gi[1]; g_img[2 * plane + pix] = gi[2];
      if (g_depth != nullptr) g_depth[pix] = gd;
    }
    loss_block_reduce(acc, loss_partials + (size_t)tile * 4);
  }
}

int launch_blend_forward(const ViewConst& v, void* ws, const WsLayout& L, float* out_color, float* out_depth,
                         float* out_opacity, int32_t* n_touched, cudaStream_t s, const FusedLoss* fl) {
  const int tiles = v.gx * v.gy;
  prof_begin("blend_forward", s);
  if (fl == nullptr) {
    LossK none{};
    blend_forward_kernel<false, false><<<tiles, 256, 0, s>>>(
        v.W, v.H, v.gx, ws_ptr<uint2>(ws, L.ranges), ws_ptr<SplatRec>(ws, L.records), v.bg, ws_ptr<float>(ws, L.final_T),
        ws_ptr<uint32_t>(ws, L.n_contrib), out_color, out_depth, out_opacity, n_touched, none, 0.f, nullptr, nullptr, nullptr);
  } else {
#define MGS_FWD_ARGS v.W, v.H, v.gx, ws_ptr<uint2>(ws, L.ranges), ws_ptr<SplatRec>(ws, L.records), v.bg,                  \
        ws_ptr<float>(ws, L.final_T), ws_ptr<uint32_t>(ws, L.n_contrib), out_color, out_depth, out_opacity, n_touched, \
        *fl->k, fl->grad_scale, fl->g_img, fl->g_depth, fl->partials
    if (fl->fast) blend_forward_kernel<true, true><<<tiles, 256, 0, s>>>(MGS_FWD_ARGS);
    else blend_forward_kernel<true, false><<<tiles, 256, 0, s>>>(MGS_FWD_ARGS);
#undef MGS_FWD_ARGS
  }
  prof_end(s);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

}  // namespace mgs
