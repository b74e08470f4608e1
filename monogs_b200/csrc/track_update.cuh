// The serial tail of a first-order tracking iteration: loss and its norm from the per-tile partial sums, tau / dab
// scaling, Adam on (rho, theta, exposure a, b), T <- Exp(tau) T and the view / projection matrices
// (utils/slam_frontend.py:596-615, utils/pose_utils.py:88-98).  Shared by track_fused.cu and blend_track.cu.
#pragma once
#include "common.cuh"
#include "pose_update.cuh"

namespace mgs {

struct TrackUpdateK {
  int mode, use_depth, npix, tiles;
  float alpha, scale;
  const double* partials; const float* tau_raw;
  float *loss, *tau, *dab;
  float *params, *exp_avg, *exp_avg_sq; int* step; float lr[4]; float beta1, beta2, eps;
  float *T, *view_out, *proj_out; const float* proj_raw; int* converged; float thr;
  float* lm_state;   // optional (mgs_track_args.lm_state): scalar ||r||_1 and the best-pose record
  float* tau_clear;  // optional: the 6-float tangent accumulator to leave zero for the next iteration (read above as tau_raw)
};

// lm_state (device f32[64]) — the tracker's cross-iteration record, shared by the first- and second-order iterations
enum : int {
  LM_LAMBDA = 0, LM_PREV_SCALAR = 1, LM_HAVE_PREV = 2, LM_PENDING = 3, LM_X = 4 /* 8 floats */, LM_BEST_SCALAR = 12,
  LM_BEST_T = 13 /* 16 floats */, LM_BEST_A = 29, LM_BEST_B = 30, LM_CONVERGED = 31, LM_ITERS = 32, LM_SCALAR = 33,
  LM_STEP_NORM = 34, LM_HAVE_BEST = 35
};

// best_viewpoint_params of the reference (utils/slam_frontend.py:523-531): the pose and exposure the frame was RENDERED
// with whenever the scalar ||residual||_1 improves.  T: pre-update pose; a, b: pre-update exposure.
static __device__ __forceinline__ void lm_track_best(float* __restrict__ st, const float scalar, const float* __restrict__ T,
                                              const float a, const float b) {
  st[LM_SCALAR] = scalar;
  if (st[LM_HAVE_BEST] == 0.f || scalar < st[LM_BEST_SCALAR]) {
    st[LM_HAVE_BEST] = 1.f; st[LM_BEST_SCALAR] = scalar;
#pragma unroll
    for (int i = 0; i < 16; i++) st[LM_BEST_T + i] = T[i];
    st[LM_BEST_A] = a; st[LM_BEST_B] = b;
  }
}

// One CTA.  Everything the serial tail needs is staged in shared memory by separate threads while the per-tile partials
// are being reduced (a single thread chasing ~60 dependent global loads and stores costs an L2 round trip each), the
// double-precision bias corrections are computed by another warp meanwhile, and the results go out in parallel.
// Callable from a CTA of 128 or 256 threads (the roles below use tid < 96): the stand-alone track_update_kernel, or the LAST CTA of blend_track_kernel (ticket in
// blend_track.cu), which saves a launch and the tau / counter memsets.  Inputs produced by other CTAs (partials, tau_raw)
// are read with ld.global.cg: they were written through L2 and this SM's L1 may not be trusted.
// beta^t for an integer step count by repeated squaring in double: ~2 log2(t) dependent multiplies instead of the generic
// pow() (which cost the serial tail 2 us); relative error < 1e-14, invisible after the rounding to float the bias
// corrections go through.
static __device__ __forceinline__ double ipow(double b, int t) {
  double r = 1.0;
  while (t > 0) { if (t & 1) r *= b; b *= b; t >>= 1; }
  return r;
}

static __device__ __noinline__ void track_update_body(const TrackUpdateK& k) {
  __shared__ double red[8][4];
  __shared__ float s_in[64];     // [0,6) tau_raw  [8,16) exp_avg  [16,24) exp_avg_sq  [24,32) params  [32,48) T  [48,64) proj_raw
  __shared__ float s_out[72];    // [0,6) tau [6,8) dab [8,16) exp_avg [16,24) exp_avg_sq [24,32) params [32,48) T [48,64) view [64] loss
  __shared__ float s_proj[16];
  __shared__ float s_bc[2];
  __shared__ float s_best[2];    // LM_HAVE_BEST, LM_BEST_SCALAR of the tracker record, fetched while the partials are summed
  __shared__ int s_conv;
  const int tid = threadIdx.x;
  const int nwarp = (int)(blockDim.x >> 5);
  const bool track_best = k.lm_state != nullptr && k.mode == MGS_LOSS_TRACK_HUBER;
  if (tid < 6) s_in[tid] = __ldcg(k.tau_raw + tid);
  else if (tid >= 8 && tid < 16) s_in[tid] = k.exp_avg[tid - 8];
  else if (tid >= 16 && tid < 24) s_in[tid] = k.exp_avg_sq[tid - 16];
  else if (tid >= 24 && tid < 32) s_in[tid] = k.params[tid - 24];
  else if (tid >= 32 && tid < 48) s_in[tid] = k.T[tid - 32];
  else if (tid >= 48 && tid < 64) s_in[tid] = k.proj_raw ? k.proj_raw[tid - 48] : 0.f;
  else if (tid == 64) {
    const int t = k.step[0] + 1;
    k.step[0] = t;
    s_bc[0] = (float)(1.0 - ipow((double)k.beta1, t));
    s_bc[1] = (float)sqrt(1.0 - ipow((double)k.beta2, t));
  } else if (tid >= 66 && tid < 68) {
    if (track_best) s_best[tid - 66] = k.lm_state[tid == 66 ? LM_HAVE_BEST : LM_BEST_SCALAR];
  }
  double acc[4] = {0, 0, 0, 0};
  const double2* p2 = reinterpret_cast<const double2*>(k.partials);
#pragma unroll 8
  for (int i = tid; i < k.tiles; i += (int)blockDim.x) {
    const double2 a = __ldcg(p2 + 2 * i), b = __ldcg(p2 + 2 * i + 1);
    acc[0] += a.x; acc[1] += a.y; acc[2] += b.x; acc[3] += b.y;
  }
#pragma unroll
  for (int q = 0; q < 4; q++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[q] += __shfl_xor_sync(0xffffffffu, acc[q], o);
    if ((tid & 31) == 0) red[tid >> 5][q] = acc[q];
  }
  __syncthreads();
  if (tid < 32) {
    // warp 0: lane i < 8 owns parameter i of [rho(3), theta(3), a, b]; every lane forms the four totals itself
    double tot[4];
#pragma unroll
    for (int q = 0; q < 4; q++) {
      double t = 0;
      for (int w = 0; w < nwarp; w++) t += red[w][q];
      tot[q] = t;
    }
    float gscale = 1.f, loss;
    if (k.mode == MGS_LOSS_TRACK_HUBER) {
      const float nrm = (float)sqrt(tot[0]);
      loss = k.scale * nrm;
      gscale = nrm > 0.f ? k.scale / nrm : 0.f;
    } else {
      const double rgb = tot[0] / (3.0 * k.npix), dep = tot[3] / (double)k.npix;
      loss = (float)(k.scale * (k.use_depth ? k.alpha * rgb + (1.0 - k.alpha) * dep : rgb));
    }
    if (tid == 0) s_out[64] = loss;
    if (tid < 8) {
      const int i = tid;
      const float g = (i < 6 ? s_in[i] : (float)(i == 6 ? tot[1] : tot[2])) * gscale;      // tau | d/da, d/db
      s_out[i] = g;
      // Adam (torch.optim.Adam, eps outside the bias-corrected sqrt)
      const float bc1 = s_bc[0], bc2s = s_bc[1];
      const float lr = k.lr[i < 3 ? 0 : (i < 6 ? 1 : i - 4)];
      float m = s_in[8 + i], v = s_in[16 + i];
      m = m + (g - m) * (1.f - k.beta1);
      v = v * k.beta2 + (1.f - k.beta2) * g * g;
      const float denom = sqrtf(v) / bc2s + k.eps;
      s_out[24 + i] = s_in[24 + i] - (lr / bc1) * (m / denom);
      s_out[8 + i] = m; s_out[16 + i] = v;
    } else if (tid < 24) s_out[32 + tid - 8] = s_in[32 + tid - 8];
    __syncwarp();
    pose_update_warp(s_out + 32, s_out + 24, s_out + 27, k.proj_raw ? s_in + 48 : nullptr, k.view_out ? s_out + 48 : nullptr,
                     k.proj_out ? s_proj : nullptr, k.converged ? &s_conv : nullptr, k.thr, tid);
  } else if (tid < 64 && track_best) {
    // warp 1: best_viewpoint_params (lm_track_best) beside the update — the PRE-update pose and exposure
    double t3 = 0;
    for (int w = 0; w < nwarp; w++) t3 += red[w][3];
    const float scalar = (float)t3;
    const int l = tid - 32;
    float* st = k.lm_state;
    if (l == 0) st[LM_SCALAR] = scalar;
    if (s_best[0] == 0.f || scalar < s_best[1]) {
      if (l < 16) st[LM_BEST_T + l] = s_in[32 + l];
      else if (l == 16) st[LM_BEST_A] = s_in[30];
      else if (l == 17) st[LM_BEST_B] = s_in[31];
      else if (l == 18) st[LM_HAVE_BEST] = 1.f;
      else if (l == 19) st[LM_BEST_SCALAR] = scalar;
    }
  }
  __syncthreads();
  if (tid < 6) k.tau[tid] = s_out[tid];
  else if (tid < 8) k.dab[tid - 6] = s_out[tid];
  else if (tid < 16) k.exp_avg[tid - 8] = s_out[tid];
  else if (tid < 24) k.exp_avg_sq[tid - 16] = s_out[tid];
  else if (tid < 32) k.params[tid - 24] = s_out[tid];
  else if (tid < 48) k.T[tid - 32] = s_out[tid];
  else if (tid < 64) { if (k.view_out) k.view_out[tid - 48] = s_out[tid]; }
  else if (tid < 80) { if (k.view_out && k.proj_out && k.proj_raw) k.proj_out[tid - 64] = s_proj[tid - 64]; }
  else if (tid == 80) k.loss[0] = s_out[64];
  else if (tid == 81) { if (k.converged) k.converged[0] = s_conv; }
  else if (tid >= 96 && tid < 102) { if (k.tau_clear) k.tau_clear[tid - 96] = 0.f; }
}

}  // namespace mgs
