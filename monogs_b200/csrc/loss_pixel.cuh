// Per-pixel loss term + its gradient, shared by the standalone loss kernel (loss_adam.cu) and the fused epilogue
// of the tracking forward blend (track_fused.cu), so both produce identical numbers.
// Semantics: MonoGS utils/slam_utils.py:83-113,188-253 and HuberLoss :58-75 (see loss_adam.cu).
#pragma once
#include "common.cuh"

namespace mgs {

struct LossK {
  int mode, npix, use_depth;
  float thr, alpha, delta, scale, eps;
  const float *image, *depth, *opacity, *gt, *gt_depth, *mask, *a, *b;
};

__device__ __forceinline__ float sgn(float x) { return (float)((x > 0.f) - (x < 0.f)); }

// One pixel.  img[3], depth, op are the rendered values; grad_scale multiplies the colour gradient:
//   L1 modes   : scale * (use_depth ? alpha : 1) / (3 npix)
//   Huber mode : scale / ||h||  (or 1 when the caller normalises afterwards — the loss is linear in it)
// acc[0] += |r| (L1) or h^2 (Huber); acc[1] += dL/da; acc[2] += dL/db; acc[3] += |depth residual|.
// g_img[3] / g_depth receive the gradients when WRITE_GRADS.
template <bool WRITE_GRADS, bool REDUCE>
__device__ __forceinline__ void loss_pixel(const LossK& k, const int p, const float (&img)[3], const float depth,
                                           const float op, const float grad_scale, float (&acc)[4],
                                           float (&g_img)[3], float& g_depth) {
  const float a_raw = k.a ? k.a[0] : 1.f;
  const float ea = k.a ? fabsf(a_raw) + k.eps : 1.f;
  const float eb = k.b ? k.b[0] : 0.f;
  const float sa = k.a ? sgn(a_raw) : 0.f;
  const float gt3[3] = {k.gt[p], k.gt[k.npix + p], k.gt[2 * k.npix + p]};
  float m = (gt3[0] + gt3[1] + gt3[2]) > k.thr ? 1.f : 0.f;
  if (k.mask) m *= k.mask[p];
  const float w = (k.mode == MGS_LOSS_MAP_L1) ? 1.f : op;   // mapping loss has no opacity weight
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const float iab = ea * img[c] + eb;
    const float r = w * (iab * m - gt3[c] * m);
    float dr;  // dLoss/dr up to grad_scale
    if (k.mode == MGS_LOSS_TRACK_HUBER) {
      const float ar = fabsf(r);
      float h, dh;
      if (ar < k.delta) { h = r; dh = 1.f; }
      else { const float sq = sqrtf(2.f * k.delta * ar - k.delta * k.delta); h = sq * sgn(r); dh = k.delta / sq; }
      if (REDUCE) { acc[0] += h * h; acc[3] += ar; }   // [3]: ||r||_1, the scalar the reference keeps for its best-pose test
      dr = h * dh;
    } else {
      if (REDUCE) acc[0] += fabsf(r);
      dr = sgn(r);
    }
    if (WRITE_GRADS) {
      const float gi = grad_scale * dr * w * m;   // d r / d iab = w*m
      g_img[c] = gi * ea;
      acc[1] += gi * img[c] * sa;
      acc[2] += gi;
    }
  }
  g_depth = 0.f;
  if (k.use_depth && k.mode != MGS_LOSS_TRACK_HUBER) {
    const float gd = k.gt_depth[p];
    float dm = gd > 0.01f ? 1.f : 0.f;
    if (k.mode == MGS_LOSS_TRACK_L1) dm *= (op > 0.95f) ? 1.f : 0.f;
    const float rd = depth * dm - gd * dm;
    if (REDUCE) acc[3] += fabsf(rd);
    if (WRITE_GRADS) g_depth = k.scale * (1.f - k.alpha) / (float)k.npix * sgn(rd) * dm;
  }
}

// One pixel of the SECOND-ORDER (sketched Gauss-Newton) tracking objective, MonoGS utils/slam_frontend.py:637-665 with
// ApplyExposure.backward (utils/slam_utils.py:132-185):
//     r_c = opacity * (image_ab_c * m - gt_c * m),  h_c = Huber_delta(r_c),  f = w * sum_c h_c / (npix / d_total)
// where w = rand_weights[p] (+-1) and d_total = stack_dim * sketch_dim.  d f / d image_ab_c = g_ab_c; the gradient that
// reaches the rasterizer is |a| * g_ab_c — the reference's ApplyExposure.backward multiplies by torch.abs(exposure_a)
// WITHOUT the eps of the forward — and its sketched exposure columns are sum_c g_ab_c * image_c (no sign(a) factor) and
// sum_c g_ab_c.  Quirks reproduced, not fixed.
// Returns f; g_img[3] = gradient for the rasterizer; da, db = this pixel's exposure columns; l1 = sum_c |r_c| (the
// un-Hubered scalar the LM accept test uses, slam_frontend.py:510).
__device__ __forceinline__ float sketch_pixel(const LossK& k, const int p, const float (&img)[3], const float op,
                                              const float wscale /* w * d_total / npix */, float (&g_img)[3],
                                              float& da, float& db, float& l1) {
  const float a_raw = k.a[0], eb = k.b[0];
  const float aa = fabsf(a_raw), ea = aa + k.eps;
  const float gt3[3] = {k.gt[p], k.gt[k.npix + p], k.gt[2 * k.npix + p]};
  float m = (gt3[0] + gt3[1] + gt3[2]) > k.thr ? 1.f : 0.f;
  if (k.mask) m *= k.mask[p];
  float f = 0.f;
  da = 0.f; db = 0.f; l1 = 0.f;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const float iab = ea * img[c] + eb;
    const float r = op * (iab * m - gt3[c] * m);
    const float ar = fabsf(r);
    float h, dh;
    if (ar < k.delta) { h = r; dh = 1.f; }
    else { const float sq = sqrtf(2.f * k.delta * ar - k.delta * k.delta); h = sq * sgn(r); dh = k.delta / sq; }
    f += h;
    l1 += ar;
    const float g_ab = wscale * dh * op * m;
    g_img[c] = aa * g_ab;
    da += g_ab * img[c];
    db += g_ab;
  }
  return f * wscale;
}

// block-level reduction of acc[4] into one double4 slot (deterministic: shuffle tree -> shared -> fixed order)
__device__ __forceinline__ void loss_block_reduce(const float (&acc)[4], double* __restrict__ slot) {
  __shared__ float part[8][4];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < 4; q++) {
    float x = acc[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) part[warp][q] = x;
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    double x = 0.0;
    const int nw = (blockDim.x + 31) >> 5;
    for (int w8 = 0; w8 < nw; w8++) x += (double)part[w8][threadIdx.x];
    slot[threadIdx.x] = x;
  }
}

inline LossK make_loss_k(const mgs_loss_args* a) {
  LossK k;
  k.mode = a->mode; k.npix = a->H * a->W; k.use_depth = a->use_depth;
  k.thr = a->rgb_boundary_threshold; k.alpha = a->alpha; k.delta = a->huber_delta; k.scale = a->loss_scale;
  k.eps = 1e-8f;  // Camera.exposure_eps (utils/camera_utils.py:55)
  k.image = a->image; k.depth = a->depth; k.opacity = a->opacity; k.gt = a->gt_image;
  k.gt_depth = a->gt_depth; k.mask = a->grad_mask; k.a = a->exposure_a; k.b = a->exposure_b;
  return k;
}

}  // namespace mgs
