// Builder of the per-Gaussian linear map screen-space gradients -> pose tangent (see pose_backward.cu).
#pragma once
#include "gaussian_backward.cuh"

namespace mgs {

// B_g layout: NCOL columns (mx, my, cxx, cxy, cyy, depth[, r, g, b]) of 6 floats each, padded to float4s.
template <int NCOL> struct PoseJac { static constexpr int kFloats = (NCOL * 6 + 3) / 4 * 4; };

// The single-pass tracking blend (blend_track.cu) consumes the map in a folded form: 32 floats = one 128-byte line per
// Gaussian.  With d = mean2D - pixel, G = exp(power) and s = opacity * G / (1 - alpha), a pixel's pose-tangent
// contribution through this splat is (backward.cu:986-1010 chained with the columns of B)
//     vp = s * ( d.x * (U + d.x * A + d.y * Bm)  +  d.y * (V + d.y * Cm) )
//     U  = -(conic.xx * W/2 * Bx + conic.xy * H/2 * By)      V  = -(conic.xy * W/2 * Bx + conic.yy * H/2 * By)
//     A  = -1/2 Bxx        Bm = -1/2 Bxy        Cm = -1/2 Byy
// so the five per-pixel scalars q_k of the column form never have to be formed: 15 packed multiply-adds instead of
// 15 + ~14 scalar instructions per (pixel, splat).  Floats 30, 31 = y_C, -x_C: the depth column (0, 0, 1, y_C, -x_C, 0).
constexpr int POSE_BASIS_FLOATS = 32;
// Packed per-Gaussian row of the fused tracker (blend_track_kernel<ROWS>): the 12 floats of a SplatRec (x, y, depth, id |
// conic xyz, opacity | r, g, b, extent) followed by the 32-float basis — one 176-byte gather by id replaces the per-instance
// 48-byte record copy of the sort kernel plus the separate basis gather.
constexpr int TRACK_ROW_FLOATS = 12 + POSE_BASIS_FLOATS;

template <int NCOL>
__device__ __forceinline__ void compute_pose_B(const ViewConst& v, const int idx, const float3& p,
                                               const float* __restrict__ shs, const float* __restrict__ scales,
                                               const float* __restrict__ rotations, const float* __restrict__ cov3D_precomp,
                                               const uint32_t clamp_mask, float (&B)[PoseJac<NCOL>::kFloats]);

__device__ __forceinline__ void build_pose_basis(const ViewConst& v, const int idx, const float3& p, const float3& conic,
                                                 const float* __restrict__ scales, const float* __restrict__ rotations,
                                                 const float* __restrict__ cov3D_precomp, float* __restrict__ basis,
                                                 const int row_floats = POSE_BASIS_FLOATS, const int row_offset = 0,
                                                 const int out_row = -1 /* row to write; default: idx */) {
  float B[PoseJac<6>::kFloats];
  compute_pose_B<6>(v, idx, p, nullptr, scales, rotations, cov3D_precomp, 0u, B);
  const float wx = 0.5f * (float)v.W, wy = 0.5f * (float)v.H;
  float o[POSE_BASIS_FLOATS];
#pragma unroll
  for (int k = 0; k < 6; k++) {
    const float bx = wx * B[0 * 6 + k], by = wy * B[1 * 6 + k];
    o[k] = -(conic.x * bx + conic.y * by);
    o[6 + k] = -(conic.y * bx + conic.z * by);
    o[12 + k] = -0.5f * B[2 * 6 + k];
    o[18 + k] = -0.5f * B[3 * 6 + k];
    o[24 + k] = -0.5f * B[4 * 6 + k];
  }
  o[30] = B[5 * 6 + 3]; o[31] = B[5 * 6 + 4];
  float4* out = reinterpret_cast<float4*>(basis + (size_t)(out_row >= 0 ? out_row : idx) * row_floats + row_offset);
#pragma unroll
  for (int k = 0; k < POSE_BASIS_FLOATS / 4; k++) out[k] = make_float4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]);
}

template <int NCOL>
__device__ __forceinline__ void compute_pose_B(const ViewConst& v, const int idx, const float3& p,
                                               const float* __restrict__ shs, const float* __restrict__ scales,
                                               const float* __restrict__ rotations, const float* __restrict__ cov3D_precomp,
                                               const uint32_t clamp_mask, float (&B)[PoseJac<NCOL>::kFloats]) {
  constexpr int NF = PoseJac<NCOL>::kFloats;
#pragma unroll
  for (int k = 0; k < NF; k++) B[k] = 0.f;
  GaussCtx g;
  if (cov3D_precomp != nullptr) {
#pragma unroll
    for (int k = 0; k < 6; k++) g.cov6[k] = cov3D_precomp[6 * (size_t)idx + k];
  } else cov3d_from_scale_rot_b(scales, rotations, idx, v.scale_modifier, g.cov6);
  gauss_ctx_init(g, v, p);
  ProjCtx q;
  proj_ctx_init(q, v, p);
  float jx[6], jy[6];
  mean2d_tau_jac(q, v.proj_raw, jx, jy);
#pragma unroll
  for (int k = 0; k < 6; k++) { B[0 * 6 + k] = jx[k]; B[1 * 6 + k] = jy[k]; }
#pragma unroll
  for (int c = 0; c < 3; c++) {
    ConicChainOut cc;
    conic_chain(g, v, c == 0 ? 1.f : 0.f, c == 1 ? 1.f : 0.f, c == 2 ? 1.f : 0.f, cc);
#pragma unroll
    for (int k = 0; k < 6; k++) B[(2 + c) * 6 + k] = cc.tau[k];
  }
  B[5 * 6 + 2] = 1.f; B[5 * 6 + 3] = q.pC.y; B[5 * 6 + 4] = -q.pC.x;
  if (NCOL == 9) {
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const float on = ((clamp_mask >> c) & 1u) ? 0.f : 1.f;
      const float3 dcol = {c == 0 ? on : 0.f, c == 1 ? on : 0.f, c == 2 ? on : 0.f};
      const float3 dm = sh_backward(v.D, v.M, p, v.campos, shs + (size_t)idx * v.M * 3, dcol, nullptr);
      B[(6 + c) * 6 + 0] = -dm.x; B[(6 + c) * 6 + 1] = -dm.y; B[(6 + c) * 6 + 2] = -dm.z;
    }
  }
}

template <int NCOL>
__device__ __forceinline__ void build_pose_jac(const ViewConst& v, const int idx, const float3& p,
                                               const float* __restrict__ shs, const float* __restrict__ scales,
                                               const float* __restrict__ rotations, const float* __restrict__ cov3D_precomp,
                                               const uint32_t clamp_mask, float* __restrict__ jac) {
  constexpr int NF = PoseJac<NCOL>::kFloats;
  float B[NF];
  compute_pose_B<NCOL>(v, idx, p, shs, scales, rotations, cov3D_precomp, clamp_mask, B);
  float4* out = reinterpret_cast<float4*>(jac + (size_t)idx * NF);
#pragma unroll
  for (int k = 0; k < NF / 4; k++) out[k] = make_float4(B[4 * k], B[4 * k + 1], B[4 * k + 2], B[4 * k + 3]);
}

}  // namespace mgs
