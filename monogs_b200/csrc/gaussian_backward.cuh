// Per-Gaussian backward chain shared by the full backward and the pose-Jacobian builder.
// Restates computeCov2DCUDA (backward.cu:154-361), preprocessCUDA (backward.cu:433-567),
// computeColorFromSH backward (backward.cu:25-149) and computeCov3D backward (backward.cu:365-428),
// fused so that one thread owns one Gaussian end to end and dL_dtau never round-trips through HBM.
#pragma once
#include "common.cuh"

namespace mgs {

// Everything that depends only on the Gaussian and the view (not on incoming gradients).
struct GaussCtx {
  float3 p;            // world mean
  float3 t;            // view-space mean with the FoV clamp applied to x,y (backward.cu:184-191)
  float xmul, ymul;    // 0 when the clamp was active (Q8)
  float cov6[6];       // world covariance (recomputed from scale/rot, or precomputed)
  M3 J, Wm, T, Vrk;
  float a, b, c;       // cov2D after the +0.3 low-pass
  float hx, hy;
};

__device__ __forceinline__ void cov3d_from_scale_rot_b(const float* __restrict__ scales, const float* __restrict__ rots,
                                                       int idx, float mod, float* cov6) {
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

__device__ __forceinline__ void gauss_ctx_init(GaussCtx& g, const ViewConst& v, const float3& p) {
  g.p = p; g.hx = v.focal_x; g.hy = v.focal_y;
  float3 t = xform4x3(p, v.view);
  const float limx = 1.3f * v.tanfovx, limy = 1.3f * v.tanfovy;
  const float txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = min(limx, max(-limx, txtz)) * t.z;
  t.y = min(limy, max(-limy, tytz)) * t.z;
  g.xmul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
  g.ymul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
  g.t = t;
  g.J = m3_cols(g.hx / t.z, 0.0f, -(g.hx * t.x) / (t.z * t.z), 0.0f, g.hy / t.z, -(g.hy * t.y) / (t.z * t.z), 0.f, 0.f, 0.f);
  g.Wm = m3_cols(v.view[0], v.view[4], v.view[8], v.view[1], v.view[5], v.view[9], v.view[2], v.view[6], v.view[10]);
  g.Vrk = m3_cols(g.cov6[0], g.cov6[1], g.cov6[2], g.cov6[1], g.cov6[3], g.cov6[4], g.cov6[2], g.cov6[4], g.cov6[5]);
  g.T = m3_mul(g.Wm, g.J);
  M3 cov2D = m3_mul(m3_mul(m3_t(g.T), m3_t(g.Vrk)), g.T);
  g.a = cov2D[0][0] + 0.3f; g.b = cov2D[0][1]; g.c = cov2D[1][1] + 0.3f;
}

struct ConicChainOut {
  float dcov6[6];   // dL/d(cov3D)
  float3 dmean;     // dL/d(mean) through the covariance projection
  float tau[6];     // pose tangent contribution (rho, theta)
};

// dL/dconic (xx, xy, yy) -> cov3D, mean and pose tangent  (backward.cu:213-358)
__device__ __forceinline__ void conic_chain(const GaussCtx& g, const ViewConst& v, float dcx, float dcy, float dcz,
                                            ConicChainOut& o) {
  const float a = g.a, b = g.b, c = g.c;
  const M3& T = g.T; const M3& Vrk = g.Vrk; const M3& W = g.Wm; const M3& J = g.J;
  const float denom = a * c - b * b;
  float dL_da = 0, dL_db = 0, dL_dc = 0;
  const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
  if (denom2inv != 0) {
    dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
    dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
    dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
    o.dcov6[0] = (T[0][0] * T[0][0] * dL_da + T[0][0] * T[1][0] * dL_db + T[1][0] * T[1][0] * dL_dc);
    o.dcov6[3] = (T[0][1] * T[0][1] * dL_da + T[0][1] * T[1][1] * dL_db + T[1][1] * T[1][1] * dL_dc);
    o.dcov6[5] = (T[0][2] * T[0][2] * dL_da + T[0][2] * T[1][2] * dL_db + T[1][2] * T[1][2] * dL_dc);
    o.dcov6[1] = 2 * T[0][0] * T[0][1] * dL_da + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][1] * dL_dc;
    o.dcov6[2] = 2 * T[0][0] * T[0][2] * dL_da + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][2] * dL_dc;
    o.dcov6[4] = 2 * T[0][2] * T[0][1] * dL_da + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_db + 2 * T[1][1] * T[1][2] * dL_dc;
  } else {
#pragma unroll
    for (int i = 0; i < 6; i++) o.dcov6[i] = 0;
  }
  // rows of T times Vrk, reused by the six dL/dT entries
  float tv0[3], tv1[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    tv0[k] = T[0][0] * Vrk[k][0] + T[0][1] * Vrk[k][1] + T[0][2] * Vrk[k][2];
    tv1[k] = T[1][0] * Vrk[k][0] + T[1][1] * Vrk[k][1] + T[1][2] * Vrk[k][2];
  }
  const float dT00 = 2 * tv0[0] * dL_da + tv1[0] * dL_db;
  const float dT01 = 2 * tv0[1] * dL_da + tv1[1] * dL_db;
  const float dT02 = 2 * tv0[2] * dL_da + tv1[2] * dL_db;
  const float dT10 = 2 * tv1[0] * dL_dc + tv0[0] * dL_db;
  const float dT11 = 2 * tv1[1] * dL_dc + tv0[1] * dL_db;
  const float dT12 = 2 * tv1[2] * dL_dc + tv0[2] * dL_db;

  const float dJ00 = W[0][0] * dT00 + W[0][1] * dT01 + W[0][2] * dT02;
  const float dJ02 = W[2][0] * dT00 + W[2][1] * dT01 + W[2][2] * dT02;
  const float dJ11 = W[1][0] * dT10 + W[1][1] * dT11 + W[1][2] * dT12;
  const float dJ12 = W[2][0] * dT10 + W[2][1] * dT11 + W[2][2] * dT12;

  const float3 t = g.t;
  const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
  const float dtx = g.xmul * -g.hx * tz2 * dJ02;
  const float dty = g.ymul * -g.hy * tz2 * dJ12;
  const float dtz = -g.hx * tz2 * dJ00 - g.hy * tz2 * dJ11 + (2 * g.hx * t.x) * tz3 * dJ02 + (2 * g.hy * t.y) * tz3 * dJ12;

  // d p_C / d rho = I ; d p_C / d theta = -[t]x, with the clamped t (backward.cu:288-303)
  o.tau[0] = dtx; o.tau[1] = dty; o.tau[2] = dtz;
  o.tau[3] = dty * -t.z + dtz * t.y;
  o.tau[4] = dtx * t.z + dtz * -t.x;
  o.tau[5] = dtx * -t.y + dty * t.x;

  o.dmean = make_float3(v.view[0] * dtx + v.view[1] * dty + v.view[2] * dtz,
                        v.view[4] * dtx + v.view[5] * dty + v.view[6] * dtz,
                        v.view[8] * dtx + v.view[9] * dty + v.view[10] * dtz);

  // rotation part of the pose tangent through W (backward.cu:314-358)
  const float dW00 = J[0][0] * dT00, dW01 = J[0][0] * dT01, dW02 = J[0][0] * dT02;
  const float dW10 = J[1][1] * dT10, dW11 = J[1][1] * dT11, dW12 = J[1][1] * dT12;
  const float dW20 = J[0][2] * dT00 + J[1][2] * dT10;
  const float dW21 = J[0][2] * dT01 + J[1][2] * dT11;
  const float dW22 = J[0][2] * dT02 + J[1][2] * dT12;
  // columns of R_cw and of dL/dW (column k = (dW0k, dW1k, dW2k))
  const float3 c1 = {v.view[0], v.view[1], v.view[2]}, c2 = {v.view[4], v.view[5], v.view[6]}, c3 = {v.view[8], v.view[9], v.view[10]};
  const float3 g1 = {dW00, dW10, dW20}, g2 = {dW01, dW11, dW21}, g3 = {dW02, dW12, dW22};
  // dot(g, col_j(-[c]x)): col0 = (0,-c.z,c.y), col1 = (c.z,0,-c.x), col2 = (-c.y,c.x,0)
  o.tau[3] += (-g1.y * c1.z + g1.z * c1.y) + (-g2.y * c2.z + g2.z * c2.y) + (-g3.y * c3.z + g3.z * c3.y);
  o.tau[4] += (g1.x * c1.z - g1.z * c1.x) + (g2.x * c2.z - g2.z * c2.x) + (g3.x * c3.z - g3.z * c3.x);
  o.tau[5] += (-g1.x * c1.y + g1.y * c1.x) + (-g2.x * c2.y + g2.y * c2.x) + (-g3.x * c3.y + g3.y * c3.x);
}

// Projection constants of one Gaussian for the mean2D / depth chains (backward.cu:474-556).
struct ProjCtx {
  float m_w, mul1, mul2;     // 1/(w+eps), and the two w-derivative factors
  float alpha, beta, gamma;  // d(ndc)/d(p_C) factors
  float3 pC;                 // camera-space point (R m + t)
};
__device__ __forceinline__ void proj_ctx_init(ProjCtx& q, const ViewConst& v, const float3& m) {
  const float* proj = v.proj;
  const float4 mh = xform4x4(m, proj);
  q.m_w = 1.0f / (mh.w + 0.0000001f);
  q.mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * q.m_w * q.m_w;
  q.mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * q.m_w * q.m_w;
  q.alpha = 1.0f * q.m_w;
  q.beta = -mh.x * q.m_w * q.m_w;
  q.gamma = -mh.y * q.m_w * q.m_w;
  const float* vm = v.view;
  q.pC = make_float3(vm[0] * m.x + vm[4] * m.y + vm[8] * m.z + vm[12],
                     vm[1] * m.x + vm[5] * m.y + vm[9] * m.z + vm[13],
                     vm[2] * m.x + vm[6] * m.y + vm[10] * m.z + vm[14]);
}
// dL/dmean2D (x,y) -> dL/dmean3D
__device__ __forceinline__ float3 mean2d_to_mean3d(const ProjCtx& q, const float* __restrict__ proj, float dmx, float dmy) {
  float3 d;
  d.x = (proj[0] * q.m_w - proj[3] * q.mul1) * dmx + (proj[1] * q.m_w - proj[3] * q.mul2) * dmy;
  d.y = (proj[4] * q.m_w - proj[7] * q.mul1) * dmx + (proj[5] * q.m_w - proj[7] * q.mul2) * dmy;
  d.z = (proj[8] * q.m_w - proj[11] * q.mul1) * dmx + (proj[9] * q.m_w - proj[11] * q.mul2) * dmy;
  return d;
}
// d(mean2D.x)/d tau and d(mean2D.y)/d tau: two 6-vectors (backward.cu:497-540)
__device__ __forceinline__ void mean2d_tau_jac(const ProjCtx& q, const float* __restrict__ proj_raw, float jx[6], float jy[6]) {
  const float a = proj_raw[0], b = proj_raw[5], e = proj_raw[11];
  const float3 v1 = {q.alpha * a, 0.f, q.beta * e};
  const float3 v2 = {0.f, q.alpha * b, q.gamma * e};
  const float3 p = q.pC;
  jx[0] = v1.x; jx[1] = v1.y; jx[2] = v1.z;
  jy[0] = v2.x; jy[1] = v2.y; jy[2] = v2.z;
  // (-[p]x)^T v = p x v
  jx[3] = p.y * v1.z - p.z * v1.y; jx[4] = p.z * v1.x - p.x * v1.z; jx[5] = p.x * v1.y - p.y * v1.x;
  jy[3] = p.y * v2.z - p.z * v2.y; jy[4] = p.z * v2.x - p.x * v2.z; jy[5] = p.x * v2.y - p.y * v2.x;
}

// SH backward (backward.cu:25-149).  Writes dL/dsh for the active coefficients (inactive ones are zeroed by
// the caller) and returns dL/dmean through the view direction.  dcol must already carry the clamp mask.
// sh and dsh may be the same row (shared-memory staging, sh_stage.cuh): every coefficient is read before dsh is written.
__device__ __forceinline__ float3 sh_backward(int deg, int M, const float3& pos, const float* __restrict__ campos,
                                              const float* sh, const float3 dcol,
                                              float* dsh /* may be null */, const int accumulate = 0 /* 1: +=, 2: atomicAdd */) {
  const float3 dir_orig = {pos.x - campos[0], pos.y - campos[1], pos.z - campos[2]};
  const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
  const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
  float w[16];
  float3 dx3 = {0, 0, 0}, dy3 = {0, 0, 0}, dz3 = {0, 0, 0};  // dRGB/dx, /dy, /dz per channel
  auto S = [&](int k) { return make_float3(sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]); };
  auto axpy = [](float3& acc, float s, const float3& vv) { acc.x += s * vv.x; acc.y += s * vv.y; acc.z += s * vv.z; };
  w[0] = kSH0;
  if (deg > 0) {
    w[1] = -kSH1 * y; w[2] = kSH1 * z; w[3] = -kSH1 * x;
    axpy(dx3, -kSH1, S(3)); axpy(dy3, -kSH1, S(1)); axpy(dz3, kSH1, S(2));
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      w[4] = kSH2[0] * xy; w[5] = kSH2[1] * yz; w[6] = kSH2[2] * (2.f * zz - xx - yy);
      w[7] = kSH2[3] * xz; w[8] = kSH2[4] * (xx - yy);
      axpy(dx3, kSH2[0] * y, S(4)); axpy(dx3, kSH2[2] * 2.f * -x, S(6)); axpy(dx3, kSH2[3] * z, S(7)); axpy(dx3, kSH2[4] * 2.f * x, S(8));
      axpy(dy3, kSH2[0] * x, S(4)); axpy(dy3, kSH2[1] * z, S(5)); axpy(dy3, kSH2[2] * 2.f * -y, S(6)); axpy(dy3, kSH2[4] * 2.f * -y, S(8));
      axpy(dz3, kSH2[1] * y, S(5)); axpy(dz3, kSH2[2] * 2.f * 2.f * z, S(6)); axpy(dz3, kSH2[3] * x, S(7));
      if (deg > 2) {
        w[9] = kSH3[0] * y * (3.f * xx - yy);
        w[10] = kSH3[1] * xy * z;
        w[11] = kSH3[2] * y * (4.f * zz - xx - yy);
        w[12] = kSH3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
        w[13] = kSH3[4] * x * (4.f * zz - xx - yy);
        w[14] = kSH3[5] * z * (xx - yy);
        w[15] = kSH3[6] * x * (xx - 3.f * yy);
        axpy(dx3, kSH3[0] * 3.f * 2.f * xy, S(9)); axpy(dx3, kSH3[1] * yz, S(10)); axpy(dx3, kSH3[2] * -2.f * xy, S(11));
        axpy(dx3, kSH3[3] * -3.f * 2.f * xz, S(12)); axpy(dx3, kSH3[4] * (-3.f * xx + 4.f * zz - yy), S(13));
        axpy(dx3, kSH3[5] * 2.f * xz, S(14)); axpy(dx3, kSH3[6] * 3.f * (xx - yy), S(15));
        axpy(dy3, kSH3[0] * 3.f * (xx - yy), S(9)); axpy(dy3, kSH3[1] * xz, S(10)); axpy(dy3, kSH3[2] * (-3.f * yy + 4.f * zz - xx), S(11));
        axpy(dy3, kSH3[3] * -3.f * 2.f * yz, S(12)); axpy(dy3, kSH3[4] * -2.f * xy, S(13));
        axpy(dy3, kSH3[5] * -2.f * yz, S(14)); axpy(dy3, kSH3[6] * -3.f * 2.f * xy, S(15));
        axpy(dz3, kSH3[1] * xy, S(10)); axpy(dz3, kSH3[2] * 4.f * 2.f * yz, S(11)); axpy(dz3, kSH3[3] * 3.f * (2.f * zz - xx - yy), S(12));
        axpy(dz3, kSH3[4] * 4.f * 2.f * xz, S(13)); axpy(dz3, kSH3[5] * (xx - yy), S(14));
      }
    }
  }
  const int n = (deg + 1) * (deg + 1);
  if (dsh != nullptr && accumulate == 2) {   // views of a mapping window running concurrently
    for (int k = 0; k < n; k++) { atomicAdd(dsh + 3 * k, w[k] * dcol.x); atomicAdd(dsh + 3 * k + 1, w[k] * dcol.y); atomicAdd(dsh + 3 * k + 2, w[k] * dcol.z); }
  } else if (dsh != nullptr && accumulate) {   // mapping window: gradients of several views summed in place
    for (int k = 0; k < n; k++) { dsh[3 * k] += w[k] * dcol.x; dsh[3 * k + 1] += w[k] * dcol.y; dsh[3 * k + 2] += w[k] * dcol.z; }
  } else if (dsh != nullptr) {
    for (int k = 0; k < n; k++) { dsh[3 * k] = w[k] * dcol.x; dsh[3 * k + 1] = w[k] * dcol.y; dsh[3 * k + 2] = w[k] * dcol.z; }
    for (int k = n; k < M; k++) { dsh[3 * k] = 0.f; dsh[3 * k + 1] = 0.f; dsh[3 * k + 2] = 0.f; }
  }
  const float3 ddir = {dx3.x * dcol.x + dx3.y * dcol.y + dx3.z * dcol.z,
                       dy3.x * dcol.x + dy3.y * dcol.y + dy3.z * dcol.z,
                       dz3.x * dcol.x + dz3.y * dcol.y + dz3.z * dcol.z};
  // gradient through v / |v| (auxiliary.h:101-112)
  const float3 vv = dir_orig;
  const float sum2 = vv.x * vv.x + vv.y * vv.y + vv.z * vv.z;
  const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  float3 dm;
  dm.x = ((+sum2 - vv.x * vv.x) * ddir.x - vv.y * vv.x * ddir.y - vv.z * vv.x * ddir.z) * invsum32;
  dm.y = (-vv.x * vv.y * ddir.x + (sum2 - vv.y * vv.y) * ddir.y - vv.z * vv.y * ddir.z) * invsum32;
  dm.z = (-vv.x * vv.z * ddir.x - vv.y * vv.z * ddir.y + (sum2 - vv.z * vv.z) * ddir.z) * invsum32;
  return dm;
}

// cov3D -> (scale, quaternion) backward (backward.cu:365-428); gradient w.r.t. the quaternion as given (Q2).
__device__ __forceinline__ void cov3d_backward(const float* __restrict__ scales, const float* __restrict__ rots, int idx,
                                               float mod, const float* dcov6, float3& dscale, float4& drot) {
  const float4 q = reinterpret_cast<const float4*>(rots)[idx];
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  M3 R = m3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                 2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                 2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
  const float3 s = {mod * scales[3 * idx], mod * scales[3 * idx + 1], mod * scales[3 * idx + 2]};
  M3 S = m3_cols(s.x, 0.f, 0.f, 0.f, s.y, 0.f, 0.f, 0.f, s.z);
  M3 Mm = m3_mul(S, R);
  M3 dSigma = m3_cols(dcov6[0], 0.5f * dcov6[1], 0.5f * dcov6[2], 0.5f * dcov6[1], dcov6[3], 0.5f * dcov6[4],
                      0.5f * dcov6[2], 0.5f * dcov6[4], dcov6[5]);
  M3 dM = m3_mul(Mm, dSigma);
#pragma unroll
  for (int cc = 0; cc < 3; cc++)
#pragma unroll
    for (int rr = 0; rr < 3; rr++) dM.m[cc][rr] *= 2.0f;
  M3 Rt = m3_t(R);
  M3 dMt = m3_t(dM);
  dscale.x = Rt[0][0] * dMt[0][0] + Rt[0][1] * dMt[0][1] + Rt[0][2] * dMt[0][2];
  dscale.y = Rt[1][0] * dMt[1][0] + Rt[1][1] * dMt[1][1] + Rt[1][2] * dMt[1][2];
  dscale.z = Rt[2][0] * dMt[2][0] + Rt[2][1] * dMt[2][1] + Rt[2][2] * dMt[2][2];
#pragma unroll
  for (int k = 0; k < 3; k++) { dMt.m[0][k] *= s.x; dMt.m[1][k] *= s.y; dMt.m[2][k] *= s.z; }
  drot.x = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
  drot.y = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
  drot.z = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
  drot.w = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
}

}  // namespace mgs
