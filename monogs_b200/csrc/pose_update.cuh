// T <- Exp(rho, theta) T on the device (utils/pose_utils.py:26-98, math.h:217-339); single-thread device function
// shared by mgs_pose_update and the fused tracking update kernel.
#pragma once
#include "common.cuh"

namespace mgs {

__device__ __forceinline__ void pose_update_dev(float* __restrict__ T /* [4,4] row-major T_w2c */, float* __restrict__ rho,
                                   float* __restrict__ theta, const float* __restrict__ proj_raw,
                                   float* __restrict__ view_out, float* __restrict__ proj_out,
                                   int32_t* __restrict__ converged, float thr) {
  const double r[3] = {rho[0], rho[1], rho[2]}, th[3] = {theta[0], theta[1], theta[2]};
  double Wm[3][3] = {{0, -th[2], th[1]}, {th[2], 0, -th[0]}, {-th[1], th[0], 0}};
  double W2[3][3];
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { double s = 0; for (int k = 0; k < 3; k++) s += Wm[i][k] * Wm[k][j]; W2[i][j] = s; }
  const double ang = sqrt(th[0] * th[0] + th[1] * th[1] + th[2] * th[2]);
  double A, B, Cc;
  if (ang < 1e-5) { A = 1.0; B = 0.5; Cc = 1.0 / 6.0; }
  else { A = sin(ang) / ang; B = (1.0 - cos(ang)) / (ang * ang); Cc = (ang - sin(ang)) / (ang * ang * ang); }
  double E[4][4] = {{0}};
  for (int i = 0; i < 3; i++) {
    double tv = 0;
    for (int j = 0; j < 3; j++) {
      E[i][j] = (i == j ? 1.0 : 0.0) + A * Wm[i][j] + B * W2[i][j];
      tv += ((i == j ? 1.0 : 0.0) + B * Wm[i][j] + Cc * W2[i][j]) * r[j];
    }
    E[i][3] = tv;
  }
  E[3][3] = 1.0;
  double Tn[4][4];
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { double s = 0; for (int k = 0; k < 4; k++) s += E[i][k] * (double)T[4 * k + j]; Tn[i][j] = s; }
  float Tf[16];
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { Tf[4 * i + j] = (float)Tn[i][j]; T[4 * i + j] = Tf[4 * i + j]; }
  if (converged) {
    const double n2 = r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + th[0] * th[0] + th[1] * th[1] + th[2] * th[2];
    converged[0] = n2 < (double)thr * thr ? 1 : 0;
  }
  for (int k = 0; k < 3; k++) { rho[k] = 0.f; theta[k] = 0.f; }
  if (view_out) {  // world_view_transform = T^T (camera_utils.py:94-96): memory order [4*col+row] of T
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) view_out[4 * i + j] = Tf[4 * j + i];
    if (proj_out && proj_raw) {  // full_proj_transform = view @ projection (camera_utils.py:98-104)
      for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) {
        float s = 0.f;
        for (int k = 0; k < 4; k++) s += Tf[4 * k + i] * proj_raw[4 * k + j];
        proj_out[4 * i + j] = s;
      }
    }
  }
}

// The same update by ONE WARP (all 32 lanes call it; T, rho, theta must be in shared or global memory the warp can see):
// lane 4i+j < 16 owns element (i, j).  The coefficients come from the series in theta^2 when |theta| < 0.5 (9 terms,
// remainder < 1e-19: a tracking step is ~1e-3 rad) instead of double-precision sin / cos, the matrix products are one
// dot product per lane, and nothing is indexed dynamically — the single-thread version spent ~8 us of the tracking
// iteration's serial tail in ~1500 dependent double-precision instructions and local-memory arrays.
__device__ __forceinline__ void pose_update_warp(float* __restrict__ T, float* __restrict__ rho, float* __restrict__ theta,
                                                 const float* __restrict__ proj_raw, float* __restrict__ view_out,
                                                 float* __restrict__ proj_out, int32_t* __restrict__ converged, float thr,
                                                 const int lane) {
  const double r0 = rho[0], r1 = rho[1], r2 = rho[2], t0 = theta[0], t1 = theta[1], t2 = theta[2];
  const int i = (lane >> 2) & 3, j = lane & 3;
  const double Tk0 = T[j], Tk1 = T[4 + j], Tk2 = T[8 + j], Tk3 = T[12 + j];     // column j of T
  __syncwarp();                                                                 // every lane has read its inputs
  const double a2 = t0 * t0 + t1 * t1 + t2 * t2;
  double A, B, Cc;
  if (a2 < 0.25) {
    // A = sin a / a, B = (1 - cos a) / a^2, C = (a - sin a) / a^3 as alternating series in a^2 (Horner, three independent chains)
    A = 1.0 / 355687428096000.0; B = 1.0 / 6402373705728000.0; Cc = 1.0 / 121645100408832000.0;       // 1/17!, 1/18!, 1/19!
    A = 1.0 / 1307674368000.0 - a2 * A;  B = 1.0 / 20922789888000.0 - a2 * B;  Cc = 1.0 / 355687428096000.0 - a2 * Cc;   // 15! 16! 17!
    A = 1.0 / 6227020800.0 - a2 * A;     B = 1.0 / 87178291200.0 - a2 * B;     Cc = 1.0 / 1307674368000.0 - a2 * Cc;     // 13! 14! 15!
    A = 1.0 / 39916800.0 - a2 * A;       B = 1.0 / 479001600.0 - a2 * B;       Cc = 1.0 / 6227020800.0 - a2 * Cc;        // 11! 12! 13!
    A = 1.0 / 362880.0 - a2 * A;         B = 1.0 / 3628800.0 - a2 * B;         Cc = 1.0 / 39916800.0 - a2 * Cc;          // 9! 10! 11!
    A = 1.0 / 5040.0 - a2 * A;           B = 1.0 / 40320.0 - a2 * B;           Cc = 1.0 / 362880.0 - a2 * Cc;            // 7! 8! 9!
    A = 1.0 / 120.0 - a2 * A;            B = 1.0 / 720.0 - a2 * B;             Cc = 1.0 / 5040.0 - a2 * Cc;              // 5! 6! 7!
    A = 1.0 / 6.0 - a2 * A;              B = 1.0 / 24.0 - a2 * B;              Cc = 1.0 / 120.0 - a2 * Cc;               // 3! 4! 5!
    A = 1.0 - a2 * A;                    B = 0.5 - a2 * B;                     Cc = 1.0 / 6.0 - a2 * Cc;
  } else {
    const double ang = sqrt(a2);
    double sn, cs;
    sincos(ang, &sn, &cs);
    A = sn / ang; B = (1.0 - cs) / a2; Cc = (ang - sn) / (a2 * ang);
  }
  // row i of Exp: R = I + A W + B W^2, t = (I + B W + C W^2) rho, with W^2 = theta theta^T - |theta|^2 I
  const double ti = i == 0 ? t0 : (i == 1 ? t1 : t2);
  // W[i][0..2]
  const double w0 = i == 0 ? 0.0 : (i == 1 ? t2 : -t1), w1 = i == 0 ? -t2 : (i == 1 ? 0.0 : t0), w2 = i == 0 ? t1 : (i == 1 ? -t0 : 0.0);
  const double q0 = ti * t0 - (i == 0 ? a2 : 0.0), q1 = ti * t1 - (i == 1 ? a2 : 0.0), q2 = ti * t2 - (i == 2 ? a2 : 0.0);
  double e0 = (i == 0 ? 1.0 : 0.0) + A * w0 + B * q0, e1 = (i == 1 ? 1.0 : 0.0) + A * w1 + B * q1,
         e2 = (i == 2 ? 1.0 : 0.0) + A * w2 + B * q2;
  double e3 = ((i == 0 ? 1.0 : 0.0) + B * w0 + Cc * q0) * r0 + ((i == 1 ? 1.0 : 0.0) + B * w1 + Cc * q1) * r1 +
              ((i == 2 ? 1.0 : 0.0) + B * w2 + Cc * q2) * r2;
  if (i == 3) { e0 = 0.0; e1 = 0.0; e2 = 0.0; e3 = 1.0; }
  const float tf = (float)(e0 * Tk0 + e1 * Tk1 + e2 * Tk2 + e3 * Tk3);          // element (i, j) of Exp(tau) T
  if (lane < 16) T[4 * i + j] = tf;
  if (lane < 3) rho[lane] = 0.f;
  else if (lane < 6) theta[lane - 3] = 0.f;
  if (lane == 0 && converged) converged[0] = (r0 * r0 + r1 * r1 + r2 * r2 + a2) < (double)thr * thr ? 1 : 0;
  if (view_out) {
    if (lane < 16) view_out[4 * j + i] = tf;      // world_view_transform = T^T (camera_utils.py:94-96)
    if (proj_out && proj_raw) {                   // full_proj_transform = view @ projection (camera_utils.py:98-104)
      __syncwarp();
      if (lane < 16) {
        float sacc = 0.f;
#pragma unroll
        for (int k = 0; k < 4; k++) sacc += T[4 * k + i] * proj_raw[4 * k + j];
        proj_out[4 * i + j] = sacc;
      }
    }
  }
}

}  // namespace mgs
