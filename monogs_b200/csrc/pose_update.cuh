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

}  // namespace mgs
