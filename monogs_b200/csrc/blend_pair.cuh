// Two pixels per thread with packed fp32 (sm_100 FADD2 / FMUL2 / FFMA2, CUDA's __f*2_rn intrinsics).
//
// The blend kernels are bound by instruction issue, not by memory: one 48-B record feeds a whole patch of pixels and
// every pixel runs the same ~60 scalar instructions on it.  Blackwell's packed fp32 pipe executes one IEEE operation on
// each half of a 64-bit register pair per issue slot, so a thread that owns TWO pixels (same column, rows y and y+4 of
// an 8x8 warp patch) evaluates both for little more than the issue cost of one, and the warp-uniform part of a hit
// (list walk, record loads, votes) is paid once per 64 pixels instead of once per 32.
//
// Every packed operation below is the same single-rounded IEEE operation the scalar code performs (a packed FMA of
// component k equals fmaf on component k), and the operation sequence is the one nvcc emits for the reference's
// expression  power = -0.5f*(con.x*d.x*d.x + con.z*d.y*d.y) - con.y*d.x*d.y  (forward.cu:341-345), namely
//     a = (con.z*dy)*dy;  s = fma(dx, con.x*dx, a);  power = fma(s, -0.5, -((con.y*dx)*dy))
// so the per-pixel bits — and with them n_contrib, n_touched and the early-termination decisions — are those of the
// scalar kernels and of the reference.
#pragma once
#include "common.cuh"

namespace mgs {

__device__ __forceinline__ float2 bcast2(const float s) { return make_float2(s, s); }
__device__ __forceinline__ float2 neg2(const float2 a) { return make_float2(-a.x, -a.y); }

template <bool FAST> __device__ __forceinline__ float2 blend_exp2(const float2 power) {
  if (FAST) {
    const float2 t = __fmul2_rn(power, bcast2(1.4426950408889634f));
    float2 r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r.x) : "f"(t.x));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r.y) : "f"(t.y));
    return r;
  }
  return make_float2(expf(power.x), expf(power.y));
}

// g0 = (x, y, depth, id), g1 = (conic xx, xy, yy, opacity) of one record; npx / npy = minus the pixel coordinates of
// the thread's two pixels.  Returns d = mean - pixel, the exponent, G = exp(power) and alpha = min(0.99, opacity * G).
template <bool FAST>
__device__ __forceinline__ void pair_alpha(const float4 g0, const float4 g1, const float2 npx, const float2 npy,
                                           float2& dx, float2& dy, float2& power, float2& G, float2& alpha) {
  dx = __fadd2_rn(bcast2(g0.x), npx);
  dy = __fadd2_rn(bcast2(g0.y), npy);
  const float2 a = __fmul2_rn(dy, __fmul2_rn(bcast2(g1.z), dy));
  const float2 b = __fmul2_rn(bcast2(g1.x), dx);
  const float2 c = __fmul2_rn(dy, __fmul2_rn(bcast2(g1.y), dx));
  const float2 s = __ffma2_rn(dx, b, a);
  power = __ffma2_rn(s, bcast2(-0.5f), neg2(c));
  G = blend_exp2<FAST>(power);
  alpha = __fmul2_rn(bcast2(g1.w), G);
  alpha.x = fminf(0.99f, alpha.x);
  alpha.y = fminf(0.99f, alpha.y);
}

// 1 - alpha with the rounding of the scalar subtraction (the product alpha * -1 is exact)
__device__ __forceinline__ float2 one_minus2(const float2 alpha) { return __ffma2_rn(alpha, bcast2(-1.f), bcast2(1.f)); }

}  // namespace mgs
