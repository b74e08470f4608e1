// Element update of torch.optim.Adam (eps outside the bias-corrected sqrt), shared by the multi-tensor kernel
// (loss_adam.cu) and the peer-memory reduce + Adam kernel (peer_exchange.cu) so both produce the same bits.
#pragma once

namespace mgs {

// coefficients are passed by value: writing into the __grid_constant__-like parameter struct would make the compiler
// spill the whole 700-byte AdamPack to local memory in every thread (measured: 9x DRAM write amplification)
struct AdamCoef { float beta1, beta2, eps, bc2_sqrt; };
__device__ __forceinline__ void adam_elem(float& p, const float g, float& m, float& v, const float step_size,
                                          const AdamCoef a) {
  // explicit roundings and contractions: the two kernels that inline this must produce the same bits whatever the
  // compiler would fuse in their different surroundings
  m = __fmaf_rn(__fsub_rn(g, m), 1.f - a.beta1, m);                                  // exp_avg.lerp_(grad, 1 - beta1)
  v = __fmaf_rn(__fmul_rn(1.f - a.beta2, g), g, __fmul_rn(v, a.beta2));              // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), a.bc2_sqrt), a.eps);
  p = __fsub_rn(p, __fmul_rn(step_size, __fdiv_rn(m, denom)));                       // param.addcdiv_(exp_avg, denom, value=-step_size)
}

}  // namespace mgs
