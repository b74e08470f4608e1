// Block-wide bitonic sort of 64-bit words held in registers (the per-tile depth sort of the binning, raster_forward.cu).
#pragma once
#include <cuda_runtime.h>

namespace mgs {

// Block-wide bitonic sort of THREADS * IPT 64-bit words held IPT per thread in blocked order (element tid * IPT + i), ascending.
// A compare-exchange at distance j stays inside the thread (j < IPT: registers, compile-time indices), inside the warp
// (lane ^ j / IPT: two shuffles per word) or crosses warps (j >= 32 IPT: one round through shared memory, [i][tid] layout,
// two barriers).  For 512 words on 256 threads that is 45 stages of which 6 touch shared memory — about half the warp
// instructions and a third of the shared-memory traffic of the merge sort it replaces for power-of-two IPT
// (cub::BlockMergeSort: 8 merge rounds through shared memory with a binary search each, 1.3 M bank conflicts per launch
// at C2, profiles/r02_ncu_full_binning_kernels.txt).  Every real word is unique; the ~0 padding words are equal to each
// other, which a compare-exchange network does not mind.
#ifndef MGS_SORT_BITONIC
#define MGS_SORT_BITONIC 1
#endif
template <int IPT, int THREADS>
__device__ __forceinline__ void block_bitonic_sort(unsigned long long (&k)[IPT], unsigned long long* xch /* NOT __restrict__:
    other threads write what this one reads; with the qualifier the compiler kept one load of xch[tid ^ lm] across the
    barriers of later rounds (IPT = 1 sorted nothing) */) {
  static_assert((IPT & (IPT - 1)) == 0 && (THREADS & (THREADS - 1)) == 0, "power-of-two network");
  constexpr int N = THREADS * IPT;
  const int tid = threadIdx.x;
#pragma unroll
  for (int size = 2; size <= N; size <<= 1) {
#pragma unroll
    for (int j = size >> 1; j > 0; j >>= 1) {
      if (j < IPT) {
#pragma unroll
        for (int i = 0; i < IPT; i++) {
          if ((i & j) == 0) {
            const bool asc = ((tid * IPT + i) & size) == 0;
            const int p = (i + j) & (IPT - 1);   // = i + j; the mask keeps the constant index in bounds where this branch is dead (IPT = 1)
            const unsigned long long a = k[i], b = k[p];
            const bool sw = (a > b) == asc;
            k[i] = sw ? b : a; k[p] = sw ? a : b;
          }
        }
      } else {
        const int lm = j / IPT;                      // partner thread = tid ^ lm, same i
        const bool lower = (tid & lm) == 0;
        if (lm < 32) {
#pragma unroll
          for (int i = 0; i < IPT; i++) {
            const bool asc = ((tid * IPT + i) & size) == 0;
            const unsigned long long o = __shfl_xor_sync(0xffffffffu, k[i], lm);
            k[i] = ((k[i] < o) == (lower == asc)) ? k[i] : o;     // the lower index keeps the minimum of an ascending pair
          }
        } else {
          __syncthreads();                           // previous round's reads are done
#pragma unroll
          for (int i = 0; i < IPT; i++) xch[i * THREADS + tid] = k[i];
          __syncthreads();
#pragma unroll
          for (int i = 0; i < IPT; i++) {
            const bool asc = ((tid * IPT + i) & size) == 0;
            const unsigned long long o = xch[i * THREADS + (tid ^ lm)];
            k[i] = ((k[i] < o) == (lower == asc)) ? k[i] : o;
          }
        }
      }
    }
  }
}


}  // namespace mgs
