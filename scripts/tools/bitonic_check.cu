// Stand-alone check of block_bitonic_sort (monogs_b200/csrc/bitonic_sort.cuh) against std::sort on random unique words with ~0
// padding: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scripts/tools/bitonic_check scripts/tools/bitonic_check.cu
#include "../../monogs_b200/csrc/bitonic_sort.cuh"
#include <algorithm>
#include <cstdio>
#include <random>
#include <vector>
using namespace mgs;
template <int IPT, int THREADS>
__global__ void k_sort(const unsigned long long* in, unsigned long long* out, const int* ns) {
  __shared__ unsigned long long xch[THREADS * IPT + 1];
  const int n = ns[blockIdx.x];
  const unsigned long long* seg = in + (size_t)blockIdx.x * THREADS * IPT;
  unsigned long long k[IPT];
  for (int i = 0; i < IPT; i++) { const int e = threadIdx.x * IPT + i; k[i] = e < n ? seg[e] : ~0ull; }
  block_bitonic_sort<IPT, THREADS>(k, xch);
  for (int i = 0; i < IPT; i++) out[(size_t)blockIdx.x * THREADS * IPT + threadIdx.x * IPT + i] = k[i];
}
template <int IPT> int run(int blocks) {
  constexpr int T = 256, N = T * IPT;
  std::mt19937_64 rng(7 + IPT);
  std::vector<unsigned long long> h((size_t)blocks * N), ref;
  std::vector<int> ns(blocks);
  for (int b = 0; b < blocks; b++) {
    ns[b] = 1 + (int)(rng() % N);
    for (int e = 0; e < N; e++) h[(size_t)b * N + e] = ((rng() >> 20) << 20) | (unsigned long long)(b * N + e);   // unique
  }
  unsigned long long *din, *dout; int* dn;
  cudaMalloc(&din, h.size() * 8); cudaMalloc(&dout, h.size() * 8); cudaMalloc(&dn, blocks * 4);
  cudaMemcpy(din, h.data(), h.size() * 8, cudaMemcpyHostToDevice); cudaMemcpy(dn, ns.data(), blocks * 4, cudaMemcpyHostToDevice);
  k_sort<IPT, T><<<blocks, T>>>(din, dout, dn);
  std::vector<unsigned long long> o(h.size());
  cudaError_t e = cudaMemcpy(o.data(), dout, h.size() * 8, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) { printf("IPT %d: CUDA error %s\n", IPT, cudaGetErrorString(e)); return 1; }
  int bad = 0;
  for (int b = 0; b < blocks; b++) {
    std::vector<unsigned long long> r(h.begin() + (size_t)b * N, h.begin() + (size_t)b * N + ns[b]);
    std::sort(r.begin(), r.end());
    r.resize(N, ~0ull);
    for (int e2 = 0; e2 < N; e2++) if (r[e2] != o[(size_t)b * N + e2]) { if (bad < 3) printf("IPT %d block %d n %d: first mismatch at %d\n", IPT, b, ns[b], e2); bad++; break; }
  }
  printf("IPT %d: %d of %d blocks wrong\n", IPT, bad, blocks);
  cudaFree(din); cudaFree(dout); cudaFree(dn);
  return bad;
}
int main() { int bad = run<1>(200) + run<2>(200) + run<4>(200) + run<8>(100) + run<16>(100); printf(bad ? "FAILED\n" : "OK\n"); return bad != 0; }
