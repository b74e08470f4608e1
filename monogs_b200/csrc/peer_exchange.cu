// Peer-memory exchange of the data-parallel mapping step: gradient reduce-scatter + Adam +
// parameter all-gather in ONE kernel over NVLink / NVSwitch peer mappings.
//
// The reference has no multi-GPU path; the step this replaces in this repo's own design is
//     dist.all_reduce(grad_arena) ; adam_kernel on every rank          (monogs_b200/mapping.py, MapEngine.step)
// which moves 2 x arena bytes per rank through NCCL and then streams 28 B per parameter element through HBM on
// EVERY rank.  Here every rank owns 1/world of the flat arena: it reads that slice of every rank's gradient arena
// through the peer mappings (summed in rank order, so the result does not depend on who computes it), applies
// Adam to the slice (the moments of a slice live only on its owner), and stores the new parameters into every
// rank's parameter arena.  Per rank: (world-1)/world x arena bytes in and the same out over NVLink, 1/world of the
// Adam traffic, two launches, no host involvement.  The trailing loss element of the gradient arena is summed by
// its owner and written to every rank's loss_out.
//
// Ordering: flags[phase][source rank] in every rank's block, written with st.release.sys by the source and polled
// with ld.acquire.sys by the owner.  Phase 0 ("my gradients are final") is signalled by CTA 0 of the kernel, which
// is stream-ordered after the rank's backward; every CTA waits for all ranks' phase-0 flags before it reads.
// Phase 1 ("my slice is stored everywhere and I no longer read your gradients") is signalled by the last CTA to
// finish; peer_wait_kernel, next on the stream, holds the rank's following work (mgs_map_begin zeroes the gradient
// arena and reads the parameters) until every rank's phase-1 flag has arrived.  Flags carry the caller's epoch
// (strictly increasing), so they are never reset.  A wait that exceeds the caller's timeout (60 s by default) sets the
// block's error word instead of hanging the GPU for good.
#include "launchers.h"
#include "adam_elem.cuh"
#include "tma.cuh"
#include <cstdlib>
#include <cstring>
#include "../../include/monogs_b200.h"

namespace mgs {

constexpr int FLAG_PHASE1 = MGS_PEER_MAX, FLAG_DONE = 2 * MGS_PEER_MAX, FLAG_ERROR = 2 * MGS_PEER_MAX + 1;

struct PeerPack {
  float* param[MGS_PEER_MAX]; const float* grad[MGS_PEER_MAX]; float* loss_out[MGS_PEER_MAX]; uint32_t* flags[MGS_PEER_MAX];
  float *m, *v;
  const float* grad_mc; float* param_mc; float* loss_mc;   // NVSwitch multicast mappings of the same arenas (null: unicast path)
  long long q_begin, q_end;                       // this rank's slice of the arena, in float4 units
  long long seg_end4[MGS_PEER_MAX_SEGMENTS];      // segment ends, float4 units
  float seg_lr[MGS_PEER_MAX_SEGMENTS];            // < 0: sum only (the trailing side segment: loss, window poses)
  long long sum_begin4;                           // first float4 of the trailing sum-only segment(s); its sums land at
                                                  // loss_out + (q - sum_begin4) * 4 on every rank
  float seg_step[MGS_PEER_MAX_SEGMENTS];          // lr / bias correction 1, the step size of the segment
  int n_seg, world, rank;
  uint32_t epoch;
  unsigned long long timeout_ns;
  float beta1, beta2, eps, bc1, bc2_sqrt;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// spin until *flag has reached `epoch` (wrap-safe); gives up after timeout_ns and records it in the local error word
__device__ __forceinline__ void wait_flag(const uint32_t* flag, uint32_t epoch, uint32_t* error_word,
                                          unsigned long long timeout_ns) {
  const unsigned long long t0 = global_ns();
  while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
    __nanosleep(64);
    if (global_ns() - t0 > timeout_ns) { atomicExch(error_word, epoch); break; }
  }
}

// shared -> global bulk store (TMA, async proxy), tracked by the issuing thread's bulk-group
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }

// Persistent CTAs walk the rank's slice in chunks of PEER_CHUNK4 float4.  Per chunk one elected thread pulls that chunk
// of EVERY rank's gradient arena into a shared-memory stage with 1-D bulk copies (TMA reads over NVLink need no
// registers and stay in flight PEER_STAGES-1 chunks ahead: the ~2 us peer latency is covered by tens of KB per CTA),
// the CTA sums the stage in rank order, applies Adam to its own parameters / moments (plain coalesced float4 traffic
// on local HBM), stages the new parameters in shared memory and the elected thread pushes them to every rank's
// parameter arena with bulk stores.
template <int PEER_CHUNK4, int PEER_STAGES>   // float4 per chunk and rank; gradient stages in flight per CTA
__global__ void __launch_bounds__(256) peer_reduce_adam_kernel(const __grid_constant__ PeerPack a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full_bar[PEER_STAGES];
  const int tid = threadIdx.x;
  const uint32_t stage_bytes = (uint32_t)a.world * PEER_CHUNK4 * 16;
  float4* out_buf = reinterpret_cast<float4*>(smem_raw + (size_t)PEER_STAGES * stage_bytes);   // 3 x PEER_CHUNK4 float4
  uint32_t* my_flags = a.flags[a.rank];
  if (tid == 0) {
    for (int st = 0; st < PEER_STAGES; st++) mbar_init(&full_bar[st], 1);
    mbar_init_fence();
  }
  if (blockIdx.x == 0 && tid < a.world)     // phase 0: this rank's gradients are final (stream order); the release
    st_release_sys(a.flags[tid] + a.rank, a.epoch);   // store carries the system-scope fence

  __syncthreads();                          // the mbarriers are initialised for everybody
  // only the elected thread has to see the other ranks' phase-0 flags: it issues every load of a peer's gradients; the
  // rest of the CTA goes on to its local loads and meets it at the stage's mbarrier
  if (tid == 0)
    for (int r = 0; r < a.world; r++) wait_flag(my_flags + r, a.epoch, my_flags + FLAG_ERROR, a.timeout_ns);

  const long long n4 = a.q_end - a.q_begin;
  const long long n_chunks = (n4 + PEER_CHUNK4 - 1) / PEER_CHUNK4;
  // chunk k of this CTA is global chunk blockIdx.x + k * gridDim.x
  const long long my_chunks = n_chunks > (long long)blockIdx.x ? (n_chunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  auto chunk_q0 = [&](long long k) { return a.q_begin + ((long long)blockIdx.x + k * gridDim.x) * PEER_CHUNK4; };
  auto chunk_len = [&](long long q0) { const long long r = a.q_end - q0; return (int)(r < PEER_CHUNK4 ? r : PEER_CHUNK4); };
  auto issue = [&](long long k) {   // elected thread: stage k % PEER_STAGES <- chunk k of every rank's gradient arena
    const int st = (int)(k % PEER_STAGES);
    const long long q0 = chunk_q0(k);
    const uint32_t bytes = (uint32_t)chunk_len(q0) * 16;
    mbar_expect_tx(&full_bar[st], bytes * a.world);
    for (int r = 0; r < a.world; r++)
      bulk_g2s(smem_raw + (size_t)st * stage_bytes + (size_t)r * PEER_CHUNK4 * 16, a.grad[r] + q0 * 4, bytes, &full_bar[st]);
  };
  if (tid == 0)
    for (long long k = 0; k < PEER_STAGES - 1 && k < my_chunks; k++) issue(k);

  const AdamCoef co{a.beta1, a.beta2, a.eps, a.bc2_sqrt};
  float* __restrict__ p_own = a.param[a.rank];
  for (long long k = 0; k < my_chunks; k++) {
    const int st = (int)(k % PEER_STAGES);
    if (tid == 0 && k + PEER_STAGES - 1 < my_chunks) issue(k + PEER_STAGES - 1);   // its stage was drained in iteration k-1
    const long long q0 = chunk_q0(k);
    const int len = chunk_len(q0);
    float4 P4[PEER_CHUNK4 / 256], M4[PEER_CHUNK4 / 256], V4[PEER_CHUNK4 / 256];
#pragma unroll
    for (int it = 0; it < PEER_CHUNK4 / 256; it++) {   // local HBM traffic goes out before the wait on the stage
      const int j = it * 256 + tid;
      if (j < len) {
        const long long i = (q0 + j) * 4;
        P4[it] = *reinterpret_cast<const float4*>(p_own + i);
        M4[it] = *reinterpret_cast<const float4*>(a.m + i); V4[it] = *reinterpret_cast<const float4*>(a.v + i);
      }
    }
    mbar_wait(&full_bar[st], (uint32_t)((k / PEER_STAGES) & 1));
    const float4* stage = reinterpret_cast<const float4*>(smem_raw + (size_t)st * stage_bytes);
    float4* out = out_buf + (size_t)(k % 3) * PEER_CHUNK4;
#pragma unroll
    for (int it = 0; it < PEER_CHUNK4 / 256; it++) {
      const int j = it * 256 + tid;
      if (j >= len) break;
      const long long q = q0 + j, i = q * 4;
      float4 S = stage[j];
      for (int r = 1; r < a.world; r++) {    // fixed rank order: the sum does not depend on the owner
        const float4 G = stage[(size_t)r * PEER_CHUNK4 + j];
        S.x += G.x; S.y += G.y; S.z += G.z; S.w += G.w;
      }
      int sg = 0;
      while (sg + 1 < a.n_seg && q >= a.seg_end4[sg]) sg++;
      const float lr = a.seg_lr[sg];
      if (lr < 0.f) {                        // side segment: every rank gets the sum; its parameter slots are unused
        for (int r = 0; r < a.world; r++) *reinterpret_cast<float4*>(a.loss_out[r] + (q - a.sum_begin4) * 4) = S;
      } else {
        const float step_size = a.seg_step[sg];
        adam_elem(P4[it].x, S.x, M4[it].x, V4[it].x, step_size, co); adam_elem(P4[it].y, S.y, M4[it].y, V4[it].y, step_size, co);
        adam_elem(P4[it].z, S.z, M4[it].z, V4[it].z, step_size, co); adam_elem(P4[it].w, S.w, M4[it].w, V4[it].w, step_size, co);
        *reinterpret_cast<float4*>(a.m + i) = M4[it]; *reinterpret_cast<float4*>(a.v + i) = V4[it];
      }
      out[j] = P4[it];
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the staged parameters become visible to the bulk stores
    __syncthreads();   // stage st is drained, out[k % 3] is complete; also orders thread 0's wait of iteration k-1
    if (tid == 0) {
      for (int r = 0; r < a.world; r++) bulk_s2g(a.param[r] + q0 * 4, out, (uint32_t)len * 16);   // parameter all-gather
      bulk_commit();
      bulk_wait_read<1>();   // the stores of iteration k-1 have read their buffer: out[(k+2) % 3] is free for k+2
    }
  }

  // phase 1: the last CTA of this rank tells every rank that the slice is stored and the gradients are released.
  // One lane per destination: every lane fences once and stores its flag, so the eight signals leave together instead
  // of eight release stores each waiting for the acknowledgement of the one before.
  __shared__ int s_last;
  if (tid == 0) {
    bulk_wait<0>();                                  // this CTA's parameter stores are complete
    asm volatile("fence.proxy.async;" ::: "memory");
    __threadfence_system();
    const unsigned int done = atomicAdd(my_flags + FLAG_DONE, 1u);
    s_last = (done == gridDim.x - 1);
    if (s_last) my_flags[FLAG_DONE] = 0;
  }
  __syncthreads();
  if (s_last && tid < a.world) {
    __threadfence_system();
    st_relaxed_sys(a.flags[tid] + FLAG_PHASE1 + a.rank, a.epoch);
  }
}

// ---- NVSwitch multicast variant (NVLS) -------------------------------------------------------------------------------
// The gradient arenas are read through the multicast mapping with multimem.ld_reduce: the switch fetches the addressed
// 16 bytes from EVERY rank and returns their fp32 sum, so the owner receives its slice once instead of once per peer;
// multimem.st through the multicast mapping of the parameter arenas hands one copy to the switch, which writes all
// ranks.  Link volume per rank and direction drops from 2(N-1)/N x arena to ~1 x arena (serve (N-1)/N of the own
// gradients + push 1/N of the parameters one way, receive 1/N reduced + (N-1)/N of the parameters the other way).
// The sum's association order is the switch's; every rank still ends with identical parameters because only the owner
// of a slice computes it.  Register path: PEER_MC_UNROLL independent reductions per thread in flight.
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc) : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st(float* mc, const float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

template <int PEER_MC_UNROLL>
__global__ void __launch_bounds__(256) peer_reduce_adam_mc_kernel(const __grid_constant__ PeerPack a) {
  const int tid = threadIdx.x;
  uint32_t* my_flags = a.flags[a.rank];
  if (blockIdx.x == 0 && tid < a.world) st_release_sys(a.flags[tid] + a.rank, a.epoch);   // phase 0, as in the unicast kernel
  if (tid < a.world) wait_flag(my_flags + tid, a.epoch, my_flags + FLAG_ERROR, a.timeout_ns);
  __syncthreads();
  const AdamCoef co{a.beta1, a.beta2, a.eps, a.bc2_sqrt};
  float* p_own = a.param[a.rank];
  constexpr int U = PEER_MC_UNROLL;
  for (long long base = a.q_begin + (long long)blockIdx.x * 256 * U; base < a.q_end; base += (long long)gridDim.x * 256 * U) {
    float4 S[U], P4[U], M4[U], V4[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const long long q = base + u * 256 + tid;
      if (q < a.q_end) {
        S[u] = multimem_ld_reduce_add(a.grad_mc + q * 4);
        P4[u] = *reinterpret_cast<const float4*>(p_own + q * 4);
        M4[u] = *reinterpret_cast<const float4*>(a.m + q * 4); V4[u] = *reinterpret_cast<const float4*>(a.v + q * 4);
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const long long q = base + u * 256 + tid;
      if (q >= a.q_end) break;
      int sg = 0;
      while (sg + 1 < a.n_seg && q >= a.seg_end4[sg]) sg++;
      const float lr = a.seg_lr[sg];
      if (lr < 0.f) { multimem_st(a.loss_mc + (q - a.sum_begin4) * 4, S[u]); continue; }   // side segment: every rank gets the sum
      const float step_size = lr / a.bc1;
      adam_elem(P4[u].x, S[u].x, M4[u].x, V4[u].x, step_size, co); adam_elem(P4[u].y, S[u].y, M4[u].y, V4[u].y, step_size, co);
      adam_elem(P4[u].z, S[u].z, M4[u].z, V4[u].z, step_size, co); adam_elem(P4[u].w, S[u].w, M4[u].w, V4[u].w, step_size, co);
      *reinterpret_cast<float4*>(a.m + q * 4) = M4[u]; *reinterpret_cast<float4*>(a.v + q * 4) = V4[u];
      multimem_st(a.param_mc + q * 4, P4[u]);   // parameter all-gather: one store, the switch writes every rank
    }
  }
  // phase 1
  __shared__ int s_last;
  __threadfence_system();
  __syncthreads();
  if (tid == 0) {
    const unsigned int done = atomicAdd(my_flags + FLAG_DONE, 1u);
    s_last = (done == gridDim.x - 1);
    if (s_last) my_flags[FLAG_DONE] = 0;
  }
  __syncthreads();
  if (s_last && tid < a.world) {
    __threadfence_system();
    st_relaxed_sys(a.flags[tid] + FLAG_PHASE1 + a.rank, a.epoch);
  }
}

__global__ void peer_wait_kernel(uint32_t* my_flags, int world, uint32_t epoch, unsigned long long timeout_ns) {
  if ((int)threadIdx.x < world) wait_flag(my_flags + FLAG_PHASE1 + threadIdx.x, epoch, my_flags + FLAG_ERROR, timeout_ns);
  __threadfence_system();
}

}  // namespace mgs

using namespace mgs;

// tuning knob (MGS_PEER_VARIANT, read once): chunk size x stage count of the kernel; 0 = the measured best
static int peer_variant() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("MGS_PEER_VARIANT"); v = e ? atoi(e) : 0; }
  return v;
}

static int peer_mc_unroll() {   // tuning knob (MGS_PEER_MC_UNROLL, read once): reductions in flight per thread
  static int v = -1;
  if (v < 0) { const char* e = getenv("MGS_PEER_MC_UNROLL"); v = e ? atoi(e) : 4; }
  return v;
}

template <int U>
static int launch_peer_mc(const PeerPack& a, cudaStream_t s) {
  static int resident_by_dev[64] = {0};   // per device: the attribute and the occupancy belong to the current device
  int dev = 0;
  MGS_CUDA_TRY(cudaGetDevice(&dev));
  int& resident_mc = resident_by_dev[dev & 63];
  if (resident_mc == 0) {
    int sms = 0, per_sm = 0;
    MGS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    MGS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, peer_reduce_adam_mc_kernel<U>, 256, 0));
    resident_mc = sms * (per_sm > 0 ? per_sm : 1);
  }
  long long ctas = (a.q_end - a.q_begin + 256 * U - 1) / (256 * U);
  if (ctas > resident_mc) ctas = resident_mc;
  if (ctas < 1) ctas = 1;
  MGS_PROF("peer_reduce_adam_mc", s, peer_reduce_adam_mc_kernel<U><<<(unsigned)ctas, 256, 0, s>>>(a));
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

template <int PEER_CHUNK4, int PEER_STAGES>
static int launch_peer(const PeerPack& a, cudaStream_t s) {
  const size_t smem = (size_t)PEER_STAGES * a.world * PEER_CHUNK4 * 16 + 3 * PEER_CHUNK4 * 16;
  static int resident_by_dev[64][MGS_PEER_MAX + 1] = {{0}};   // persistent grid: resident CTAs per device and world size
  int dev = 0;
  MGS_CUDA_TRY(cudaGetDevice(&dev));
  int* resident = resident_by_dev[dev & 63];
  if (resident[a.world] == 0) {
    int sms = 0, per_sm = 0;
    MGS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    MGS_CUDA_TRY(cudaFuncSetAttribute(peer_reduce_adam_kernel<PEER_CHUNK4, PEER_STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    MGS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, peer_reduce_adam_kernel<PEER_CHUNK4, PEER_STAGES>, 256, smem));
    if (per_sm < 1) { set_error("peer_reduce_adam: %zu bytes of shared memory do not fit", smem); return MGS_ENOSPC; }
    resident[a.world] = sms * per_sm;
  }
  long long ctas = (a.q_end - a.q_begin + PEER_CHUNK4 - 1) / PEER_CHUNK4;
  if (ctas > resident[a.world]) ctas = resident[a.world];
  if (ctas < 1) ctas = 1;   // an empty slice still takes part in both phases
  MGS_PROF("peer_reduce_adam", s, peer_reduce_adam_kernel<PEER_CHUNK4, PEER_STAGES><<<(unsigned)ctas, 256, smem, s>>>(a));
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

extern "C" {

size_t mgs_peer_flags_bytes(void) { return 256; }   // u32[64]: phase-0 flags, phase-1 flags, done counter, error word

int mgs_peer_alloc(size_t bytes, void** out_ptr) {
  if (!out_ptr || bytes == 0) { set_error("peer_alloc: bad arguments"); return MGS_EINVAL; }
  MGS_CUDA_TRY(cudaMalloc(out_ptr, bytes));
  MGS_CUDA_TRY(cudaMemset(*out_ptr, 0, bytes));
  MGS_CUDA_TRY(cudaDeviceSynchronize());
  return MGS_OK;
}

int mgs_peer_free(void* ptr) {
  if (ptr) MGS_CUDA_TRY(cudaFree(ptr));
  return MGS_OK;
}

int mgs_peer_export(const void* ptr, uint8_t* handle) {
  static_assert(sizeof(cudaIpcMemHandle_t) == MGS_PEER_HANDLE_BYTES, "handle size");
  if (!ptr || !handle) { set_error("peer_export: null argument"); return MGS_EINVAL; }
  cudaIpcMemHandle_t h;
  MGS_CUDA_TRY(cudaIpcGetMemHandle(&h, const_cast<void*>(ptr)));
  memcpy(handle, &h, sizeof(h));
  return MGS_OK;
}

int mgs_peer_open(const uint8_t* handle, void** out_ptr) {
  if (!handle || !out_ptr) { set_error("peer_open: null argument"); return MGS_EINVAL; }
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  MGS_CUDA_TRY(cudaIpcOpenMemHandle(out_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return MGS_OK;
}

int mgs_peer_close(void* ptr) {
  if (ptr) MGS_CUDA_TRY(cudaIpcCloseMemHandle(ptr));
  return MGS_OK;
}

int mgs_peer_reduce_adam(const mgs_peer_exchange* x, int32_t step, uint32_t epoch, float beta1, float beta2, float eps,
                         void* stream) {
  if (!x) { set_error("peer_reduce_adam: null exchange"); return MGS_EINVAL; }
  if (x->world < 1 || x->world > MGS_PEER_MAX || x->rank < 0 || x->rank >= x->world) {
    set_error("peer_reduce_adam: world %d / rank %d out of range (max %d ranks)", x->world, x->rank, MGS_PEER_MAX); return MGS_EINVAL;
  }
  if (step < 1 || epoch == 0) { set_error("peer_reduce_adam: step and epoch start at 1"); return MGS_EINVAL; }
  if (x->numel <= 0 || (x->numel & 3) || x->n_segments < 1 || x->n_segments > MGS_PEER_MAX_SEGMENTS) {
    set_error("peer_reduce_adam: numel must be a positive multiple of 4 and 1..%d segments given", MGS_PEER_MAX_SEGMENTS); return MGS_EINVAL;
  }
  if (!x->exp_avg || !x->exp_avg_sq) { set_error("peer_reduce_adam: null moments"); return MGS_EINVAL; }
  PeerPack a;
  for (int r = 0; r < x->world; r++) {
    if (!x->param_arena[r] || !x->grad_arena[r] || !x->loss_out[r] || !x->flags[r]) { set_error("peer_reduce_adam: rank %d has null pointers", r); return MGS_EINVAL; }
    if (((uintptr_t)x->param_arena[r] | (uintptr_t)x->grad_arena[r] | (uintptr_t)x->loss_out[r]) & 15) { set_error("peer_reduce_adam: arenas must be 16-byte aligned"); return MGS_EINVAL; }
    a.param[r] = x->param_arena[r]; a.grad[r] = x->grad_arena[r]; a.loss_out[r] = x->loss_out[r]; a.flags[r] = x->flags[r];
  }
  for (int r = x->world; r < MGS_PEER_MAX; r++) { a.param[r] = nullptr; a.grad[r] = nullptr; a.loss_out[r] = nullptr; a.flags[r] = nullptr; }
  if (((uintptr_t)x->exp_avg | (uintptr_t)x->exp_avg_sq) & 15) { set_error("peer_reduce_adam: moments must be 16-byte aligned"); return MGS_EINVAL; }
  long long prev = 0;
  for (int sgi = 0; sgi < x->n_segments; sgi++) {
    const long long e = x->segment_end[sgi];
    if (e <= prev || (e & 3) || e > x->numel) { set_error("peer_reduce_adam: segment %d ends at %lld (ends must increase in multiples of 4)", sgi, e); return MGS_EINVAL; }
    a.seg_end4[sgi] = e / 4; a.seg_lr[sgi] = x->segment_lr[sgi]; prev = e;
  }
  if (prev != x->numel) { set_error("peer_reduce_adam: the segments must cover the arena"); return MGS_EINVAL; }
  a.sum_begin4 = x->numel / 4;
  for (int sgi = x->n_segments - 1; sgi >= 0 && x->segment_lr[sgi] < 0.f; sgi--) a.sum_begin4 = sgi > 0 ? x->segment_end[sgi - 1] / 4 : 0;
  for (int sgi = 0; sgi < x->n_segments; sgi++)
    if (x->segment_lr[sgi] < 0.f && x->segment_end[sgi] / 4 <= a.sum_begin4) { set_error("peer_reduce_adam: sum-only segments must be the trailing ones"); return MGS_EINVAL; }
  a.n_seg = x->n_segments; a.world = x->world; a.rank = x->rank; a.epoch = epoch;
  a.timeout_ns = (unsigned long long)(x->timeout_ms ? x->timeout_ms : 60000u) * 1000000ull;
  a.m = x->exp_avg; a.v = x->exp_avg_sq;
  a.grad_mc = x->grad_multicast; a.param_mc = x->param_multicast; a.loss_mc = x->loss_multicast;
  const bool mc = a.grad_mc || a.param_mc || a.loss_mc;
  if (mc && (!a.grad_mc || !a.param_mc || !a.loss_mc || (((uintptr_t)a.grad_mc | (uintptr_t)a.param_mc | (uintptr_t)a.loss_mc) & 15))) {
    set_error("peer_reduce_adam: the multicast mappings come as three 16-byte aligned pointers or not at all"); return MGS_EINVAL;
  }
  const long long n4 = x->numel / 4, per = (n4 + x->world - 1) / x->world;
  a.q_begin = per * x->rank < n4 ? per * x->rank : n4;
  a.q_end = a.q_begin + per < n4 ? a.q_begin + per : n4;
  a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
  a.bc1 = (float)(1.0 - pow((double)beta1, (double)step));
  a.bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
  for (int sgi = 0; sgi < a.n_seg; sgi++) a.seg_step[sgi] = a.seg_lr[sgi] / a.bc1;   // the same fp32 division adam_kernel does
  cudaStream_t s = (cudaStream_t)stream;
  int rc = MGS_OK;
  if (mc) {
    switch (peer_mc_unroll()) {
      case 2: rc = launch_peer_mc<2>(a, s); break;
      case 8: rc = launch_peer_mc<8>(a, s); break;
      default: rc = launch_peer_mc<4>(a, s); break;
    }
  } else
  switch (peer_variant()) {
    case 1: rc = launch_peer<1024, 3>(a, s); break;
    case 2: rc = launch_peer<512, 4>(a, s); break;
    case 3: rc = launch_peer<1024, 2>(a, s); break;
    case 4: rc = launch_peer<256, 4>(a, s); break;
    case 5: rc = launch_peer<256, 6>(a, s); break;
    default: rc = launch_peer<512, 3>(a, s); break;
  }
  if (rc) return rc;
  MGS_PROF("peer_wait", s, peer_wait_kernel<<<1, 32, 0, s>>>(a.flags[a.rank], a.world, epoch, a.timeout_ns));
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

int mgs_peer_error(const uint32_t* flags, uint32_t* out_epoch, void* stream) {
  if (!flags || !out_epoch) { set_error("peer_error: null argument"); return MGS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  MGS_CUDA_TRY(cudaMemcpyAsync(out_epoch, flags + FLAG_ERROR, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  MGS_CUDA_TRY(cudaStreamSynchronize(s));
  return MGS_OK;
}

}  // extern "C"
