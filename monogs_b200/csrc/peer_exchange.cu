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
// (strictly increasing), so they are never reset.  A wait that exceeds ~4 s sets the block's error word instead of
// hanging the GPU.
#include "launchers.h"
#include "adam_elem.cuh"
#include "../../include/monogs_b200.h"

namespace mgs {

constexpr int PEER_CHUNK4 = 512;   // float4 per CTA (256 threads x 2)
constexpr int FLAG_PHASE1 = MGS_PEER_MAX, FLAG_DONE = 2 * MGS_PEER_MAX, FLAG_ERROR = 2 * MGS_PEER_MAX + 1;

struct PeerPack {
  float* param[MGS_PEER_MAX]; const float* grad[MGS_PEER_MAX]; float* loss_out[MGS_PEER_MAX]; uint32_t* flags[MGS_PEER_MAX];
  float *m, *v;
  long long q_begin, q_end;                       // this rank's slice of the arena, in float4 units
  long long seg_end4[MGS_PEER_MAX_SEGMENTS];      // segment ends, float4 units
  float seg_lr[MGS_PEER_MAX_SEGMENTS];            // < 0: sum only (the loss element)
  int n_seg, world, rank;
  uint32_t epoch;
  float beta1, beta2, eps, bc1, bc2_sqrt;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
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
// spin until *flag has reached `epoch` (wrap-safe); gives up after ~4 s and records it in the local error word
__device__ __forceinline__ void wait_flag(const uint32_t* flag, uint32_t epoch, uint32_t* error_word) {
  const unsigned long long t0 = global_ns();
  while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
    __nanosleep(64);
    if (global_ns() - t0 > 4000000000ull) { atomicExch(error_word, epoch); break; }
  }
}

__global__ void __launch_bounds__(256) peer_reduce_adam_kernel(const __grid_constant__ PeerPack a) {
  uint32_t* my_flags = a.flags[a.rank];
  if (blockIdx.x == 0 && threadIdx.x < a.world) {   // phase 0: this rank's gradients are final (stream order)
    __threadfence_system();
    st_release_sys(a.flags[threadIdx.x] + a.rank, a.epoch);
  }
  if (threadIdx.x < a.world) wait_flag(my_flags + threadIdx.x, a.epoch, my_flags + FLAG_ERROR);
  __syncthreads();

  const AdamCoef co{a.beta1, a.beta2, a.eps, a.bc2_sqrt};
  float* __restrict__ p_own = a.param[a.rank];
#pragma unroll
  for (int it = 0; it < PEER_CHUNK4 / 256; it++) {
    const long long q = a.q_begin + (long long)blockIdx.x * PEER_CHUNK4 + it * 256 + threadIdx.x;
    if (q >= a.q_end) break;
    const long long i = q * 4;
    float4 G[MGS_PEER_MAX];
#pragma unroll
    for (int r = 0; r < MGS_PEER_MAX; r++)   // all loads in flight first; L1 bypassed (peer lines must not be reused)
      if (r < a.world) G[r] = __ldcg(reinterpret_cast<const float4*>(a.grad[r] + i));
    float4 S = G[0];
#pragma unroll
    for (int r = 1; r < MGS_PEER_MAX; r++)   // fixed rank order: the sum does not depend on the owner
      if (r < a.world) { S.x += G[r].x; S.y += G[r].y; S.z += G[r].z; S.w += G[r].w; }
    int sg = 0;
    while (sg + 1 < a.n_seg && q >= a.seg_end4[sg]) sg++;
    const float lr = a.seg_lr[sg];
    if (lr < 0.f) {                          // the loss element: every rank gets the sum
#pragma unroll
      for (int r = 0; r < MGS_PEER_MAX; r++)
        if (r < a.world) *reinterpret_cast<float4*>(a.loss_out[r]) = S;
      continue;
    }
    const float step_size = lr / a.bc1;
    float4 P4 = *reinterpret_cast<const float4*>(p_own + i);
    float4 M4 = *reinterpret_cast<const float4*>(a.m + i), V4 = *reinterpret_cast<const float4*>(a.v + i);
    adam_elem(P4.x, S.x, M4.x, V4.x, step_size, co); adam_elem(P4.y, S.y, M4.y, V4.y, step_size, co);
    adam_elem(P4.z, S.z, M4.z, V4.z, step_size, co); adam_elem(P4.w, S.w, M4.w, V4.w, step_size, co);
    *reinterpret_cast<float4*>(a.m + i) = M4; *reinterpret_cast<float4*>(a.v + i) = V4;
#pragma unroll
    for (int r = 0; r < MGS_PEER_MAX; r++)   // parameter all-gather: the owner stores its slice into every arena
      if (r < a.world) *reinterpret_cast<float4*>(a.param[r] + i) = P4;
  }

  // phase 1: the last CTA of this rank tells every rank that the slice is stored and the gradients are released
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int done = atomicAdd(my_flags + FLAG_DONE, 1u);
    if (done == gridDim.x - 1) {
      my_flags[FLAG_DONE] = 0;
      __threadfence_system();
      for (int r = 0; r < a.world; r++) st_release_sys(a.flags[r] + FLAG_PHASE1 + a.rank, a.epoch);
    }
  }
}

__global__ void peer_wait_kernel(uint32_t* my_flags, int world, uint32_t epoch) {
  if ((int)threadIdx.x < world) wait_flag(my_flags + FLAG_PHASE1 + threadIdx.x, epoch, my_flags + FLAG_ERROR);
  __threadfence_system();
}

}  // namespace mgs

using namespace mgs;

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
  a.n_seg = x->n_segments; a.world = x->world; a.rank = x->rank; a.epoch = epoch;
  a.m = x->exp_avg; a.v = x->exp_avg_sq;
  const long long n4 = x->numel / 4, per = (n4 + x->world - 1) / x->world;
  a.q_begin = per * x->rank < n4 ? per * x->rank : n4;
  a.q_end = a.q_begin + per < n4 ? a.q_begin + per : n4;
  a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
  a.bc1 = (float)(1.0 - pow((double)beta1, (double)step));
  a.bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
  cudaStream_t s = (cudaStream_t)stream;
  long long ctas = (a.q_end - a.q_begin + PEER_CHUNK4 - 1) / PEER_CHUNK4;
  if (ctas < 1) ctas = 1;   // an empty slice still takes part in both phases
  MGS_PROF("peer_reduce_adam", s, peer_reduce_adam_kernel<<<(unsigned)ctas, 256, 0, s>>>(a));
  MGS_LAUNCH_CHECK();
  MGS_PROF("peer_wait", s, peer_wait_kernel<<<1, 32, 0, s>>>(a.flags[a.rank], a.world, epoch));
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
