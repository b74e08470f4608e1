// Fused tracking / mapping loss (forward + backward in one sweep) and fused multi-tensor Adam.
// Loss semantics: MonoGS utils/slam_utils.py:83-113 (tracking L1 rgb / rgbd), :188-205 + :58-75 +
// utils/slam_frontend.py:596-598 (per-pixel residual -> Huber -> L2 norm), :224-253 (mapping L1).
// Adam semantics: torch.optim.Adam(eps=1e-15, betas=(0.9,0.999)) as set up in
// gaussian_splatting/scene/gaussian_model.py:285 and utils/slam_frontend.py:453.
#include "launchers.h"
#include <cstring>
#include <mutex>
#include "loss_pixel.cuh"
#include "adam_elem.cuh"

namespace mgs {

// partial sums per block: [0] loss (L1 modes) or sum of squares (Huber), [1] dL/da, [2] dL/db, [3] depth term
// pass 0: reductions only (needed first for the Huber norm); pass 1: gradients.  L1 modes do both in one launch.
template <bool WRITE_GRADS, bool REDUCE>
__global__ void __launch_bounds__(256)
loss_kernel(const LossK k, const float* __restrict__ norm_in, double* __restrict__ partials,
            float* __restrict__ dL_dimage, float* __restrict__ dL_ddepth) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  if (p < k.npix) {
    float grad_scale = k.scale;
    if (k.mode == MGS_LOSS_TRACK_HUBER) {
      if (WRITE_GRADS) { const float nrm = norm_in[0]; grad_scale = nrm > 0.f ? k.scale / nrm : 0.f; }
    } else {
      grad_scale *= (k.use_depth ? k.alpha : 1.f) / (3.f * (float)k.npix);
    }
    const float img[3] = {k.image[p], k.image[k.npix + p], k.image[2 * k.npix + p]};
    const float op = k.opacity ? k.opacity[p] : 1.f;
    const float d = (k.use_depth && k.depth) ? k.depth[p] : 0.f;
    float g_img[3], g_d;
    loss_pixel<WRITE_GRADS, REDUCE>(k, p, img, d, op, grad_scale, acc, g_img, g_d);
    if (WRITE_GRADS) {
      dL_dimage[p] = g_img[0]; dL_dimage[k.npix + p] = g_img[1]; dL_dimage[2 * k.npix + p] = g_img[2];
      if (dL_ddepth != nullptr) dL_ddepth[p] = g_d;
    }
  }
  loss_block_reduce(acc, partials + (size_t)blockIdx.x * 4);
}

// stage: 0 = Huber norm only (writes norm to out_loss), 1 = final (loss + dab)
__global__ void __launch_bounds__(256)
loss_finalize_kernel(const LossK k, const double* __restrict__ partials, const int nblocks, const int stage,
                     float* __restrict__ out_loss, float* __restrict__ out_dab) {
  __shared__ double sh[256][4];
  double acc[4] = {0, 0, 0, 0};
  for (int i = threadIdx.x; i < nblocks; i += 256)
#pragma unroll
    for (int q = 0; q < 4; q++) acc[q] += partials[(size_t)i * 4 + q];
#pragma unroll
  for (int q = 0; q < 4; q++) sh[threadIdx.x][q] = acc[q];
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o)
#pragma unroll
      for (int q = 0; q < 4; q++) sh[threadIdx.x][q] += sh[threadIdx.x + o][q];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    if (k.mode == MGS_LOSS_TRACK_HUBER) {
      if (stage == 0) out_loss[0] = (float)sqrt(sh[0][0]);   // un-scaled norm, consumed by the gradient pass
      else { out_loss[0] = k.scale * out_loss[0]; if (out_dab) { out_dab[0] = (float)sh[0][1]; out_dab[1] = (float)sh[0][2]; } }
    } else {
      const double rgb = sh[0][0] / (3.0 * k.npix), dep = sh[0][3] / (double)k.npix;
      const double l = k.use_depth ? k.alpha * rgb + (1.0 - k.alpha) * dep : rgb;
      out_loss[0] = (float)(k.scale * l);
      if (out_dab) { out_dab[0] = (float)sh[0][1]; out_dab[1] = (float)sh[0][2]; }
    }
  }
}

// ---------------------------------------------------------------------------------------------------
constexpr int ADAM_MAX_GROUPS = 16;
constexpr int ADAM_CHUNK = 2048;   // elements per block (256 threads x 2 float4)
struct AdamPack {
  float* p[ADAM_MAX_GROUPS]; const float* g[ADAM_MAX_GROUPS]; float* m[ADAM_MAX_GROUPS]; float* v[ADAM_MAX_GROUPS];
  long long numel[ADAM_MAX_GROUPS]; float lr[ADAM_MAX_GROUPS]; int chunk_start[ADAM_MAX_GROUPS + 1];
  const float* lr_dev[ADAM_MAX_GROUPS];   // non-null: the group's learning rate is read from device memory at run time
  int n;
  float beta1, beta2, eps, bc1, bc2_sqrt;
  const int* step_dev;   // when non-null the step number is *step_dev + 1 and bc1/bc2_sqrt are derived on device
};

__global__ void adam_step_inc_kernel(int* step_dev) { step_dev[0] += 1; }

__global__ void __launch_bounds__(256) adam_kernel(const __grid_constant__ AdamPack a) {
  float bc1 = a.bc1, bc2_sqrt = a.bc2_sqrt;
  if (a.step_dev != nullptr) {   // uniform branch: device-side step counter (CUDA-graph friendly)
    __shared__ float s_bc[2];
    if (threadIdx.x == 0) {
      // beta^t for the integer step number by repeated squaring (every CTA waits for this thread: the generic double pow()
      // is several hundred dependent instructions, twice)
      const int t = a.step_dev[0] + 1;
      double p1 = 1.0, p2 = 1.0, b1 = (double)a.beta1, b2 = (double)a.beta2;
      for (int e = t; e > 0; e >>= 1) { if (e & 1) { p1 *= b1; p2 *= b2; } b1 *= b1; b2 *= b2; }
      s_bc[0] = (float)(1.0 - p1);
      s_bc[1] = (float)sqrt(1.0 - p2);
    }
    __syncthreads();
    bc1 = s_bc[0]; bc2_sqrt = s_bc[1];
  }
  const AdamCoef co{a.beta1, a.beta2, a.eps, bc2_sqrt};
  int gi = 0;
  while (gi + 1 < a.n && (int)blockIdx.x >= a.chunk_start[gi + 1]) gi++;
  const long long base = (long long)(blockIdx.x - a.chunk_start[gi]) * ADAM_CHUNK;
  const long long n = a.numel[gi];
  float* __restrict__ p = a.p[gi]; const float* __restrict__ g = a.g[gi];
  float* __restrict__ m = a.m[gi]; float* __restrict__ v = a.v[gi];
  const float step_size = (a.lr_dev[gi] != nullptr ? __ldg(a.lr_dev[gi]) : a.lr[gi]) / bc1;
  const bool vec = ((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0);
#pragma unroll
  for (int it = 0; it < ADAM_CHUNK / 1024; it++) {
    const long long i = base + (long long)it * 1024 + threadIdx.x * 4;
    if (i >= n) break;
    if (vec && i + 4 <= n) {
      float4 P4 = *reinterpret_cast<float4*>(p + i);
      const float4 G4 = *reinterpret_cast<const float4*>(g + i);
      float4 M4 = *reinterpret_cast<float4*>(m + i), V4 = *reinterpret_cast<float4*>(v + i);
      adam_elem(P4.x, G4.x, M4.x, V4.x, step_size, co); adam_elem(P4.y, G4.y, M4.y, V4.y, step_size, co);
      adam_elem(P4.z, G4.z, M4.z, V4.z, step_size, co); adam_elem(P4.w, G4.w, M4.w, V4.w, step_size, co);
      *reinterpret_cast<float4*>(p + i) = P4; *reinterpret_cast<float4*>(m + i) = M4; *reinterpret_cast<float4*>(v + i) = V4;
    } else {
      for (long long j = i; j < min(i + 4, n); j++) {
        float pj = p[j], mj = m[j], vj = v[j];
        adam_elem(pj, g[j], mj, vj, step_size, co);
        p[j] = pj; m[j] = mj; v[j] = vj;
      }
    }
  }
}

}  // namespace mgs

using namespace mgs;

extern "C" {

size_t mgs_loss_scratch_bytes(int32_t H, int32_t W) {
  const size_t nblocks = ((size_t)H * W + 255) / 256;
  return nblocks * 4 * sizeof(double) + 256;
}

int mgs_loss_fwd_bwd(const mgs_loss_args* args, float* out_loss, float* out_dab, float* dL_dimage, float* dL_ddepth,
                     void* scratch, size_t scratch_bytes, void* stream) {
  if (!args || !out_loss || !scratch) { set_error("loss: null argument"); return MGS_EINVAL; }
  if (args->H <= 0 || args->W <= 0 || !args->image || !args->gt_image) { set_error("loss: bad image arguments"); return MGS_EINVAL; }
  if (args->mode < 0 || args->mode > 2) { set_error("loss: unknown mode %d", args->mode); return MGS_EINVAL; }
  if (args->mode != MGS_LOSS_MAP_L1 && !args->opacity) { set_error("loss: tracking modes need the opacity image"); return MGS_EINVAL; }
  if (args->use_depth && (!args->depth || !args->gt_depth)) { set_error("loss: use_depth needs depth and gt_depth"); return MGS_EINVAL; }
  if (args->use_depth && args->mode == MGS_LOSS_TRACK_HUBER) { set_error("loss: the per-pixel rgbd variant raises NotImplementedError in the reference (slam_utils.py:219)"); return MGS_EINVAL; }
  if (scratch_bytes < mgs_loss_scratch_bytes(args->H, args->W)) { set_error("loss: scratch too small"); return MGS_ENOSPC; }
  cudaStream_t s = (cudaStream_t)stream;
  const LossK k = make_loss_k(args);
  double* partials = reinterpret_cast<double*>(scratch);
  const int nblocks = (k.npix + 255) / 256;
  const bool grads = dL_dimage != nullptr;
  if (grads && k.use_depth && !dL_ddepth) { set_error("loss: dL_ddepth required with use_depth"); return MGS_EINVAL; }
  prof_begin("loss", s);
  if (k.mode == MGS_LOSS_TRACK_HUBER) {
    loss_kernel<false, true><<<nblocks, 256, 0, s>>>(k, nullptr, partials, nullptr, nullptr);
    MGS_LAUNCH_CHECK();
    loss_finalize_kernel<<<1, 256, 0, s>>>(k, partials, nblocks, 0, out_loss, nullptr);
    MGS_LAUNCH_CHECK();
    if (grads) {
      loss_kernel<true, false><<<nblocks, 256, 0, s>>>(k, out_loss, partials, dL_dimage, dL_ddepth);
      MGS_LAUNCH_CHECK();
    } else MGS_CUDA_TRY(memset_async(partials, 0, (size_t)nblocks * 4 * sizeof(double), s));
    loss_finalize_kernel<<<1, 256, 0, s>>>(k, partials, nblocks, 1, out_loss, out_dab);
    MGS_LAUNCH_CHECK();
  } else {
    if (grads) loss_kernel<true, true><<<nblocks, 256, 0, s>>>(k, nullptr, partials, dL_dimage, dL_ddepth);
    else loss_kernel<false, true><<<nblocks, 256, 0, s>>>(k, nullptr, partials, nullptr, nullptr);
    MGS_LAUNCH_CHECK();
    loss_finalize_kernel<<<1, 256, 0, s>>>(k, partials, nblocks, 1, out_loss, out_dab);
    MGS_LAUNCH_CHECK();
  }
  prof_end(s);
  return MGS_OK;
}

// ---- mgs_loss_fwd_bwd_static: the same loss with its scratch in the library's own device memory (no allocation
//      anywhere on the call path; see include/monogs_b200.h).  The array below is a file-scope device symbol: in HOST
//      code its name is only a stand-in, so its device address is fetched once per device with cudaGetSymbolAddress,
//      with that device current, and that pointer is what the kernels receive.
constexpr size_t kLossStaticBlocks = ((size_t)MGS_LOSS_STATIC_MAX_PIXELS + 255) / 256;
constexpr size_t kLossStaticBytes = kLossStaticBlocks * 4 * sizeof(double) + 256;
static __device__ __align__(256) unsigned char g_loss_static_scratch[kLossStaticBytes];
constexpr int kLossStaticMaxDevices = 64;
static void* g_loss_static_ptr[kLossStaticMaxDevices];     // per device, null until that device's first call
static std::mutex g_loss_static_mutex;

int mgs_loss_fwd_bwd_static(const mgs_loss_args* args, float* out_loss, float* out_dab, float* dL_dimage,
                            float* dL_ddepth, void* stream) {
  if (!args) { set_error("loss_static: null argument"); return MGS_EINVAL; }
  if (args->H <= 0 || args->W <= 0) { set_error("loss_static: bad image size %d x %d", args->W, args->H); return MGS_EINVAL; }
  if ((long long)args->H * (long long)args->W > (long long)MGS_LOSS_STATIC_MAX_PIXELS) {
    set_error("loss_static: %d x %d = %lld pixels exceed the built-in scratch (MGS_LOSS_STATIC_MAX_PIXELS = %d); use "
              "mgs_loss_fwd_bwd with caller-owned scratch", args->W, args->H, (long long)args->H * args->W,
              (int)MGS_LOSS_STATIC_MAX_PIXELS);
    return MGS_ENOSPC;
  }
  int dev = -1;
  MGS_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kLossStaticMaxDevices) { set_error("loss_static: device ordinal %d not supported (max %d)", dev, kLossStaticMaxDevices - 1); return MGS_EINVAL; }
  void* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_loss_static_mutex);
    if (g_loss_static_ptr[dev] == nullptr) {     // set-up, once per device, with that device current
      void* p = nullptr;
      const cudaError_t e = cudaGetSymbolAddress(&p, g_loss_static_scratch);
      if (e != cudaSuccess || p == nullptr) {
        set_error("loss_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        return MGS_ECUDA;
      }
      g_loss_static_ptr[dev] = p;
    }
    scratch = g_loss_static_ptr[dev];
  }
  // Every per-block partial slot the finalize kernels read is written earlier in the SAME call (the Huber path without
  // gradients clears its own second-stage slots), so the contents the previous call left behind are never read.
  return mgs_loss_fwd_bwd(args, out_loss, out_dab, dL_dimage, dL_ddepth, scratch, kLossStaticBytes, stream);
}

static int adam_impl(const mgs_adam_group* groups, int32_t n_groups, int32_t step, int32_t* step_dev, float beta1,
                     float beta2, float eps, void* stream);

int mgs_adam_step(const mgs_adam_group* groups, int32_t n_groups, int32_t step, float beta1, float beta2, float eps,
                  void* stream) {
  if (step < 1) { set_error("adam: step must be >= 1"); return MGS_EINVAL; }
  return adam_impl(groups, n_groups, step, nullptr, beta1, beta2, eps, stream);
}

int mgs_adam_step_device(const mgs_adam_group* groups, int32_t n_groups, int32_t* step_counter, float beta1,
                         float beta2, float eps, void* stream) {
  if (!step_counter) { set_error("adam: step_counter is null"); return MGS_EINVAL; }
  return adam_impl(groups, n_groups, 1, step_counter, beta1, beta2, eps, stream);
}

static int adam_impl(const mgs_adam_group* groups, int32_t n_groups, int32_t step, int32_t* step_dev, float beta1,
                     float beta2, float eps, void* stream) {
  if (!groups || n_groups <= 0) { set_error("adam: bad arguments"); return MGS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  for (int g0 = 0; g0 < n_groups; g0 += ADAM_MAX_GROUPS) {
    AdamPack a;
    memset(&a, 0, sizeof(a));
    a.n = 0; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.step_dev = step_dev;
    a.bc1 = (float)(1.0 - pow((double)beta1, (double)step));
    a.bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
    int chunks = 0;
    for (int gi = g0; gi < n_groups && gi < g0 + ADAM_MAX_GROUPS; gi++) {
      const mgs_adam_group& G = groups[gi];
      if (G.numel < 0 || (G.numel > 0 && (!G.param || !G.grad || !G.exp_avg || !G.exp_avg_sq))) { set_error("adam: group %d has null pointers", gi); return MGS_EINVAL; }
      if (G.numel == 0) continue;
      const int j = a.n++;
      a.p[j] = G.param; a.g[j] = G.grad; a.m[j] = G.exp_avg; a.v[j] = G.exp_avg_sq; a.numel[j] = G.numel; a.lr[j] = G.lr; a.lr_dev[j] = G.lr_device;
      a.chunk_start[j] = chunks;
      chunks += (int)((G.numel + ADAM_CHUNK - 1) / ADAM_CHUNK);
    }
    a.chunk_start[a.n] = chunks;
    if (chunks == 0) continue;
    MGS_PROF("adam", s, adam_kernel<<<chunks, 256, 0, s>>>(a));
    MGS_LAUNCH_CHECK();
  }
  if (step_dev) { adam_step_inc_kernel<<<1, 1, 0, s>>>(step_dev); MGS_LAUNCH_CHECK(); }
  return MGS_OK;
}

}  // extern "C"
