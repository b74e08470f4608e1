// Pose-only backward (tracking path) and the sketch-Jacobian reduction.
//
// The map from a Gaussian's screen-space gradients (dL/dmean2D, dL/dconic, dL/ddepth[, dL/dcolor]) to its
// pose tangent dL/dtau is LINEAR (backward.cu:154-361, 433-567).  So instead of reducing screen-space
// gradients over the 256 pixels of every tile for every Gaussian (the reference's dominant cost) and then
// running the per-Gaussian chain, we
//   1. build that linear map B_g (6 x NCOL floats) once per visible Gaussian (pose_jacobian_kernel),
//   2. let every PIXEL apply B_g to its own contribution and accumulate its own 6-vector while it walks its
//      tile list back to front (blend_backward_pose_kernel) — no cross-thread reduction, no atomics per
//      Gaussian, no per-Gaussian gradient buffers,
//   3. sum the per-pixel 6-vectors: over the image for grad_rho/grad_theta (__init__.py:339-342), or per
//      sketch row for the sketched Jacobian (__init__.py:216-253, backward_sketch_jacobian.cu), where the
//      SAME tau image serves every stack because only the pixel->row routing differs between stacks.
#include "launchers.h"
#include <cstdlib>
#include "pose_jac.cuh"
#include "tma.cuh"

namespace mgs {

constexpr int POSE_BATCH = 128;

template <int NCOL>
__global__ void __launch_bounds__(256)
pose_jacobian_kernel(const ViewConst v, const int P, const float* __restrict__ means3D, const float* __restrict__ shs,
                     const float* __restrict__ scales, const float* __restrict__ rotations,
                     const float* __restrict__ cov3D_precomp, const int32_t* __restrict__ radii,
                     const float4* __restrict__ rgb4, float* __restrict__ jac) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P || !(radii[idx] > 0)) return;
  const float3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
  const uint32_t cm = (NCOL == 9) ? __float_as_uint(rgb4[idx].w) : 0u;
  build_pose_jac<NCOL>(v, idx, p, shs, scales, rotations, cov3D_precomp, cm, jac);
}

template <int NCOL, bool HAS_DEPTH, bool FAST>
__global__ void __launch_bounds__(256)
blend_backward_pose_kernel(const int W, const int H, const int gx, const uint2* __restrict__ ranges,
                           const SplatRec* __restrict__ recs, const float* __restrict__ jac,
                           const float* __restrict__ bg, const float* __restrict__ final_T,
                           const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix,
                           const float* __restrict__ dL_dpix_depth, float* __restrict__ tau_image /* [6,H,W] or null */,
                           float* __restrict__ tau_sum /* [6] or null */) {
  constexpr int NF = PoseJac<NCOL>::kFloats;
  __shared__ __align__(128) SplatRec stage[2][POSE_BATCH];
  __shared__ __align__(16) float jstage[POSE_BATCH * NF];
  __shared__ __align__(8) uint64_t full_bar[2];
  __shared__ int s_max_contrib;
  __shared__ float s_part[8][6];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile = blockIdx.x;
  const int tx = tile % gx, ty = tile / gx;
  const int px = tx * TILE_X + (warp & 1) * 8 + (lane & 7);
  const int py = ty * TILE_Y + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = px < W && py < H;
  const size_t pix = (size_t)py * W + px, plane = (size_t)W * H;
  const float pxf = (float)px, pyf = (float)py;
  const float patch_x0 = (float)(tx * TILE_X + (warp & 1) * 8), patch_y0 = (float)(ty * TILE_Y + (warp >> 1) * 4);
  const uint2 range = ranges[tile];

  const float T_final = inside ? final_T[pix] : 0.f;
  const int last_contributor = inside ? (int)n_contrib[pix] : 0;
  float dpx0 = 0.f, dpx1 = 0.f, dpx2 = 0.f, dpd = 0.f;
  if (inside) {
    dpx0 = dL_dpix[pix]; dpx1 = dL_dpix[plane + pix]; dpx2 = dL_dpix[2 * plane + pix];
    if (HAS_DEPTH) dpd = dL_dpix_depth[pix];
  }
  if (tid == 0) {
    s_max_contrib = 0;
    mbar_init(&full_bar[0], 1); mbar_init(&full_bar[1], 1);
    mbar_init_fence();
  }
  __syncthreads();
  {
    int m = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0 && m > 0) atomicMax(&s_max_contrib, m);
  }
  __syncthreads();
  const int n = min(s_max_contrib, (int)(range.y - range.x));
  const int rounds = (n + POSE_BATCH - 1) / POSE_BATCH;
  const SplatRec* src = recs + range.x;
  auto issue = [&](int i) {
    const int hi = n - i * POSE_BATCH, lo = max(0, hi - POSE_BATCH);
    const uint32_t bytes = (uint32_t)(hi - lo) * sizeof(SplatRec);
    mbar_expect_tx(&full_bar[i & 1], bytes);
    bulk_g2s(&stage[i & 1][0], src + lo, bytes, &full_bar[i & 1]);
  };
  if (tid == 0 && rounds > 0) issue(0);

  float T = T_final;
  float ar0 = 0.f, ar1 = 0.f, ar2 = 0.f, ard = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, ld = 0.f, last_alpha = 0.f;
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
  const float bg_dot_dpixel = bg[0] * dpx0 + bg[1] * dpx1 + bg[2] * dpx2;
  float tau[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};

  for (int i = 0; i < rounds; i++) {
    __syncthreads();  // all warps are done with jstage and with the stage the next copy overwrites
    if (tid == 0 && i + 1 < rounds) issue(i + 1);
    mbar_wait(&full_bar[i & 1], (i >> 1) & 1);
    const int hi = n - i * POSE_BATCH, lo = max(0, hi - POSE_BATCH);
    const int cnt = hi - lo;
    const float4* rec4 = reinterpret_cast<const float4*>(&stage[i & 1][0]);
    // gather the batch's Jacobians (L2-resident per-Gaussian array) into shared memory, float4 per thread
    for (int e = tid; e < cnt * (NF / 4); e += 256) {
      const int slot = e / (NF / 4), k = e - slot * (NF / 4);
      const uint32_t id = __float_as_uint(rec4[3 * slot + 0].w);
      reinterpret_cast<float4*>(jstage)[e] = __ldg(reinterpret_cast<const float4*>(jac + (size_t)id * NF) + k);
    }
    __syncthreads();
    const float* recf = reinterpret_cast<const float*>(&stage[i & 1][0]);
    for (int c0 = ((cnt - 1) >> 5) << 5; c0 >= 0; c0 -= 32) {
     // lane-parallel exact cull of 32 splats against this warp's 8x4 patch (see splat_extent)
     const int jj = c0 + lane;
     unsigned hits = __ballot_sync(0xffffffffu, jj < cnt && splat_hits(recf + 12 * jj, patch_x0, patch_x0 + 7.f, patch_y0, patch_y0 + 3.f));
     while (hits) {
      const int j = c0 + 31 - __clz(hits);
      hits &= ~(1u << (j - c0));
      const int contributor = lo + j;
      const float4 g0 = rec4[3 * j + 0];
      const float4 g1 = rec4[3 * j + 1];
      const float dx = g0.x - pxf, dy = g0.y - pyf;
      const float power = -0.5f * (g1.x * dx * dx + g1.z * dy * dy) - g1.y * dx * dy;
      const float G = blend_exp<FAST>(power);
      const float alpha = min(0.99f, g1.w * G);
      const bool skip = !inside || contributor >= last_contributor || power > 0.0f || alpha < 1.0f / 255.0f;
      if (__ballot_sync(0xffffffffu, !skip) == 0u) continue;
      if (!skip) {
        const float4 g2 = rec4[3 * j + 2];
        const float inv_1ma = rcp_unit(1.f - alpha);   // 1-alpha in [0.01,1]: MUFU.RCP, 1 ulp vs the IEEE divide
        T = T * inv_1ma;
        const float dchannel_dcolor = alpha * T;
        float dL_dalpha = 0.0f;
        ar0 = last_alpha * lc0 + (1.f - last_alpha) * ar0; lc0 = g2.x; dL_dalpha += (g2.x - ar0) * dpx0;
        ar1 = last_alpha * lc1 + (1.f - last_alpha) * ar1; lc1 = g2.y; dL_dalpha += (g2.y - ar1) * dpx1;
        ar2 = last_alpha * lc2 + (1.f - last_alpha) * ar2; lc2 = g2.z; dL_dalpha += (g2.z - ar2) * dpx2;
        if (HAS_DEPTH) { ard = last_alpha * ld + (1.f - last_alpha) * ard; ld = g0.z; dL_dalpha += (g0.z - ard) * dpd; }
        dL_dalpha *= T;
        last_alpha = alpha;
        dL_dalpha += (-T_final * inv_1ma) * bg_dot_dpixel;
        const float dL_dG = g1.w * dL_dalpha;
        const float gdx = G * dx, gdy = G * dy;
        float s[NCOL];
        s[0] = dL_dG * (-gdx * g1.x - gdy * g1.y) * ddelx_dx;
        s[1] = dL_dG * (-gdy * g1.z - gdx * g1.y) * ddely_dy;
        s[2] = -0.5f * gdx * dx * dL_dG;
        s[3] = -0.5f * gdx * dy * dL_dG;
        s[4] = -0.5f * gdy * dy * dL_dG;
        s[5] = HAS_DEPTH ? dchannel_dcolor * dpd : 0.f;
        if (NCOL == 9) { s[6] = dchannel_dcolor * dpx0; s[7] = dchannel_dcolor * dpx1; s[8] = dchannel_dcolor * dpx2; }
        const float4* J4 = reinterpret_cast<const float4*>(jstage + j * NF);
        float Bf[NF];
#pragma unroll
        for (int k = 0; k < NF / 4; k++) { const float4 t4 = J4[k]; Bf[4 * k] = t4.x; Bf[4 * k + 1] = t4.y; Bf[4 * k + 2] = t4.z; Bf[4 * k + 3] = t4.w; }
#pragma unroll
        for (int c = 0; c < NCOL; c++) {
          if (!HAS_DEPTH && c == 5) continue;   // no depth gradient (mono tracking): skip the depth column
#pragma unroll
          for (int k = 0; k < 6; k++) tau[k] = fmaf(s[c], Bf[c * 6 + k], tau[k]);
        }
      }
     }
    }
  }

  if (tau_image != nullptr && inside) {
#pragma unroll
    for (int k = 0; k < 6; k++) tau_image[k * plane + pix] = tau[k];
  }
  if (tau_sum != nullptr) {
#pragma unroll
    for (int k = 0; k < 6; k++) {
      float x = tau[k];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      if (lane == 0) s_part[warp][k] = x;
    }
    __syncthreads();
    if (tid < 6) {
      float x = 0.f;
#pragma unroll
      for (int w = 0; w < 8; w++) x += s_part[w][tid];
      if (x != 0.f) atomicAdd(&tau_sum[tid], x);
    }
  }
}

// SJ[j, i, :] = sum over pixels with sketch_indices[j, pix] == i of rows[:, pix]   (NV values per pixel: 6 = the pose
// tangent image of the drop-in sketch backward; 9 = tangent + the two exposure columns + f of the fused second-order step)
template <int NV>
__global__ void __launch_bounds__(256)
sketch_reduce_kernel(const float* __restrict__ rows, const int npix, const int d,
                     const int32_t* __restrict__ sketch_indices, float* __restrict__ out) {
  extern __shared__ float acc[];  // [d][NV]
  const int stack = blockIdx.y;
  for (int k = threadIdx.x; k < d * NV; k += blockDim.x) acc[k] = 0.f;
  __syncthreads();
  const int32_t* idx = sketch_indices + (size_t)stack * npix;
  // Eight pixels per thread and step, their routing words loaded together: a stack routes 1 / n_stacks of the pixels, so the
  // loop is a stream of index loads with little behind them, and one load in flight per thread left it latency-bound
  // (52 -> 41 us for 16 stacks of 640x480; what remains are the shared-memory float atomics, compare-and-swap loops in
  // hardware terms — the fused second-order iteration routes from the blend kernel's epilogue instead, SketchRoute)
  constexpr int U = 8;
  for (int p0 = blockIdx.x * blockDim.x * U + threadIdx.x; p0 < npix; p0 += gridDim.x * blockDim.x * U) {
    int row[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int p = p0 + u * (int)blockDim.x;
      row[u] = p < npix ? __ldg(idx + p) : -1;
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      if (row[u] < 0 || row[u] >= d) continue;
      const int p = p0 + u * (int)blockDim.x;
      float x[NV];
#pragma unroll
      for (int k = 0; k < NV; k++) x[k] = rows[(size_t)k * npix + p];
#pragma unroll
      for (int k = 0; k < NV; k++)
        if (x[k] != 0.f) atomicAdd(&acc[row[u] * NV + k], x[k]);
    }
  }
  __syncthreads();
  float* o = out + (size_t)stack * d * NV;
  for (int k = threadIdx.x; k < d * NV; k += blockDim.x) {
    const float x = acc[k];
    if (x != 0.f) atomicAdd(&o[k], x);
  }
}

size_t backward_pose_scratch_bytes(int P, int W, int H) {
  (void)W; (void)H;
  return ((size_t)(P > 0 ? P : 1) * PoseJac<9>::kFloats * sizeof(float) + 255) / 256 * 256;
}

template <int NC, bool HD, bool FAST>
static int launch_blend_pose_t(const BackwardArgs& a, const float* jac, float* tau_image, cudaStream_t s) {
  const ViewConst& v = a.v;
  prof_begin("blend_backward_pose", s);
  blend_backward_pose_kernel<NC, HD, FAST><<<v.gx * v.gy, 256, 0, s>>>(
      v.W, v.H, v.gx, ws_ptr<uint2>(a.ws, a.L.ranges), ws_ptr<SplatRec>(a.ws, a.L.records), jac, v.bg,
      ws_ptr<float>(a.ws, a.L.final_T), ws_ptr<uint32_t>(a.ws, a.L.n_contrib), a.dL_dcolor_img, a.dL_ddepth_img,
      tau_image, a.dL_dtau_sum);
  prof_end(s);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

// blend pass only: `jac` already holds the per-Gaussian maps (built by pose_jacobian_kernel or by the fused
// tracking preprocess); a null dL_ddepth_img means "no depth gradient" and drops the depth column entirely;
// `fast` must match the exp variant the forward pass of this iteration used
int launch_blend_backward_pose(const BackwardArgs& a, const float* jac, bool with_color, cudaStream_t s, bool fast) {
  float* tau_image = a.sketch_dtau;  // reused field: per-pixel tau image [6,H,W] (may be null)
  const bool hd = a.dL_ddepth_img != nullptr;
  if (fast) {
    if (with_color) return hd ? launch_blend_pose_t<9, true, true>(a, jac, tau_image, s) : launch_blend_pose_t<9, false, true>(a, jac, tau_image, s);
    return hd ? launch_blend_pose_t<6, true, true>(a, jac, tau_image, s) : launch_blend_pose_t<6, false, true>(a, jac, tau_image, s);
  }
  if (with_color) return hd ? launch_blend_pose_t<9, true, false>(a, jac, tau_image, s) : launch_blend_pose_t<9, false, false>(a, jac, tau_image, s);
  return hd ? launch_blend_pose_t<6, true, false>(a, jac, tau_image, s) : launch_blend_pose_t<6, false, false>(a, jac, tau_image, s);
}

int launch_backward_pose(const BackwardArgs& a, cudaStream_t s) {
  const ViewConst& v = a.v;
  const int P = a.P;
  const size_t need = backward_pose_scratch_bytes(P, v.W, v.H);
  if (a.scratch == nullptr || a.scratch_bytes < need) { set_error("pose backward scratch too small: %zu < %zu", a.scratch_bytes, need); return MGS_ENOSPC; }
  float* jac = reinterpret_cast<float*>(a.scratch);
  if (a.dL_dtau_sum) MGS_CUDA_TRY(memset_async(a.dL_dtau_sum, 0, 6 * sizeof(float), s));
  const bool with_color = (a.shs != nullptr && v.D > 0);
  prof_begin("pose_jacobian", s);
  if (with_color)
    pose_jacobian_kernel<9><<<(P + 255) / 256, 256, 0, s>>>(v, P, a.means3D, a.shs, a.scales, a.rotations, a.cov3D_precomp,
                                                            a.radii, ws_ptr<float4>(a.ws, a.L.rgb4), jac);
  else
    pose_jacobian_kernel<6><<<(P + 255) / 256, 256, 0, s>>>(v, P, a.means3D, a.shs, a.scales, a.rotations, a.cov3D_precomp,
                                                            a.radii, ws_ptr<float4>(a.ws, a.L.rgb4), jac);
  prof_end(s);
  MGS_LAUNCH_CHECK();
  return launch_blend_backward_pose(a, jac, with_color, s);
}

int launch_sketch_reduce(const float* rows, int H, int W, int n_stacks, int d, const int32_t* sketch_indices,
                         float* out, cudaStream_t s, int nv) {
  MGS_CUDA_TRY(memset_async(out, 0, (size_t)n_stacks * d * nv * sizeof(float), s));
  const int npix = W * H;
  // every CTA ends with d * nv global atomics, so the grid stays small: ~4 CTAs per SM over all stacks
  const int per_stack = max(1, min((npix + 255) / 256, (148 * 4 + n_stacks - 1) / n_stacks));
  dim3 grid(per_stack, n_stacks);
  prof_begin("sketch_reduce", s);
  if (nv == 9) sketch_reduce_kernel<9><<<grid, 256, (size_t)d * 9 * sizeof(float), s>>>(rows, npix, d, sketch_indices, out);
  else sketch_reduce_kernel<6><<<grid, 256, (size_t)d * 6 * sizeof(float), s>>>(rows, npix, d, sketch_indices, out);
  prof_end(s);
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

}  // namespace mgs
