// Forward pipeline of the rasterizer: per-Gaussian preprocess, scan, key generation, sort, tile ranges +
// packed records, alpha blend.  Replaces CudaRasterizer::Rasterizer::forward (rasterizer_impl.cu:198-344),
// FORWARD::preprocess / FORWARD::render (forward.cu:157-258, 263-392).
#include "launchers.h"
#include "pose_jac.cuh"
#include <cub/device/device_scan.cuh>
#include <cub/device/device_radix_sort.cuh>
#include <cub/iterator/transform_input_iterator.cuh>
#include <cstdarg>
#include <cstdio>
#include <map>
#include <string>
#include <vector>

namespace mgs {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}
const char* get_error() { return g_err; }

// ---- kernel timer ----
struct ProfRec { const char* name; cudaEvent_t e0, e1; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof;
static ProfRec g_cur;
void prof_begin(const char* name, cudaStream_t s) {
  if (!g_prof_on) return;
  g_cur.name = name;
  cudaEventCreate(&g_cur.e0); cudaEventCreate(&g_cur.e1);
  cudaEventRecord(g_cur.e0, s);
}
void prof_end(cudaStream_t s) {
  if (!g_prof_on) return;
  cudaEventRecord(g_cur.e1, s);
  g_prof.push_back(g_cur);
}
void prof_enable(bool on) {
  for (auto& r : g_prof) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  g_prof.clear();
  g_prof_on = on;
}
int prof_report(char* buf, size_t n) {
  cudaDeviceSynchronize();
  std::map<std::string, std::pair<int, double>> agg;
  for (auto& r : g_prof) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) { auto& a = agg[r.name]; a.first++; a.second += ms; }
  }
  size_t off = 0;
  for (auto& kv : agg) {
    int w = snprintf(buf + off, off < n ? n - off : 0, "%s %d %.6f\n", kv.first.c_str(), kv.second.first, kv.second.second);
    if (w < 0 || off + (size_t)w >= n) break;
    off += (size_t)w;
  }
  if (n) buf[off < n ? off : n - 1] = 0;
  return (int)agg.size();
}

static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// tiles_touched read through the depth-order permutation (input iterator of the scan)
struct PermLoad {
  const uint32_t* tiles_touched;
  __host__ __device__ __forceinline__ uint32_t operator()(const uint32_t& g) const { return tiles_touched[g]; }
};
using PermIter = cub::TransformInputIterator<uint32_t, PermLoad, const uint32_t*>;

int ws_layout(int32_t P, int32_t W, int32_t H, int64_t R_cap, WsLayout* L) {
  if (P < 0 || W <= 0 || H <= 0 || R_cap < 0) { set_error("ws_layout: bad sizes"); return MGS_EINVAL; }
  const size_t Pp = (size_t)(P > 0 ? P : 1), Rp = (size_t)(R_cap > 0 ? R_cap : 1);
  const size_t npix = (size_t)W * H;
  const size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
  size_t scan_b = 0, sort_b = 0, dsort_b = 0;
  cudaError_t e = cub::DeviceScan::InclusiveSum(nullptr, scan_b, PermIter(nullptr, PermLoad{nullptr}), (uint32_t*)nullptr, (int)Pp);
  if (e == cudaSuccess)
    e = cub::DeviceRadixSort::SortPairs(nullptr, sort_b, (uint32_t*)nullptr, (uint32_t*)nullptr,
                                        (uint32_t*)nullptr, (uint32_t*)nullptr, (int)Rp);
  if (e == cudaSuccess)
    e = cub::DeviceRadixSort::SortPairs(nullptr, dsort_b, (uint32_t*)nullptr, (uint32_t*)nullptr,
                                        (uint32_t*)nullptr, (uint32_t*)nullptr, (int)Pp);
  if (e != cudaSuccess) { set_error("CUB size query failed: %s", cudaGetErrorString(e)); return MGS_ECUDA; }
  L->scan_tmp_bytes = scan_b; L->sort_tmp_bytes = sort_b; L->dsort_tmp_bytes = dsort_b;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes); return o; };
  // ---- R_cap-independent part (stage 1 and backward use only this) ----
  L->header = take(sizeof(WsHeader));
  L->depths = take(Pp * 4);
  L->means2D = take(Pp * 8);
  L->conic_opacity = take(Pp * 16);
  L->rgb4 = take(Pp * 16);
  L->tiles_touched = take(Pp * 4);
  L->point_offsets = take(Pp * 4);
  L->dkey_in = take(Pp * 4);
  L->dkey_out = take(Pp * 4);
  L->perm_in = take(Pp * 4);
  L->perm = take(Pp * 4);
  L->scan_tmp = take(scan_b);
  L->dsort_tmp = take(dsort_b);
  L->final_T = take(npix * 4);
  L->n_contrib = take(npix * 4);
  L->ranges = take(tiles * 8);
  // ---- per-instance part ----
  L->records = take(Rp * sizeof(SplatRec));
  L->keys_in = take(Rp * 4);
  L->keys_out = take(Rp * 4);
  L->vals_in = take(Rp * 4);
  L->vals_out = take(Rp * 4);
  L->sort_tmp = take(sort_b);
  L->total = off;
  return MGS_OK;
}

// ------------------------------------------------------------------------------------------------
// K1: per-Gaussian preprocess.  One thread per Gaussian; identical arithmetic order to forward.cu:157-258
// so that radii, tile rects and depth bits match the reference bit for bit.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cov3d_from_scale_rot(const float* __restrict__ scales, const float* __restrict__ rots,
                                                     int idx, float mod, float* cov6) {
  // forward.cu:120-154 — quaternion is used as given (not normalised), order (r,x,y,z) = q[0..3]
  const float sx = mod * scales[3 * idx + 0], sy = mod * scales[3 * idx + 1], sz = mod * scales[3 * idx + 2];
  const float4 q = reinterpret_cast<const float4*>(rots)[idx];
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  M3 S = m3_cols(sx, 0.f, 0.f, 0.f, sy, 0.f, 0.f, 0.f, sz);
  M3 R = m3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                 2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                 2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
  M3 Mm = m3_mul(S, R);
  M3 Sg = m3_mul(m3_t(Mm), Mm);
  cov6[0] = Sg[0][0]; cov6[1] = Sg[0][1]; cov6[2] = Sg[0][2];
  cov6[3] = Sg[1][1]; cov6[4] = Sg[1][2]; cov6[5] = Sg[2][2];
}

struct Cov2DOut { float a, b, c; };
// forward.cu:76-115 — EWA projection of the 3D covariance, FoV clamp 1.3x, +0.3 low-pass
__device__ __forceinline__ Cov2DOut cov2d_project(const float3& mean, const ViewConst& v, const float* cov6) {
  float3 t = xform4x3(mean, v.view);
  const float limx = 1.3f * v.tanfovx, limy = 1.3f * v.tanfovy;
  const float txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = min(limx, max(-limx, txtz)) * t.z;
  t.y = min(limy, max(-limy, tytz)) * t.z;
  M3 J = m3_cols(v.focal_x / t.z, 0.0f, -(v.focal_x * t.x) / (t.z * t.z),
                 0.0f, v.focal_y / t.z, -(v.focal_y * t.y) / (t.z * t.z),
                 0.f, 0.f, 0.f);
  M3 Wm = m3_cols(v.view[0], v.view[4], v.view[8], v.view[1], v.view[5], v.view[9],
                  v.view[2], v.view[6], v.view[10]);
  M3 T = m3_mul(Wm, J);
  M3 Vrk = m3_cols(cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]);
  M3 cov = m3_mul(m3_mul(m3_t(T), m3_t(Vrk)), T);
  Cov2DOut o;
  o.a = cov[0][0] + 0.3f; o.b = cov[0][1]; o.c = cov[1][1] + 0.3f;
  return o;
}

// SH -> RGB (forward.cu:22-73).  Returns the clamped colour and a 3-bit mask of channels clamped at 0.
__device__ __forceinline__ float3 sh_to_rgb(int idx, int deg, int M, const float3& pos, const float* __restrict__ campos,
                                            const float* __restrict__ shs, uint32_t& clamp_mask) {
  const float* sh = shs + (size_t)idx * M * 3;
  float w[16];
  w[0] = kSH0;
  if (deg > 0) {
    float dx = pos.x - campos[0], dy = pos.y - campos[1], dz = pos.z - campos[2];
    const float len = sqrtf(dx * dx + dy * dy + dz * dz);
    const float x = dx / len, y = dy / len, z = dz / len;
    w[1] = -(kSH1 * y); w[2] = kSH1 * z; w[3] = -(kSH1 * x);
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      w[4] = kSH2[0] * xy; w[5] = kSH2[1] * yz; w[6] = kSH2[2] * (2.0f * zz - xx - yy);
      w[7] = kSH2[3] * xz; w[8] = kSH2[4] * (xx - yy);
      if (deg > 2) {
        w[9] = kSH3[0] * y * (3.0f * xx - yy);
        w[10] = kSH3[1] * xy * z;
        w[11] = kSH3[2] * y * (4.0f * zz - xx - yy);
        w[12] = kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
        w[13] = kSH3[4] * x * (4.0f * zz - xx - yy);
        w[14] = kSH3[5] * z * (xx - yy);
        w[15] = kSH3[6] * x * (xx - 3.0f * yy);
      }
    }
  }
  const int n = (deg + 1) * (deg + 1);
  float c[3];
#pragma unroll
  for (int ch = 0; ch < 3; ch++) c[ch] = w[0] * sh[ch];
  for (int k = 1; k < n; k++) {
#pragma unroll
    for (int ch = 0; ch < 3; ch++) c[ch] = fmaf(w[k], sh[3 * k + ch], c[ch]);
  }
  clamp_mask = 0;
#pragma unroll
  for (int ch = 0; ch < 3; ch++) {
    c[ch] += 0.5f;
    if (c[ch] < 0.f) clamp_mask |= (1u << ch);
    c[ch] = fmaxf(c[ch], 0.f);
  }
  return make_float3(c[0], c[1], c[2]);
}

template <int JAC_NCOL>   // 0: plain; 6 / 9: also build the pose Jacobian of visible Gaussians (tracking path)
__global__ void __launch_bounds__(256)
preprocess_kernel(const ViewConst v, const int P,
                  const float* __restrict__ means3D, const float* __restrict__ opacities,
                  const float* __restrict__ shs, const float* __restrict__ colors_precomp,
                  const float* __restrict__ scales, const float* __restrict__ rotations,
                  const float* __restrict__ cov3D_precomp,
                  int32_t* __restrict__ radii, float* __restrict__ depths, float2* __restrict__ means2D,
                  float4* __restrict__ conic_opacity, float4* __restrict__ rgb4,
                  uint32_t* __restrict__ tiles_touched, uint32_t* __restrict__ dkey, uint32_t* __restrict__ iota,
                  float* __restrict__ jac) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  int my_radius_i = 0;
  uint32_t touched = 0, depth_key = 0xFFFFFFFFu;   // culled Gaussians sort to the end and emit nothing
  do {
    const float3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    // near cull only (auxiliary.h:139-164)
    const float4 ph = xform4x4(p, v.proj);
    const float pw = 1.0f / (ph.w + 0.0000001f);
    const float3 pproj = {ph.x * pw, ph.y * pw, ph.z * pw};
    const float3 pview = xform4x3(p, v.view);
    if (pview.z <= 0.2f) {
      if (v.prefiltered) { printf("Point is filtered although prefiltered is set. This shouldn't happen!"); __trap(); }
      break;
    }
    float cov6_local[6];
    const float* cov6;
    if (cov3D_precomp != nullptr) cov6 = cov3D_precomp + 6 * (size_t)idx;
    else { cov3d_from_scale_rot(scales, rotations, idx, v.scale_modifier, cov6_local); cov6 = cov6_local; }
    const Cov2DOut cv = cov2d_project(p, v, cov6);
    const float det = (cv.a * cv.c - cv.b * cv.b);
    if (det == 0.0f) break;
    const float det_inv = 1.f / det;
    const float3 conic = {cv.c * det_inv, -cv.b * det_inv, cv.a * det_inv};
    const float mid = 0.5f * (cv.a + cv.c);
    const float lambda1 = mid + sqrtf(max(0.1f, mid * mid - det));
    const float lambda2 = mid - sqrtf(max(0.1f, mid * mid - det));
    const float my_radius = ceilf(3.f * sqrtf(max(lambda1, lambda2)));
    const float2 pix = {ndc_to_pix(pproj.x, v.W), ndc_to_pix(pproj.y, v.H)};
    uint2 rmin, rmax;
    tile_rect(pix, (int)my_radius, v.gx, v.gy, rmin, rmax);
    const uint32_t area = (rmax.x - rmin.x) * (rmax.y - rmin.y);
    if (area == 0) break;
    uint32_t cm = 0;
    if (colors_precomp == nullptr) {
      const float3 c = sh_to_rgb(idx, v.D, v.M, p, v.campos, shs, cm);
      rgb4[idx] = make_float4(c.x, c.y, c.z, __uint_as_float(cm));
    }
    if (JAC_NCOL > 0) build_pose_jac<(JAC_NCOL > 0 ? JAC_NCOL : 6)>(v, idx, p, shs, scales, rotations, cov3D_precomp, cm, jac);
    depths[idx] = pview.z;
    means2D[idx] = pix;
    conic_opacity[idx] = make_float4(conic.x, conic.y, conic.z, opacities[idx]);
    my_radius_i = (int)my_radius;
    touched = area;
    depth_key = __float_as_uint(pview.z);   // > 0.2, so the bit pattern orders like the float (rasterizer_impl.cu:99-101)
  } while (0);
  radii[idx] = my_radius_i;
  tiles_touched[idx] = touched;
  dkey[idx] = depth_key;
  iota[idx] = (uint32_t)idx;
}

// ------------------------------------------------------------------------------------------------
// Binning.  bin_header: device-side instance count (no host sync, rasterizer_impl.cu:284 removed).
// ------------------------------------------------------------------------------------------------
__global__ void bin_header_kernel(const uint32_t* __restrict__ point_offsets, int P, uint32_t R_cap, WsHeader* hdr) {
  const uint32_t n = P > 0 ? point_offsets[P - 1] : 0u;
  hdr->num_rendered = n;
  hdr->overflow = n > R_cap ? 1u : 0u;
  hdr->r_used = n > R_cap ? R_cap : n;
}

// K3: emit tile keys + Gaussian ids (rasterizer_impl.cu:70-111).  The reference sorts 64-bit (tile|depth) keys in
// one go; here the Gaussians were first stably sorted by depth bits (P items, 32-bit keys), instances are emitted in
// that order, and only the tile bits remain to be sorted (stable LSD radix => identical final permutation, ties keep
// ascending Gaussian index, at a third of the sort passes on 8-byte instead of 12-byte pairs).  Slots
// [r_used, R_cap) get a sentinel tile id that is never a real tile so the sort length stays a host constant.
__global__ void __launch_bounds__(256)
emit_keys_kernel(const int P, const uint32_t R_cap, const int gx, const int gy, const uint32_t sentinel,
                 const uint32_t* __restrict__ perm, const float2* __restrict__ means2D,
                 const uint32_t* __restrict__ offsets, const int32_t* __restrict__ radii,
                 const WsHeader* __restrict__ hdr,
                 uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (uint32_t)P) {
    const uint32_t g = perm[i];
    const int rad = radii[g];
    if (rad > 0) {
      uint32_t off = (i == 0) ? 0u : offsets[i - 1];
      uint2 rmin, rmax;
      tile_rect(means2D[g], rad, gx, gy, rmin, rmax);
      for (uint32_t y = rmin.y; y < rmax.y; y++)
        for (uint32_t x = rmin.x; x < rmax.x; x++) {
          if (off < R_cap) { keys[off] = y * gx + x; vals[off] = g; }
          off++;
        }
    }
  }
  if (i < R_cap && i >= hdr->r_used) { keys[i] = sentinel; vals[i] = 0xFFFFFFFFu; }
}

// K5 + record packing: tile [start,end) ranges (rasterizer_impl.cu:116-138) and, in the same pass over the
// sorted list, the 48-B records the blend kernels stream.
__global__ void __launch_bounds__(256)
ranges_records_kernel(const uint32_t R_cap, const WsHeader* __restrict__ hdr,
                      const uint32_t* __restrict__ keys, const uint32_t* __restrict__ ids,
                      const float2* __restrict__ means2D, const float* __restrict__ depths,
                      const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb4,
                      const float* __restrict__ colors_precomp,
                      uint2* __restrict__ ranges, SplatRec* __restrict__ recs) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t L = hdr->r_used;
  if (i >= L) return;
  const uint32_t tile = keys[i];
  if (i == 0) ranges[tile].x = 0;
  else {
    const uint32_t prev = keys[i - 1];
    if (tile != prev) { ranges[prev].y = i; ranges[tile].x = i; }
  }
  if (i == L - 1) ranges[tile].y = L;
  const uint32_t id = ids[i];
  const float2 xy = means2D[id];
  const float4 co = conic_opacity[id];
  float4 c;
  if (colors_precomp != nullptr) c = make_float4(colors_precomp[3 * id], colors_precomp[3 * id + 1], colors_precomp[3 * id + 2], 0.f);
  else c = rgb4[id];
  float4* out = reinterpret_cast<float4*>(recs + i);
  out[0] = make_float4(xy.x, xy.y, depths[id], __uint_as_float(id));
  out[1] = co;
  out[2] = make_float4(c.x, c.y, c.z, __uint_as_float(splat_extent(co.x, co.y, co.z, co.w)));
}


int launch_preprocess(const ViewConst& v, int P, const float* means3D, const float* opacities, const float* shs,
                      const float* colors_precomp, const float* scales, const float* rotations,
                      const float* cov3D_precomp, int32_t* radii, void* ws, const WsLayout& L, cudaStream_t s,
                      float* jac, int jac_ncol) {
  if (P == 0) return MGS_OK;
#define MGS_PRE_ARGS v, P, means3D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp, radii,           \
      ws_ptr<float>(ws, L.depths), ws_ptr<float2>(ws, L.means2D), ws_ptr<float4>(ws, L.conic_opacity),               \
      ws_ptr<float4>(ws, L.rgb4), ws_ptr<uint32_t>(ws, L.tiles_touched), ws_ptr<uint32_t>(ws, L.dkey_in),            \
      ws_ptr<uint32_t>(ws, L.perm_in), jac
  prof_begin("preprocess", s);
  if (jac == nullptr) preprocess_kernel<0><<<(P + 255) / 256, 256, 0, s>>>(MGS_PRE_ARGS);
  else if (jac_ncol == 9) preprocess_kernel<9><<<(P + 255) / 256, 256, 0, s>>>(MGS_PRE_ARGS);
  else preprocess_kernel<6><<<(P + 255) / 256, 256, 0, s>>>(MGS_PRE_ARGS);
  prof_end(s);
#undef MGS_PRE_ARGS
  MGS_LAUNCH_CHECK();
  // stable sort of the Gaussians by depth bits -> perm (depth order, ties by index)
  size_t dtmp = L.dsort_tmp_bytes;
  prof_begin("cub_sort_depth", s);
  MGS_CUDA_TRY(cub::DeviceRadixSort::SortPairs(ws_ptr<void>(ws, L.dsort_tmp), dtmp, ws_ptr<uint32_t>(ws, L.dkey_in),
                                               ws_ptr<uint32_t>(ws, L.dkey_out), ws_ptr<uint32_t>(ws, L.perm_in),
                                               ws_ptr<uint32_t>(ws, L.perm), P, 0, 32, s));
  prof_end(s);
  // inclusive scan of tiles_touched in depth order
  size_t tmp = L.scan_tmp_bytes;
  prof_begin("cub_scan", s);
  MGS_CUDA_TRY(cub::DeviceScan::InclusiveSum(ws_ptr<void>(ws, L.scan_tmp), tmp,
                                             PermIter(ws_ptr<uint32_t>(ws, L.perm), PermLoad{ws_ptr<uint32_t>(ws, L.tiles_touched)}),
                                             ws_ptr<uint32_t>(ws, L.point_offsets), P, s));
  prof_end(s);
  return MGS_OK;
}

int launch_binning(const ViewConst& v, int P, const float* colors_precomp, const int32_t* radii, void* ws,
                   const WsLayout& L, int64_t R_cap, cudaStream_t s) {
  WsHeader* hdr = ws_ptr<WsHeader>(ws, L.header);
  const uint32_t tiles = (uint32_t)v.gx * v.gy;
  bin_header_kernel<<<1, 1, 0, s>>>(ws_ptr<uint32_t>(ws, L.point_offsets), P, (uint32_t)R_cap, hdr);
  MGS_LAUNCH_CHECK();
  MGS_CUDA_TRY(cudaMemsetAsync(ws_ptr<void>(ws, L.ranges), 0, (size_t)tiles * sizeof(uint2), s));
  if (R_cap == 0 || P == 0) return MGS_OK;
  const uint32_t bit = tile_id_bits(tiles);
  const uint32_t sentinel = (1u << bit) - 1u;
  const int64_t n_thr = P > R_cap ? P : R_cap;
  MGS_PROF("emit_keys", s, emit_keys_kernel<<<(unsigned)((n_thr + 255) / 256), 256, 0, s>>>(
      P, (uint32_t)R_cap, v.gx, v.gy, sentinel, ws_ptr<uint32_t>(ws, L.perm), ws_ptr<float2>(ws, L.means2D),
      ws_ptr<uint32_t>(ws, L.point_offsets), radii, hdr, ws_ptr<uint32_t>(ws, L.keys_in),
      ws_ptr<uint32_t>(ws, L.vals_in)));
  MGS_LAUNCH_CHECK();
  size_t tmp = L.sort_tmp_bytes;
  prof_begin("cub_sort", s);
  MGS_CUDA_TRY(cub::DeviceRadixSort::SortPairs(ws_ptr<void>(ws, L.sort_tmp), tmp, ws_ptr<uint32_t>(ws, L.keys_in),
                                               ws_ptr<uint32_t>(ws, L.keys_out), ws_ptr<uint32_t>(ws, L.vals_in),
                                               ws_ptr<uint32_t>(ws, L.vals_out), (int)R_cap, 0, (int)bit, s));
  prof_end(s);
  MGS_PROF("ranges_records", s, ranges_records_kernel<<<(unsigned)((R_cap + 255) / 256), 256, 0, s>>>(
      (uint32_t)R_cap, hdr, ws_ptr<uint32_t>(ws, L.keys_out), ws_ptr<uint32_t>(ws, L.vals_out),
      ws_ptr<float2>(ws, L.means2D), ws_ptr<float>(ws, L.depths), ws_ptr<float4>(ws, L.conic_opacity),
      ws_ptr<float4>(ws, L.rgb4), colors_precomp, ws_ptr<uint2>(ws, L.ranges), ws_ptr<SplatRec>(ws, L.records)));
  MGS_LAUNCH_CHECK();
  return MGS_OK;
}

}  // namespace mgs
