// Input producers next to the hot path, one thread per pixel, HBM-bound:
//   gen_rays      reference src/util/cam_geometry.py:36-79   32 B written per ray, nothing read but the camera
//   gen_rays_at   the same for selected pixels (diner.py:229-258): 8 B index read + 32 B written per ray
//   depth2normal  reference src/util/depth2normal.py:6-87    4 B read (+ cached neighbours), 12 B written per pixel
// Every fp32 operation of the reference is one rounding here (explicit __f*_rn, no contraction), so the maps agree
// with torch to the last bits except where torch's own reduction order is free (3-term sums, the 3x3 rotation).
#include "common.cuh"

namespace dnr {

// ray of pixel (x, y) of camera (e, k): pixel centre -> camera plane -> unit direction (cam_geometry.py:57-63),
// world direction = R^T d_cam, camera centre = -R^T t (:66-70)
__device__ __forceinline__ void ray_of_pixel(const float* __restrict__ e, const float* __restrict__ k, int x, int y,
                                             float z_near, float z_far, float4* __restrict__ out) {
  const float px = __fdiv_rn(__fsub_rn((float)x + 0.5f, k[2]), k[0]);
  const float py = __fdiv_rn(__fsub_rn((float)y + 0.5f, k[5]), k[4]);
  const float nrm = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(px, px), __fmul_rn(py, py)), 1.0f));
  const float dx = __fdiv_rn(px, nrm), dy = __fdiv_rn(py, nrm), dz = __fdiv_rn(1.0f, nrm);
  const float wx = fmaf(e[8], dz, fmaf(e[4], dy, __fmul_rn(e[0], dx)));
  const float wy = fmaf(e[9], dz, fmaf(e[5], dy, __fmul_rn(e[1], dx)));
  const float wz = fmaf(e[10], dz, fmaf(e[6], dy, __fmul_rn(e[2], dx)));
  const float ox = -fmaf(e[8], e[11], fmaf(e[4], e[7], __fmul_rn(e[0], e[3])));
  const float oy = -fmaf(e[9], e[11], fmaf(e[5], e[7], __fmul_rn(e[1], e[3])));
  const float oz = -fmaf(e[10], e[11], fmaf(e[6], e[7], __fmul_rn(e[2], e[3])));
  out[0] = make_float4(ox, oy, oz, wx);
  out[1] = make_float4(wy, wz, z_near, z_far);
}

__global__ void gen_rays_kernel(const float* __restrict__ E, const float* __restrict__ K, int B, int W, int H,
                                const float* __restrict__ z_near, const float* __restrict__ z_far,
                                float4* __restrict__ rays) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long per = (long long)H * W;
  if (idx >= per * B) return;
  const int b = (int)(idx / per);
  const int rem = (int)(idx % per);
  ray_of_pixel(E + (size_t)b * 16, K + (size_t)b * 9, rem % W, rem / W, z_near[b], z_far[b], rays + idx * 2);
}

// rays of SELECTED pixels: rays[b, i] = gen_rays(...)[b].view(H*W, 8)[pix[b, i]]  (the caller's gather, diner.py:229-258)
__global__ void gen_rays_at_kernel(const float* __restrict__ E, const float* __restrict__ K, int B, int W, int H,
                                   const float* __restrict__ z_near, const float* __restrict__ z_far,
                                   const long long* __restrict__ pix, int n, float4* __restrict__ rays,
                                   int* __restrict__ bad) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)n * B) return;
  const int b = (int)(idx / n);
  long long p = pix[idx];
  const long long per = (long long)H * W;
  if (p < 0) p += per;                                  // torch indexing accepts negative indices
  if (p < 0 || p >= per) { if (bad) atomicExch(bad, 1); p = 0; }
  ray_of_pixel(E + (size_t)b * 16, K + (size_t)b * 9, (int)(p % W), (int)(p / W), z_near[b], z_far[b], rays + idx * 2);
}

struct D2NParams {
  const float* dmap;   // [N,1,H,W]
  const float* K;      // [N,3,3]
  int N, H, W;
  int batch_axis;      // 1: the reference's torch.cross (no dim) runs over the batch axis (N == 3)
  float* normal;       // [N,3,H,W]
};

// back-projected point of pixel (y, x) of view n, replicate padding (depth2normal.py:23-32)
__device__ __forceinline__ float3 d2n_point(const D2NParams& P, int n, int y, int x) {
  y = min(max(y, 0), P.H - 1); x = min(max(x, 0), P.W - 1);
  const float* k = P.K + (size_t)n * 9;
  const float d = __ldg(P.dmap + ((size_t)n * P.H + y) * P.W + x);
  const float rx = __fdiv_rn(__fsub_rn((float)x + 0.5f, k[2]), k[0]);
  const float ry = __fdiv_rn(__fsub_rn((float)y + 0.5f, k[5]), k[4]);
  return make_float3(__fmul_rn(rx, d), __fmul_rn(ry, d), d);
}

__device__ __forceinline__ void d2n_diffs(const D2NParams& P, int n, int y, int x, float3& v, float3& h) {
  const float3 dn = d2n_point(P, n, y + 1, x), up = d2n_point(P, n, y - 1, x);
  const float3 rt = d2n_point(P, n, y, x + 1), lf = d2n_point(P, n, y, x - 1);
  v = make_float3(__fsub_rn(dn.x, up.x), __fsub_rn(dn.y, up.y), __fsub_rn(dn.z, up.z));      // :51
  h = make_float3(__fsub_rn(rt.x, lf.x), __fsub_rn(rt.y, lf.y), __fsub_rn(rt.z, lf.z));      // :52
}

__device__ __forceinline__ float d2n_cross1(float a1, float b2, float a2, float b1) {
  return __fsub_rn(__fmul_rn(a1, b2), __fmul_rn(a2, b1));
}

// un-cleaned unit normal at (n, y, x) (:53-54)
__device__ __forceinline__ float3 d2n_raw(const D2NParams& P, int n, int y, int x) {
  float3 c;
  if (!P.batch_axis) {
    float3 v, h;
    d2n_diffs(P, n, y, x, v, h);
    c = make_float3(d2n_cross1(v.y, h.z, v.z, h.y), d2n_cross1(v.z, h.x, v.x, h.z), d2n_cross1(v.x, h.y, v.y, h.x));
  } else {
    // N == 3: cross product over the VIEW axis, component by component: c[n] = v[n+1] * h[n+2] - v[n+2] * h[n+1]
    float3 v1, h1, v2, h2;
    d2n_diffs(P, (n + 1) % 3, y, x, v1, h1);
    d2n_diffs(P, (n + 2) % 3, y, x, v2, h2);
    c = make_float3(d2n_cross1(v1.x, h2.x, v2.x, h1.x), d2n_cross1(v1.y, h2.y, v2.y, h1.y),
                    d2n_cross1(v1.z, h2.z, v2.z, h1.z));
  }
  const float nrm = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(c.x, c.x), __fmul_rn(c.y, c.y)), __fmul_rn(c.z, c.z)));
  return make_float3(__fdiv_rn(c.x, nrm), __fdiv_rn(c.y, nrm), __fdiv_rn(c.z, nrm));         // 0/0 = NaN, as torch
}

__global__ void depth2normal_kernel(const D2NParams P) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long per = (long long)P.H * P.W;
  if (idx >= per * P.N) return;
  const int n = (int)(idx / per);
  const int rem = (int)(idx % per);
  const int y = rem / P.W, x = rem % P.W;
  float3 out = make_float3(0.f, 0.f, 0.f);
  if (__ldg(P.dmap + idx) != 0.f) {                                                           // :78
    // cleaning (:57-77): step away from neighbours that are holes (x == 0), then clip
    const int oy = -(int)(d2n_point(P, n, y + 1, x).x == 0.f) + (int)(d2n_point(P, n, y - 1, x).x == 0.f);
    const int ox = -(int)(d2n_point(P, n, y, x + 1).x == 0.f) + (int)(d2n_point(P, n, y, x - 1).x == 0.f);
    out = d2n_raw(P, n, min(max(y + oy, 0), P.H - 1), min(max(x + ox, 0), P.W - 1));
  }
  float* o = P.normal + (size_t)n * 3 * per + rem;
  o[0] = out.x; o[per] = out.y; o[2 * per] = out.z;
}

}  // namespace dnr

// Pixel-centre rays [origin 3 | direction 3 | near | far] of B cameras; replaces gen_rays (cam_geometry.py:36-79).
DNR_EXPORT int dnr_gen_rays(const float* extrinsics, const float* intrinsics, int B, int W, int H, const float* z_near,
                            const float* z_far, float* rays, cudaStream_t stream) {
  DNR_REQUIRE(B >= 0 && W >= 1 && H >= 1, "dnr_gen_rays: bad sizes");
  DNR_REQUIRE(extrinsics && intrinsics && z_near && z_far && rays, "dnr_gen_rays: null argument");
  const long long n = (long long)B * W * H;
  if (n == 0) return 0;
  dnr::gen_rays_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(extrinsics, intrinsics, B, W, H, z_near, z_far,
                                                                        (float4*)rays);
  DNR_CHECK_LAUNCH("dnr_gen_rays");
  return 0;
}

// Rays of n selected pixels per camera (flat indices y*W+x, int64 [B,n]); same arithmetic as dnr_gen_rays, so
// rays[b,i] == dnr_gen_rays(...)[b, pix[b,i]] bit for bit.  Replaces gen_rays + the advanced-index gather of
// DINER.calc_losses (diner.py:222-258).  `bad_flag` (device int, nullable) is set to 1 if an index is out of range.
DNR_EXPORT int dnr_gen_rays_at(const float* extrinsics, const float* intrinsics, int B, int W, int H, const float* z_near,
                               const float* z_far, const long long* pix_idx, int n, float* rays, int* bad_flag,
                               cudaStream_t stream) {
  DNR_REQUIRE(B >= 0 && W >= 1 && H >= 1 && n >= 0, "dnr_gen_rays_at: bad sizes");
  const long long tot = (long long)B * n;
  if (tot == 0) return 0;
  DNR_REQUIRE(extrinsics && intrinsics && z_near && z_far && rays && pix_idx, "dnr_gen_rays_at: null argument");
  dnr::gen_rays_at_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>(extrinsics, intrinsics, B, W, H, z_near, z_far,
                                                                          pix_idx, n, (float4*)rays, bad_flag);
  DNR_CHECK_LAUNCH("dnr_gen_rays_at");
  return 0;
}

// Camera-space normal maps from depth maps; replaces depth2normal (depth2normal.py:6-87), including the axis the
// reference's torch.cross picks when N == 3.
DNR_EXPORT int dnr_depth2normal(const float* dmap, const float* K, int N, int H, int W, float* normal,
                                cudaStream_t stream) {
  DNR_REQUIRE(N >= 0 && H >= 1 && W >= 1, "dnr_depth2normal: bad sizes");
  DNR_REQUIRE(H != 3 && W != 3, "dnr_depth2normal: H == 3 or W == 3 is not supported (axis of the reference's cross)");
  DNR_REQUIRE(dmap && K && normal, "dnr_depth2normal: null argument");
  const long long n = (long long)N * W * H;
  if (n == 0) return 0;
  dnr::D2NParams P;
  P.dmap = dmap; P.K = K; P.N = N; P.H = H; P.W = W; P.batch_axis = N == 3; P.normal = normal;
  dnr::depth2normal_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(P);
  DNR_CHECK_LAUNCH("dnr_depth2normal");
  return 0;
}
