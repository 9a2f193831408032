// Alpha compositing of (rgb, sigma) samples into colour / depth / weights, forward and backward.
// Restates reference nerf_renderer.py:299-301 (deltas) and :344-360 (blending); one warp per ray,
// exclusive transmittance by a warp-level product scan, single pass over HBM.
//
// Algorithmic bytes per ray: fwd 32 + K*(16+4+4) + 16 = 1008 B at K=40,
// bwd 16 + K*(16+4+4 [+4]) + K*16 = 1616 B (1776 B with d_weights).
#include "common.cuh"

namespace dnr {

__device__ __forceinline__ float warp_excl_prod(float v, int lane, float& total) {
  // inclusive product scan, then shift
  float x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x *= y;
  }
  total = __shfl_sync(0xffffffffu, x, 31);
  float e = __shfl_up_sync(0xffffffffu, x, 1);
  return lane == 0 ? 1.0f : e;
}

__device__ __forceinline__ float warp_excl_suffix_sum(float v, int lane, float& total) {
  float x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float y = __shfl_down_sync(0xffffffffu, x, o);
    if (lane + o < 32) x += y;
  }
  total = __shfl_sync(0xffffffffu, x, 0);
  float e = __shfl_down_sync(0xffffffffu, x, 1);
  return lane == 31 ? 0.0f : e;
}

__global__ void __launch_bounds__(256) composite_fwd_kernel(
    const float4* __restrict__ rgbsigma,  // [n_rays, K]
    const float* __restrict__ z,          // [n_rays, K] ascending
    const float* __restrict__ rays,       // [n_rays, 8]
    int n_rays, int K, int white_bkgd,
    float* __restrict__ weights,          // [n_rays, K] or null
    float* __restrict__ rgb,              // [n_rays, 3]
    float* __restrict__ depth)            // [n_rays]
{
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  for (long long ray = (long long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); ray < n_rays;
       ray += (long long)gridDim.x * warps_per_block) {
    const float far = __ldg(rays + ray * 8 + 7);
    const float* zr = z + ray * K;
    const float4* cr = rgbsigma + ray * K;
    float T_carry = 1.0f;
    float ar = 0.f, ag = 0.f, ab = 0.f, ad = 0.f, aw = 0.f;
    for (int k0 = 0; k0 < K; k0 += 32) {
      const int k = k0 + lane;
      const bool valid = k < K;
      float zk = 0.f, znext = 0.f;
      float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
      if (valid) {
        zk = zr[k];
        znext = (k + 1 < K) ? zr[k + 1] : far;
        c = cr[k];
      }
      const float delta = __fsub_rn(znext, zk);
      const float alpha = valid ? __fsub_rn(1.0f, expf(-__fmul_rn(delta, fmaxf(c.w, 0.f)))) : 0.f;
      const float t = valid ? __fadd_rn(__fsub_rn(1.0f, alpha), 1e-10f) : 1.0f;
      float total;
      const float T = T_carry * warp_excl_prod(t, lane, total);
      T_carry *= total;
      const float w = alpha * T;
      if (valid && weights) weights[ray * K + k] = w;
      ar += w * c.x; ag += w * c.y; ab += w * c.z; ad += w * zk; aw += w;
    }
    ar = warp_sum(ar); ag = warp_sum(ag); ab = warp_sum(ab); ad = warp_sum(ad); aw = warp_sum(aw);
    if (lane == 0) {
      if (white_bkgd) { ar = ar + 1.f - aw; ag = ag + 1.f - aw; ab = ab + 1.f - aw; }
      rgb[ray * 3 + 0] = ar; rgb[ray * 3 + 1] = ag; rgb[ray * 3 + 2] = ab;
      depth[ray] = ad;
    }
  }
}

// d_rgbsigma is w.r.t. the (already activated) rgb and sigma the forward consumed.
__global__ void __launch_bounds__(256) composite_bwd_kernel(
    const float4* __restrict__ rgbsigma, const float* __restrict__ z, const float* __restrict__ rays,
    const float* __restrict__ d_rgb,      // [n_rays,3]
    const float* __restrict__ d_depth,    // [n_rays] or null
    const float* __restrict__ d_weights,  // [n_rays,K] or null
    int n_rays, int K, int white_bkgd,
    float4* __restrict__ d_rgbsigma)      // [n_rays,K]
{
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int nchunk = (K + 31) >> 5;
  for (long long ray = (long long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); ray < n_rays;
       ray += (long long)gridDim.x * warps_per_block) {
    const float far = __ldg(rays + ray * 8 + 7);
    const float* zr = z + ray * K;
    const float4* cr = rgbsigma + ray * K;
    const float gr = d_rgb[ray * 3 + 0], gg = d_rgb[ray * 3 + 1], gb = d_rgb[ray * 3 + 2];
    const float gd = d_depth ? d_depth[ray] : 0.f;
    const float gbg = white_bkgd ? (gr + gg + gb) : 0.f;

    // pass 1 (forward order): transmittance carried into every 32-sample chunk (<= 8 chunks: K <= 256)
    float T_in[8];
    {
      float T_carry = 1.0f;
      for (int ch = 0; ch < nchunk; ++ch) {
        const int k = ch * 32 + lane;
        float t = 1.0f;
        if (k < K) {
          const float zk = zr[k];
          const float znext = (k + 1 < K) ? zr[k + 1] : far;
          const float alpha = 1.0f - expf(-(znext - zk) * fmaxf(cr[k].w, 0.f));
          t = 1.0f - alpha + 1e-10f;
        }
        float total;
        (void)warp_excl_prod(t, lane, total);
        T_in[ch] = T_carry;
        T_carry *= total;
      }
    }
    // pass 2 (reverse order): suffix sums of g_m * w_m
    float S_carry = 0.f;
    for (int ch = nchunk - 1; ch >= 0; --ch) {
      const int k = ch * 32 + lane;
      const bool valid = k < K;
      float zk = 0.f, znext = 0.f;
      float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
      if (valid) { zk = zr[k]; znext = (k + 1 < K) ? zr[k + 1] : far; c = cr[k]; }
      const float delta = znext - zk;
      const float sig = fmaxf(c.w, 0.f);
      const float alpha = valid ? 1.0f - expf(-delta * sig) : 0.f;
      const float t = valid ? 1.0f - alpha + 1e-10f : 1.0f;
      float total;
      const float T = T_in[ch] * warp_excl_prod(t, lane, total);
      const float w = alpha * T;
      float g = gr * c.x + gg * c.y + gb * c.z + gd * zk - gbg;
      if (valid && d_weights) g += d_weights[ray * K + k];
      const float gw = valid ? g * w : 0.f;
      float tot2;
      const float S = S_carry + warp_excl_suffix_sum(gw, lane, tot2);
      S_carry += tot2;
      if (valid) {
        const float d_alpha = g * T - S / t;
        const float d_sigma = (c.w > 0.f) ? d_alpha * delta * (1.0f - alpha) : 0.f;
        d_rgbsigma[ray * K + k] = make_float4(w * gr, w * gg, w * gb, d_sigma);
      }
    }
  }
}

// ---- K even and <= 64: two adjacent samples per lane, every load of a ray issued before the first use -----------
// (the generic kernels below walk 32-sample chunks one after the other: two dependent HBM round trips at K = 40)

__device__ __forceinline__ float warp_excl_prod1(float v, int lane) {
  float x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x *= y;
  }
  float e = __shfl_up_sync(0xffffffffu, x, 1);
  return lane == 0 ? 1.0f : e;
}

__device__ __forceinline__ float warp_excl_suffix_sum1(float v, int lane) {
  float x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float y = __shfl_down_sync(0xffffffffu, x, o);
    if (lane + o < 32) x += y;
  }
  float e = __shfl_down_sync(0xffffffffu, x, 1);
  return lane == 31 ? 0.0f : e;
}

struct PairRaw { float4 c0, c1; float2 zz; float far; };
struct PairSamples {
  float4 c0, c1;
  float z0, z1, delta0, delta1, alpha0, alpha1, t0, t1;
};

// samples (2*lane, 2*lane+1) of one ray: the loads ...
__device__ __forceinline__ PairRaw load_pair(const float4* __restrict__ rgbsigma, const float* __restrict__ z,
                                             const float* __restrict__ rays, long long ray, int K, int lane) {
  PairRaw r;
  r.far = __ldg(rays + ray * 8 + 7);
  r.zz = make_float2(0.f, 0.f);
  r.c0 = make_float4(0.f, 0.f, 0.f, 0.f); r.c1 = r.c0;
  if (2 * lane < K) {
    r.zz = __ldg(reinterpret_cast<const float2*>(z + ray * K) + lane);
    r.c0 = __ldg(rgbsigma + ray * K + 2 * lane);
    r.c1 = __ldg(rgbsigma + ray * K + 2 * lane + 1);
  }
  return r;
}

// ... and their per-sample terms (nerf_renderer.py:299-301, :344-349)
__device__ __forceinline__ PairSamples derive_pair(const PairRaw& r, int K, int lane) {
  PairSamples s;
  const bool valid = 2 * lane < K;
  float zn1 = __shfl_down_sync(0xffffffffu, r.zz.x, 1);
  if (2 * lane + 2 >= K) zn1 = r.far;
  s.c0 = r.c0; s.c1 = r.c1;
  s.z0 = r.zz.x; s.z1 = r.zz.y;
  s.delta0 = __fsub_rn(r.zz.y, r.zz.x);
  s.delta1 = __fsub_rn(zn1, r.zz.y);
  s.alpha0 = valid ? __fsub_rn(1.0f, expf(-__fmul_rn(s.delta0, fmaxf(s.c0.w, 0.f)))) : 0.f;
  s.alpha1 = valid ? __fsub_rn(1.0f, expf(-__fmul_rn(s.delta1, fmaxf(s.c1.w, 0.f)))) : 0.f;
  s.t0 = valid ? __fadd_rn(__fsub_rn(1.0f, s.alpha0), 1e-10f) : 1.0f;
  s.t1 = valid ? __fadd_rn(__fsub_rn(1.0f, s.alpha1), 1e-10f) : 1.0f;
  return s;
}

constexpr int kRaysInFlight = 2;   // rays per warp whose loads are issued together (bytes in flight per SM)

__global__ void __launch_bounds__(256) composite_fwd2_kernel(
    const float4* __restrict__ rgbsigma, const float* __restrict__ z, const float* __restrict__ rays,
    int n_rays, int K, int white_bkgd, float* __restrict__ weights, float* __restrict__ rgb, float* __restrict__ depth)
{
  const int lane = threadIdx.x & 31;
  const long long stride = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long ray0 = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); ray0 < n_rays;
       ray0 += stride * kRaysInFlight) {
    PairRaw raw[kRaysInFlight];
#pragma unroll
    for (int u = 0; u < kRaysInFlight; ++u)
      if (ray0 + u * stride < n_rays) raw[u] = load_pair(rgbsigma, z, rays, ray0 + u * stride, K, lane);
#pragma unroll
    for (int u = 0; u < kRaysInFlight; ++u) {
      const long long ray = ray0 + u * stride;
      if (ray >= n_rays) break;
      const PairSamples s = derive_pair(raw[u], K, lane);
      const float T0 = warp_excl_prod1(s.t0 * s.t1, lane);
      const float T1 = T0 * s.t0;
      const float w0 = s.alpha0 * T0, w1 = s.alpha1 * T1;
      if (weights && 2 * lane < K)
        reinterpret_cast<float2*>(weights + ray * K)[lane] = make_float2(w0, w1);
      float ar = fmaf(w1, s.c1.x, w0 * s.c0.x), ag = fmaf(w1, s.c1.y, w0 * s.c0.y), ab = fmaf(w1, s.c1.z, w0 * s.c0.z);
      float ad = fmaf(w1, s.z1, w0 * s.z0), aw = w0 + w1;
      ar = warp_sum(ar); ag = warp_sum(ag); ab = warp_sum(ab); ad = warp_sum(ad);
      if (white_bkgd) aw = warp_sum(aw);
      if (lane == 0) {
        if (white_bkgd) { ar = ar + 1.f - aw; ag = ag + 1.f - aw; ab = ab + 1.f - aw; }
        rgb[ray * 3 + 0] = ar; rgb[ray * 3 + 1] = ag; rgb[ray * 3 + 2] = ab;
        depth[ray] = ad;
      }
    }
  }
}

// ---- K <= 48 (the reference's K = 40): one ray per HALF warp, three adjacent samples per lane ---------------------
// At K = 40 the two-samples-per-lane kernel leaves 12 of 32 lanes idle and spends ~31 shuffles per ray (one 32-lane
// scan + five 32-lane sums): it is issue-bound (ncu: 68 % issue slots, 194 warp instructions per ray, 58 % of the HBM
// roofline at 1 M rays).  Here a warp blends two rays at once: 16-lane scan (4 shuffles), the four outputs reduced
// together by a butterfly that halves the number of live values per step (5 shuffles for 4 sums), so a pair of rays
// costs 14 shuffles.
struct TriRaw { float4 c0, c1, c2; float z0, z1, z2; float far; };

__device__ __forceinline__ TriRaw load_tri(const float4* __restrict__ rgbsigma, const float* __restrict__ z,
                                           const float* __restrict__ rays, long long ray, int K, int l16) {
  TriRaw r;
  r.far = __ldg(rays + ray * 8 + 7);
  const int k = 3 * l16;
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  r.c0 = zero; r.c1 = zero; r.c2 = zero; r.z0 = 0.f; r.z1 = 0.f; r.z2 = 0.f;
  const float* zr = z + ray * K;
  const float4* cr = rgbsigma + ray * K;
  if (k < K) { r.z0 = __ldg(zr + k); r.c0 = __ldg(cr + k); }
  if (k + 1 < K) { r.z1 = __ldg(zr + k + 1); r.c1 = __ldg(cr + k + 1); }
  if (k + 2 < K) { r.z2 = __ldg(zr + k + 2); r.c2 = __ldg(cr + k + 2); }
  return r;
}

__device__ __forceinline__ void tri_term(bool valid, float zk, float znext, float sigma, float& alpha, float& t) {
  const float delta = __fsub_rn(znext, zk);
  alpha = valid ? __fsub_rn(1.0f, expf(-__fmul_rn(delta, fmaxf(sigma, 0.f)))) : 0.f;      // nerf_renderer.py:344-346
  t = valid ? __fadd_rn(__fsub_rn(1.0f, alpha), 1e-10f) : 1.0f;                             // :347
}

__global__ void __launch_bounds__(256) composite_fwd3_kernel(
    const float4* __restrict__ rgbsigma, const float* __restrict__ z, const float* __restrict__ rays,
    int n_rays, int K, int white_bkgd, float* __restrict__ weights, float* __restrict__ rgb, float* __restrict__ depth)
{
  const int lane = threadIdx.x & 31, l16 = lane & 15, half = lane >> 4;
  const long long n_pairs = ((long long)n_rays + 1) >> 1;
  const long long stride = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long p0 = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); p0 < n_pairs;
       p0 += stride * kRaysInFlight) {
    TriRaw raw[kRaysInFlight];
#pragma unroll
    for (int u = 0; u < kRaysInFlight; ++u) {
      const long long ray = 2 * (p0 + u * stride) + half;
      if (ray < n_rays) raw[u] = load_tri(rgbsigma, z, rays, ray, K, l16);
    }
#pragma unroll
    for (int u = 0; u < kRaysInFlight; ++u) {
      if (p0 + u * stride >= n_pairs) break;                    // warp-uniform
      const long long ray = 2 * (p0 + u * stride) + half;
      const bool live = ray < n_rays;                           // odd n_rays: the last pair has one ray
      TriRaw r = raw[u];
      if (!live) { r.c0 = make_float4(0.f, 0.f, 0.f, 0.f); r.c1 = r.c0; r.c2 = r.c0; r.z0 = r.z1 = r.z2 = 0.f; r.far = 0.f; }
      const int k = 3 * l16;
      float zn = __shfl_down_sync(0xffffffffu, r.z0, 1, 16);    // first sample of the next lane of this ray
      const bool v0 = live && k < K, v1 = live && k + 1 < K, v2 = live && k + 2 < K;
      const float n0 = (k + 1 < K) ? r.z1 : r.far, n1 = (k + 2 < K) ? r.z2 : r.far, n2 = (k + 3 < K) ? zn : r.far;
      float a0, a1, a2, t0, t1, t2;
      tri_term(v0, r.z0, n0, r.c0.w, a0, t0);
      tri_term(v1, r.z1, n1, r.c1.w, a1, t1);
      tri_term(v2, r.z2, n2, r.c2.w, a2, t2);
      // exclusive transmittance: 16-lane product scan of the per-lane products (nerf_renderer.py:348-349)
      float x = (t0 * t1) * t2;
#pragma unroll
      for (int o = 1; o < 16; o <<= 1) {
        const float y = __shfl_up_sync(0xffffffffu, x, o, 16);
        if (l16 >= o) x *= y;
      }
      const float e = __shfl_up_sync(0xffffffffu, x, 1, 16);
      const float T0 = l16 == 0 ? 1.0f : e;
      const float T1 = T0 * t0, T2 = T1 * t1;
      const float w0 = a0 * T0, w1 = a1 * T1, w2 = a2 * T2;
      if (weights) {
        float* wr = weights + ray * K + k;
        if (v0) wr[0] = w0;
        if (v1) wr[1] = w1;
        if (v2) wr[2] = w2;
      }
      float ar = fmaf(w2, r.c2.x, fmaf(w1, r.c1.x, w0 * r.c0.x));
      float ag = fmaf(w2, r.c2.y, fmaf(w1, r.c1.y, w0 * r.c0.y));
      float ab = fmaf(w2, r.c2.z, fmaf(w1, r.c1.z, w0 * r.c0.z));
      float ad = fmaf(w2, r.z2, fmaf(w1, r.z1, w0 * r.z0));
      // butterfly: (ar, ab) and (ag, ad) fold over bit 3, the two results over bit 2, then bits 1, 0:
      // lanes 0-3 of the half warp end with sum(ar), 4-7 sum(ag), 8-11 sum(ab), 12-15 sum(ad)
      const bool b3 = l16 & 8, b2 = l16 & 4;
      float A = (b3 ? ab : ar) + __shfl_xor_sync(0xffffffffu, b3 ? ar : ab, 8, 16);
      float B = (b3 ? ad : ag) + __shfl_xor_sync(0xffffffffu, b3 ? ag : ad, 8, 16);
      float V = (b2 ? B : A) + __shfl_xor_sync(0xffffffffu, b2 ? A : B, 4, 16);
      V += __shfl_xor_sync(0xffffffffu, V, 2, 16);
      V += __shfl_xor_sync(0xffffffffu, V, 1, 16);
      if (white_bkgd) {
        float aw = (w0 + w1) + w2;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) aw += __shfl_xor_sync(0xffffffffu, aw, o, 16);
        if (l16 < 12) V = V + 1.f - aw;                         // nerf_renderer.py:361-363
      }
      if (live && (l16 & 3) == 0) {
        if (l16 < 12) rgb[ray * 3 + (l16 >> 2)] = V; else depth[ray] = V;
      }
    }
  }
}

__global__ void __launch_bounds__(256) composite_bwd2_kernel(
    const float4* __restrict__ rgbsigma, const float* __restrict__ z, const float* __restrict__ rays,
    const float* __restrict__ d_rgb, const float* __restrict__ d_depth, const float* __restrict__ d_weights,
    int n_rays, int K, int white_bkgd, float4* __restrict__ d_rgbsigma)
{
  const int lane = threadIdx.x & 31;
  const bool valid = 2 * lane < K;
  const long long stride = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long ray0 = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); ray0 < n_rays;
       ray0 += stride * kRaysInFlight) {
    PairRaw raw[kRaysInFlight];
    float2 dws[kRaysInFlight];
    float4 gs[kRaysInFlight];            // d_rgb (3) and d_depth of the ray
#pragma unroll
    for (int u = 0; u < kRaysInFlight; ++u) {
      const long long ray = ray0 + u * stride;
      if (ray >= n_rays) continue;
      raw[u] = load_pair(rgbsigma, z, rays, ray, K, lane);
      dws[u] = make_float2(0.f, 0.f);
      if (d_weights && valid) dws[u] = __ldg(reinterpret_cast<const float2*>(d_weights + ray * K) + lane);
      gs[u] = make_float4(__ldg(d_rgb + ray * 3 + 0), __ldg(d_rgb + ray * 3 + 1), __ldg(d_rgb + ray * 3 + 2),
                          d_depth ? __ldg(d_depth + ray) : 0.f);
    }
#pragma unroll
    for (int u = 0; u < kRaysInFlight; ++u) {
      const long long ray = ray0 + u * stride;
      if (ray >= n_rays) break;
      const PairSamples s = derive_pair(raw[u], K, lane);
      const float gr = gs[u].x, gg = gs[u].y, gb = gs[u].z, gd = gs[u].w;
      const float2 dw = dws[u];
      const float gbg = white_bkgd ? (gr + gg + gb) : 0.f;
      const float T0 = warp_excl_prod1(s.t0 * s.t1, lane);
      const float T1 = T0 * s.t0;
      const float w0 = s.alpha0 * T0, w1 = s.alpha1 * T1;
      const float g0 = gr * s.c0.x + gg * s.c0.y + gb * s.c0.z + gd * s.z0 - gbg + dw.x;
      const float g1 = gr * s.c1.x + gg * s.c1.y + gb * s.c1.z + gd * s.z1 - gbg + dw.y;
      const float gw0 = valid ? g0 * w0 : 0.f, gw1 = valid ? g1 * w1 : 0.f;
      const float S1 = warp_excl_suffix_sum1(gw0 + gw1, lane);   // sum over the samples after 2*lane+1
      const float S0 = S1 + gw1;
      if (valid) {
        const float da0 = g0 * T0 - S0 / s.t0, da1 = g1 * T1 - S1 / s.t1;
        const float ds0 = (s.c0.w > 0.f) ? da0 * s.delta0 * (1.0f - s.alpha0) : 0.f;
        const float ds1 = (s.c1.w > 0.f) ? da1 * s.delta1 * (1.0f - s.alpha1) : 0.f;
        float4* o = d_rgbsigma + ray * K + 2 * lane;
        o[0] = make_float4(w0 * gr, w0 * gg, w0 * gb, ds0);
        o[1] = make_float4(w1 * gr, w1 * gg, w1 * gb, ds1);
      }
    }
  }
}

}  // namespace dnr

// C ABI -- replaces NeRFRendererDGS.composite's blending tail (nerf_renderer.py:344-365).
DNR_EXPORT int dnr_composite_fwd(const float* rgbsigma, const float* z, const float* rays, int n_rays, int K,
                                 int white_bkgd, float* weights, float* rgb, float* depth, cudaStream_t stream) {
  DNR_REQUIRE(K >= 1 && K <= 256, "dnr_composite_fwd: K must be in [1,256]");
  if (n_rays <= 0) return 0;
  const int wpb = 8;
  dnr::DeviceInfo* di = dnr::device_info();
  DNR_REQUIRE(di != nullptr, "dnr_composite_fwd: no current CUDA device");
  const int sms = di->num_sms;
  int blocks = (int)((n_rays + wpb - 1) / wpb);
  if (blocks > sms * 8) blocks = sms * 8;      // one resident wave (2 048 threads per SM), grid-stride over the rays
  // K <= 48: one ray per half warp, three samples per lane (fwd3); 48 < K <= 64, K even: two samples per lane (fwd2);
  // otherwise the generic one-warp-per-ray kernel
  auto kern = K <= 48 ? dnr::composite_fwd3_kernel
                      : (K % 2 == 0 && K <= 64) ? dnr::composite_fwd2_kernel : dnr::composite_fwd_kernel;
  if (K <= 48) {                                 // two rays per warp, three resident blocks per SM (80 registers)
    blocks = (int)(((n_rays + 1) / 2 + wpb - 1) / wpb);
    if (blocks > sms * 3) blocks = sms * 3;
  }
  kern<<<blocks, wpb * 32, 0, stream>>>((const float4*)rgbsigma, z, rays, n_rays, K, white_bkgd, weights, rgb, depth);
  DNR_CHECK_LAUNCH("dnr_composite_fwd");
  return 0;
}

DNR_EXPORT int dnr_composite_bwd(const float* rgbsigma, const float* z, const float* rays, const float* d_rgb,
                                 const float* d_depth, const float* d_weights, int n_rays, int K, int white_bkgd,
                                 float* d_rgbsigma, cudaStream_t stream) {
  DNR_REQUIRE(K >= 1 && K <= 256, "dnr_composite_bwd: K must be in [1,256]");
  if (n_rays <= 0) return 0;
  const int wpb = 8;
  dnr::DeviceInfo* di = dnr::device_info();
  DNR_REQUIRE(di != nullptr, "dnr_composite_bwd: no current CUDA device");
  int blocks = (int)((n_rays + wpb - 1) / wpb);
  if (blocks > di->num_sms * 8) blocks = di->num_sms * 8;
  auto kern = (K % 2 == 0 && K <= 64) ? dnr::composite_bwd2_kernel : dnr::composite_bwd_kernel;
  kern<<<blocks, wpb * 32, 0, stream>>>((const float4*)rgbsigma, z, rays, d_rgb, d_depth, d_weights, n_rays, K, white_bkgd,
                                        (float4*)d_rgbsigma);
  DNR_CHECK_LAUNCH("dnr_composite_bwd");
  return 0;
}
