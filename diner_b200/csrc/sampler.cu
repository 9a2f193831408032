// Depth-guided ray sampler: stratified candidates -> projection into the NV source views ->
// nearest taps of depth / padded depth-std / normal -> erf surface likelihood (max over views) ->
// occlusion-aware likelihood -> top-(K-G) shortlist + G gaussian samples -> uniform fill-up -> sort.
//
// One warp per ray.  Restates reference nerf_renderer.py:39-63 (sample_coarse), :65-190
// (sample_depthguided), :367-397 (fill_up_uniform_samples), image_encoder.py:129-204 (index_depth /
// index_depth_std / index_normal) and torch_helpers.py:125-160,294-302.  Compiled with --fmad=false:
// every fp32 operation between the ray and the texel index rounds exactly like the corresponding
// torch op (nearest-neighbour taps are discontinuous); the only fused operations are the 3x3
// rotation, written as the explicit FMA chain torch's matmul uses.
#include "common.cuh"

namespace dnr {

constexpr int kMaxViews = 8;
constexpr int kSamplerWarps = 4;

struct SamplerParams {
  const float* rays;      // [SB*NR, 8]
  const float* poses;     // [SB, NV, 4, 4] world->cam
  const float* focal;     // [SB, NV, 2]
  const float* cxy;       // [SB, NV, 2]
  const float* depth;     // [SB, NV, H, W]
  const float* stdpad;    // [SB, NV, H+2P, W+2P]  exponentially padded depth std
  const float* normal;    // [SB, NV, 3, H, W]
  const float* lin_tab;   // [C]  torch.linspace(0, 1-1/C, C)
  const float* u_coarse;  // [SB*NR, C] or null (-> Philox)
  const float* n_gauss;   // packed [n_masked, G] or null
  const float* u_fill;    // packed [n_missing] or null
  const int* gauss_off;   // [SB*NR] exclusive prefix of ray_mask (with n_gauss)
  const int* fill_off;    // [SB*NR] exclusive prefix of n_missing (with u_fill)
  const float* cand_off;  // [SB*NR, C, 3] or null: NOVEL variant, per-candidate deformation (dnr_novel_deform)
  unsigned long long seed;
  int SB, NV, NR, H, W, C, K, G, P;
  float depth_diff_max, img_w, img_h, std_sx, std_sy, inv_c_step;
  float inv_w, inv_h;     // exact reciprocals when img_w / img_h are powers of two (x / 2^k == x * 2^-k bit for bit), else 0
  int count_only;
  int unpacked;           // n_gauss is [SB*NR, G] and u_fill is [SB*NR, K] (indexed by ray / sorted slot)
  float* z_out;           // [SB*NR, K]
  int* ray_mask_out;      // [SB*NR] or null
  int* n_missing_out;     // [SB*NR] or null
  int* sel_idx_out;       // [SB*NR, K-G] or null (-1 = empty slot)
  float* p_out;           // [SB*NR, C] or null: per-candidate likelihood (debug / parity tests)
};

__device__ __forceinline__ int nearest_border(float u, int size) {
  float x = ((u + 1.0f) * (float)size - 1.0f) / 2.0f;
  x = fminf(fmaxf(x, 0.0f), (float)(size - 1));
  return (int)nearbyintf(x);
}
// returns -1 when outside (zeros padding)
__device__ __forceinline__ int nearest_zeros(float u, int size) {
  float x = ((u + 1.0f) * (float)size - 1.0f) / 2.0f;
  float r = nearbyintf(x);
  if (!(r >= 0.0f && r <= (float)(size - 1))) return -1;
  return (int)r;
}

__global__ void __launch_bounds__(kSamplerWarps * 32) sampler_kernel(SamplerParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = P.C, K = P.K, NV = P.NV;
  // per-warp carve-up
  const size_t per_warp = (size_t)C * 10 + (size_t)K * 8 + kMaxViews * 20 * 4 + 16;
  unsigned char* base = smem_raw + (size_t)warp * ((per_warp + 15) / 16 * 16);
  float* lp = reinterpret_cast<float*>(base);                 // [C] likelihood (later: opaque likelihood)
  float* lz = lp + C;                                         // [C] candidate depth
  float* slot = lz + C;                                       // [K]
  float* sorted = slot + K;                                   // [K]
  float* cam = sorted + K;                                    // [kMaxViews][20]
  unsigned short* lidx = reinterpret_cast<unsigned short*>(cam + kMaxViews * 20);  // [C]

  const long long n_rays = (long long)P.SB * P.NR;
  const long long ray = (long long)blockIdx.x * kSamplerWarps + warp;
  if (ray >= n_rays) return;
  const int sb = (int)(ray / P.NR);

  const float* rp = P.rays + ray * 8;
  const float o0 = rp[0], o1 = rp[1], o2 = rp[2], d0 = rp[3], d1 = rp[4], d2 = rp[5];
  const float near = rp[6], far = rp[7];
  const float step = (far - near) / (float)C;          // nerf_renderer.py:95
  const float half_step = step / 2.0f;

  // per-view constants: R (9), t (3), dir_cam (3), focal (2), c (2)
  if (lane < NV) {
    const float* E = P.poses + ((size_t)sb * NV + lane) * 16;
    float* cv = cam + lane * 20;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      cv[i * 3 + 0] = E[i * 4 + 0]; cv[i * 3 + 1] = E[i * 4 + 1]; cv[i * 3 + 2] = E[i * 4 + 2];
      cv[9 + i] = E[i * 4 + 3];
      cv[12 + i] = fmaf(E[i * 4 + 2], d2, fmaf(E[i * 4 + 1], d1, E[i * 4 + 0] * d0));   // :103
    }
    cv[15] = P.focal[((size_t)sb * NV + lane) * 2 + 0];
    cv[16] = P.focal[((size_t)sb * NV + lane) * 2 + 1];
    cv[17] = P.cxy[((size_t)sb * NV + lane) * 2 + 0];
    cv[18] = P.cxy[((size_t)sb * NV + lane) * 2 + 1];
  }
  __syncwarp();

  const int H = P.H, W = P.W, Wp = P.W + 2 * P.P, Hp = P.H + 2 * P.P;
  const size_t map_sz = (size_t)H * W, pad_sz = (size_t)Hp * Wp;
  const float sqrt2 = 1.41421356237309504880f;
  Philox rng(P.seed);

  // ---- 1. candidates and their likelihoods, compacted to the non-zero ones (in index order) ----
  // A lane scores four consecutive candidates per pass: one Philox call yields the four uniforms of candidates
  // 4j..4j+3 (counter i >> 2, component i & 3), and the compaction keeps index order (lane-major, q-minor).
  // Per view the four candidates go through the taps stage by stage (std -> depth -> normal -> erf), so that the four
  // map loads of a stage are in flight together instead of one dependent L2 round trip per candidate; the float
  // operations of a candidate are the same, in the same order, as in the one-candidate form.
  int n_nz = 0;
  const unsigned lt_mask = (1u << lane) - 1u;
  for (int i0 = 0; i0 < C; i0 += 128) {
    const int ib = i0 + 4 * lane;
    float pq[4], zq[4], x0[4], x1[4], x2[4];
    uint4 r = make_uint4(0u, 0u, 0u, 0u);
    if (!P.u_coarse && ib < C) r = rng((uint32_t)ray, (uint32_t)(ray >> 32), (uint32_t)(ib >> 2), 0u);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int i = ib + q;
      pq[q] = 0.0f; zq[q] = 0.0f; x0[q] = x1[q] = x2[q] = 0.0f;
      if (i < C) {
        const float u = P.u_coarse ? P.u_coarse[ray * C + i] : u01(q == 0 ? r.x : q == 1 ? r.y : q == 2 ? r.z : r.w);
        float s = P.lin_tab[i] + u * P.inv_c_step;       // :55-57
        const float zc = near * (1.0f - s) + far * s;     // :60
        zq[q] = zc;
        x0[q] = o0 + zc * d0; x1[q] = o1 + zc * d1; x2[q] = o2 + zc * d2;   // :96
        if (P.cand_off) {                                 // novel/nerf_novel_renderer.py:104-107 (deform_points)
          const float* co = P.cand_off + ((size_t)ray * C + i) * 3;
          x0[q] += co[0]; x1[q] += co[1]; x2[q] += co[2];
        }
      }
    }
    for (int v = 0; v < NV; ++v) {
      const float* cv = cam + v * 20;
      const float* std_v = P.stdpad + ((size_t)sb * NV + v) * pad_sz;
      const float* depth_v = P.depth + ((size_t)sb * NV + v) * map_sz;
      const float* normal_v = P.normal + ((size_t)sb * NV + v) * 3 * map_sz;
      float zcam[4], uu[4], vv[4], sref[4], dref[4];
      bool ok[4];
      // stage A: project, tap the padded std map (background candidates, std == 0, need nothing else)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float xc = fmaf(cv[2], x2[q], fmaf(cv[1], x1[q], cv[0] * x0[q])) + cv[9];    // :100-101
        const float yc = fmaf(cv[5], x2[q], fmaf(cv[4], x1[q], cv[3] * x0[q])) + cv[10];
        zcam[q] = fmaf(cv[8], x2[q], fmaf(cv[7], x1[q], cv[6] * x0[q])) + cv[11];
        float a = xc / zcam[q], b = yc / zcam[q];      // :107
        a = a * cv[15]; b = b * cv[16];                // :108
        a = a + cv[17]; b = b + cv[18];                // :109
        uu[q] = (P.inv_w != 0.0f ? a * P.inv_w : a / P.img_w) * 2.0f - 1.0f;   // :110
        vv[q] = (P.inv_h != 0.0f ? b * P.inv_h : b / P.img_h) * 2.0f - 1.0f;
        const int sx = nearest_zeros(uu[q] * P.std_sx, Wp), sy = nearest_zeros(vv[q] * P.std_sy, Hp);
        ok[q] = (ib + q < C) && sx >= 0 && sy >= 0;
        sref[q] = ok[q] ? __ldg(std_v + (size_t)sy * Wp + sx) : 0.0f;
      }
      // stage B: depth tap and the depth-difference gate
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        ok[q] = ok[q] && sref[q] != 0.0f;              // bg_mask :123
        const int ix = nearest_border(uu[q], W), iy = nearest_border(vv[q], H);
        dref[q] = ok[q] ? __ldg(depth_v + (size_t)iy * W + ix) : 0.0f;
      }
      bool any_ok = false;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        ok[q] = ok[q] && (fabsf(dref[q] - zcam[q]) < P.depth_diff_max);   // :122
        any_ok = any_ok || ok[q];
      }
      if (!__any_sync(0xffffffffu, any_ok)) continue;
      // stage C: normal tap, facing test; stage D: likelihood
      float n0[4], n1[4], n2[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int nx = nearest_zeros(uu[q], W), ny = nearest_zeros(vv[q], H);
        const bool tap = ok[q] && nx >= 0 && ny >= 0;
        const float* np_ = normal_v + (size_t)ny * W + nx;
        n0[q] = tap ? __ldg(np_) : 0.f; n1[q] = tap ? __ldg(np_ + map_sz) : 0.f; n2[q] = tap ? __ldg(np_ + 2 * map_sz) : 0.f;
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float cosd = cv[12] * n0[q] + cv[13] * n1[q] + cv[14] * n2[q];   // :119
        if (ok[q] && cosd <= 0.0f) {                   // :121
          const float den = sref[q] * sqrt2;
          const float a = erff((zcam[q] + half_step - dref[q]) / den);
          const float b = erff((zcam[q] - half_step - dref[q]) / den);
          const float pv = fabsf(0.5f * (a - b));      // :125-128
          pq[q] = fmaxf(pq[q], pv);                    // :129
        }
      }
    }
    if (P.p_out) {
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (ib + q < C) P.p_out[ray * C + ib + q] = pq[q];
    }
    const unsigned b0 = __ballot_sync(0xffffffffu, pq[0] != 0.0f), b1 = __ballot_sync(0xffffffffu, pq[1] != 0.0f);
    const unsigned b2 = __ballot_sync(0xffffffffu, pq[2] != 0.0f), b3 = __ballot_sync(0xffffffffu, pq[3] != 0.0f);
    int pos = n_nz + __popc(b0 & lt_mask) + __popc(b1 & lt_mask) + __popc(b2 & lt_mask) + __popc(b3 & lt_mask);
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (pq[q] != 0.0f) { lp[pos] = pq[q]; lz[pos] = zq[q]; lidx[pos] = (unsigned short)(ib + q); ++pos; }
    n_nz += __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
  }
  __syncwarp();

  // ---- 2. shortlist: the M = K-G most likely candidates (ties -> lower index), :172-178,190 ----
  const int G = P.G;
  const int M = K - G;
  for (int k = lane; k < K; k += 32) slot[k] = 0.0f;
  __syncwarp();
  for (int e = lane; e < n_nz; e += 32) {
    const float pe = lp[e];
    int rank = 0;
    for (int j = 0; j < n_nz; ++j) {
      const float pj = lp[j];
      rank += (pj > pe) || (pj == pe && j < e);
    }
    if (rank < M) {
      slot[rank] = lz[e];
      if (P.sel_idx_out) P.sel_idx_out[ray * M + rank] = lidx[e];
    }
  }
  if (P.sel_idx_out)
    for (int k = n_nz + lane; k < M; k += 32) P.sel_idx_out[ray * M + k] = -1;
  __syncwarp();

  // ---- 3. occlusion-aware likelihood (sequential product, same order as torch.cumprod), :131-132 ----
  if (lane == 0) {
    float run = 1.0f;
    for (int j = 0; j < n_nz; ++j) {
      const float pj = lp[j];
      lp[j] = (j == 0) ? pj : pj * run;
      run = run * (1.0f - pj);
    }
  }
  __syncwarp();

  // ---- 4. gaussian samples around the weighted mean depth, :181-190 + torch_helpers.py:294-302 ----
  float wsum = 0.f;
  int any_nz = 0;
  for (int e = lane; e < n_nz; e += 32) { const float w = lp[e]; wsum += w; any_nz |= (w != 0.0f); }
  wsum = warp_sum(wsum);
  const bool ray_mask = __any_sync(0xffffffffu, any_nz) != 0;
  if (G > 0 && ray_mask && !P.count_only) {
    float m = 0.f;
    for (int e = lane; e < n_nz; e += 32) m += lz[e] * (lp[e] / wsum);
    m = warp_sum(m);
    float var = 0.f;
    for (int e = lane; e < n_nz; e += 32) { const float dz = lz[e] - m; var += dz * dz * (lp[e] / wsum); }
    const float sd = sqrtf(warp_sum(var));
    for (int g = lane; g < G; g += 32) {
      float n;
      if (P.n_gauss) n = P.n_gauss[(size_t)(P.unpacked ? ray : (long long)P.gauss_off[ray]) * G + g];
      else {
        uint4 r = rng((uint32_t)ray, (uint32_t)(ray >> 32), (uint32_t)g, 1u);
        const float u1 = fmaxf(u01(r.x), 5.9604645e-8f), u2 = u01(r.y);
        n = sqrtf(-2.0f * logf(u1)) * cosf(6.283185307179586f * u2);
      }
      slot[M + g] = n * sd + m;
    }
  }
  __syncwarp();

  // ---- 5. fill-up, :367-397 ----
  int n_zero = 0, n_neg = 0;
  for (int k = lane; k < K; k += 32) { const float v = slot[k]; n_zero += (v == 0.0f); n_neg += (v < 0.0f); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { n_zero += __shfl_xor_sync(0xffffffffu, n_zero, o); n_neg += __shfl_xor_sync(0xffffffffu, n_neg, o); }
  if (P.count_only) {
    // a masked ray's G gaussian draws are non-zero almost surely
    if (lane == 0) {
      if (P.ray_mask_out) P.ray_mask_out[ray] = (G > 0 && ray_mask) ? 1 : 0;
      if (P.n_missing_out) P.n_missing_out[ray] = n_zero - ((G > 0 && ray_mask) ? G : 0);
    }
    return;
  }
  if (lane == 0) {
    if (P.ray_mask_out) P.ray_mask_out[ray] = (G > 0 && ray_mask) ? 1 : 0;
    if (P.n_missing_out) P.n_missing_out[ray] = n_zero;
  }
  // first sort (ascending): zeros end up right after the negative values
  for (int e = lane; e < K; e += 32) {
    const float ve = slot[e];
    int rank = 0;
    for (int j = 0; j < K; ++j) { const float vj = slot[j]; rank += (vj < ve) || (vj == ve && j < e); }
    sorted[rank] = ve;
  }
  __syncwarp();
  if (n_zero > 0) {
    const float mstep = (far - near) / (float)n_zero;         // :388
    for (int i = lane; i < K; i += 32) {
      if (sorted[i] == 0.0f) {
        float u;
        if (P.u_fill) u = P.unpacked ? P.u_fill[(size_t)ray * K + i] : P.u_fill[(size_t)P.fill_off[ray] + (i - n_neg)];
        else { uint4 r = rng((uint32_t)ray, (uint32_t)(ray >> 32), (uint32_t)i, 2u); u = u01(r.x); }
        float zm = near + (float)i * mstep;                   // :389
        zm = zm + u * mstep;                                  // :390
        sorted[i] = zm;
      }
    }
    __syncwarp();
  }
  // second sort
  for (int e = lane; e < K; e += 32) {
    const float ve = sorted[e];
    int rank = 0;
    for (int j = 0; j < K; ++j) { const float vj = sorted[j]; rank += (vj < ve) || (vj == ve && j < e); }
    P.z_out[ray * K + rank] = ve;
  }
}

static size_t sampler_smem_bytes(int C, int K) {
  size_t per_warp = (size_t)C * 10 + (size_t)K * 8 + kMaxViews * 20 * 4 + 16;
  per_warp = (per_warp + 15) / 16 * 16;
  return per_warp * kSamplerWarps;
}

// Exponentially padded depth-std map, closed form of torch_helpers.py:100-122 (see oracle).
__global__ void pad_std_kernel(const float* __restrict__ std, const float* __restrict__ gain_tab, int n_maps, int H,
                               int W, int P, float* __restrict__ out) {
  const int Wp = W + 2 * P, Hp = H + 2 * P;
  const size_t total = (size_t)n_maps * Hp * Wp;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int x = (int)(idx % Wp), y = (int)((idx / Wp) % Hp);
    const size_t m = idx / ((size_t)Wp * Hp);
    const int dx = max(max(P - x, x - (W + P - 1)), 0), dy = max(max(P - y, y - (H + P - 1)), 0);
    const int e = max(max(dx, dy) - 1, 0);
    const int sx = min(max(x - P, 0), W - 1), sy = min(max(y - P, 0), H - 1);
    out[idx] = std[m * H * W + (size_t)sy * W + sx] * gain_tab[e];
  }
}

}  // namespace dnr

// Replaces the per-call rebuild of the padded std map inside SpatialEncoder.index_depth_std
// (image_encoder.py:153-180 -> torch_helpers.py:100-122); gain_tab[e] = exp(e/12*ln2), e in [0,P).
DNR_EXPORT int dnr_pad_depth_std(const float* depth_std, const float* gain_tab, int n_maps, int H, int W, int pad,
                                 float* out, cudaStream_t stream) {
  DNR_REQUIRE(n_maps > 0 && H > 0 && W > 0 && pad >= 0, "dnr_pad_depth_std: bad shape");
  const size_t total = (size_t)n_maps * (H + 2 * pad) * (W + 2 * pad);
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  dnr::pad_std_kernel<<<blocks, 256, 0, stream>>>(depth_std, gain_tab, n_maps, H, W, pad, out);
  DNR_CHECK_LAUNCH("dnr_pad_depth_std");
  return 0;
}

// Replaces NeRFRendererDGS.sample_depthguided + fill_up_uniform_samples (nerf_renderer.py:65-284, 367-397).
DNR_EXPORT int dnr_sampler_fwd(const float* rays, const float* poses, const float* focal, const float* cxy,
                               const float* depth, const float* depth_std_padded, const float* normal,
                               const float* lin_tab, const float* u_coarse, const float* n_gauss, const float* u_fill,
                               const int* gauss_off, const int* fill_off, const float* cand_offset,
                               unsigned long long seed, int SB, int NV,
                               int NR, int H, int W, int C, int K, int G, int pad, float depth_diff_max, float img_w,
                               float img_h, float std_sx, float std_sy, int count_only, int unpacked_noise,
                               float* z_out,
                               int* ray_mask_out, int* n_missing_out, int* sel_idx_out, float* p_out,
                               cudaStream_t stream) {
  DNR_REQUIRE(NV >= 1 && NV <= dnr::kMaxViews, "dnr_sampler_fwd: NV must be in [1,8]");
  DNR_REQUIRE(K >= 1 && K <= 256 && G >= 0 && G <= K, "dnr_sampler_fwd: need 1 <= K <= 256, 0 <= G <= K");
  DNR_REQUIRE(C >= 1 && C <= 4096, "dnr_sampler_fwd: C must be in [1,4096]");
  DNR_REQUIRE((n_gauss == nullptr) || unpacked_noise || gauss_off != nullptr, "dnr_sampler_fwd: n_gauss needs gauss_off");
  DNR_REQUIRE((u_fill == nullptr) || unpacked_noise || fill_off != nullptr, "dnr_sampler_fwd: u_fill needs fill_off");
  const long long n_rays = (long long)SB * NR;
  if (n_rays <= 0) return 0;
  dnr::SamplerParams P;
  P.rays = rays; P.poses = poses; P.focal = focal; P.cxy = cxy; P.depth = depth; P.stdpad = depth_std_padded;
  P.normal = normal; P.lin_tab = lin_tab; P.u_coarse = u_coarse; P.n_gauss = n_gauss; P.u_fill = u_fill;
  P.gauss_off = gauss_off; P.fill_off = fill_off; P.cand_off = cand_offset; P.seed = seed;
  P.SB = SB; P.NV = NV; P.NR = NR; P.H = H; P.W = W; P.C = C; P.K = K; P.G = G; P.P = pad;
  P.depth_diff_max = depth_diff_max; P.img_w = img_w; P.img_h = img_h; P.std_sx = std_sx; P.std_sy = std_sy;
  {
    int ew = 0, eh = 0;
    P.inv_w = (img_w >= 1.0f && frexpf(img_w, &ew) == 0.5f) ? 1.0f / img_w : 0.0f;
    P.inv_h = (img_h >= 1.0f && frexpf(img_h, &eh) == 0.5f) ? 1.0f / img_h : 0.0f;
  }
  P.inv_c_step = (float)(1.0 / (double)C);
  P.count_only = count_only; P.unpacked = unpacked_noise; P.z_out = z_out; P.ray_mask_out = ray_mask_out; P.n_missing_out = n_missing_out;
  P.sel_idx_out = sel_idx_out; P.p_out = p_out;
  const size_t smem = dnr::sampler_smem_bytes(C, K);
  dnr::DeviceInfo* di = dnr::device_info();
  DNR_REQUIRE(di != nullptr, "dnr_sampler_fwd: no current CUDA device");
  if (smem > di->sampler_smem) {
    DNR_CHECK(cudaFuncSetAttribute(dnr::sampler_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
              "dnr_sampler_fwd/attr");
    di->sampler_smem = smem;
  }
  const int blocks = (int)((n_rays + dnr::kSamplerWarps - 1) / dnr::kSamplerWarps);
  dnr::sampler_kernel<<<blocks, dnr::kSamplerWarps * 32, smem, stream>>>(P);
  DNR_CHECK_LAUNCH("dnr_sampler_fwd");
  return 0;
}
