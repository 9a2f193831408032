// NOVEL renderer variant (reference src/models/novel/nerf_novel_renderer.py, src/models/novel/novel_pixelnerf.py):
// nearest-vertex deformation of every sampler CANDIDATE (nerf_novel_renderer.py:104-107) and of every SAMPLE point
// (twice: tgt_in_offsets / tgt_gen_offsets, :441-442) -- K = 1 nearest neighbour against the target mesh
// (pytorch3d.ops.knn_points: squared L2, first minimum), offset gather, add.
//
// A frame has SB*NR*C = 65.5 M candidates against ~5 000 vertices: brute force (csrc/deform.cu, 3e12 pairs/s) would cost
// 110 ms per FaceScape frame, three times the whole DINER render.  Here the vertices arrive sorted along a Morton curve
// in groups of 32 with a bounding sphere per group, and 8 consecutive groups form a super-group with an axis-aligned box
// (diner_b200/novel.py builds both per mesh, ~0.1 ms).  A query scans the super-group whose box is nearest first, which
// gives a real distance to prune with, then visits only the super-groups / groups / runs of 8 vertices inside a group whose
// lower bound (box distance, |x - c| - r) does not exceed the best distance so far (a candidate far from the mesh sees a
// whole cap of the surface at nearly the same distance: the 8-vertex spheres cut the vertices it scans from ~900 to ~150).  The pruning is conservative (every vertex that can be the
// minimum or tie with it is scanned) and the scanned distances use the brute-force kernel's arithmetic (x, y, z accumulated with fused
// multiply-adds) with ties going to the lowest ORIGINAL index, so the result equals the brute force bit for bit.
//
// The candidate positions are recomputed exactly as csrc/sampler.cu computes them (same jitter, same op order; this file
// is compiled with --fmad=false like the sampler), so the offsets land on the very candidates the sampler scores.
#include "common.cuh"
#include <stdlib.h>

namespace dnr {

struct NovelParams {
  const float* rays;        // [SB*NR, 8]
  const float* z;           // samples mode: [SB*NR, K]
  const float* lin_tab;     // candidates mode: [C]
  const float* u_coarse;    // candidates mode: [SB*NR, C] or null (Philox)
  unsigned long long seed;
  const float4* verts;      // [SB][G*32] Morton-sorted (x, y, z, original index as int bits); padding: +huge, INT_MAX
  const float4* spheres;    // [SB][G] (cx, cy, cz, r)
  const float4* supers;     // [SB][S][2] (lo.xyz, -), (hi.xyz, -): boxes of 8 consecutive groups, S = ceil(G / 8)
  const float4* subs;       // [SB][G][4] (cx, cy, cz, r): spheres of the four runs of 8 vertices inside a group, or null
  const float* off_a;       // [SB][V][3] offsets indexed by ORIGINAL vertex index (tgt_in_offsets)
  const float* off_b;       // samples mode: tgt_gen_offsets, or null
  int SB, NR, per_ray, V, G, S;
  int tile;                 // candidates mode: warps take 8 candidates x 4 neighbouring rays where that is more compact
  float inv_c_step;
  float* out_a;             // candidates: offsets [SB*NR, C, 3]; samples: deformed points [SB*NR*K, 3] (in)
  float* out_b;             // samples: deformed points (gen), or null
  int* idx_out;             // optional nearest-vertex index per query
};

constexpr int kNovelThreads = 256;

template <bool CANDIDATES>
__global__ void __launch_bounds__(kNovelThreads) novel_deform_kernel(const NovelParams P) {
  extern __shared__ float4 sph[];                      // [G] spheres, then [S][2] boxes
  float4* box = sph + P.G;
  const int sb = blockIdx.y;
  const float4* sp = P.spheres + (size_t)sb * P.G;
  const float4* bx = P.supers + (size_t)sb * P.S * 2;
  for (int g = threadIdx.x; g < P.G; g += blockDim.x) sph[g] = sp[g];
  for (int g = threadIdx.x; g < 2 * P.S; g += blockDim.x) box[g] = bx[g];
  __syncthreads();
  const long long per_scene = (long long)P.NR * P.per_ray;
  long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;           // query inside the scene
  if (CANDIDATES && P.tile) {
    // A warp's union of visited vertex runs grows with the extent of its 32 queries.  32 consecutive candidates of ONE ray
    // span 31 steps of (far - near) / C; 8 candidates of 4 NEIGHBOURING rays (adjacent pixels: the callers hand over
    // frames, rows of frames or patches) span 7 steps by 3 pixel footprints -- a third of the extent.  Rays are taken in
    // groups of four; a group whose rays are not neighbours keeps the plain order (one decision per group, from its first
    // and last ray, so every warp that works on the group sees the same mapping; the results do not depend on it).
    const long long per_group = 4LL * P.per_ray;                             // a multiple of 32 (host: per_ray % 8 == 0)
    const long long r0 = q / per_group * 4;
    const int rem = (int)(q % per_group);
    if (r0 >= P.NR) return;
    const long long r3 = min(r0 + 3, (long long)P.NR - 1);
    const float* ra = P.rays + ((long long)sb * P.NR + r0) * 8;
    const float* rb = P.rays + ((long long)sb * P.NR + r3) * 8;
    const float la = ra[7], lb = rb[7];                                      // at the far end, where neighbours are farthest
    const float e0 = (rb[0] + lb * rb[3]) - (ra[0] + la * ra[3]), e1 = (rb[1] + lb * rb[4]) - (ra[1] + la * ra[4]);
    const float e2 = (rb[2] + lb * rb[5]) - (ra[2] + la * ra[5]);
    const float step = (ra[7] - ra[6]) * P.inv_c_step * sqrtf(ra[3] * ra[3] + ra[4] * ra[4] + ra[5] * ra[5]);
    const float lateral2 = e0 * e0 + e1 * e1 + e2 * e2;
    long long rr; int ii;
    if (lateral2 + 49.f * step * step < 961.f * step * step) {               // hypot(lateral, 7 steps) < 31 steps
      const int l = rem & 31;
      ii = (rem >> 5) * 8 + (l & 7);
      rr = r0 + (l >> 3);
    } else {
      rr = r0 + rem / P.per_ray;
      ii = rem % P.per_ray;
    }
    if (rr >= P.NR) return;
    q = rr * P.per_ray + ii;
  }
  if (q >= per_scene) return;
  const long long ray = (long long)sb * P.NR + q / P.per_ray;
  const int i = (int)(q % P.per_ray);
  const float* rp = P.rays + ray * 8;
  const float o0 = rp[0], o1 = rp[1], o2 = rp[2], d0 = rp[3], d1 = rp[4], d2 = rp[5];
  float zc;
  if (CANDIDATES) {
    // sampler.cu, stage 1: u from the recorded draws or from Philox counter (ray, i >> 2), component i & 3
    float u;
    if (P.u_coarse) u = P.u_coarse[ray * P.per_ray + i];
    else {
      Philox rng(P.seed);
      const uint4 r = rng((uint32_t)ray, (uint32_t)(ray >> 32), (uint32_t)(i >> 2), 0u);
      const int c = i & 3;
      u = u01(c == 0 ? r.x : c == 1 ? r.y : c == 2 ? r.z : r.w);
    }
    const float near = rp[6], far = rp[7];
    const float s = P.lin_tab[i] + u * P.inv_c_step;       // nerf_novel_renderer.py:70-73
    zc = near * (1.0f - s) + far * s;                      // :76
  } else {
    zc = P.z[ray * P.per_ray + i];
  }
  const float x0 = o0 + zc * d0, x1 = o1 + zc * d1, x2 = o2 + zc * d2;      // :93 / :416
  float best = 3.0e38f;                                    // squared distance, the brute-force kernel's arithmetic
  int best_i = 0x7fffffff;
  const float4* vs = P.verts + (size_t)sb * P.G * 32;
  const float4* sub = P.subs ? P.subs + (size_t)sb * P.G * 4 : nullptr;
  // a lower bound may be a rounding above the true distance: prune only beyond a relative + absolute slack, and never on
  // equality (ties go to the lowest original index, so every equal-distance vertex must be seen)
  auto beyond = [&](float lb2) { return lb2 > best * 1.00002f + 1e-12f; };
  // sphere test without a square root per sphere: |x - c| - r > sqrt(best)  <=>  |x - c|^2 > (sqrt(best) + r)^2.  `reach`
  // is sqrt(best) with the same slack as `beyond`, refreshed after every scanned run of vertices; the radii are inflated by
  // 1e-5 relative when they are built (diner_b200/novel.py), which covers the roundings of the large terms
  float reach = 3.0e18f;
  auto outside = [&](const float4& sp) {
    const float a = x0 - sp.x, b = x1 - sp.y, c = x2 - sp.z;
    const float t = reach + sp.w;
    return fmaf(c, c, fmaf(b, b, a * a)) > t * t;
  };
  auto box_lb2 = [&](int sidx) {
    const float4 lo = box[2 * sidx], hi = box[2 * sidx + 1];
    const float a = fmaxf(fmaxf(lo.x - x0, x0 - hi.x), 0.f), b = fmaxf(fmaxf(lo.y - x1, x1 - hi.y), 0.f);
    const float c = fmaxf(fmaxf(lo.z - x2, x2 - hi.z), 0.f);
    return a * a + b * b + c * c;
  };
  auto scan_super = [&](int sidx) {
    const int g1 = min(P.G, sidx * 8 + 8);
    for (int g = sidx * 8; g < g1; ++g) {
      if (outside(sph[g])) continue;
      const float4* vg = vs + (size_t)g * 32;
      const float4* sg4 = sub ? sub + (size_t)g * 4 : nullptr;
#pragma unroll 1
      for (int q = 0; q < 4; ++q) {
        if (sg4) {                                           // third level: the run of 8 vertices
          if (outside(__ldg(sg4 + q))) continue;
        }
        const float before = best;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 v = __ldg(vg + q * 8 + j);
          const float e0 = x0 - v.x, e1 = x1 - v.y, e2 = x2 - v.z;
          const float dd = fmaf(e2, e2, fmaf(e1, e1, e0 * e0));        // the brute-force kernel's arithmetic (deform.cu)
          const int vi = __float_as_int(v.w);
          if (dd < best || (dd == best && vi < best_i)) { best = dd; best_i = vi; }
        }
        if (best < before) reach = sqrtf(best * 1.00002f + 1e-12f) * 1.000001f;
      }
    }
  };
  // seed: the super-group whose box is nearest
  int s0 = 0;
  float lb0 = 3.0e38f;
  for (int sidx = 0; sidx < P.S; ++sidx) {
    const float lb = box_lb2(sidx);
    if (lb < lb0) { lb0 = lb; s0 = sidx; }
  }
  scan_super(s0);
  for (int sidx = 0; sidx < P.S; ++sidx) {
    if (sidx == s0 || beyond(box_lb2(sidx))) continue;
    scan_super(sidx);
  }
  const size_t qq = (size_t)sb * per_scene + q;
  if (P.idx_out) P.idx_out[qq] = best_i;
  const float* oa = P.off_a + ((size_t)sb * P.V + best_i) * 3;
  if (CANDIDATES) {
    P.out_a[qq * 3 + 0] = oa[0]; P.out_a[qq * 3 + 1] = oa[1]; P.out_a[qq * 3 + 2] = oa[2];
  } else {
    P.out_a[qq * 3 + 0] = x0 + oa[0]; P.out_a[qq * 3 + 1] = x1 + oa[1]; P.out_a[qq * 3 + 2] = x2 + oa[2];     // :48-49
    if (P.out_b) {
      const float* ob = P.off_b + ((size_t)sb * P.V + best_i) * 3;
      P.out_b[qq * 3 + 0] = x0 + ob[0]; P.out_b[qq * 3 + 1] = x1 + ob[1]; P.out_b[qq * 3 + 2] = x2 + ob[2];
    }
  }
}

}  // namespace dnr

// Offsets of the sampler's candidates (mode = candidates: per_ray = C, z from lin_tab + jitter exactly as dnr_sampler_fwd
// draws it; out_a = offsets [SB*NR, C, 3] for dnr_sampler_fwd's cand_offset) or deformed sample points (mode = samples:
// per_ray = K, z given; out_a / out_b = points + in / gen offsets, [SB*NR*K, 3]) of the NOVEL renderer variant
// (nerf_novel_renderer.py:40-50,104-107,441-442).  verts_sorted [SB][G*32] float4 (x, y, z, original index bits) and
// spheres [SB][G] float4 (centre, radius) and supers [SB][ceil(G/8)][2] float4 (box lo, hi of 8 consecutive groups) come
// from diner_b200.novel.build_vertex_groups, as does subs [SB][G][4] float4 (spheres of the four runs of 8 vertices of a
// group; optional third pruning level, null = scan whole groups); offsets are indexed by original vertex index [SB][V][3].
DNR_EXPORT int dnr_novel_deform(const float* rays, const float* z, const float* lin_tab, const float* u_coarse,
                                unsigned long long seed, const void* verts_sorted, const void* spheres,
                                const void* supers, const void* subs, const float* offsets_a, const float* offsets_b, int SB, int NR,
                                int per_ray, int V, int G,
                                int candidates, float* out_a, float* out_b, int* idx_out, cudaStream_t stream) {
  DNR_REQUIRE(rays && verts_sorted && spheres && supers && offsets_a && out_a, "dnr_novel_deform: null argument");
  DNR_REQUIRE(SB >= 1 && NR >= 0 && per_ray >= 1 && V >= 1 && G >= 1 && G <= 2048, "dnr_novel_deform: bad sizes (G <= 2048)");
  DNR_REQUIRE(candidates ? (lin_tab != nullptr) : (z != nullptr), "dnr_novel_deform: candidates need lin_tab, samples need z");
  DNR_REQUIRE(SB <= 65535, "dnr_novel_deform: at most 65535 scenes per call");
  if (NR == 0) return 0;
  dnr::NovelParams P;
  P.rays = rays; P.z = z; P.lin_tab = lin_tab; P.u_coarse = u_coarse; P.seed = seed;
  P.verts = (const float4*)verts_sorted; P.spheres = (const float4*)spheres; P.supers = (const float4*)supers; P.subs = (const float4*)subs; P.off_a = offsets_a; P.off_b = offsets_b;
  P.SB = SB; P.NR = NR; P.per_ray = per_ray; P.V = V; P.G = G; P.S = (G + 7) / 8;
  P.inv_c_step = (float)(1.0 / (double)per_ray);
  P.out_a = out_a; P.out_b = out_b; P.idx_out = idx_out;
  const char* tile_env = getenv("DNR_NOVEL_TILE");                          // 0: plain order (A/B and tests; same results)
  P.tile = (candidates && per_ray % 8 == 0 && !(tile_env && tile_env[0] == '0')) ? 1 : 0;
  const long long per_scene = (long long)(P.tile ? (NR + 3) / 4 * 4 : NR) * per_ray;      // whole groups of four rays
  dim3 grid((unsigned)((per_scene + dnr::kNovelThreads - 1) / dnr::kNovelThreads), SB);
  const size_t smem = ((size_t)G + 2 * (size_t)((G + 7) / 8)) * sizeof(float4);
  if (candidates) dnr::novel_deform_kernel<true><<<grid, dnr::kNovelThreads, smem, stream>>>(P);
  else dnr::novel_deform_kernel<false><<<grid, dnr::kNovelThreads, smem, stream>>>(P);
  DNR_CHECK_LAUNCH("dnr_novel_deform");
  return 0;
}
