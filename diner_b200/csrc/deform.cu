// Nearest-vertex deformation of sample points (first piece): the NOVEL renderer variants move every
// candidate / sample point by the offset of its nearest target-mesh vertex before projecting it into the source views
// (reference src/models/novel/nerf_novel_renderer.py:40-50: pytorch3d.ops.knn_points(points, vertices, K=1) + a gather of
// offsets + an add).  pytorch3d is not installed in the build image, so this kernel restates the published brute-force
// K = 1 search: squared L2 distance accumulated over x, y, z in fp32, strict `<` against the running minimum (the lowest
// vertex index wins a tie).  Parity with the reference variant is UNPINNED (it cannot be imported here); the checker is the
// numpy restatement in oracle/deform_oracle.py.
//
// One thread per point, vertices staged through shared memory in tiles of 1024 (float4 x, y, z, -): compute-bound,
// ~9 instructions per (point, vertex) pair; the points stream once (12 B in, 12 + 4 B out per point).
#include "common.cuh"

namespace dnr {

constexpr int kDfThreads = 256;
constexpr int kDfTile = 1024;

__global__ void __launch_bounds__(kDfThreads) deform_points_kernel(const float* __restrict__ points,     // [SB,B,3]
                                                                   const float* __restrict__ vertices,   // [SB,V,3]
                                                                   const float* __restrict__ offsets,    // [SB,V,3]
                                                                   long long B, int V,
                                                                   float* __restrict__ out,              // [SB,B,3]
                                                                   int* __restrict__ idx_out,            // [SB,B] or null
                                                                   float* __restrict__ dist_out) {       // [SB,B] or null
  __shared__ float4 tile[kDfTile];
  const int sb = blockIdx.y;
  const long long i = (long long)blockIdx.x * kDfThreads + threadIdx.x;
  const bool live = i < B;
  float px = 0.f, py = 0.f, pz = 0.f;
  if (live) {
    const float* p = points + ((size_t)sb * B + i) * 3;
    px = p[0]; py = p[1]; pz = p[2];
  }
  const float* vb = vertices + (size_t)sb * V * 3;
  float best = 3.402823466e38f;
  int best_i = 0;
  for (int v0 = 0; v0 < V; v0 += kDfTile) {
    const int n = min(kDfTile, V - v0);
    __syncthreads();
    for (int t = threadIdx.x; t < n; t += kDfThreads) {
      const float* q = vb + (size_t)(v0 + t) * 3;
      tile[t] = make_float4(q[0], q[1], q[2], 0.f);
    }
    __syncthreads();
    if (live) {
#pragma unroll 8
      for (int t = 0; t < n; ++t) {
        const float4 q = tile[t];
        const float dx = px - q.x, dy = py - q.y, dz = pz - q.z;
        float d = dx * dx;
        d = fmaf(dy, dy, d);
        d = fmaf(dz, dz, d);
        if (d < best) { best = d; best_i = v0 + t; }
      }
    }
  }
  if (!live) return;
  const float* o = offsets + ((size_t)sb * V + best_i) * 3;
  float* r = out + ((size_t)sb * B + i) * 3;
  r[0] = px + o[0]; r[1] = py + o[1]; r[2] = pz + o[2];
  if (idx_out) idx_out[(size_t)sb * B + i] = best_i;
  if (dist_out) dist_out[(size_t)sb * B + i] = best;
}

}  // namespace dnr

// deformed[sb, i] = points[sb, i] + offsets[sb, argmin_v |points[sb, i] - vertices[sb, v]|^2]; idx_out / dist_out (nullable)
// receive the vertex index and the squared distance.  Replaces NeRFRendererDGS.deform_points of the NOVEL variants
// (src/models/novel/nerf_novel_renderer.py:40-50).
DNR_EXPORT int dnr_deform_points(const float* points, const float* vertices, const float* offsets, int SB, long long B,
                                 int V, float* out, int* idx_out, float* dist_out, cudaStream_t stream) {
  DNR_REQUIRE(SB >= 0 && B >= 0 && SB <= 65535, "dnr_deform_points: bad batch sizes");
  if (SB == 0 || B == 0) return 0;
  DNR_REQUIRE(V >= 1, "dnr_deform_points: needs at least one vertex");
  DNR_REQUIRE(points && vertices && offsets && out, "dnr_deform_points: null argument");
  dim3 grid((unsigned)((B + dnr::kDfThreads - 1) / dnr::kDfThreads), (unsigned)SB);
  dnr::deform_points_kernel<<<grid, dnr::kDfThreads, 0, stream>>>(points, vertices, offsets, B, V, out, idx_out, dist_out);
  DNR_CHECK_LAUNCH("dnr_deform_points");
  return 0;
}
