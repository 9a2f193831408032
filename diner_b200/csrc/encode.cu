// Field input encoding: every sample point is projected into the NV source views; per (sample, view)
// row the kernel gathers the 512-channel encoder feature (bilinear, border) from the NHWC bf16 latent,
// taps the predicted depth (nearest, border) and builds the 55-wide geometric code
// [x_cam 3 | posenc 36 | dir_cam 3 | depth-dist 1 | posenc 12].  Restates reference
// pixelnerf.py:91-128, image_encoder.py:97-151, positional_encoding.py:33-53.
//
// Rows are written straight into the MLP's packed A-operand layouts (PK<128>, see common.cuh), staged
// through shared memory so that every global store is a full 128-byte line.  The backward kernel
// scatters d(feature) back into the NHWC fp32 latent gradient with vector reductions.
//
// Algorithmic bytes per (sample, view) row, fwd: 4 taps * 512 ch * 2 B + 4 B depth read,
// 512*2 + 64*2 B written.
#include "common.cuh"

namespace dnr {

struct EncodeParams {
  const float* rays;     // [SB*NR, 8]
  const float* z;        // [SB*NR, K]
  const float* poses;    // [SB,NV,4,4]
  const float* focal;    // [SB,NV,2]
  const float* cxy;      // [SB,NV,2]
  const float* depth;    // [SB,NV,H,W]
  const __nv_bfloat16* latent;  // [SB,NV,Hl,Wl,L] channels-last
  const float* freqs;    // [2*NF] (f0,f0,f1,f1,...)
  const float* phases;   // [2*NF] (0,pi/2,...)
  int SB, NV, NR, K, H, W, Hl, Wl, NF;
  float img_w, img_h, lat_sx, lat_sy;
  long long n_samples;   // rays_in_chunk*K
  long long ray_offset;  // index of the chunk's first ray in the (SB*NR) batch
  long long n_rays;      // rays in this chunk
  unsigned char* lat_pk; // PK<128> [rows_pad x 512] bf16
  unsigned char* in_pk;  // PK<128> [rows_pad x 64] bf16
  float* d_latent;       // bwd: [SB,NV,Hl,Wl,L] fp32 (accumulated into)
  const float* d_lat_xr; // bwd: XR [rows_pad x 512] fp32
  // NOVEL variant (src/models/novel/novel_pixelnerf.py:143-196), all null / 0 otherwise
  const float* pts_in;   // [n_samples, 3] `in`-deformed sample points: replace o + z d in the projection chain
  const float* pts_gen;  // [n_samples, 3] `gen`-deformed sample points, projected by the ONE generic camera
  const float* gen_poses;   // [SB,1,4,4]
  const float* gen_focal;   // [SB,1,2]
  const float* gen_cxy;     // [SB,1,2]
  const __nv_bfloat16* gen_latent;  // [Hg,Wg,L] channels-last, shared by all scenes and views
  int Hg, Wg;
  float gen_img_w, gen_img_h, gen_sx, gen_sy;
  float* d_gen_latent;   // bwd, NOVEL: [Hg,Wg,L] fp32 (accumulated into; one map for all scenes and views)
  const float* d_gen_xr; // bwd, NOVEL: XR gradient rows scattered into d_gen_latent (null: d_lat_xr)
  // NOVEL_PE variant (src/models/novel_pe/pe_novel_pixelnerf.py:200-275), all null / 0 otherwise
  const float* pe_src;   // [SB,NV,3,Hp,Wp] fp32 positional maps of the source views, tapped at the source uv (:133-166,260)
  const float* pe_tgt;   // [SB,1,3,Ht,Wt] fp32 positional map of the target view, tapped at the target-camera uv of the
                         // UNDEFORMED point (:168-198,223-226,248-251,262)
  const float* tgt_poses;   // [SB,1,4,4]
  const float* tgt_focal;   // [SB,1,2]
  const float* tgt_cxy;     // [SB,1,2]
  int Hp, Wp, Ht, Wt;
  float pe_sx, pe_sy, tgt_img_w, tgt_img_h, tgt_sx, tgt_sy;
  unsigned char* gen_pk;  // PK<128> [rows_pad x 512] bf16: the gen_latent taps, kept apart from lat_pk (the deformation
                          // layer sits between the encoder feature and the sum, :265-268)
};

struct RowGeom {
  float xc, yc, zc, dx, dy, dz, dd;
  float u, v;   // normalised image coordinates in the source view, before the feature-padding rescale
  int x0, y0, x1, y1;
  float wx, wy;
  size_t map;   // (sb*NV+v)
};

// projection chain shared by fwd and bwd (pixelnerf.py:91-93,99-101,105-116)
__device__ __forceinline__ void row_geometry(const EncodeParams& P, long long sample, int view, RowGeom& g) {
  const long long ray = sample / P.K;
  const int sb = (int)((ray + P.ray_offset) / P.NR);
  const float* rp = P.rays + ray * 8;
  float x0, x1, x2;
  if (P.pts_in) {          // NOVEL variant: the (deformed) point is given
    x0 = P.pts_in[sample * 3 + 0]; x1 = P.pts_in[sample * 3 + 1]; x2 = P.pts_in[sample * 3 + 2];
  } else {
    const float zs = P.z[sample];
    x0 = __fadd_rn(rp[0], __fmul_rn(zs, rp[3]));
    x1 = __fadd_rn(rp[1], __fmul_rn(zs, rp[4]));
    x2 = __fadd_rn(rp[2], __fmul_rn(zs, rp[5]));
  }
  g.map = (size_t)sb * P.NV + view;
  const float* E = P.poses + g.map * 16;
  g.xc = __fadd_rn(fmaf(E[2], x2, fmaf(E[1], x1, __fmul_rn(E[0], x0))), E[3]);
  g.yc = __fadd_rn(fmaf(E[6], x2, fmaf(E[5], x1, __fmul_rn(E[4], x0))), E[7]);
  g.zc = __fadd_rn(fmaf(E[10], x2, fmaf(E[9], x1, __fmul_rn(E[8], x0))), E[11]);
  g.dx = fmaf(E[2], rp[5], fmaf(E[1], rp[4], __fmul_rn(E[0], rp[3])));
  g.dy = fmaf(E[6], rp[5], fmaf(E[5], rp[4], __fmul_rn(E[4], rp[3])));
  g.dz = fmaf(E[10], rp[5], fmaf(E[9], rp[4], __fmul_rn(E[8], rp[3])));
  float u = __fdiv_rn(g.xc, g.zc), v = __fdiv_rn(g.yc, g.zc);
  u = __fadd_rn(__fmul_rn(u, P.focal[g.map * 2 + 0]), P.cxy[g.map * 2 + 0]);
  v = __fadd_rn(__fmul_rn(v, P.focal[g.map * 2 + 1]), P.cxy[g.map * 2 + 1]);
  u = __fsub_rn(__fmul_rn(__fdiv_rn(u, P.img_w), 2.0f), 1.0f);
  v = __fsub_rn(__fmul_rn(__fdiv_rn(v, P.img_h), 2.0f), 1.0f);
  g.u = u; g.v = v;
  // nearest depth tap, border padding
  float px = __fdiv_rn(__fsub_rn(__fmul_rn(__fadd_rn(u, 1.0f), (float)P.W), 1.0f), 2.0f);
  float py = __fdiv_rn(__fsub_rn(__fmul_rn(__fadd_rn(v, 1.0f), (float)P.H), 1.0f), 2.0f);
  px = fminf(fmaxf(px, 0.0f), (float)(P.W - 1));
  py = fminf(fmaxf(py, 0.0f), (float)(P.H - 1));
  const int ix = (int)nearbyintf(px), iy = (int)nearbyintf(py);
  const float dref = __ldg(P.depth + g.map * (size_t)P.H * P.W + (size_t)iy * P.W + ix);
  g.dd = __fsub_rn(dref, g.zc);
  // bilinear latent tap, border padding; feature-padding rescale (image_encoder.py:113-123)
  const float ul = __fmul_rn(u, P.lat_sx), vl = __fmul_rn(v, P.lat_sy);
  float fx = fmaf(__fadd_rn(ul, 1.0f), (float)P.Wl, -1.0f) * 0.5f;   // torch's bilinear path fuses here
  float fy = fmaf(__fadd_rn(vl, 1.0f), (float)P.Hl, -1.0f) * 0.5f;
  fx = fminf(fmaxf(fx, 0.0f), (float)(P.Wl - 1));
  fy = fminf(fmaxf(fy, 0.0f), (float)(P.Hl - 1));
  const float flx = floorf(fx), fly = floorf(fy);
  g.x0 = (int)flx; g.y0 = (int)fly;
  g.x1 = min(g.x0 + 1, P.Wl - 1); g.y1 = min(g.y0 + 1, P.Hl - 1);
  g.wx = fx - flx; g.wy = fy - fly;     // weight of the +1 neighbour (0 when it is out of range)
}

// NOVEL variant: bilinear footprint of the `gen`-deformed point in the learned map (novel_pixelnerf.py:163-165,186-189,
// index :108-141): one generic camera per scene, the same for every source view.
struct GenTaps {
  uint32_t p00, p01, p10, p11;      // element offsets inside gen_latent
  float w00, w01, w10, w11;
};
__device__ __forceinline__ void gen_geometry(const EncodeParams& P, long long sample, int sb, GenTaps& t) {
  const float x0 = P.pts_gen[sample * 3 + 0], x1 = P.pts_gen[sample * 3 + 1], x2 = P.pts_gen[sample * 3 + 2];
  const float* E = P.gen_poses + (size_t)sb * 16;
  const float xc = __fadd_rn(fmaf(E[2], x2, fmaf(E[1], x1, __fmul_rn(E[0], x0))), E[3]);
  const float yc = __fadd_rn(fmaf(E[6], x2, fmaf(E[5], x1, __fmul_rn(E[4], x0))), E[7]);
  const float zc = __fadd_rn(fmaf(E[10], x2, fmaf(E[9], x1, __fmul_rn(E[8], x0))), E[11]);
  float u = __fdiv_rn(xc, zc), v = __fdiv_rn(yc, zc);
  u = __fadd_rn(__fmul_rn(u, P.gen_focal[sb * 2 + 0]), P.gen_cxy[sb * 2 + 0]);
  v = __fadd_rn(__fmul_rn(v, P.gen_focal[sb * 2 + 1]), P.gen_cxy[sb * 2 + 1]);
  u = __fsub_rn(__fmul_rn(__fdiv_rn(u, P.gen_img_w), 2.0f), 1.0f);
  v = __fsub_rn(__fmul_rn(__fdiv_rn(v, P.gen_img_h), 2.0f), 1.0f);
  const float ul = __fmul_rn(u, P.gen_sx), vl = __fmul_rn(v, P.gen_sy);
  float fx = fmaf(__fadd_rn(ul, 1.0f), (float)P.Wg, -1.0f) * 0.5f;
  float fy = fmaf(__fadd_rn(vl, 1.0f), (float)P.Hg, -1.0f) * 0.5f;
  fx = fminf(fmaxf(fx, 0.0f), (float)(P.Wg - 1));
  fy = fminf(fmaxf(fy, 0.0f), (float)(P.Hg - 1));
  const float flx = floorf(fx), fly = floorf(fy);
  const int gx0 = (int)flx, gy0 = (int)fly;
  const int gx1 = min(gx0 + 1, P.Wg - 1), gy1 = min(gy0 + 1, P.Hg - 1);
  const float wx = fx - flx, wy = fy - fly;
  t.p00 = (uint32_t)(((size_t)gy0 * P.Wg + gx0) * 512); t.p01 = (uint32_t)(((size_t)gy0 * P.Wg + gx1) * 512);
  t.p10 = (uint32_t)(((size_t)gy1 * P.Wg + gx0) * 512); t.p11 = (uint32_t)(((size_t)gy1 * P.Wg + gx1) * 512);
  t.w00 = (1.f - wx) * (1.f - wy); t.w01 = wx * (1.f - wy); t.w10 = (1.f - wx) * wy; t.w11 = wx * wy;
}

// NOVEL_PE: bilinear tap (border padding, align_corners = False) of a 3-channel fp32 map [3][H][W] at normalised (u, v)
// after the feature-padding rescale (pe_novel_pixelnerf.py:150-163); one lane per row
__device__ __forceinline__ void tap3_bilinear(const float* map, int H, int W, float u, float v, float* out) {
  float fx = fmaf(__fadd_rn(u, 1.0f), (float)W, -1.0f) * 0.5f;
  float fy = fmaf(__fadd_rn(v, 1.0f), (float)H, -1.0f) * 0.5f;
  fx = fminf(fmaxf(fx, 0.0f), (float)(W - 1));
  fy = fminf(fmaxf(fy, 0.0f), (float)(H - 1));
  const float flx = floorf(fx), fly = floorf(fy);
  const int x0 = (int)flx, y0 = (int)fly;
  const int x1 = min(x0 + 1, W - 1), y1 = min(y0 + 1, H - 1);
  const float wx = fx - flx, wy = fy - fly;
  const float w00 = (1.f - wx) * (1.f - wy), w01 = wx * (1.f - wy), w10 = (1.f - wx) * wy, w11 = wx * wy;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float* m = map + (size_t)c * H * W;
    out[c] = fmaf(__ldg(m + (size_t)y1 * W + x1), w11, fmaf(__ldg(m + (size_t)y1 * W + x0), w10,
             fmaf(__ldg(m + (size_t)y0 * W + x1), w01, __ldg(m + (size_t)y0 * W + x0) * w00)));
  }
}
// NOVEL_PE: the six positional values of a row -- the source view's map at the row's own uv, the target view's map at
// the projection of the undeformed sample point o + z d by the target camera (pe_novel_pixelnerf.py:223-226,248-262)
__device__ __forceinline__ void pe_values(const EncodeParams& P, long long sample, const RowGeom& g, float* out) {
  tap3_bilinear(P.pe_src + g.map * 3 * (size_t)P.Hp * P.Wp, P.Hp, P.Wp, __fmul_rn(g.u, P.pe_sx), __fmul_rn(g.v, P.pe_sy), out);
  const long long ray = sample / P.K;
  const int sb = (int)(g.map / P.NV);
  const float* rp = P.rays + ray * 8;
  const float zs = P.z[sample];
  const float x0 = __fadd_rn(rp[0], __fmul_rn(zs, rp[3])), x1 = __fadd_rn(rp[1], __fmul_rn(zs, rp[4]));
  const float x2 = __fadd_rn(rp[2], __fmul_rn(zs, rp[5]));
  const float* E = P.tgt_poses + (size_t)sb * 16;
  const float xc = __fadd_rn(fmaf(E[2], x2, fmaf(E[1], x1, __fmul_rn(E[0], x0))), E[3]);
  const float yc = __fadd_rn(fmaf(E[6], x2, fmaf(E[5], x1, __fmul_rn(E[4], x0))), E[7]);
  const float zc = __fadd_rn(fmaf(E[10], x2, fmaf(E[9], x1, __fmul_rn(E[8], x0))), E[11]);
  float u = __fdiv_rn(xc, zc), v = __fdiv_rn(yc, zc);
  u = __fadd_rn(__fmul_rn(u, P.tgt_focal[sb * 2 + 0]), P.tgt_cxy[sb * 2 + 0]);
  v = __fadd_rn(__fmul_rn(v, P.tgt_focal[sb * 2 + 1]), P.tgt_cxy[sb * 2 + 1]);
  u = __fsub_rn(__fmul_rn(__fdiv_rn(u, P.tgt_img_w), 2.0f), 1.0f);
  v = __fsub_rn(__fmul_rn(__fdiv_rn(v, P.tgt_img_h), 2.0f), 1.0f);
  tap3_bilinear(P.pe_tgt + (size_t)sb * 3 * P.Ht * P.Wt, P.Ht, P.Wt, __fmul_rn(u, P.tgt_sx), __fmul_rn(v, P.tgt_sy), out + 3);
}

// packed fp32 pairs (Blackwell FMUL2 / FFMA2): the two channels of a bf16x2 word travel together
__device__ __forceinline__ uint64_t splat_f32x2(float w) {
  uint64_t r; asm("mov.b64 %0, {%1, %1};" : "=l"(r) : "f"(w)); return r;
}
__device__ __forceinline__ uint64_t bf16x2_to_f32x2(uint32_t v) {
  uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(v << 16), "r"(v & 0xffff0000u)); return r;
}
__device__ __forceinline__ uint64_t mul_f32x2(uint64_t a, uint64_t b) {
  uint64_t r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r;
}
__device__ __forceinline__ uint64_t fma_f32x2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r;
}

constexpr int kEncWarps = 4;
// staging of one 32-row slab in the PK<128> order: [feature chunk of 8][32 rows] x 16 B, i.e. the 512 contiguous bytes a
// slab owns of every chunk in global memory; + 16 B pad per chunk -> conflict-free 8-byte stores and 16-byte loads
constexpr int kLatChunkStride = 32 * 16 + 16;
constexpr int kEncLatBytes = 64 * kLatChunkStride;       // 33 792
constexpr int kEncInBytes = 8 * kLatChunkStride;         //  4 224

// per-row tap description, produced by one lane per row and read back by the whole CTA (broadcast LDS)
struct __align__(16) RowTaps {
  uint32_t p00, p01, p10, p11;      // element offsets of the four taps inside the whole latent (all maps)
  float w00, w01, w10, w11;
  float xc, yc, zc, dx, dy, dz, dd;
  uint32_t map_plus1;               // (sb*NV + view) + 1, 0 = padding row
};

__device__ __forceinline__ float blend4_lo(uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint64_t W00, uint64_t W01,
                                           uint64_t W10, uint64_t W11, float& hi) {
  // both channels of a bf16x2 word in one packed fp32 operation (FMUL2 / FFMA2): a*w00, then fma by fma in tap order --
  // the same roundings as the scalar chain
  uint64_t acc = mul_f32x2(bf16x2_to_f32x2(a), W00);
  acc = fma_f32x2(bf16x2_to_f32x2(b), W01, acc);
  acc = fma_f32x2(bf16x2_to_f32x2(c), W10, acc);
  acc = fma_f32x2(bf16x2_to_f32x2(d), W11, acc);
  float lo;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc));
  return lo;
}

// One CTA per 32-row slab (floor(32/NV) samples x NV views); warp w owns feature channels [128 w, 128 w + 128), four per
// lane.  The warp walks the slab's rows VIEW BY VIEW in sample order: consecutive samples of a ray project into the same
// bilinear footprint of a source view most of the time (the depth-guided samples cluster within a few std of the
// surface, and the latent has half the image resolution), so the four taps stay in registers across such a run and only
// the blend weights change -- the L2 -> SM tap traffic falls from 4 KB per row to 4 KB per footprint run.  Same blend
// order as before (mul, fma, fma, fma in tap order): bit-identical outputs.
// SPLIT (verification mode): the latent is a bf16 hi + lo pair, the blended feature and the geometric code leave as
// hi + lo pairs of operand images (v_hi = bf16(v), v_lo = bf16(v - v_hi)).
// NOVEL (src/models/novel/novel_pixelnerf.py:143-196): the sample points are given (deformed), and the feature row is
// gen_latent[gen_uv] + latent[uv] -- a second bilinear gather from the learned map, with its own footprint cache.
// PE (NOVEL_PE, src/models/novel_pe/pe_novel_pixelnerf.py:260-275; with NOVEL): the encoder feature and the gen_latent
// taps leave as two operand images (a linear layer sits between them), and the six positional values of the row take the
// first six padding columns (55..60) of the geometric code.
template <bool SPLIT, bool NOVEL, bool PE = false>
__global__ void __launch_bounds__(kEncWarps * 32, (SPLIT || PE) ? 1 : (NOVEL ? 4 : 5)) encode_fwd_kernel(EncodeParams P, const __nv_bfloat16* latent_lo,
                                                                    unsigned char* lat_pk_lo, unsigned char* in_pk_lo) {
  extern __shared__ __align__(16) unsigned char enc_smem[];
  unsigned char* sl = enc_smem;                                   // [64][kLatChunkStride]
  unsigned char* si = enc_smem + kEncLatBytes;                    // [8][kLatChunkStride]
  RowTaps* taps_s = reinterpret_cast<RowTaps*>(si + kEncInBytes); // [32]
  unsigned char* sl_lo = reinterpret_cast<unsigned char*>(taps_s + 32);     // SPLIT only
  unsigned char* si_lo = sl_lo + kEncLatBytes;
  GenTaps* gen_s = reinterpret_cast<GenTaps*>(taps_s + 32);                 // NOVEL only (never together with SPLIT)
  unsigned char* sg = reinterpret_cast<unsigned char*>(gen_s + 32);         // PE only: [64][kLatChunkStride] gen taps
  float* pe_s = reinterpret_cast<float*>(sg + kEncLatBytes);                // PE only: [32][8]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const RowMap rm = RowMap::make(P.NV);
  const long long row_base = (long long)blockIdx.x * 32;
  const int L = 512;

  // phase 1: warp 0 runs the projection chains of the slab's 32 rows, one per lane (the other warps would only repeat
  // the same instruction stream for eight rows each)
  if (warp == 0) {
    long long sample; int view;
    RowTaps tp;
    tp.map_plus1 = 0;
    tp.p00 = tp.p01 = tp.p10 = tp.p11 = 0u;
    tp.w00 = tp.w01 = tp.w10 = tp.w11 = 0.f;
    tp.xc = tp.yc = tp.zc = tp.dx = tp.dy = tp.dz = tp.dd = 0.f;
    if (rm.decode(row_base + lane, P.n_samples, sample, view)) {
      RowGeom g;
      row_geometry(P, sample, view, g);
      const size_t mb = g.map * (size_t)P.Hl * P.Wl;          // texels before this (scene, view) map; the whole latent holds < 2^32 elements
      tp.p00 = (uint32_t)((mb + (size_t)g.y0 * P.Wl + g.x0) * L); tp.p01 = (uint32_t)((mb + (size_t)g.y0 * P.Wl + g.x1) * L);
      tp.p10 = (uint32_t)((mb + (size_t)g.y1 * P.Wl + g.x0) * L); tp.p11 = (uint32_t)((mb + (size_t)g.y1 * P.Wl + g.x1) * L);
      tp.w00 = (1.f - g.wx) * (1.f - g.wy); tp.w01 = g.wx * (1.f - g.wy);
      tp.w10 = (1.f - g.wx) * g.wy; tp.w11 = g.wx * g.wy;
      tp.xc = g.xc; tp.yc = g.yc; tp.zc = g.zc; tp.dx = g.dx; tp.dy = g.dy; tp.dz = g.dz; tp.dd = g.dd;
      tp.map_plus1 = (uint32_t)g.map + 1u;
      if (NOVEL) {
        GenTaps gt;
        gen_geometry(P, sample, (int)(g.map / P.NV), gt);
        gen_s[lane] = gt;
      }
      if (PE) pe_values(P, sample, g, pe_s + lane * 8);
    } else if (PE) {
#pragma unroll
      for (int j = 0; j < 6; ++j) pe_s[lane * 8 + j] = 0.f;
    }
    taps_s[lane] = tp;
  }
  __syncthreads();

  // phase 1b: the 55-wide geometric code of the warp's 8 rows (zero-padded to 64).  The 48 sines of a row are spread
  // over the warp as 8 x 48 = 384 = 12 full-warp evaluations, one branch-free sinf each; the 16 plain elements follow.
  {
    constexpr int NF = 6;                    // required by fill_encode_params (55-wide code)
    constexpr int npos = 3 + 6 * NF;         // 39
    constexpr int nsin_pos = 6 * NF;         // 36
    // 3 x 32 lanes cover the 2 x 48 sines of a row pair, so a lane meets only three (element, frequency, coordinate)
    // combinations: q = lane (row 2m), lane + 32 or lane - 16 (row 2m or 2m + 1), lane + 16 (row 2m + 1).  Everything
    // that depends on q alone is hoisted out of the loop over the four row pairs.
    int q3[3] = {lane, lane < 16 ? lane + 32 : lane - 16, lane + 16};
    const int r3[3] = {0, lane < 16 ? 0 : 1, 1};                    // row of the pair
    int foff[3], sto[3]; float fr[3], ph[3];
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int q = q3[u];
      int e, fs;
      if (q < nsin_pos) { fs = q / 3; e = 3 + q; foff[u] = q % 3; }            // xc, yc, zc
      else { fs = q - nsin_pos; e = npos + 4 + fs; foff[u] = 6; }              // dd
      fr[u] = P.freqs[fs]; ph[u] = P.phases[fs];
      sto[u] = (e >> 3) * kLatChunkStride + (e & 7) * 2;                        // byte offset inside the staged code slab
    }
    auto put = [&](int off, float val) {
      const __nv_bfloat16 hi = __float2bfloat16_rn(val);
      *reinterpret_cast<__nv_bfloat16*>(si + off) = hi;
      if (SPLIT) *reinterpret_cast<__nv_bfloat16*>(si_lo + off) = __float2bfloat16_rn(val - __bfloat162float(hi));
    };
#pragma unroll
    for (int m = 0; m < 4; ++m) {
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const int i = warp * 8 + 2 * m + r3[u];
        const RowTaps& tp = taps_s[i];
        const float x = (&tp.xc)[foff[u]];
        float val = sinf(__fadd_rn(ph[u], __fmul_rn(x, fr[u])));
        if (!tp.map_plus1) val = 0.f;
        put(sto[u] + i * 16, val);
      }
    }
    // the 16 plain elements of a row (x_cam 3 | dir_cam 3 | depth distance 1 | 9 zeros): lane & 15 fixes the element, two
    // rows per iteration
    {
      const int r = lane & 15;
      int e, fo = -1;                                              // fo: field offset from RowTaps::xc, -1 = zero padding
      if (r < 3) { e = r; fo = r; }                                // xc, yc, zc
      else if (r < 6) { e = npos + (r - 3); fo = r; }              // dx, dy, dz
      else if (r == 6) { e = npos + 3; fo = 6; }                   // dd
      else e = npos + 4 + 2 * NF + (r - 7);
      const int so = (e >> 3) * kLatChunkStride + (e & 7) * 2;
#pragma unroll
      for (int m = 0; m < 4; ++m) {
        const int i = warp * 8 + 2 * m + (lane >> 4);
        const RowTaps& tp = taps_s[i];
        float val = fo >= 0 ? (&tp.xc)[fo] : 0.f;
        if (PE && r >= 7 && r < 13) val = pe_s[i * 8 + (r - 7)];    // columns 55..60: src_pos_encodings | tgt_pos_encodings
        if (!tp.map_plus1) val = 0.f;
        put(so + i * 16, val);
      }
    }
  }
  __syncthreads();

  // phase 2: gather.  Lane's channels: 128 warp + 4 lane .. + 3 = 8 bytes of feature chunk c = 16 warp + lane / 2.
  // Rows are taken four at a time (same view, consecutive samples): the taps of every NEW footprint among the four are
  // requested first -- up to four footprints (4 x 1 KB per warp) in flight -- and only then are the rows blended, each
  // from the footprint it shares (a warp-uniform choice, so the selection is a branch, not a register copy).
  {
    const int ch = warp * 128 + lane * 4;
    unsigned char* dst0 = sl + (warp * 16 + (lane >> 1)) * kLatChunkStride + (lane & 1) * 8;
    unsigned char* dst0_lo = sl_lo + (warp * 16 + (lane >> 1)) * kLatChunkStride + (lane & 1) * 8;
    unsigned char* dstg = sg + (warp * 16 + (lane >> 1)) * kLatChunkStride + (lane & 1) * 8;
    const __nv_bfloat16* lat = P.latent + ch;
    const __nv_bfloat16* lat_lo = SPLIT ? latent_lo + ch : nullptr;
    struct Foot { uint2 a, b, c, d, al, bl, cl, dl; };            // (the lo taps exist in the SPLIT instantiation only)
    uint32_t prev00 = 0xffffffffu, prev11 = 0xffffffffu;
    uint2 ga = make_uint2(0, 0), gb = ga, gc = ga, gd = ga;       // NOVEL: the four taps of the current gen_latent footprint
    uint32_t gcur00 = 0xffffffffu, gcur11 = 0xffffffffu;
    const __nv_bfloat16* glat = NOVEL ? P.gen_latent + ch : nullptr;
    const int n_live = rm.SPW * rm.NV;
    auto fetch = [&](Foot& f, const uint4& pp) {
      f.a = __ldg(reinterpret_cast<const uint2*>(lat + pp.x)); f.b = __ldg(reinterpret_cast<const uint2*>(lat + pp.y));
      f.c = __ldg(reinterpret_cast<const uint2*>(lat + pp.z)); f.d = __ldg(reinterpret_cast<const uint2*>(lat + pp.w));
      if (SPLIT) {
        f.al = __ldg(reinterpret_cast<const uint2*>(lat_lo + pp.x)); f.bl = __ldg(reinterpret_cast<const uint2*>(lat_lo + pp.y));
        f.cl = __ldg(reinterpret_cast<const uint2*>(lat_lo + pp.z)); f.dl = __ldg(reinterpret_cast<const uint2*>(lat_lo + pp.w));
      }
    };
    // the current footprint, unpacked ONCE to fp32 pairs: rows that share it only redo the four weighted sums
    uint64_t ua0 = 0, ub0 = 0, uc0 = 0, ud0 = 0, ua1 = 0, ub1 = 0, uc1 = 0, ud1 = 0;
    uint64_t la0 = 0, lb0 = 0, lc0 = 0, ld0 = 0, la1 = 0, lb1 = 0, lc1 = 0, ld1 = 0;          // SPLIT: the lo taps
    auto unpack = [&](const Foot& f) {
      ua0 = bf16x2_to_f32x2(f.a.x); ub0 = bf16x2_to_f32x2(f.b.x); uc0 = bf16x2_to_f32x2(f.c.x); ud0 = bf16x2_to_f32x2(f.d.x);
      ua1 = bf16x2_to_f32x2(f.a.y); ub1 = bf16x2_to_f32x2(f.b.y); uc1 = bf16x2_to_f32x2(f.c.y); ud1 = bf16x2_to_f32x2(f.d.y);
      if (SPLIT) {
        la0 = bf16x2_to_f32x2(f.al.x); lb0 = bf16x2_to_f32x2(f.bl.x); lc0 = bf16x2_to_f32x2(f.cl.x); ld0 = bf16x2_to_f32x2(f.dl.x);
        la1 = bf16x2_to_f32x2(f.al.y); lb1 = bf16x2_to_f32x2(f.bl.y); lc1 = bf16x2_to_f32x2(f.cl.y); ld1 = bf16x2_to_f32x2(f.dl.y);
      }
    };
    auto wsum = [&](uint64_t a, uint64_t b, uint64_t c, uint64_t d, uint64_t W00, uint64_t W01, uint64_t W10, uint64_t W11,
                    float& lo, float& hi) {
      // a*w00, then fma by fma in tap order -- the same roundings as the scalar chain (and as blend4_lo)
      uint64_t acc = mul_f32x2(a, W00);
      acc = fma_f32x2(b, W01, acc);
      acc = fma_f32x2(c, W10, acc);
      acc = fma_f32x2(d, W11, acc);
      asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc));
    };
    auto blend_store = [&](const float4& ww, int l) {
      const uint64_t W00 = splat_f32x2(ww.x), W01 = splat_f32x2(ww.y), W10 = splat_f32x2(ww.z), W11 = splat_f32x2(ww.w);
      float v0, v1, v2, v3;
      wsum(ua0, ub0, uc0, ud0, W00, W01, W10, W11, v0, v1);
      wsum(ua1, ub1, uc1, ud1, W00, W01, W10, W11, v2, v3);
      if (SPLIT) {
        float u0, u1, u2, u3;
        wsum(la0, lb0, lc0, ld0, W00, W01, W10, W11, u0, u1);
        wsum(la1, lb1, lc1, ld1, W00, W01, W10, W11, u2, u3);
        v0 += u0; v1 += u1; v2 += u2; v3 += u3;
      }
      if (NOVEL) {                                                 // final_latent = gen_latent + latent (novel_pixelnerf.py:196)
        const uint4 gp = *reinterpret_cast<const uint4*>(&gen_s[l].p00);
        const float4 gw = *reinterpret_cast<const float4*>(&gen_s[l].w00);
        if (gp.x != gcur00 || gp.w != gcur11) {
          ga = __ldg(reinterpret_cast<const uint2*>(glat + gp.x)); gb = __ldg(reinterpret_cast<const uint2*>(glat + gp.y));
          gc = __ldg(reinterpret_cast<const uint2*>(glat + gp.z)); gd = __ldg(reinterpret_cast<const uint2*>(glat + gp.w));
          gcur00 = gp.x; gcur11 = gp.w;
        }
        const uint64_t G00 = splat_f32x2(gw.x), G01 = splat_f32x2(gw.y), G10 = splat_f32x2(gw.z), G11 = splat_f32x2(gw.w);
        float u0, u1, u2, u3;
        u0 = blend4_lo(ga.x, gb.x, gc.x, gd.x, G00, G01, G10, G11, u1);
        u2 = blend4_lo(ga.y, gb.y, gc.y, gd.y, G00, G01, G10, G11, u3);
        if (PE) *reinterpret_cast<uint2*>(dstg + l * 16) = make_uint2(pack_bf16x2(u0, u1), pack_bf16x2(u2, u3));
        else { v0 = u0 + v0; v1 = u1 + v1; v2 = u2 + v2; v3 = u3 + v3; }
      }
      const uint2 o = make_uint2(pack_bf16x2(v0, v1), pack_bf16x2(v2, v3));
      *reinterpret_cast<uint2*>(dst0 + l * 16) = o;
      if (SPLIT)
        *reinterpret_cast<uint2*>(dst0_lo + l * 16) =
            make_uint2(pack_bf16x2(v0 - bf16lo(o.x), v1 - bf16hi(o.x)), pack_bf16x2(v2 - bf16lo(o.y), v3 - bf16hi(o.y)));
    };
    for (int v = 0; v < rm.NV; ++v) {
      for (int s0 = 0; s0 < rm.SPW; s0 += 4) {
        uint4 pp[4]; float4 ww[4]; bool in[4], live[4], need[4];
        Foot f[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          in[j] = s0 + j < rm.SPW;
          const int l = in[j] ? (s0 + j) * rm.NV + v : 0;
          pp[j] = *reinterpret_cast<const uint4*>(&taps_s[l].p00);
          ww[j] = *reinterpret_cast<const float4*>(&taps_s[l].w00);
          live[j] = in[j] && taps_s[l].map_plus1 != 0u;
          need[j] = live[j] && (pp[j].x != prev00 || pp[j].w != prev11);      // p00 fixes (map, x0, y0), p11 the clamps
          if (live[j]) { prev00 = pp[j].x; prev11 = pp[j].w; }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (need[j]) fetch(f[j], pp[j]);                                    // (warp-uniform)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (!in[j]) continue;
          const int l = (s0 + j) * rm.NV + v;
          if (!live[j]) {
            *reinterpret_cast<uint2*>(dst0 + l * 16) = make_uint2(0, 0);
            if (SPLIT) *reinterpret_cast<uint2*>(dst0_lo + l * 16) = make_uint2(0, 0);
            if (PE) *reinterpret_cast<uint2*>(dstg + l * 16) = make_uint2(0, 0);
            continue;
          }
          // a row that opens a new footprint unpacks it (its taps were requested with the batch); the others reuse it
          if (j == 0 && need[0]) unpack(f[0]);
          if (j == 1 && need[1]) unpack(f[1]);
          if (j == 2 && need[2]) unpack(f[2]);
          if (j == 3 && need[3]) unpack(f[3]);
          blend_store(ww[j], l);
        }
      }
    }
    for (int l = n_live; l < 32; ++l) {                            // padding rows of the slab (NV = 3: rows 30, 31)
      *reinterpret_cast<uint2*>(dst0 + l * 16) = make_uint2(0, 0);
      if (SPLIT) *reinterpret_cast<uint2*>(dst0_lo + l * 16) = make_uint2(0, 0);
      if (PE) *reinterpret_cast<uint2*>(dstg + l * 16) = make_uint2(0, 0);
    }
  }
  __syncthreads();
  // write-out: a warp moves the slab's 512 contiguous bytes of one feature chunk per instruction
  const long long tile = row_base / 128;
  const int rr = (int)(row_base % 128);
  {
    unsigned char* lat_tile = P.lat_pk + (size_t)tile * (512 / 8) * 2048 + (size_t)(rr + lane) * 16;
    unsigned char* in_tile = P.in_pk + (size_t)tile * (64 / 8) * 2048 + (size_t)(rr + lane) * 16;
#pragma unroll 4
    for (int it = 0; it < 16; ++it) {
      const int c = it * 4 + warp;
      *reinterpret_cast<uint4*>(lat_tile + (size_t)c * 2048) = *reinterpret_cast<const uint4*>(sl + c * kLatChunkStride + lane * 16);
    }
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int c = it * 4 + warp;
      *reinterpret_cast<uint4*>(in_tile + (size_t)c * 2048) = *reinterpret_cast<const uint4*>(si + c * kLatChunkStride + lane * 16);
    }
  }
  if (PE) {
    unsigned char* gen_tile = P.gen_pk + (size_t)tile * (512 / 8) * 2048 + (size_t)(rr + lane) * 16;
#pragma unroll 4
    for (int it = 0; it < 16; ++it) {
      const int c = it * 4 + warp;
      *reinterpret_cast<uint4*>(gen_tile + (size_t)c * 2048) = *reinterpret_cast<const uint4*>(sg + c * kLatChunkStride + lane * 16);
    }
  }
  if (SPLIT) {
    unsigned char* lat_tile = lat_pk_lo + (size_t)tile * (512 / 8) * 2048 + (size_t)(rr + lane) * 16;
    unsigned char* in_tile = in_pk_lo + (size_t)tile * (64 / 8) * 2048 + (size_t)(rr + lane) * 16;
    for (int it = 0; it < 16; ++it) {
      const int c = it * 4 + warp;
      *reinterpret_cast<uint4*>(lat_tile + (size_t)c * 2048) = *reinterpret_cast<const uint4*>(sl_lo + c * kLatChunkStride + lane * 16);
    }
    for (int it = 0; it < 2; ++it) {
      const int c = it * 4 + warp;
      *reinterpret_cast<uint4*>(in_tile + (size_t)c * 2048) = *reinterpret_cast<const uint4*>(si_lo + c * kLatChunkStride + lane * 16);
    }
  }
}

// ---- production gather (no SPLIT / NOVEL): one warp per 8 rows, 16 channels per lane, every row loads its four taps.
// Round 2 measured three footprint-reusing variants of the slab kernel above against it on the same B200
// (profiles/r2_encode.md): they cut the L2 -> SM tap traffic by a third (2.7 instead of 4.1 KB per row) but none was
// faster -- the gather is bound by ALU issue (bf16 -> fp32 unpacking, address and footprint bookkeeping), not by tap
// bytes: 2.20 ms per FaceScape frame here, 2.30 / 2.79 ms for the 2- / 4-warp slab kernels.  The slab kernel stays for
// the SPLIT and NOVEL instantiations, where clarity matters more than the last 20 %.
constexpr int kEnc8Warps = 4;
constexpr int kRows8LatStride = 144;   // 8 rows * 16 B + 16 B pad -> conflict-free staging

// per-row tap description, produced by one lane per row and read back by the whole warp (broadcast LDS)
struct __align__(16) RowTaps8 {
  uint32_t p00, p01, p10, p11;      // element offsets of the four taps inside the whole latent (all maps)
  float w00, w01, w10, w11;
  float xc, yc, zc, dx, dy, dz, dd;
  uint32_t map_plus1;               // (sb*NV + view) + 1, 0 = padding row
};

__global__ void __launch_bounds__(kEnc8Warps * 32) encode_fwd_rows8_kernel(EncodeParams P) {
  __shared__ __align__(16) unsigned char stage_lat[kEnc8Warps][64 * kRows8LatStride];
  __shared__ __align__(16) unsigned char stage_in[kEnc8Warps][8 * 128];
  __shared__ RowTaps8 taps_s[kEnc8Warps][8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const RowMap rm = RowMap::make(P.NV);
  const long long group = (long long)blockIdx.x * kEnc8Warps + warp;   // 8-row group index
  const long long row_base = group * 8;
  unsigned char* sl = stage_lat[warp];
  unsigned char* si = stage_in[warp];
  const int L = 512;

  // phase 1: lanes 0..7 run the projection chain of one row each (the eight chains overlap instead of running
  // back to back on all 32 lanes)
  if (lane < 8) {
    long long sample; int view;
    RowTaps8 tp;
    tp.map_plus1 = 0;
    if (rm.decode(row_base + lane, P.n_samples, sample, view)) {
      RowGeom g;
      row_geometry(P, sample, view, g);
      const size_t mb = g.map * (size_t)P.Hl * P.Wl;          // texels before this (scene, view) map; the whole latent holds < 2^32 elements
      tp.p00 = (uint32_t)((mb + (size_t)g.y0 * P.Wl + g.x0) * L); tp.p01 = (uint32_t)((mb + (size_t)g.y0 * P.Wl + g.x1) * L);
      tp.p10 = (uint32_t)((mb + (size_t)g.y1 * P.Wl + g.x0) * L); tp.p11 = (uint32_t)((mb + (size_t)g.y1 * P.Wl + g.x1) * L);
      tp.w00 = (1.f - g.wx) * (1.f - g.wy); tp.w01 = g.wx * (1.f - g.wy);
      tp.w10 = (1.f - g.wx) * g.wy; tp.w11 = g.wx * g.wy;
      tp.xc = g.xc; tp.yc = g.yc; tp.zc = g.zc; tp.dx = g.dx; tp.dy = g.dy; tp.dz = g.dz; tp.dd = g.dd;
      tp.map_plus1 = (uint32_t)g.map + 1u;
    }
    taps_s[warp][lane] = tp;
  }
  __syncwarp();

  // phase 1b: the 55-wide geometric code of the 8 rows (zero-padded to 64).  The 48 sines of a row are spread over the
  // warp as 8 x 48 = 384 = 12 full-warp evaluations, one branch-free sinf each; the 16 plain elements follow.
  {
    constexpr int NF = 6;                    // required by fill_encode_params (55-wide code)
    constexpr int npos = 3 + 6 * NF;         // 39
    constexpr int nsin_pos = 6 * NF;      // 36, 48
    // 3 x 32 lanes cover the 2 x 48 sines of a row pair, so a lane meets only three (element, frequency, coordinate)
    // combinations: q = lane (row 2m), lane + 32 or lane - 16 (row 2m or 2m + 1), lane + 16 (row 2m + 1).  Everything
    // that depends on q alone is hoisted out of the loop over the four row pairs.
    int q3[3] = {lane, lane < 16 ? lane + 32 : lane - 16, lane + 16};
    const int r3[3] = {0, lane < 16 ? 0 : 1, 1};                    // row of the pair
    int foff[3], sto[3]; float fr[3], ph[3];
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int q = q3[u];
      int e, fs;
      if (q < nsin_pos) { fs = q / 3; e = 3 + q; foff[u] = q % 3; }            // xc, yc, zc
      else { fs = q - nsin_pos; e = npos + 4 + fs; foff[u] = 6; }              // dd
      fr[u] = P.freqs[fs]; ph[u] = P.phases[fs];
      sto[u] = (e >> 3) * 128 + (e & 7) * 2;                                    // byte offset inside the staged code tile
    }
#pragma unroll
    for (int m = 0; m < 4; ++m) {
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const int i = 2 * m + r3[u];
        const RowTaps8& tp = taps_s[warp][i];
        const float x = (&tp.xc)[foff[u]];
        float val = sinf(__fadd_rn(ph[u], __fmul_rn(x, fr[u])));
        if (!tp.map_plus1) val = 0.f;
        *reinterpret_cast<__nv_bfloat16*>(si + sto[u] + i * 16) = __float2bfloat16_rn(val);
      }
    }
    // the 16 plain elements of a row (x_cam 3 | dir_cam 3 | depth distance 1 | 9 zeros): lane & 15 fixes the element, two
    // rows per iteration
    {
      const int r = lane & 15;
      int e, fo = -1;                                              // fo: field offset from RowTaps::xc, -1 = zero padding
      if (r < 3) { e = r; fo = r; }                                // xc, yc, zc
      else if (r < 6) { e = npos + (r - 3); fo = r; }              // dx, dy, dz
      else if (r == 6) { e = npos + 3; fo = 6; }                   // dd
      else e = npos + 4 + 2 * NF + (r - 7);
      const int so = (e >> 3) * 128 + (e & 7) * 2;
#pragma unroll
      for (int m = 0; m < 4; ++m) {
        const int i = 2 * m + (lane >> 4);
        const RowTaps8& tp = taps_s[warp][i];
        float val = fo >= 0 ? (&tp.xc)[fo] : 0.f;
        if (!tp.map_plus1) val = 0.f;
        *reinterpret_cast<__nv_bfloat16*>(si + so + i * 16) = __float2bfloat16_rn(val);
      }
    }
  }

  // phase 2: the warp gathers row after row, two rows' loads in flight
#pragma unroll 2
  for (int i = 0; i < 8; ++i) {
    const RowTaps8 tp = taps_s[warp][i];
    uint4 o0 = make_uint4(0, 0, 0, 0), o1 = o0;
    if (tp.map_plus1) {
      const __nv_bfloat16* lat = P.latent;
      const uint64_t W00 = splat_f32x2(tp.w00), W01 = splat_f32x2(tp.w01), W10 = splat_f32x2(tp.w10), W11 = splat_f32x2(tp.w11);
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int ch = half * 256 + lane * 8;
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(lat + tp.p00 + ch));
        const uint4 b = __ldg(reinterpret_cast<const uint4*>(lat + tp.p01 + ch));
        const uint4 c = __ldg(reinterpret_cast<const uint4*>(lat + tp.p10 + ch));
        const uint4 d = __ldg(reinterpret_cast<const uint4*>(lat + tp.p11 + ch));
        const uint32_t av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
        const uint32_t cv[4] = {c.x, c.y, c.z, c.w}, dv[4] = {d.x, d.y, d.z, d.w};
        uint32_t r[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          // both channels of a bf16x2 word in one packed fp32 operation (FMUL2 / FFMA2): a*w00, then fma by fma in
          // tap order -- the same roundings as the scalar chain
          uint64_t acc = mul_f32x2(bf16x2_to_f32x2(av[q]), W00);
          acc = fma_f32x2(bf16x2_to_f32x2(bv[q]), W01, acc);
          acc = fma_f32x2(bf16x2_to_f32x2(cv[q]), W10, acc);
          acc = fma_f32x2(bf16x2_to_f32x2(dv[q]), W11, acc);
          float lo, hi;
          asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc));
          r[q] = pack_bf16x2(lo, hi);
        }
        if (half == 0) o0 = make_uint4(r[0], r[1], r[2], r[3]); else o1 = make_uint4(r[0], r[1], r[2], r[3]);
      }
    }
    *reinterpret_cast<uint4*>(sl + lane * kRows8LatStride + i * 16) = o0;
    *reinterpret_cast<uint4*>(sl + (32 + lane) * kRows8LatStride + i * 16) = o1;
  }
  __syncwarp();
  // write-out: every 8 lanes move one 128-byte line (8 rows x 16 B of one feature chunk)
  const long long tile = row_base / 128;
  const int rr = (int)(row_base % 128);
  unsigned char* lat_tile = P.lat_pk + (size_t)tile * (512 / 8) * 2048;
  unsigned char* in_tile = P.in_pk + (size_t)tile * (64 / 8) * 2048;
#pragma unroll 4
  for (int it = 0; it < 16; ++it) {
    const int c = it * 4 + (lane >> 3), i = lane & 7;
    const uint4 v = *reinterpret_cast<const uint4*>(sl + c * kRows8LatStride + i * 16);
    *reinterpret_cast<uint4*>(lat_tile + (size_t)c * 2048 + (size_t)(rr + i) * 16) = v;
  }
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const int c = it * 4 + (lane >> 3), i = lane & 7;
    const uint4 v = *reinterpret_cast<const uint4*>(si + c * 128 + i * 16);
    *reinterpret_cast<uint4*>(in_tile + (size_t)c * 2048 + (size_t)(rr + i) * 16) = v;
  }
}


__device__ __forceinline__ void red_add_v4(float* addr, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// Backward scatter d_latent[tap] += w_tap * d_feature.  One CTA per 32-row slab; warp w owns channels
// [128 w, 128 w + 128).  The slab's gradient rows arrive in XR order (one row per lane, coalesced), are transposed
// through shared memory, and the warp then walks the rows view by view in sample order with the channels across
// lanes: consecutive samples of a ray that fall into the same bilinear footprint of a view (most of the
// depth-guided samples do: they cluster within a few std of the surface) are summed in registers and leave as
// ONE set of four 512-byte-contiguous vector reductions per warp instead of one 16-byte reduction per lane and row.
constexpr int kBwdRowStride = 33;      // float4 per staged row: 32 + 1 pad -> conflict-free 128-bit stores
constexpr int kBwdSmemBytes = 4 * 32 * kBwdRowStride * 16;

// NOVEL (novel_pixelnerf.py:196, final_latent = gen_latent + latent): the same gradient row also belongs to the learned
// map, at the footprint of the `gen`-deformed point, which is the same for the NV views of a sample -- the views' rows
// are summed first, then runs of samples inside one footprint, and leave as vector reductions into d_gen_latent.
template <bool NOVEL>
__global__ void __launch_bounds__(128) encode_bwd_kernel(EncodeParams P) {
  extern __shared__ float4 bwd_stage[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const RowMap rm = RowMap::make(P.NV);
  const long long slab = blockIdx.x;
  const long long row = slab * 32 + lane;
  long long sample; int view;
  const bool valid = rm.decode(row, P.n_samples, sample, view);
  if (!__any_sync(0xffffffffu, valid)) return;
  const int L = 512;
  uint32_t kx = 0, ky = 0, km = 0;
  float wx = 0.f, wy = 0.f;
  if (valid) {
    RowGeom g;
    row_geometry(P, sample, view, g);
    kx = (uint32_t)g.x0 | ((uint32_t)g.x1 << 16);
    ky = (uint32_t)g.y0 | ((uint32_t)g.y1 << 16);
    km = (uint32_t)g.map + 1u;
    wx = g.wx; wy = g.wy;
  }
  // stage this warp's 128 channels of the 32 rows: st[row][n4]
  float4* st = bwd_stage + warp * 32 * kBwdRowStride;
  auto stage = [&](const float* xr) {
    const long long tile = row / 128;
    const int rr = (int)(row % 128);
    const float4* src = reinterpret_cast<const float4*>(xr) + ((size_t)tile * (L / 4) + warp * 32) * 128 + rr;
#pragma unroll 8
    for (int n4 = 0; n4 < 32; ++n4) st[lane * kBwdRowStride + n4] = src[(size_t)n4 * 128];
  };
  const float* gen_xr = (NOVEL && P.d_gen_xr) ? P.d_gen_xr : P.d_lat_xr;
  const float* first_xr = P.d_latent ? P.d_lat_xr : gen_xr;
  stage(first_xr);
  __syncwarp();
  float4 a00 = make_float4(0.f, 0.f, 0.f, 0.f), a01 = a00, a10 = a00, a11 = a00;
  uint32_t cx = 0, cy = 0, cm = 0;
  const int ch = warp * 128 + lane * 4;
  auto flush = [&]() {
    if (cm == 0) return;
    float* dl = P.d_latent + (size_t)(cm - 1) * P.Hl * P.Wl * L + ch;
    const size_t x0 = cx & 0xffffu, x1 = cx >> 16, y0 = cy & 0xffffu, y1 = cy >> 16;
    red_add_v4(dl + (y0 * P.Wl + x0) * L, a00);
    if (x1 != x0) red_add_v4(dl + (y0 * P.Wl + x1) * L, a01);
    if (y1 != y0) {
      red_add_v4(dl + (y1 * P.Wl + x0) * L, a10);
      if (x1 != x0) red_add_v4(dl + (y1 * P.Wl + x1) * L, a11);
    }
  };
  for (int v = 0; v < (P.d_latent ? rm.NV : 0); ++v) {
    for (int s = 0; s < rm.SPW; ++s) {
      const int l = s * rm.NV + v;
      const uint32_t m = __shfl_sync(0xffffffffu, km, l);
      if (m == 0) continue;                              // padding row (warp-uniform)
      const uint32_t x = __shfl_sync(0xffffffffu, kx, l), y = __shfl_sync(0xffffffffu, ky, l);
      const float fx = __shfl_sync(0xffffffffu, wx, l), fy = __shfl_sync(0xffffffffu, wy, l);
      if (m != cm || x != cx || y != cy) {
        flush();
        a00 = make_float4(0.f, 0.f, 0.f, 0.f); a01 = a00; a10 = a00; a11 = a00;
        cm = m; cx = x; cy = y;
      }
      const float4 d = st[l * kBwdRowStride + lane];
      const float w00 = (1.f - fx) * (1.f - fy), w01 = fx * (1.f - fy), w10 = (1.f - fx) * fy, w11 = fx * fy;
      a00.x = fmaf(d.x, w00, a00.x); a00.y = fmaf(d.y, w00, a00.y); a00.z = fmaf(d.z, w00, a00.z); a00.w = fmaf(d.w, w00, a00.w);
      a01.x = fmaf(d.x, w01, a01.x); a01.y = fmaf(d.y, w01, a01.y); a01.z = fmaf(d.z, w01, a01.z); a01.w = fmaf(d.w, w01, a01.w);
      a10.x = fmaf(d.x, w10, a10.x); a10.y = fmaf(d.y, w10, a10.y); a10.z = fmaf(d.z, w10, a10.z); a10.w = fmaf(d.w, w10, a10.w);
      a11.x = fmaf(d.x, w11, a11.x); a11.y = fmaf(d.y, w11, a11.y); a11.z = fmaf(d.z, w11, a11.z); a11.w = fmaf(d.w, w11, a11.w);
    }
  }
  flush();
  if (NOVEL && P.d_gen_latent) {
    if (gen_xr != first_xr) {            // NOVEL_PE: the learned map has its own gradient rows
      __syncwarp();
      stage(gen_xr);
      __syncwarp();
    }
    GenTaps gt;
    gt.p00 = gt.p01 = gt.p10 = gt.p11 = 0u; gt.w00 = gt.w01 = gt.w10 = gt.w11 = 0.f;
    uint32_t gok = 0;
    const long long smp = slab * rm.SPW + lane;
    if (lane < rm.SPW && smp < P.n_samples) {
      gen_geometry(P, smp, (int)((smp / P.K + P.ray_offset) / P.NR), gt);
      gok = 1;
    }
    uint32_t c00 = 0, c01 = 0, c10 = 0, c11 = 0;
    bool open = false;
    a00 = make_float4(0.f, 0.f, 0.f, 0.f); a01 = a00; a10 = a00; a11 = a00;
    auto gflush = [&]() {
      if (!open) return;
      float* dg = P.d_gen_latent + ch;
      red_add_v4(dg + c00, a00);
      if (c01 != c00) red_add_v4(dg + c01, a01);
      if (c10 != c00) {
        red_add_v4(dg + c10, a10);
        if (c11 != c10) red_add_v4(dg + c11, a11);
      }
    };
    for (int s = 0; s < rm.SPW; ++s) {
      if (__shfl_sync(0xffffffffu, gok, s) == 0) continue;           // past the last sample (warp-uniform)
      const uint32_t q00 = __shfl_sync(0xffffffffu, gt.p00, s), q01 = __shfl_sync(0xffffffffu, gt.p01, s);
      const uint32_t q10 = __shfl_sync(0xffffffffu, gt.p10, s), q11 = __shfl_sync(0xffffffffu, gt.p11, s);
      const float w00 = __shfl_sync(0xffffffffu, gt.w00, s), w01 = __shfl_sync(0xffffffffu, gt.w01, s);
      const float w10 = __shfl_sync(0xffffffffu, gt.w10, s), w11 = __shfl_sync(0xffffffffu, gt.w11, s);
      if (!open || q00 != c00 || q01 != c01 || q10 != c10 || q11 != c11) {
        gflush();
        a00 = make_float4(0.f, 0.f, 0.f, 0.f); a01 = a00; a10 = a00; a11 = a00;
        c00 = q00; c01 = q01; c10 = q10; c11 = q11; open = true;
      }
      float4 d = st[(s * rm.NV) * kBwdRowStride + lane];              // sum over the sample's views
      for (int v = 1; v < rm.NV; ++v) {
        const float4 e = st[(s * rm.NV + v) * kBwdRowStride + lane];
        d.x += e.x; d.y += e.y; d.z += e.z; d.w += e.w;
      }
      a00.x = fmaf(d.x, w00, a00.x); a00.y = fmaf(d.y, w00, a00.y); a00.z = fmaf(d.z, w00, a00.z); a00.w = fmaf(d.w, w00, a00.w);
      a01.x = fmaf(d.x, w01, a01.x); a01.y = fmaf(d.y, w01, a01.y); a01.z = fmaf(d.z, w01, a01.z); a01.w = fmaf(d.w, w01, a01.w);
      a10.x = fmaf(d.x, w10, a10.x); a10.y = fmaf(d.y, w10, a10.y); a10.z = fmaf(d.z, w10, a10.z); a10.w = fmaf(d.w, w10, a10.w);
      a11.x = fmaf(d.x, w11, a11.x); a11.y = fmaf(d.y, w11, a11.y); a11.z = fmaf(d.z, w11, a11.z); a11.w = fmaf(d.w, w11, a11.w);
    }
    gflush();
  }
}

// latent (N,C,H,W) fp32 -> (N,H,W,C) bf16, tiled transpose through shared memory
__global__ void nchw_to_nhwc_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int C, int HW) {
  __shared__ float tile[32][33];
  const size_t n = blockIdx.z;
  const int c0 = blockIdx.y * 32, p0 = blockIdx.x * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;   // 32 x 8
  for (int j = ty; j < 32; j += 8) {
    const int c = c0 + j, p = p0 + tx;
    tile[j][tx] = (c < C && p < HW) ? src[(n * C + c) * (size_t)HW + p] : 0.f;
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {
    const int p = p0 + j, c = c0 + tx;
    if (c < C && p < HW) dst[(n * (size_t)HW + p) * C + c] = __float2bfloat16_rn(tile[tx][j]);
  }
}
// d_latent (N,H,W,C) fp32 -> (N,C,H,W) fp32
__global__ void nhwc_to_nchw_f32_kernel(const float* __restrict__ src, float* __restrict__ dst, int C, int HW) {
  __shared__ float tile[32][33];
  const size_t n = blockIdx.z;
  const int c0 = blockIdx.y * 32, p0 = blockIdx.x * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;
  for (int j = ty; j < 32; j += 8) {
    const int p = p0 + j, c = c0 + tx;
    tile[j][tx] = (c < C && p < HW) ? src[(n * (size_t)HW + p) * C + c] : 0.f;
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {
    const int c = c0 + j, p = p0 + tx;
    if (c < C && p < HW) dst[(n * C + c) * (size_t)HW + p] = tile[tx][j];
  }
}

}  // namespace dnr

static int fill_encode_params(dnr::EncodeParams& P, const float* rays, const float* z, const float* poses,
                              const float* focal, const float* cxy, const float* depth, const void* latent_nhwc,
                              const float* freqs, const float* phases, int SB, int NV, int NR, int K, int H, int W,
                              int Hl, int Wl, int L, int NF, long long ray_offset, long long n_rays, float img_w, float img_h,
                              float lat_sx, float lat_sy) {
  DNR_REQUIRE(L == 512, "encode: latent width must be 512");
  DNR_REQUIRE(NF == 6, "encode: num_freqs must be 6 (55-wide geometric code)");
  DNR_REQUIRE(NV >= 1 && NV <= 8, "encode: NV must be in [1,8]");
  P.rays = rays; P.z = z; P.poses = poses; P.focal = focal; P.cxy = cxy; P.depth = depth;
  P.latent = (const __nv_bfloat16*)latent_nhwc; P.freqs = freqs; P.phases = phases;
  P.SB = SB; P.NV = NV; P.NR = NR; P.K = K; P.H = H; P.W = W; P.Hl = Hl; P.Wl = Wl; P.NF = NF;
  P.img_w = img_w; P.img_h = img_h; P.lat_sx = lat_sx; P.lat_sy = lat_sy;
  P.n_samples = n_rays * K; P.ray_offset = ray_offset; P.n_rays = n_rays;
  DNR_REQUIRE(ray_offset >= 0 && ray_offset + n_rays <= (long long)SB * NR, "encode: ray chunk out of range");
  P.lat_pk = nullptr; P.in_pk = nullptr; P.d_latent = nullptr; P.d_lat_xr = nullptr;
  P.pts_in = nullptr; P.pts_gen = nullptr; P.gen_poses = nullptr; P.gen_focal = nullptr; P.gen_cxy = nullptr;
  P.gen_latent = nullptr; P.Hg = P.Wg = 0; P.gen_img_w = P.gen_img_h = P.gen_sx = P.gen_sy = 0.f;
  P.d_gen_latent = nullptr; P.d_gen_xr = nullptr;
  P.pe_src = P.pe_tgt = P.tgt_poses = P.tgt_focal = P.tgt_cxy = nullptr; P.Hp = P.Wp = P.Ht = P.Wt = 0;
  P.pe_sx = P.pe_sy = P.tgt_img_w = P.tgt_img_h = P.tgt_sx = P.tgt_sy = 0.f; P.gen_pk = nullptr;
  return 0;
}

// Number of 128-row tiles of the per-view MLP batch for n_samples points seen from NV views.
DNR_EXPORT int dnr_view_row_tiles(long long n_samples, int NV) {
  if (NV < 1 || NV > 8) return -1;
  return dnr::RowMap::make(NV).tiles_for(n_samples);
}

static int launch_encode_fwd(dnr::EncodeParams& P, int NV, const void* latent_lo, void* lat_pk_lo, void* in_pk_lo,
                             cudaStream_t stream) {
  if (P.n_samples <= 0) return 0;
  const bool split = latent_lo != nullptr, novel = P.gen_latent != nullptr, pe = P.pe_src != nullptr;
  const int tiles = dnr::RowMap::make(NV).tiles_for(P.n_samples);
  const int slabs = tiles * 4;
  const int smem = (dnr::kEncLatBytes + dnr::kEncInBytes) * (split ? 2 : 1) + 32 * (int)sizeof(dnr::RowTaps) +
                   (novel ? 32 * (int)sizeof(dnr::GenTaps) : 0) + (pe ? dnr::kEncLatBytes + 32 * 8 * 4 : 0);
  if (split) {
    DNR_CHECK(cudaFuncSetAttribute(dnr::encode_fwd_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
              "dnr_encode_fwd_split/attr");
    dnr::encode_fwd_kernel<true, false><<<slabs, dnr::kEncWarps * 32, smem, stream>>>(
        P, (const __nv_bfloat16*)latent_lo, (unsigned char*)lat_pk_lo, (unsigned char*)in_pk_lo);
  } else if (pe) {
    DNR_CHECK(cudaFuncSetAttribute(dnr::encode_fwd_kernel<false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
              "dnr_encode_novel_pe_fwd/attr");
    dnr::encode_fwd_kernel<false, true, true><<<slabs, dnr::kEncWarps * 32, smem, stream>>>(P, nullptr, nullptr, nullptr);
  } else if (novel) {
    dnr::encode_fwd_kernel<false, true><<<slabs, dnr::kEncWarps * 32, smem, stream>>>(P, nullptr, nullptr, nullptr);
  } else {
    dnr::encode_fwd_rows8_kernel<<<tiles * 16 / dnr::kEnc8Warps, dnr::kEnc8Warps * 32, 0, stream>>>(P);
  }
  return 0;
}

// Replaces PixelNeRF.forward's projection / gather / encoding (pixelnerf.py:91-128).
DNR_EXPORT int dnr_encode_fwd(const float* rays, const float* z, const float* poses, const float* focal,
                              const float* cxy, const float* depth, const void* latent_nhwc_bf16, const float* freqs,
                              const float* phases, int SB, int NV, int NR, int K, int H, int W, int Hl, int Wl, int L,
                              int NF, long long ray_offset, long long n_rays, float img_w, float img_h, float lat_sx,
                              float lat_sy, void* lat_pk, void* in_pk,
                              cudaStream_t stream) {
  dnr::EncodeParams P;
  int rc = fill_encode_params(P, rays, z, poses, focal, cxy, depth, latent_nhwc_bf16, freqs, phases, SB, NV, NR, K, H,
                              W, Hl, Wl, L, NF, ray_offset, n_rays, img_w, img_h, lat_sx, lat_sy);
  if (rc) return rc;
  DNR_REQUIRE((unsigned long long)SB * NV * Hl * Wl * L < (1ull << 32),
              "dnr_encode_fwd: the latent holds 2^32 elements or more (tap offsets are 32-bit); render fewer scenes per call");
  DNR_REQUIRE(latent_nhwc_bf16 && lat_pk && in_pk, "dnr_encode_fwd: null argument");
  P.lat_pk = (unsigned char*)lat_pk; P.in_pk = (unsigned char*)in_pk;
  rc = launch_encode_fwd(P, NV, nullptr, nullptr, nullptr, stream);
  if (rc) return rc;
  DNR_CHECK_LAUNCH("dnr_encode_fwd");
  return 0;
}

// Verification-mode variant (NeRFRendererDGS(precision="split")): the latent and both outputs are bf16 hi + lo pairs
// (x ~ hi + lo to 2^-17 relative), otherwise identical to dnr_encode_fwd.
DNR_EXPORT int dnr_encode_fwd_split(const float* rays, const float* z, const float* poses, const float* focal,
                                    const float* cxy, const float* depth, const void* latent_hi, const void* latent_lo,
                                    const float* freqs, const float* phases, int SB, int NV, int NR, int K, int H, int W,
                                    int Hl, int Wl, int L, int NF, long long ray_offset, long long n_rays, float img_w,
                                    float img_h, float lat_sx, float lat_sy, void* lat_pk_hi, void* in_pk_hi,
                                    void* lat_pk_lo, void* in_pk_lo, cudaStream_t stream) {
  dnr::EncodeParams P;
  int rc = fill_encode_params(P, rays, z, poses, focal, cxy, depth, latent_hi, freqs, phases, SB, NV, NR, K, H,
                              W, Hl, Wl, L, NF, ray_offset, n_rays, img_w, img_h, lat_sx, lat_sy);
  if (rc) return rc;
  DNR_REQUIRE((unsigned long long)SB * NV * Hl * Wl * L < (1ull << 32), "dnr_encode_fwd_split: latent too large");
  DNR_REQUIRE(latent_hi && latent_lo && lat_pk_hi && in_pk_hi && lat_pk_lo && in_pk_lo, "dnr_encode_fwd_split: null argument");
  P.lat_pk = (unsigned char*)lat_pk_hi; P.in_pk = (unsigned char*)in_pk_hi;
  rc = launch_encode_fwd(P, NV, latent_lo, lat_pk_lo, in_pk_lo, stream);
  if (rc) return rc;
  DNR_CHECK_LAUNCH("dnr_encode_fwd_split");
  return 0;
}

// NOVEL variant of dnr_encode_fwd (src/models/novel/novel_pixelnerf.py:143-241): pts_in / pts_gen [n_rays*K, 3] are the
// chunk's sample points deformed with tgt_in_offsets / tgt_gen_offsets (dnr_novel_deform); pts_in replaces o + z d in the
// projection chain, pts_gen is projected by the generic camera (gen_poses [SB,1,4,4], gen_focal / gen_cxy [SB,1,2],
// gen_img_w/h) into gen_latent_nhwc [Hg,Wg,512] bf16 (feature-padding rescale gen_sx / gen_sy), and the feature row is
// gen_latent[gen_uv] + latent[uv].  Backward: dnr_encode_novel_bwd.
DNR_EXPORT int dnr_encode_novel_fwd(const float* rays, const float* z, const float* pts_in, const float* pts_gen,
                                    const float* poses, const float* focal, const float* cxy, const float* depth,
                                    const void* latent_nhwc_bf16, const float* gen_poses, const float* gen_focal,
                                    const float* gen_cxy, const void* gen_latent_nhwc_bf16, int Hg, int Wg,
                                    float gen_img_w, float gen_img_h, float gen_sx, float gen_sy, const float* freqs,
                                    const float* phases, int SB, int NV, int NR, int K, int H, int W, int Hl, int Wl, int L,
                                    int NF, long long ray_offset, long long n_rays, float img_w, float img_h, float lat_sx,
                                    float lat_sy, void* lat_pk, void* in_pk, cudaStream_t stream) {
  dnr::EncodeParams P;
  int rc = fill_encode_params(P, rays, z, poses, focal, cxy, depth, latent_nhwc_bf16, freqs, phases, SB, NV, NR, K, H,
                              W, Hl, Wl, L, NF, ray_offset, n_rays, img_w, img_h, lat_sx, lat_sy);
  if (rc) return rc;
  DNR_REQUIRE((unsigned long long)SB * NV * Hl * Wl * L < (1ull << 32), "dnr_encode_novel_fwd: latent too large");
  DNR_REQUIRE(latent_nhwc_bf16 && lat_pk && in_pk && pts_in && pts_gen && gen_poses && gen_focal && gen_cxy &&
              gen_latent_nhwc_bf16, "dnr_encode_novel_fwd: null argument");
  DNR_REQUIRE(Hg >= 1 && Wg >= 1 && (unsigned long long)Hg * Wg * L < (1ull << 32), "dnr_encode_novel_fwd: bad gen_latent size");
  P.lat_pk = (unsigned char*)lat_pk; P.in_pk = (unsigned char*)in_pk;
  P.pts_in = pts_in; P.pts_gen = pts_gen; P.gen_poses = gen_poses; P.gen_focal = gen_focal; P.gen_cxy = gen_cxy;
  P.gen_latent = (const __nv_bfloat16*)gen_latent_nhwc_bf16; P.Hg = Hg; P.Wg = Wg;
  P.gen_img_w = gen_img_w; P.gen_img_h = gen_img_h; P.gen_sx = gen_sx; P.gen_sy = gen_sy;
  rc = launch_encode_fwd(P, NV, nullptr, nullptr, nullptr, stream);
  if (rc) return rc;
  DNR_CHECK_LAUNCH("dnr_encode_novel_fwd");
  return 0;
}

// NOVEL_PE variant of dnr_encode_novel_fwd (src/models/novel_pe/pe_novel_pixelnerf.py:200-275).  Against the NOVEL entry:
// lat_pk receives the encoder feature ALONE and gen_pk [rows_pad x 512] bf16 PK128 the gen_latent taps (the variant's
// deformation layer, a 518 -> 512 nn.Linear, is applied to [feature | 6 positional values] before the two are added,
// :265-268 -- a dnr_mlp_gemm with gen_pk as its bf16 residual operand), and columns 55..60 of in_pk carry
// [src_pos_encodings 3 | tgt_pos_encodings 3]: pe_src [SB,NV,3,Hp,Wp] fp32 tapped bilinearly at the row's source uv
// (rescaled by pe_sx / pe_sy, :133-166), pe_tgt [SB,1,3,Ht,Wt] fp32 at the projection of the UNDEFORMED point o + z d by
// the target camera tgt_poses [SB,1,4,4], tgt_focal / tgt_cxy [SB,1,2], tgt_img_w/h (rescaled by tgt_sx / tgt_sy,
// :168-198,223-226,248-251).  The MLP input of the variant, [final_latent | pe 6 | code 55] (:275), is these operands.
DNR_EXPORT int dnr_encode_novel_pe_fwd(const float* rays, const float* z, const float* pts_in, const float* pts_gen,
                                       const float* poses, const float* focal, const float* cxy, const float* depth,
                                       const void* latent_nhwc_bf16, const float* gen_poses, const float* gen_focal,
                                       const float* gen_cxy, const void* gen_latent_nhwc_bf16, int Hg, int Wg,
                                       float gen_img_w, float gen_img_h, float gen_sx, float gen_sy, const float* pe_src,
                                       int Hp, int Wp, float pe_sx, float pe_sy, const float* tgt_poses,
                                       const float* tgt_focal, const float* tgt_cxy, float tgt_img_w, float tgt_img_h,
                                       const float* pe_tgt, int Ht, int Wt, float tgt_sx, float tgt_sy,
                                       const float* freqs, const float* phases, int SB, int NV, int NR, int K, int H,
                                       int W, int Hl, int Wl, int L, int NF, long long ray_offset, long long n_rays,
                                       float img_w, float img_h, float lat_sx, float lat_sy, void* lat_pk, void* gen_pk,
                                       void* in_pk, cudaStream_t stream) {
  dnr::EncodeParams P;
  int rc = fill_encode_params(P, rays, z, poses, focal, cxy, depth, latent_nhwc_bf16, freqs, phases, SB, NV, NR, K, H,
                              W, Hl, Wl, L, NF, ray_offset, n_rays, img_w, img_h, lat_sx, lat_sy);
  if (rc) return rc;
  DNR_REQUIRE((unsigned long long)SB * NV * Hl * Wl * L < (1ull << 32), "dnr_encode_novel_pe_fwd: latent too large");
  DNR_REQUIRE(z && latent_nhwc_bf16 && lat_pk && gen_pk && in_pk && pts_in && pts_gen && gen_poses && gen_focal && gen_cxy &&
              gen_latent_nhwc_bf16 && pe_src && pe_tgt && tgt_poses && tgt_focal && tgt_cxy,
              "dnr_encode_novel_pe_fwd: null argument");
  DNR_REQUIRE(Hg >= 1 && Wg >= 1 && (unsigned long long)Hg * Wg * L < (1ull << 32), "dnr_encode_novel_pe_fwd: bad gen_latent size");
  DNR_REQUIRE(Hp >= 1 && Wp >= 1 && Ht >= 1 && Wt >= 1, "dnr_encode_novel_pe_fwd: bad positional map size");
  P.lat_pk = (unsigned char*)lat_pk; P.in_pk = (unsigned char*)in_pk; P.gen_pk = (unsigned char*)gen_pk;
  P.pts_in = pts_in; P.pts_gen = pts_gen; P.gen_poses = gen_poses; P.gen_focal = gen_focal; P.gen_cxy = gen_cxy;
  P.gen_latent = (const __nv_bfloat16*)gen_latent_nhwc_bf16; P.Hg = Hg; P.Wg = Wg;
  P.gen_img_w = gen_img_w; P.gen_img_h = gen_img_h; P.gen_sx = gen_sx; P.gen_sy = gen_sy;
  P.pe_src = pe_src; P.Hp = Hp; P.Wp = Wp; P.pe_sx = pe_sx; P.pe_sy = pe_sy;
  P.tgt_poses = tgt_poses; P.tgt_focal = tgt_focal; P.tgt_cxy = tgt_cxy; P.tgt_img_w = tgt_img_w; P.tgt_img_h = tgt_img_h;
  P.pe_tgt = pe_tgt; P.Ht = Ht; P.Wt = Wt; P.tgt_sx = tgt_sx; P.tgt_sy = tgt_sy;
  rc = launch_encode_fwd(P, NV, nullptr, nullptr, nullptr, stream);
  if (rc) return rc;
  DNR_CHECK_LAUNCH("dnr_encode_novel_pe_fwd");
  return 0;
}

// Backward of SpatialEncoder.index (grid_sample bilinear w.r.t. the latent map, image_encoder.py:117-123).
DNR_EXPORT int dnr_encode_bwd(const float* rays, const float* z, const float* poses, const float* focal,
                              const float* cxy, const float* depth, const float* freqs, const float* phases, int SB,
                              int NV, int NR, int K, int H, int W, int Hl, int Wl, int L, int NF, long long ray_offset,
                              long long n_rays, float img_w, float img_h, float lat_sx, float lat_sy,
                              const float* d_lat_xr, float* d_latent_nhwc,
                              cudaStream_t stream) {
  dnr::EncodeParams P;
  int rc = fill_encode_params(P, rays, z, poses, focal, cxy, depth, nullptr, freqs, phases, SB, NV, NR, K, H, W, Hl,
                              Wl, L, NF, ray_offset, n_rays, img_w, img_h, lat_sx, lat_sy);
  if (rc) return rc;
  P.d_lat_xr = d_lat_xr; P.d_latent = d_latent_nhwc;
  if (P.n_samples <= 0) return 0;
  const int tiles = dnr::RowMap::make(NV).tiles_for(P.n_samples);
  DNR_CHECK(cudaFuncSetAttribute(dnr::encode_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 dnr::kBwdSmemBytes), "dnr_encode_bwd");
  const long long slabs = (P.n_samples + dnr::RowMap::make(NV).SPW - 1) / dnr::RowMap::make(NV).SPW;
  (void)tiles;
  dnr::encode_bwd_kernel<false><<<(unsigned)slabs, 128, dnr::kBwdSmemBytes, stream>>>(P);
  DNR_CHECK_LAUNCH("dnr_encode_bwd");
  return 0;
}

// Backward of dnr_encode_novel_fwd w.r.t. the two feature maps (novel_pixelnerf.py:196: the gradient of final_latent goes
// to encoder.latent at the footprint of the `in`-deformed point in each source view, and to gen_latent at the footprint
// of the `gen`-deformed point under the generic camera, summed over scenes and views -- the reference repeats the one
// learned map, :110).  Either output may be null.  d_gen_latent_nhwc [Hg,Wg,512] fp32, accumulated into.  The deformed
// points carry no gradient (the offsets are inputs of the training step, novel.py:280-285,361-363).  d_gen_xr: the
// gradient rows of the gen_latent taps when they differ from those of the encoder feature (NOVEL_PE: a linear layer sits
// between the encoder feature and the sum); null = d_lat_xr.
DNR_EXPORT int dnr_encode_novel_bwd(const float* rays, const float* z, const float* pts_in, const float* pts_gen,
                                    const float* poses, const float* focal, const float* cxy, const float* depth,
                                    const float* gen_poses, const float* gen_focal, const float* gen_cxy, int Hg, int Wg,
                                    float gen_img_w, float gen_img_h, float gen_sx, float gen_sy, const float* freqs,
                                    const float* phases, int SB, int NV, int NR, int K, int H, int W, int Hl, int Wl, int L,
                                    int NF, long long ray_offset, long long n_rays, float img_w, float img_h, float lat_sx,
                                    float lat_sy, const float* d_lat_xr, const float* d_gen_xr, float* d_latent_nhwc,
                                    float* d_gen_latent_nhwc, cudaStream_t stream) {
  dnr::EncodeParams P;
  int rc = fill_encode_params(P, rays, z, poses, focal, cxy, depth, nullptr, freqs, phases, SB, NV, NR, K, H, W, Hl,
                              Wl, L, NF, ray_offset, n_rays, img_w, img_h, lat_sx, lat_sy);
  if (rc) return rc;
  DNR_REQUIRE(pts_in && pts_gen && gen_poses && gen_focal && gen_cxy, "dnr_encode_novel_bwd: null argument");
  DNR_REQUIRE(!d_latent_nhwc || d_lat_xr, "dnr_encode_novel_bwd: d_latent_nhwc needs d_lat_xr");
  DNR_REQUIRE(!d_gen_latent_nhwc || d_gen_xr || d_lat_xr, "dnr_encode_novel_bwd: d_gen_latent_nhwc needs d_gen_xr or d_lat_xr");
  DNR_REQUIRE(Hg >= 1 && Wg >= 1 && (unsigned long long)Hg * Wg * L < (1ull << 32), "dnr_encode_novel_bwd: bad gen_latent size");
  P.pts_in = pts_in; P.pts_gen = pts_gen; P.gen_poses = gen_poses; P.gen_focal = gen_focal; P.gen_cxy = gen_cxy;
  P.Hg = Hg; P.Wg = Wg; P.gen_img_w = gen_img_w; P.gen_img_h = gen_img_h; P.gen_sx = gen_sx; P.gen_sy = gen_sy;
  P.d_lat_xr = d_lat_xr; P.d_gen_xr = d_gen_xr; P.d_latent = d_latent_nhwc; P.d_gen_latent = d_gen_latent_nhwc;
  if (P.n_samples <= 0 || (!d_latent_nhwc && !d_gen_latent_nhwc)) return 0;
  DNR_CHECK(cudaFuncSetAttribute(dnr::encode_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 dnr::kBwdSmemBytes), "dnr_encode_novel_bwd");
  const long long slabs = (P.n_samples + dnr::RowMap::make(NV).SPW - 1) / dnr::RowMap::make(NV).SPW;
  dnr::encode_bwd_kernel<true><<<(unsigned)slabs, 128, dnr::kBwdSmemBytes, stream>>>(P);
  DNR_CHECK_LAUNCH("dnr_encode_novel_bwd");
  return 0;
}

// Per-scene repack of encoder.latent (SB*NV, C, Hl, Wl) fp32 -> channels-last bf16.
DNR_EXPORT int dnr_pack_latent_nhwc(const float* latent_nchw, int n_maps, int C, int Hl, int Wl, void* out_bf16,
                                    cudaStream_t stream) {
  const int HW = Hl * Wl;
  dim3 grid((HW + 31) / 32, (C + 31) / 32, n_maps), block(32, 8);
  dnr::nchw_to_nhwc_bf16_kernel<<<grid, block, 0, stream>>>(latent_nchw, (__nv_bfloat16*)out_bf16, C, HW);
  DNR_CHECK_LAUNCH("dnr_pack_latent_nhwc");
  return 0;
}

DNR_EXPORT int dnr_unpack_dlatent_nchw(const float* d_latent_nhwc, int n_maps, int C, int Hl, int Wl, float* out_nchw,
                                       cudaStream_t stream) {
  const int HW = Hl * Wl;
  dim3 grid((HW + 31) / 32, (C + 31) / 32, n_maps), block(32, 8);
  dnr::nhwc_to_nchw_f32_kernel<<<grid, block, 0, stream>>>(d_latent_nhwc, out_nchw, C, HW);
  DNR_CHECK_LAUNCH("dnr_unpack_dlatent_nchw");
  return 0;
}
