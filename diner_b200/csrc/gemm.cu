// tcgen05 GEMM for the ResnetFC field MLP (reference resnetfc.py:61-69,129-159), forward and dgrad:
//
//   D[rows x N] = sum_seg  A_seg[rows x K_seg] * W_seg[N x K_seg]^T        (bf16 in, fp32 accumulate in TMEM)
//
// with a fused epilogue (bias, ReLU mask, residual add, view mean-pool, bf16 repack).  Operands live in
// HBM in the packed PK layouts of common.cuh, i.e. already in the tensor core's canonical shared-memory
// image, so every pipeline stage is filled by two cp.async.bulk copies (TMA engine) and consumed by
// four tcgen05.mma (M=128, N=256, K=16) issued by one thread; accumulators are double-buffered in TMEM
// (2 x 256 columns) so the epilogue of tile i overlaps the main loop of tile i+1.  Persistent kernel:
// one CTA per SM, 10 warps = {bulk-copy producer, MMA issuer, 8 epilogue warps (two per TMEM lane quadrant)}.
#include "common.cuh"
#include "tc05.cuh"

namespace dnr {

constexpr int kBM = 128, kBN = 256, kBK = 64;
constexpr int kStages = 4;
constexpr int kAStage = kBM * kBK * 2;          // 16 KB
constexpr int kBStage = kBN * kBK * 2;          // 32 KB
constexpr int kGemmThreads = 320;     // producer, MMA issuer, 8 epilogue warps (two per TMEM lane quadrant)
constexpr size_t kGemmSmem = (size_t)kStages * (kAStage + kBStage) + 256;

struct GemmSeg {
  const unsigned char* A;   // PK<128> [rows_pad x K] bf16
  const unsigned char* B;   // PK<256> [N x K] bf16 (weights, K-major)
  int K;
};

struct GemmParams {
  GemmSeg seg[3];
  int nseg;
  int m_tiles;              // rows_pad / 128
  int N;                    // multiple of 256
  const float* bias;        // [N] or null
  const unsigned char* mask_pk;  // PK<128> [rows x N] bf16: keep acc where mask > 0 (ReLU backward) or null
  const float* res_in;      // XR [rows x N] fp32 added to the result, or null
  const unsigned char* res_in_pk;  // PK<128> [rows x N] bf16 added to the result instead of res_in (bf16 gradient stream), or null
  float* res_out;           // XR [rows x N] fp32, or null
  unsigned char* act_out;   // PK<128> [rows x N] bf16, or null
  unsigned char* act_lo_out;  // verification mode: PK<128> bf16 image of (v - bf16(v)) beside act_out, or null
  int relu_act;             // act_out = relu(v) instead of v
  // view mean-pool (combine layer): outputs are indexed by sample instead of by (sample, view) row
  int mean_nv;              // 0 = off, else NV
  int res_in_nv;            // > 0: res_in is indexed by SAMPLE and enters as res_in / NV (backward of the view mean)
  long long n_samples;
};

// Global operands of one 32-column epilogue chunk (residual stream, ReLU mask): loaded one chunk ahead of the
// TMEM read so that their HBM latency overlaps the previous chunk (and, for the first chunk, the wait for the MMAs).
struct EpiOperands {
  float4 r[8];
  uint4 m[4];
  uint4 rp[4];      // res_in_pk
};
__device__ __forceinline__ void epilogue_prefetch(const GemmParams& P, int m_tile, int n0, int row_in_tile,
                                                  EpiOperands& e) {
  const int N = P.N;
  if (P.mask_pk) {
    const unsigned char* mp = P.mask_pk + ((size_t)((size_t)m_tile * (N / 8) + n0 / 8) * 128 + row_in_tile) * 16;
#pragma unroll
    for (int c = 0; c < 4; ++c) e.m[c] = __ldg(reinterpret_cast<const uint4*>(mp + (size_t)c * 2048));
  }
  if (P.res_in_pk) {
    const unsigned char* rq = P.res_in_pk + ((size_t)((size_t)m_tile * (N / 8) + n0 / 8) * 128 + row_in_tile) * 16;
#pragma unroll
    for (int c = 0; c < 4; ++c) e.rp[c] = __ldg(reinterpret_cast<const uint4*>(rq + (size_t)c * 2048));
  }
  if (P.res_in && P.res_in_nv) {
    // d_x[(sample, view)] = d_mean[sample] / NV (resnetfc.py:9-14): the NV rows of a sample read the same pooled row
    const int NV = P.res_in_nv, SPW = 32 / NV, l = row_in_tile & 31;
    const long long sample = ((long long)m_tile * 4 + (row_in_tile >> 5)) * SPW + l / NV;
    const bool ok = l < SPW * NV && sample < P.n_samples;
    const float4* rp = reinterpret_cast<const float4*>(P.res_in) + ((size_t)(sample >> 7) * (N / 4) + n0 / 4) * 128 + (sample & 127);
    const float inv = 1.f / (float)NV;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      float4 r = ok ? __ldg(rp + (size_t)c * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
      r.x *= inv; r.y *= inv; r.z *= inv; r.w *= inv;
      e.r[c] = r;
    }
  } else if (P.res_in) {
    const float4* rp = reinterpret_cast<const float4*>(P.res_in) + ((size_t)m_tile * (N / 4) + n0 / 4) * 128 + row_in_tile;
#pragma unroll
    for (int c = 0; c < 8; ++c) e.r[c] = __ldg(rp + (size_t)c * 128);
  }
}

__device__ __forceinline__ void epilogue_chunk(const GemmParams& P, int m_tile, int n0, int row_in_tile, int lane,
                                               uint32_t (&acc)[32], const EpiOperands& e) {
  const int N = P.N;
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc[j]);
  if (P.bias) {
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 b = __ldg(reinterpret_cast<const float4*>(P.bias + n0 + j));
      v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
    }
  }
  if (P.mask_pk) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const uint4 m = e.m[c];
      const uint32_t mv[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (!(bf16lo(mv[q]) > 0.f)) v[c * 8 + q * 2] = 0.f;
        if (!(bf16hi(mv[q]) > 0.f)) v[c * 8 + q * 2 + 1] = 0.f;
      }
    }
  }
  if (P.res_in) {
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const float4 r = e.r[c];
      v[c * 4] += r.x; v[c * 4 + 1] += r.y; v[c * 4 + 2] += r.z; v[c * 4 + 3] += r.w;
    }
  }
  if (P.res_in_pk) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const uint4 r = e.rp[c];
      const uint32_t rv[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) { v[c * 8 + q * 2] += bf16lo(rv[q]); v[c * 8 + q * 2 + 1] += bf16hi(rv[q]); }
    }
  }
  // destination row: the (sample, view) row itself, or -- at the combine layer -- the sample
  long long out_row = (long long)m_tile * 128 + row_in_tile;
  bool write = true;
  if (P.mean_nv) {
    // rows of one sample are NV consecutive lanes of this warp (RowMap): average them (resnetfc.py:9-14, torch.mean
    // over views; same association as view_mean of the cluster kernels), the leader lane writes
    const int NV = P.mean_nv, SPW = 32 / NV;
    if (NV == 2) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = (v[j] + __shfl_xor_sync(0xffffffffu, v[j], 1)) * 0.5f;
    } else if (NV == 4) {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float t = v[j] + __shfl_xor_sync(0xffffffffu, v[j], 1);
        t += __shfl_xor_sync(0xffffffffu, t, 2);
        v[j] = t * 0.25f;
      }
    } else if (NV > 1) {
      const float nv = (float)NV;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float t = v[j];
        for (int d = 1; d < NV; ++d) t += __shfl_down_sync(0xffffffffu, v[j], d);
        v[j] = t / nv;                                               // same bits as the chain kernel (mlp_fused.cu)
      }
    }
    const long long slab = (long long)m_tile * 4 + (row_in_tile >> 5);
    const long long sample = slab * SPW + lane / NV;
    write = (lane % NV == 0) && (lane < SPW * NV) && (sample < P.n_samples);
    out_row = sample;
  }
  if (!write) return;
  const long long ot = out_row >> 7;
  const int orow = (int)(out_row & 127);
  if (P.res_out) {
    float4* op = reinterpret_cast<float4*>(P.res_out) + ((size_t)ot * (N / 4) + n0 / 4) * 128 + orow;
#pragma unroll
    for (int c = 0; c < 8; ++c) op[(size_t)c * 128] = make_float4(v[c * 4], v[c * 4 + 1], v[c * 4 + 2], v[c * 4 + 3]);
  }
  if (P.act_out) {
    const size_t aoff = ((size_t)((size_t)ot * (N / 8) + n0 / 8) * 128 + orow) * 16;
    unsigned char* ap = P.act_out + aoff;
    const bool relu = P.relu_act != 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t w[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float a = v[c * 8 + q * 2], b = v[c * 8 + q * 2 + 1];
        if (relu) { a = fmaxf(a, 0.f); b = fmaxf(b, 0.f); }
        w[q] = pack_bf16x2(a, b);
        v[c * 8 + q * 2] = a - bf16lo(w[q]); v[c * 8 + q * 2 + 1] = b - bf16hi(w[q]);      // residue (read below only)
      }
      *reinterpret_cast<uint4*>(ap + (size_t)c * 2048) = make_uint4(w[0], w[1], w[2], w[3]);
      if (P.act_lo_out)
        *reinterpret_cast<uint4*>(P.act_lo_out + aoff + (size_t)c * 2048) =
            make_uint4(pack_bf16x2(v[c * 8], v[c * 8 + 1]), pack_bf16x2(v[c * 8 + 2], v[c * 8 + 3]),
                       pack_bf16x2(v[c * 8 + 4], v[c * 8 + 5]), pack_bf16x2(v[c * 8 + 6], v[c * 8 + 7]));
    }
  }
}

__global__ void __launch_bounds__(kGemmThreads, 1) mlp_gemm_kernel(const GemmParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* sA = smem;                                   // [kStages][16 KB]
  unsigned char* sB = smem + (size_t)kStages * kAStage;       // [kStages][32 KB]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)kStages * (kAStage + kBStage));
  uint64_t* full = bars;                 // [kStages]  bulk copies landed
  uint64_t* empty = bars + kStages;      // [kStages]  MMAs that read the stage retired
  uint64_t* tfull = bars + 2 * kStages;  // [2]        accumulator complete
  uint64_t* tempty = tfull + 2;          // [2]        accumulator drained by the epilogue
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles_n = P.N / kBN;
  const int total_tiles = P.m_tiles * n_tiles_n;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kStages; ++s) { tc::mbar_init(full + s, 1); tc::mbar_init(empty + s, 1); }
    for (int b = 0; b < 2; ++b) { tc::mbar_init(tfull + b, 1); tc::mbar_init(tempty + b, 256); }
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= producer: bulk copies of packed operand K-blocks =================
    {     // the whole warp runs the loop (warp-uniform operands -> uniform registers); one elected lane issues the copies
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int m_tile = tile / n_tiles_n, n_tile = tile % n_tiles_n;
        for (int sg = 0; sg < P.nseg; ++sg) {
          const int K = P.seg[sg].K;
          const unsigned char* a = P.seg[sg].A + (size_t)m_tile * (K / 8) * 2048;
          const unsigned char* b = P.seg[sg].B + (size_t)n_tile * (K / 8) * 4096;
          for (int kb = 0; kb < K / kBK; ++kb) {
            tc::mbar_wait(empty + stage, phase ^ 1);
            if (tc::elect_one()) {
              tc::mbar_arrive_expect_tx(full + stage, kAStage + kBStage);
              tc::bulk_g2s(sA + (size_t)stage * kAStage, a + (size_t)kb * kAStage, kAStage, full + stage);
              tc::bulk_g2s(sB + (size_t)stage * kBStage, b + (size_t)kb * kBStage, kBStage, full + stage);
            }
            __syncwarp();
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    {     // whole warp, one elected lane issues (see tc::elect_one)
      constexpr uint32_t idesc = tc::make_idesc_bf16(kBM, kBN, 0, 0);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        tc::mbar_wait(tempty + buf, ((it >> 1) & 1) ^ 1);
        tc::fence_after_sync();
        const uint32_t d_tmem = tmem_base + (uint32_t)buf * kBN;
        uint32_t accumulate = 0;
        for (int sg = 0; sg < P.nseg; ++sg) {
          for (int kb = 0; kb < P.seg[sg].K / kBK; ++kb) {
            tc::mbar_wait(full + stage, phase);
            tc::fence_after_sync();
            const uint32_t a0 = tc::smem_u32(sA + (size_t)stage * kAStage);
            const uint32_t b0 = tc::smem_u32(sB + (size_t)stage * kBStage);
            if (tc::elect_one()) {
#pragma unroll
              for (int j = 0; j < kBK / 16; ++j) {
                // K-major: one MMA consumes two 8-element K chunks (A: 2 KB apart, B: 4 KB apart)
                const uint64_t ad = tc::make_smem_desc(a0 + j * 2 * 2048, 2048, 128);
                const uint64_t bd = tc::make_smem_desc(b0 + j * 2 * 4096, 4096, 128);
                tc::mma_bf16(d_tmem, ad, bd, idesc, j == 0 ? accumulate : 1u);
              }
              tc::mma_commit(empty + stage);
            }
            __syncwarp();
            accumulate = 1;
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        }
        if (tc::elect_one()) tc::mma_commit(tfull + buf);
        __syncwarp();
      }
    }
  } else {
    // ================= epilogue: TMEM -> registers -> HBM =================
    // two warps per TMEM lane quadrant; each takes four of the tile's eight 32-column chunks
    const int quad = warp & 3;                 // TMEM lane quadrant this warp may access
    const int half = (warp - 2) >> 2;
    const int row_in_tile = quad * 32 + lane;
    constexpr int kChunksPerWarp = kBN / 32 / 2;
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int m_tile = tile / n_tiles_n, n_tile = tile % n_tiles_n;
      const int buf = it & 1;
      const int n_base = n_tile * kBN + half * kChunksPerWarp * 32;
      EpiOperands e[2];
      epilogue_prefetch(P, m_tile, n_base, row_in_tile, e[0]);
      tc::mbar_wait(tfull + buf, (it >> 1) & 1);
      tc::fence_after_sync();
      const uint32_t t0 = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)buf * kBN + half * kChunksPerWarp * 32;
#pragma unroll
      for (int ch = 0; ch < kChunksPerWarp; ++ch) {
        uint32_t acc[32];
        tc::tmem_ld32(t0 + ch * 32, acc);
        if (ch + 1 < kChunksPerWarp) epilogue_prefetch(P, m_tile, n_base + (ch + 1) * 32, row_in_tile, e[(ch + 1) & 1]);
        tc::tmem_ld_wait();
        epilogue_chunk(P, m_tile, n_base + ch * 32, row_in_tile, lane, acc, e[ch & 1]);
      }
      tc::fence_before_sync();
      tc::mbar_arrive(tempty + buf);
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 1) { tc::fence_after_sync(); tc::tmem_dealloc(tmem_base, 512); }
}

}  // namespace dnr

// One fused linear layer of the field MLP (forward or dgrad), see GemmParams.
DNR_EXPORT int dnr_mlp_gemm(const void* A0, const void* B0, int K0, const void* A1, const void* B1, int K1,
                            const void* A2, const void* B2, int K2, int m_tiles, int N, const float* bias,
                            const void* mask_pk, const float* res_in, const void* res_in_pk, float* res_out, void* act_out,
                            void* act_lo_out, int relu_act, int mean_nv, long long n_samples, int max_ctas,
                            cudaStream_t stream) {
  DNR_REQUIRE(N > 0 && N % dnr::kBN == 0, "dnr_mlp_gemm: N must be a multiple of 256");
  DNR_REQUIRE(A0 && B0 && K0 > 0 && K0 % dnr::kBK == 0, "dnr_mlp_gemm: segment 0 invalid (K multiple of 64)");
  DNR_REQUIRE(K1 % dnr::kBK == 0 && K2 % dnr::kBK == 0, "dnr_mlp_gemm: K must be a multiple of 64");
  DNR_REQUIRE(mean_nv >= -8 && mean_nv <= 8, "dnr_mlp_gemm: mean_nv must be in [-8,8]");
  DNR_REQUIRE(mean_nv >= 0 || res_in != nullptr, "dnr_mlp_gemm: mean_nv < 0 (pooled residual input) needs res_in");
  if (m_tiles <= 0) return 0;
  dnr::GemmParams P;
  P.seg[0] = {(const unsigned char*)A0, (const unsigned char*)B0, K0};
  P.seg[1] = {(const unsigned char*)A1, (const unsigned char*)B1, K1};
  P.seg[2] = {(const unsigned char*)A2, (const unsigned char*)B2, K2};
  P.nseg = 1 + (A1 != nullptr && K1 > 0 ? 1 : 0);
  if (P.nseg == 2 && A2 != nullptr && K2 > 0) P.nseg = 3;
  P.m_tiles = m_tiles; P.N = N; P.bias = bias; P.mask_pk = (const unsigned char*)mask_pk; P.res_in = res_in; P.res_in_pk = (const unsigned char*)res_in_pk;
  DNR_REQUIRE(!(res_in && res_in_pk), "dnr_mlp_gemm: res_in and res_in_pk are exclusive");
  DNR_REQUIRE(!(res_in_pk && mean_nv != 0), "dnr_mlp_gemm: res_in_pk cannot be combined with a view mean");
  P.res_out = res_out; P.act_out = (unsigned char*)act_out; P.act_lo_out = (unsigned char*)act_lo_out; P.relu_act = relu_act;
  DNR_REQUIRE(act_lo_out == nullptr || act_out != nullptr, "dnr_mlp_gemm: act_lo_out needs act_out");
  P.mean_nv = mean_nv > 0 ? mean_nv : 0; P.res_in_nv = mean_nv < 0 ? -mean_nv : 0;
  P.n_samples = n_samples;
  dnr::DeviceInfo* di = dnr::device_info();
  DNR_REQUIRE(di != nullptr, "dnr_mlp_gemm: no current CUDA device");
  if (!di->gemm_configured) {          // function attributes are per device
    DNR_CHECK(cudaFuncSetAttribute(dnr::mlp_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)dnr::kGemmSmem), "dnr_mlp_gemm/attr");
    di->gemm_configured = true;
  }
  const int num_sms = di->num_sms;
  int grid = m_tiles * (N / dnr::kBN);
  int cap = (max_ctas > 0 && max_ctas < num_sms) ? max_ctas : num_sms;
  if (grid > cap) grid = cap;
  dnr::mlp_gemm_kernel<<<grid, dnr::kGemmThreads, dnr::kGemmSmem, stream>>>(P);
  DNR_CHECK_LAUNCH("dnr_mlp_gemm");
  return 0;
}
