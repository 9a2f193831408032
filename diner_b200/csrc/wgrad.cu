// Weight-gradient GEMM of the field MLP on tcgen05:
//
//   dW[n, k] += sum_rows dY[row, n] * X[row, k]        db[n] += sum_rows dY[row, n]
//
// Both operands are the packed PK<128> activations / gradients the forward and dgrad GEMMs already
// produced; read as MN-major operands (rows = the MMA's K dimension) the very same bytes are a valid
// canonical shared-memory image, so no transposed copies exist anywhere.  Each CTA owns one
// [128 x NB] tile of dW and a contiguous range of 128-row tiles (split-K over the batch), accumulates in
// TMEM and reduces into the fp32 gradient with red.global.add.  The bias gradient rides along as one
// extra N=16 MMA per K step against a constant tile of ones.
#include "common.cuh"
#include "tc05.cuh"

namespace dnr {

constexpr int kWgStages = 2;
constexpr int kWgAStage = 128 * 128 * 2;     // dY: 16 feature chunks x 2 KB
constexpr int kWgThreads = 192;

struct WgradParams {
  const unsigned char* dY;   // PK<128> [rows_pad x Fy] bf16
  const unsigned char* X;    // PK<128> [rows_pad x Fx] bf16
  int Fy, Fx;                // Fy multiple of 128; Fx = 64 or multiple of 256
  int NB;                    // B tile width: min(Fx, 256)
  int row_tiles;             // rows_pad / 128
  int splits;                // CTAs along the batch
  float* dW;                 // [Fy_valid x ldw] fp32 row-major (accumulated into)
  int ldw, n_valid, k_valid; // only n < n_valid, k < k_valid are written
  float* db;                 // [n_valid] or null
  int copy_piece;            // bytes per bulk copy of the X tile
};

__global__ void __launch_bounds__(kWgThreads, 1) mlp_wgrad_kernel(const WgradParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int b_stage = P.NB * 128 * 2;
  unsigned char* sA = smem;
  unsigned char* sB = smem + (size_t)kWgStages * kWgAStage;
  unsigned char* sOnes = sB + (size_t)kWgStages * b_stage;          // 16 features x 128 rows of 1.0 (4 KB)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sOnes + 4096);
  uint64_t* full = bars;
  uint64_t* empty = bars + kWgStages;
  uint64_t* done = bars + 2 * kWgStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k_tiles = P.Fx / P.NB;
  const int out_tile = blockIdx.x % ((P.Fy / 128) * k_tiles);
  const int split = blockIdx.x / ((P.Fy / 128) * k_tiles);
  const int n_tile = out_tile / k_tiles, k_tile = out_tile % k_tiles;
  const int per = (P.row_tiles + P.splits - 1) / P.splits;
  const int rt0 = split * per, rt1 = min(P.row_tiles, rt0 + per);
  const bool with_bias = (P.db != nullptr) && (k_tile == 0);

  for (int i = threadIdx.x; i < 4096 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sOnes)[i] = 0x3f803f80u;
  tc::fence_proxy_async();
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kWgStages; ++s) { tc::mbar_init(full + s, 1); tc::mbar_init(empty + s, 1); }
    tc::mbar_init(done, 1);
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  if (rt0 >= rt1) {   // nothing to do for this split (uniform per CTA)
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc(tmem_base, 512);
    return;
  }

  // producer and MMA issuer: the whole warp runs the loop (warp-uniform operands -> uniform registers), one elected lane
  // issues (tc::elect_one)
  if (warp == 0) {
    {
      int stage = 0; uint32_t phase = 0;
      for (int rt = rt0; rt < rt1; ++rt) {
        tc::mbar_wait(empty + stage, phase ^ 1);
        if (tc::elect_one()) {
          tc::mbar_arrive_expect_tx(full + stage, kWgAStage + b_stage);
          tc::bulk_g2s(sA + (size_t)stage * kWgAStage, P.dY + ((size_t)rt * (P.Fy / 8) + (size_t)n_tile * 16) * 2048,
                       kWgAStage, full + stage);
          const unsigned char* xsrc = P.X + ((size_t)rt * (P.Fx / 8) + (size_t)k_tile * (P.NB / 8)) * 2048;
          const int piece = P.copy_piece;
          for (int off = 0; off < b_stage; off += piece)
            tc::bulk_g2s(sB + (size_t)stage * b_stage + off, xsrc + off, piece, full + stage);
        }
        __syncwarp();
        if (++stage == kWgStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    {
      const uint32_t idesc = tc::make_idesc_bf16(128, P.NB, 1, 1);
      const uint32_t idesc_ones = tc::make_idesc_bf16(128, 16, 1, 1);
      const uint32_t ones0 = tc::smem_u32(sOnes);
      int stage = 0; uint32_t phase = 0;
      uint32_t accumulate = 0;
      for (int rt = rt0; rt < rt1; ++rt) {
        tc::mbar_wait(full + stage, phase);
        tc::fence_after_sync();
        const uint32_t a0 = tc::smem_u32(sA + (size_t)stage * kWgAStage);
        const uint32_t b0 = tc::smem_u32(sB + (size_t)stage * b_stage);
        if (tc::elect_one()) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            // MN-major: 16 batch rows per MMA = two 8-row groups 128 B apart; feature groups 2 KB apart
            const uint64_t ad = tc::make_smem_desc(a0 + j * 256, 128, 2048);
            const uint64_t bd = tc::make_smem_desc(b0 + j * 256, 128, 2048);
            tc::mma_bf16(tmem_base, ad, bd, idesc, j == 0 ? accumulate : 1u);
            if (with_bias) {
              const uint64_t od = tc::make_smem_desc(ones0 + j * 256, 128, 2048);
              tc::mma_bf16(tmem_base + 256, ad, od, idesc_ones, j == 0 ? accumulate : 1u);
            }
          }
          tc::mma_commit(empty + stage);
        }
        __syncwarp();
        accumulate = 1;
        if (++stage == kWgStages) { stage = 0; phase ^= 1; }
      }
      if (tc::elect_one()) tc::mma_commit(done);
      __syncwarp();
    }
  } else {
    const int quad = warp & 3;
    const int n = n_tile * 128 + quad * 32 + lane;
    tc::mbar_wait(done, 0);
    tc::fence_after_sync();
    const uint32_t t0 = tmem_base + ((uint32_t)(quad * 32) << 16);
    for (int ch = 0; ch < P.NB / 32; ++ch) {
      uint32_t acc[32];
      tc::tmem_ld32(t0 + ch * 32, acc);
      tc::tmem_ld_wait();
      if (n < P.n_valid) {
        const int k0 = k_tile * P.NB + ch * 32;
        float* dst = P.dW + (size_t)n * P.ldw + k0;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (k0 + j < P.k_valid) atomicAdd(dst + j, __uint_as_float(acc[j]));
      }
    }
    if (with_bias) {
      uint32_t acc[32];
      tc::tmem_ld32(t0 + 256, acc);
      tc::tmem_ld_wait();
      if (n < P.n_valid) atomicAdd(P.db + n, __uint_as_float(acc[0]));
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 1) { tc::fence_after_sync(); tc::tmem_dealloc(tmem_base, 512); }
}

// ---- CTA-pair variant (Fy, Fx multiples of 256) ---------------------------------------------------------------------
// The single-CTA kernel pulls 96 KB (32 KB of dY + 64 KB of X) from L2 per 128-row batch tile for 8 MMAs, i.e.
// 96 B/clk/SM at full tensor rate against the ~47 B/clk/SM the L2 -> SM fabric sustains chip-wide: it is
// L2-bandwidth-bound (measured 41 B/clk/SM, 1 000 TFLOP/s).  Here the two CTAs of a cluster own two adjacent 128-row
// tiles of dW (n tiles 2i, 2i+1) and share the X operand through tcgen05 cta_group::2 (M = 256): each CTA streams its
// own dY tile and only HALF of the X tile (64 KB per batch tile), one thread of the even CTA issues for both SMs, the
// odd CTA forwards "my operands landed" with one remote mbarrier arrival per stage, and the commits are multicast to
// the barriers of both CTAs.
constexpr int kW2Stages = 3;
constexpr int kW2Op = 128 * 128 * 2;          // 32 KB: own dY tile / own half of the X tile

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kWgThreads, 1) mlp_wgrad2_kernel(const WgradParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* sA = smem;
  unsigned char* sB = smem + (size_t)kW2Stages * kW2Op;
  unsigned char* sOnes = sB + (size_t)kW2Stages * kW2Op;           // 16 features x 128 rows of 1.0 (4 KB)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sOnes + 4096);
  uint64_t* full = bars;                       // [S] my operands of the stage landed
  uint64_t* empty = bars + kW2Stages;          // [S] stage consumed in both CTAs (multicast commit)
  uint64_t* peer_full = bars + 2 * kW2Stages;  // [S] (even CTA) the odd CTA's operands landed
  uint64_t* done = bars + 3 * kW2Stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cr = tc::cluster_ctarank();
  const int pair = blockIdx.x >> 1;
  const int k_tiles = P.Fx / 256, n_pairs = P.Fy / 256;
  const int out = pair % (n_pairs * k_tiles), split = pair / (n_pairs * k_tiles);
  const int n_tile = 2 * (out / k_tiles) + (int)cr, k_tile = out % k_tiles;
  const int per = (P.row_tiles + P.splits - 1) / P.splits;
  const int rt0 = split * per, rt1 = min(P.row_tiles, rt0 + per);          // the same range in both CTAs of the pair
  const bool with_bias = (P.db != nullptr) && (k_tile == 0);

  for (int i = threadIdx.x; i < 4096 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sOnes)[i] = 0x3f803f80u;
  tc::fence_proxy_async();
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kW2Stages; ++s) { tc::mbar_init(full + s, 1); tc::mbar_init(empty + s, 1); tc::mbar_init(peer_full + s, 1); }
    tc::mbar_init(done, 1);
    tc::fence_barrier_init();
  }
  __syncthreads();
  tc::cluster_sync_all();                      // the peer's barriers and ones tile exist before anything signals / reads them
  if (warp == 1) tc::tmem_alloc_2cta(tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::cluster_sync_all();
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (rt0 < rt1) {
    if (warp == 0) {
      {     // whole warp, one elected lane issues the copies (tc::elect_one)
        int stage = 0; uint32_t phase = 0;
        for (int rt = rt0; rt < rt1; ++rt) {
          tc::mbar_wait(empty + stage, phase ^ 1);
          if (tc::elect_one()) {
            tc::mbar_arrive_expect_tx(full + stage, 2 * kW2Op);
            tc::bulk_g2s(sA + (size_t)stage * kW2Op, P.dY + ((size_t)rt * (P.Fy / 8) + (size_t)n_tile * 16) * 2048, kW2Op,
                         full + stage);
            tc::bulk_g2s(sB + (size_t)stage * kW2Op,
                         P.X + ((size_t)rt * (P.Fx / 8) + (size_t)k_tile * 32 + (size_t)cr * 16) * 2048, kW2Op, full + stage);
          }
          __syncwarp();
          if (++stage == kW2Stages) { stage = 0; phase ^= 1; }
        }
      }
    } else if (warp == 1) {
      if (cr == 0) {        // whole warp, one elected lane issues
        constexpr uint32_t idesc = tc::make_idesc_bf16(256, 256, 1, 1);
        constexpr uint32_t idesc_ones = tc::make_idesc_bf16(256, 32, 1, 1);
        const uint32_t ones0 = tc::smem_u32(sOnes);
        int stage = 0; uint32_t phase = 0;
        uint32_t accumulate = 0;
        for (int rt = rt0; rt < rt1; ++rt) {
          tc::mbar_wait(full + stage, phase);
          tc::mbar_wait(peer_full + stage, phase);
          tc::fence_after_sync();
          const uint32_t a0 = tc::smem_u32(sA + (size_t)stage * kW2Op);
          const uint32_t b0 = tc::smem_u32(sB + (size_t)stage * kW2Op);
          if (tc::elect_one()) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint64_t ad = tc::make_smem_desc(a0 + j * 256, 128, 2048);
              tc::mma_bf16_2cta(tmem_base, ad, tc::make_smem_desc(b0 + j * 256, 128, 2048), idesc, j == 0 ? accumulate : 1u);
              if (with_bias)
                tc::mma_bf16_2cta(tmem_base + 256, ad, tc::make_smem_desc(ones0 + j * 256, 128, 2048), idesc_ones,
                                  j == 0 ? accumulate : 1u);
            }
            tc::mma_commit_2cta_mc(empty + stage, (uint16_t)3);
          }
          __syncwarp();
          accumulate = 1;
          if (++stage == kW2Stages) { stage = 0; phase ^= 1; }
        }
        if (tc::elect_one()) tc::mma_commit_2cta_mc(done, (uint16_t)3);
        __syncwarp();
      } else if (lane == 0) {
        int stage = 0; uint32_t phase = 0;
        for (int rt = rt0; rt < rt1; ++rt) {
          tc::mbar_wait(full + stage, phase);
          tc::mbar_arrive_remote_relaxed(peer_full + stage, 0);
          if (++stage == kW2Stages) { stage = 0; phase ^= 1; }
        }
      }
    } else {
      const int quad = warp & 3;
      const int n = n_tile * 128 + quad * 32 + lane;
      tc::mbar_wait(done, 0);
      tc::fence_after_sync();
      const uint32_t t0 = tmem_base + ((uint32_t)(quad * 32) << 16);
      for (int ch = 0; ch < 8; ++ch) {
        uint32_t acc[32];
        tc::tmem_ld32(t0 + ch * 32, acc);
        tc::tmem_ld_wait();
        if (n < P.n_valid) {
          const int k0 = k_tile * 256 + ch * 32;
          float* dst = P.dW + (size_t)n * P.ldw + k0;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (k0 + j < P.k_valid) atomicAdd(dst + j, __uint_as_float(acc[j]));
        }
      }
      if (with_bias) {
        uint32_t acc[32];
        tc::tmem_ld32(t0 + 256, acc);
        tc::tmem_ld_wait();
        if (n < P.n_valid) atomicAdd(P.db + n, __uint_as_float(acc[0]));
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::cluster_sync_all();                      // nobody exits or frees TMEM while the peer may still signal or read it
  if (warp == 1) { tc::fence_after_sync(); tc::tmem_dealloc_2cta(tmem_base, 512); }
}

}  // namespace dnr

// Weight / bias gradient of one nn.Linear of the field MLP (autograd of F.linear in resnetfc.py).
DNR_EXPORT int dnr_mlp_wgrad(const void* dY_pk, int Fy, const void* X_pk, int Fx, int row_tiles, float* dW, int ldw,
                             int n_valid, int k_valid, float* db, cudaStream_t stream) {
  DNR_REQUIRE(Fy > 0 && Fy % 128 == 0, "dnr_mlp_wgrad: Fy must be a multiple of 128");
  DNR_REQUIRE(Fx == 64 || (Fx > 0 && Fx % 256 == 0), "dnr_mlp_wgrad: Fx must be 64 or a multiple of 256");
  if (row_tiles <= 0) return 0;
  dnr::WgradParams P;
  P.dY = (const unsigned char*)dY_pk; P.X = (const unsigned char*)X_pk; P.Fy = Fy; P.Fx = Fx;
  P.NB = Fx < 256 ? Fx : 256;
  P.copy_piece = P.NB * 256;
  P.row_tiles = row_tiles; P.dW = dW; P.ldw = ldw; P.n_valid = n_valid; P.k_valid = k_valid; P.db = db;
  dnr::DeviceInfo* di = dnr::device_info();
  DNR_REQUIRE(di != nullptr, "dnr_mlp_wgrad: no current CUDA device");
  const int num_sms = di->num_sms;
  if (Fy % 256 == 0 && Fx % 256 == 0) {
    // CTA pairs sharing the X operand (cta_group::2)
    const int out_pairs = (Fy / 256) * (Fx / 256);
    int splits = num_sms / (2 * out_pairs);
    if (splits < 1) splits = 1;
    if (splits > row_tiles) splits = row_tiles;
    P.splits = splits;
    const size_t smem2 = (size_t)dnr::kW2Stages * 2 * dnr::kW2Op + 4096 + 256;
    if (!di->wgrad2_configured) {
      DNR_CHECK(cudaFuncSetAttribute(dnr::mlp_wgrad2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2),
                "dnr_mlp_wgrad/attr2");
      di->wgrad2_configured = true;
    }
    dnr::mlp_wgrad2_kernel<<<out_pairs * splits * 2, dnr::kWgThreads, smem2, stream>>>(P);
    DNR_CHECK_LAUNCH("dnr_mlp_wgrad");
    return 0;
  }
  const int out_tiles = (Fy / 128) * (Fx / P.NB);
  int splits = num_sms / out_tiles;
  if (splits < 1) splits = 1;
  if (splits > row_tiles) splits = row_tiles;
  P.splits = splits;
  const size_t smem = (size_t)dnr::kWgStages * (dnr::kWgAStage + (size_t)P.NB * 256) + 4096 + 256;
  if (smem > di->wgrad_smem) {
    DNR_CHECK(cudaFuncSetAttribute(dnr::mlp_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
              "dnr_mlp_wgrad/attr");
    di->wgrad_smem = smem;
  }
  dnr::mlp_wgrad_kernel<<<out_tiles * splits, dnr::kWgThreads, smem, stream>>>(P);
  DNR_CHECK_LAUNCH("dnr_mlp_wgrad");
  return 0;
}
