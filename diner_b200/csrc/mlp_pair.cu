// ResnetFC field MLP, inference, fully on-chip: a CLUSTER OF TWO CTAs walks one 128-row view tile through all
// layers (reference resnetfc.py:129-159, combine() :9-14, head pixelnerf.py:137-145).
//
// CTA r of the pair owns output columns [256 r, 256 r + 256) of every layer:
//   TMEM   (512 cols)  X = fp32 residual stream of its 256 columns (fc_1 / lin_z accumulate straight into it, so the
//                      residual add costs nothing and never leaves the SM) + T = fc_0 accumulator
//   smem   128 KB      the full-K bf16 A operand (relu(x) or relu(net)) in the tcgen05 canonical layout: its own
//                      256 columns are written by its epilogue warps and pushed, one 16 KB K-block at a time, into the peer's
//                      buffer by cp.async.bulk shared::cta -> shared::cluster (DSMEM), which completes the peer's
//                      landing barrier; the peer does the same, so nothing of the exchange touches L2
//          3 x 32 KB   ring for the streamed operands: weight K-blocks of its 256 output rows, and the
//                      gathered-feature / code K-blocks for lin_in / lin_z
// so per layer a CTA pulls 256 KB of weights from L2 instead of ~1.1 MB of weights + activations, and nothing of
// the layer chain depends on HBM.  lin_z[b](lat) K-steps are issued FIRST in their layer: they do not depend on the
// previous epilogue and hide it.  After the view mean every row carries its sample's mean, i.e. blocks 3-4 run on
// NV duplicates (+18 % MMA work at NV=2) - the price of keeping the tile on chip.
//
//   warp 0 : ring producer      warp 1 : tcgen05.mma issuer      warp 2 : pusher      warp 3 : receiver (credits / acks)
//   warps 4..11 : epilogue (two per TMEM lane quadrant; four 32-column pieces each, the two warps of a quadrant
//                 complete one 64-column K-block per piece so the next layer's MMAs start after a quarter epilogue)
#include "common.cuh"
#include "tc05.cuh"
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

// in-kernel cycle counters (tools/pair_prof.py): compiled in only with -DDNR_CLUSTER_PROF=1 (DNR_CLUSTER_PROF=1 in the
// environment of diner_b200._build); a production build reads no clocks
#ifndef DNR_CLUSTER_PROF
#define DNR_CLUSTER_PROF 0
#endif
#if DNR_CLUSTER_PROF
#define TIMED(acc, stmt) { long long t__ = clock64(); stmt; acc += clock64() - t__; }
#define PROF_CLOCK() clock64()
#else
#define TIMED(acc, stmt) { stmt; }
#define PROF_CLOCK() 0LL
#endif

namespace dnr {

constexpr int kPThreads = 384;
constexpr int kPRing = 3;
constexpr int kPSlot = 32768;
constexpr int kPAbuf = 131072;
constexpr int kPHeadScratch = 2048;
constexpr size_t kPSmem = (size_t)kPAbuf + kPRing * kPSlot + kPHeadScratch + 256;
constexpr int kPOps = 11;

struct PairParams {
  const unsigned char* lat_pk;    // PK128 [mv_tiles x 512]
  const unsigned char* code_pk;   // PK128 [mv_tiles x 64]
  const unsigned char* w_in;      // PK256 [512 x 64]
  const unsigned char* w_z[3];
  const unsigned char* w_zc[2];   // NOVEL_PE, else null: [512 x 64] weights of a second code K-step in fc_1 of blocks 0 / 1
                                  // (lin_z[1], lin_z[2] applied to the six positional values in the code operand)
  const unsigned char* w_0[5];
  const unsigned char* w_1[5];
  const float* cb[6];             // cumulative residual-stream biases: after lin_in+z0, block 0, 1, 2 | block 3, 4
  const float* b_0[5];
  const float* w_out;             // fp32 [4 x 512]
  float4* head_partial;           // [mv_tiles][2][128]: partial lin_out sums of each CTA's 256 columns
  float* merge_scratch;           // quad kernel, MERGE: [gridDim][128 rows][256] fp32 parked pooled rows
  int NV, SPW, mv_tiles;
  long long* prof;                // optional [gridDim][8] cycle counters (debug)
  int tma2;                       // quad kernel: operands through tensor-map TMA with cta_group::2 completion (QuadMaps)
};

// Tensor maps of the quad kernel's streamed operands (tma2 mode): every packed buffer is viewed as a 2-D bf16 tensor of
// 256-element rows (512 B); a 16 KB K-block is the box of 32 consecutive rows starting at byte offset / 512.
struct QuadMaps {
  CUtensorMap w[14];              // lin_in, lin_z x3, fc_0 x5, fc_1 x5 (PK128 images)
  CUtensorMap lat, code;
  CUtensorMap wc[2];              // PairParams::w_zc
};

enum { EPI_X = 0, EPI_T = 1, EPI_MEAN = 2, EPI_HEAD = 3 };

struct POp {
  const unsigned char* lat_w;     // lin_z weights (A = streamed lat) or null
  const unsigned char* abuf_w;    // weights applied to the resident A operand or null
  const float* bias;
  const unsigned char* code_w;    // weights of the op's code K-step (lin_in in op 0; NOVEL_PE: w_zc in ops 2, 4) or null
  int code;                       // the op starts with a K-step on the 64-wide code operand
  int acc_x;                      // accumulator: X (1) or T (0)
  int fresh;                      // first MMA overwrites
  int epi;
};

__device__ __forceinline__ void pair_get_op(const PairParams& P, int o, POp& d) {
  d.lat_w = nullptr; d.abuf_w = nullptr; d.code_w = nullptr; d.code = 0; d.acc_x = 1; d.fresh = 0; d.epi = EPI_X;
  if (o == 0) { d.code = 1; d.code_w = P.w_in; d.lat_w = P.w_z[0]; d.fresh = 1; d.bias = P.cb[0]; return; }
  const int b = (o - 1) >> 1;
  if (o & 1) {                      // fc_0 of block b
    d.abuf_w = P.w_0[b]; d.acc_x = 0; d.fresh = 1; d.epi = EPI_T; d.bias = P.b_0[b];
  } else {                          // fc_1 of block b (+ lin_z[b+1])
    d.abuf_w = P.w_1[b];
    if (b < 2) {
      d.lat_w = P.w_z[b + 1];
      if (P.w_zc[b]) { d.code = 1; d.code_w = P.w_zc[b]; }
    }
    d.bias = P.cb[b + 1];
    d.epi = b == 4 ? EPI_HEAD : EPI_X;
  }
}

// K-block order of the resident-operand GEMMs: own K-blocks first, in the order the epilogue finishes them
// ({0,2} after its first 64-column stage, {1,3} after the second), then the peer's four.
__device__ __forceinline__ int pair_kb_order(int rank, int i) {
  return i < 4 ? 4 * rank + i : 4 * (rank ^ 1) + (i - 4);      // own K-blocks first, in the order the epilogue writes them
}

__device__ __forceinline__ void bar_sync_epi() { asm volatile("bar.sync 1, 256;" ::: "memory"); }


// mean over the NV view rows of a sample (adjacent lanes of the warp; RowMap)
template <int NV>
__device__ __forceinline__ float view_mean(float v, int lane) {
  if (NV == 1) return v;
  if (NV == 2) return (v + __shfl_xor_sync(0xffffffffu, v, 1)) * 0.5f;
  if (NV == 4) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    return v * 0.25f;
  }
  // NV == 3: the two other rows of the sample with two shuffles; every lane adds in view order (a0 + a1) + a2 so the
  // three rows of a sample hold the same bits.  Lanes 30, 31 are padding rows (SPW = 10).
  const int r = lane % 3, l0 = lane - r;
  const float x = __shfl_sync(0xffffffffu, v, min(l0 + (r == 2 ? 0 : r + 1), 31));      // a_{r+1 mod 3}
  const float y = __shfl_sync(0xffffffffu, v, min(l0 + (r == 0 ? 2 : r - 1), 31));      // a_{r+2 mod 3}
  const float A = r == 2 ? x : v, B = r == 0 ? x : y, C = r == 0 ? y : (r == 1 ? x : v);
  return ((A + B) + C) * (1.0f / 3.0f);
}

// one 32-column piece of an activation epilogue: (+bias, [view mean], relu, bf16) -> resident A buffer
template <int NV, bool POOLED>
__device__ __forceinline__ void pair_act_stage(uint32_t (&acc)[32], const float4 (&bq)[8],
                                               const float* __restrict__ bias, int lane, unsigned char* dst_s,
                                               uint32_t taddr) {
  if (POOLED) {
    // x (pooled) = mean over the view rows (resnetfc.py:9-14,148-151), taken ONCE: the mean is written back over the
    // per-view rows of the residual stream in TMEM.  Blocks 3-4 then add identical updates to identical rows, so
    // every later read of the stream (fc_1 of blocks 3/4, the head) already sees the pooled value in all NV rows.
#pragma unroll
    for (int j = 0; j < 32; ++j) acc[j] = __float_as_uint(view_mean<NV>(__uint_as_float(acc[j]), lane));
    tc::tmem_st32(taddr, acc);
  }
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    // biases: prefetched into registers by the caller, except for NV = 3 whose three-way mean needs the registers
    const float4 b0 = NV == 3 ? __ldg(reinterpret_cast<const float4*>(bias) + 2 * g) : bq[2 * g];
    const float4 b1 = NV == 3 ? __ldg(reinterpret_cast<const float4*>(bias) + 2 * g + 1) : bq[2 * g + 1];
    const float v[8] = {__uint_as_float(acc[g * 8 + 0]) + b0.x, __uint_as_float(acc[g * 8 + 1]) + b0.y,
                        __uint_as_float(acc[g * 8 + 2]) + b0.z, __uint_as_float(acc[g * 8 + 3]) + b0.w,
                        __uint_as_float(acc[g * 8 + 4]) + b1.x, __uint_as_float(acc[g * 8 + 5]) + b1.y,
                        __uint_as_float(acc[g * 8 + 6]) + b1.z, __uint_as_float(acc[g * 8 + 7]) + b1.w};
    const uint32_t w0 = pack_bf16x2(fmaxf(v[0], 0.f), fmaxf(v[1], 0.f)), w1 = pack_bf16x2(fmaxf(v[2], 0.f), fmaxf(v[3], 0.f));
    const uint32_t w2 = pack_bf16x2(fmaxf(v[4], 0.f), fmaxf(v[5], 0.f)), w3 = pack_bf16x2(fmaxf(v[6], 0.f), fmaxf(v[7], 0.f));
    *reinterpret_cast<uint4*>(dst_s + (size_t)g * 2048) = make_uint4(w0, w1, w2, w3);
  }
  if (POOLED) tc::tmem_st_wait();       // before this warp signals the K-block: the next MMAs on X come after that signal
}

// one 32-column piece of the head: partial lin_out sums on the bf16-rounded relu(x) (contract of dnr_head_fwd); the
// stream is already pooled (all NV rows of a sample equal, see pair_act_stage)
__device__ __forceinline__ void pair_head_stage(const uint32_t (&acc)[32], const float4 (&bq)[8],
                                                const float* __restrict__ wo, float& h0, float& h1, float& h2,
                                                float& h3) {
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const float4 b = bq[q];
    float v[4] = {fmaxf(__uint_as_float(acc[q * 4 + 0]) + b.x, 0.f), fmaxf(__uint_as_float(acc[q * 4 + 1]) + b.y, 0.f),
                  fmaxf(__uint_as_float(acc[q * 4 + 2]) + b.z, 0.f), fmaxf(__uint_as_float(acc[q * 4 + 3]) + b.w, 0.f)};
    {   // round to bf16 two at a time (one packed convert instead of two quarter-rate scalar ones)
      const uint32_t p01 = pack_bf16x2(v[0], v[1]), p23 = pack_bf16x2(v[2], v[3]);
      v[0] = bf16lo(p01); v[1] = bf16hi(p01); v[2] = bf16lo(p23); v[3] = bf16hi(p23);
    }
    const float4 w0 = __ldg(reinterpret_cast<const float4*>(wo + q * 4));
    const float4 w1 = __ldg(reinterpret_cast<const float4*>(wo + 512 + q * 4));
    const float4 w2 = __ldg(reinterpret_cast<const float4*>(wo + 1024 + q * 4));
    const float4 w3 = __ldg(reinterpret_cast<const float4*>(wo + 1536 + q * 4));
    h0 = fmaf(v[0], w0.x, fmaf(v[1], w0.y, fmaf(v[2], w0.z, fmaf(v[3], w0.w, h0))));
    h1 = fmaf(v[0], w1.x, fmaf(v[1], w1.y, fmaf(v[2], w1.z, fmaf(v[3], w1.w, h1))));
    h2 = fmaf(v[0], w2.x, fmaf(v[1], w2.y, fmaf(v[2], w2.z, fmaf(v[3], w2.w, h2))));
    h3 = fmaf(v[0], w3.x, fmaf(v[1], w3.y, fmaf(v[2], w3.z, fmaf(v[3], w3.w, h3))));
  }
}

template <int NV>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPThreads, 1) mlp_pair_fwd_kernel(const PairParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* abuf = smem;                                   // 128 KB resident A operand
  unsigned char* ring = smem + kPAbuf;                          // 3 x 32 KB
  float* head_s = reinterpret_cast<float*>(ring + kPRing * kPSlot);      // [128][4]
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + kPRing * kPSlot + kPHeadScratch);
  uint64_t* full = bars;                  // [3]  ring slot landed
  uint64_t* empty = bars + 3;             // [3]  ring slot consumed
  uint64_t* accf = bars + 6;              // [2]  accumulator of op (parity) complete
  uint64_t* a_own = bars + 12;            // [4]  own K-block p of the A operand written (one arrival per epilogue warp)
  uint64_t* a_peer = bars + 8;            // [4]  peer K-block p landed (one 16 KB bulk copy over DSMEM each)
  uint64_t* pushed = bars + 16;           // [1]  the peer has received my whole half (remote ack): my half may be rewritten
  uint64_t* peer_free = bars + 17;        // [1]  the peer's A buffer may receive my next operand (remote credit)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = tc::cluster_ctarank();
  const uint32_t peer = rank ^ 1u;
  const int cluster = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kPRing; ++s) { tc::mbar_init(full + s, 1); tc::mbar_init(empty + s, 1); }
    tc::mbar_init(accf + 0, 1); tc::mbar_init(accf + 1, 1);
    for (int i = 0; i < 4; ++i) tc::mbar_init(a_own + i, 8);
    for (int i = 0; i < 4; ++i) tc::mbar_init(a_peer + i, 1);
    tc::mbar_init(pushed, 1); tc::mbar_init(peer_free, 1);
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::cluster_sync_all();               // the peer's barriers exist before anyone arrives on them remotely
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= ring producer =================
    {     // whole warp (warp-uniform operands -> uniform registers), one elected lane issues the copies
      const uint64_t pol_keep = tc::policy_evict_last(), pol_stream = tc::policy_evict_first();
      const uint64_t pol_norm = tc::policy_evict_normal();
      int slot = 0; uint32_t phase = 0;
      auto fill = [&](const unsigned char* src, uint32_t bytes, uint64_t pol) {
        tc::mbar_wait(empty + slot, phase ^ 1);
        if (tc::elect_one()) {
          tc::mbar_arrive_expect_tx(full + slot, bytes);
          tc::bulk_g2s_hint(ring + (size_t)slot * kPSlot, src, bytes, full + slot, pol);
        }
        __syncwarp();
        if (++slot == kPRing) { slot = 0; phase ^= 1; }
      };
      POp d;
      for (int tile = cluster; tile < P.mv_tiles; tile += n_clusters) {
        const unsigned char* lat = P.lat_pk + (size_t)tile * 131072;
        if (tile + n_clusters < P.mv_tiles && lane == 0) {     // pull the next tile's inputs from HBM into L2 while this one computes
          tc::bulk_prefetch_l2(P.lat_pk + (size_t)(tile + n_clusters) * 131072, 131072);
          tc::bulk_prefetch_l2(P.code_pk + (size_t)(tile + n_clusters) * 16384, 16384);
        }
        for (int o = 0; o < kPOps; ++o) {
          pair_get_op(P, o, d);
          if (d.code) {
            fill(P.code_pk + (size_t)tile * 16384, 16384, (o == 4 || !P.w_zc[0]) ? pol_stream : pol_norm);
            fill(d.code_w + (size_t)rank * 32768, 32768, pol_keep);
          }
          if (d.lat_w) {
            for (int kb = 0; kb < 8; ++kb) {
              fill(lat + (size_t)kb * 16384, 16384, o == 4 ? pol_stream : pol_norm);   // last use: evict first
              fill(d.lat_w + (size_t)rank * 262144 + (size_t)kb * 32768, 32768, pol_keep);
            }
          }
          if (d.abuf_w) {
            for (int i = 0; i < 8; ++i) {
              const int kb = pair_kb_order((int)rank, i);
              fill(d.abuf_w + (size_t)rank * 262144 + (size_t)kb * 32768, 32768, pol_keep);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    {     // whole warp (warp-uniform operands -> uniform registers), one elected lane issues
      constexpr uint32_t idesc = tc::make_idesc_bf16(128, 256, 0, 0);
      int slot = 0; uint32_t phase = 0;
      int opc = 0;                           // ops issued (all kinds)
      int aops = 0;                          // ops that read the resident A operand
      int titer = 0;
      POp d;
      const uint32_t abuf0 = tc::smem_u32(abuf), ring0 = tc::smem_u32(ring);
      long long w_ring = 0, w_own = 0, w_peer = 0, t_start = PROF_CLOCK();

      for (int tile = cluster; tile < P.mv_tiles; tile += n_clusters, ++titer) {
        const uint32_t x_col = (titer & 1) ? 256u : 0u;           // X / T swap regions every tile
        for (int o = 0; o < kPOps; ++o, ++opc) {
          pair_get_op(P, o, d);
          const uint32_t d_tmem = tmem_base + (d.acc_x ? x_col : (256u - x_col));
          uint32_t accumulate = d.fresh ? 0u : 1u;
          auto step_ring_ab = [&]() {        // A and B both come from consecutive ring slots
            uint64_t ad0 = tc::make_smem_desc(ring0 + (uint32_t)slot * kPSlot, 2048, 128);
            uint64_t bd0 = tc::make_smem_desc(ring0 + (uint32_t)(slot + 1 == kPRing ? 0 : slot + 1) * kPSlot, 4096, 128);
            asm volatile("" : "+l"(ad0), "+l"(bd0));     // (before the waits: off the path "operands landed" -> first MMA)
            TIMED(w_ring, tc::mbar_wait(full + slot, phase));
            const int sa = slot;
            if (++slot == kPRing) { slot = 0; phase ^= 1; }
            TIMED(w_ring, tc::mbar_wait(full + slot, phase));
            const int sb = slot;
            if (++slot == kPRing) { slot = 0; phase ^= 1; }
            tc::fence_after_sync();
            if (tc::elect_one()) {
#pragma unroll
              for (int j = 0; j < 4; ++j)        // K = 16 further: 4 KB in A (+256), 8 KB in B (+512, 256 weight rows)
                tc::mma_bf16(d_tmem, ad0 + (uint64_t)(j * 256), bd0 + (uint64_t)(j * 512), idesc, j == 0 ? accumulate : 1u);
              tc::mma_commit(empty + sa);
              tc::mma_commit(empty + sb);
            }
            __syncwarp();
            accumulate = 1;
          };
          if (d.code) step_ring_ab();
          if (d.lat_w)
            for (int kb = 0; kb < 8; ++kb) step_ring_ab();
          if (d.abuf_w) {
            const uint32_t par = (uint32_t)((aops) & 1);
            for (int i = 0; i < 8; ++i) {
              const int kb = pair_kb_order((int)rank, i);
              uint64_t ad0 = tc::make_smem_desc(abuf0 + (uint32_t)kb * 16384u, 2048, 128);
              uint64_t bd0 = tc::make_smem_desc(ring0 + (uint32_t)slot * kPSlot, 4096, 128);
              asm volatile("" : "+l"(ad0), "+l"(bd0));
              if (i < 4) TIMED(w_own, tc::mbar_wait(a_own + i, par));
              if (i >= 4) TIMED(w_peer, tc::mbar_wait(a_peer + (i - 4), par));
              TIMED(w_ring, tc::mbar_wait(full + slot, phase));
              tc::fence_after_sync();
              if (tc::elect_one()) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  tc::mma_bf16(d_tmem, ad0 + (uint64_t)(j * 256), bd0 + (uint64_t)(j * 512), idesc, j == 0 ? accumulate : 1u);
                tc::mma_commit(empty + slot);
              }
              __syncwarp();
              accumulate = 1;
              if (++slot == kPRing) { slot = 0; phase ^= 1; }
            }
            ++aops;
          }
          if (tc::elect_one()) tc::mma_commit(accf + (opc & 1));
          __syncwarp();
        }
      }
      if (P.prof && lane == 0) {
        long long* pr = P.prof + (size_t)blockIdx.x * 16;
        pr[0] = PROF_CLOCK() - t_start; pr[1] = w_ring; pr[2] = w_own; pr[3] = w_peer;
      }
    }
  } else if (warp == 2) {
    // ================= pusher: my finished K-blocks -> the peer's resident A buffer (bulk copy over DSMEM) =============
    if (lane == 0) {
      int xops = 0;
      for (int tile = cluster; tile < P.mv_tiles; tile += n_clusters) {
        for (int o = 0; o < kPOps - 1; ++o, ++xops) {          // every op but the head produces an A operand
          unsigned char* mine = abuf + (size_t)rank * 65536;
          for (int kb = 0; kb < 4; ++kb) {
            // all epilogue warps wrote (and proxy-fenced) K-block kb: push it now (the peer's MMAs need its first remote
            // K-block four K-steps after their first own one)
            tc::mbar_wait(a_own + kb, (uint32_t)(xops & 1));
            if (kb == 0) tc::mbar_wait(peer_free, (uint32_t)(xops & 1));   // peer's MMAs left its buffer, tx armed
            tc::bulk_s2s_peer(mine + (size_t)kb * 16384, 16384, a_peer + kb, peer);
          }
        }
      }
    }
  } else if (warp == 3) {
    // ================= receiver: arm the landing barriers, grant the credit, acknowledge =================
    if (lane == 0) {
      int opc = 0, xops = 0;
      for (int tile = cluster; tile < P.mv_tiles; tile += n_clusters) {
        for (int o = 0; o < kPOps; ++o, ++opc) {
          if (o == kPOps - 1) continue;
          tc::mbar_wait(accf + (opc & 1), (uint32_t)((opc >> 1) & 1));     // my MMAs of this op no longer read the buffer
          for (int e = 0; e < 4; ++e) tc::mbar_arrive_expect_tx(a_peer + e, 16384);
          tc::mbar_arrive_remote_relaxed(peer_free, peer);
          for (int e = 3; e >= 0; --e) tc::mbar_wait(a_peer + e, (uint32_t)(xops & 1));     // everything of this operand has landed
          tc::mbar_arrive_remote_relaxed(pushed, peer);
          ++xops;
        }
      }
    }
  }
  else if (warp >= 4) {
    // ================= epilogue =================
    const int quad = warp & 3;
    const int half = (warp - 4) >> 2;
    const int row = quad * 32 + lane;
    const bool leader = (warp == 4 && lane == 0);
    int opc = 0, xops = 0, titer = 0;
    long long w_acc = 0, w_fence = 0, w_sig = 0, w_push = 0, t_epi0 = PROF_CLOCK();
    long long w_kind[4] = {0, 0, 0, 0};
    POp d;
    for (int tile = cluster; tile < P.mv_tiles; tile += n_clusters, ++titer) {
      const uint32_t x_col = (titer & 1) ? 256u : 0u;
      for (int o = 0; o < kPOps; ++o, ++opc) {
        pair_get_op(P, o, d);
        TIMED(w_acc, tc::mbar_wait(accf + (opc & 1), (uint32_t)((opc >> 1) & 1)));
        tc::fence_after_sync();
        const uint32_t t0 = tmem_base + ((uint32_t)(quad * 32) << 16) + (d.acc_x ? x_col : (256u - x_col));
        const bool pooled = o == 6;                            // fc_1 of block 2: the view mean (taken once, written back)
        // the previous operand has left my half of the A buffer (pusher's bulk stores have read it)
        if (d.epi != EPI_HEAD && xops > 0) TIMED(w_push, tc::mbar_wait(pushed, (uint32_t)((xops - 1) & 1)));
        float hp0 = 0.f, hp1 = 0.f, hp2 = 0.f, hp3 = 0.f;
        const long long tl0 = PROF_CLOCK();
#pragma unroll 1
        for (int st = 0; st < 4; ++st) {                       // four 32-column pieces; the two halves share K-block st
          uint32_t acc[32];
          const int nl0 = st * 64 + half * 32;                 // first column inside the CTA's 256
          tc::tmem_ld32(t0 + nl0, acc);
          float4 bq[8];                                        // the piece's biases travel while the TMEM load does
          if (NV != 3) {                                       // (NV = 3: the three-way mean needs the registers)
#pragma unroll
            for (int g = 0; g < 8; ++g) bq[g] = __ldg(reinterpret_cast<const float4*>(d.bias + (int)rank * 256 + nl0) + g);
          }
          tc::tmem_ld_wait();
          const float* bias = d.bias + (int)rank * 256 + nl0;
          if (NV == 3 && d.epi == EPI_HEAD) {
#pragma unroll
            for (int g = 0; g < 8; ++g) bq[g] = __ldg(reinterpret_cast<const float4*>(bias) + g);
          }
          if (d.epi == EPI_HEAD) {
            const float* wo = P.w_out + (int)rank * 256 + nl0;
            pair_head_stage(acc, bq, wo, hp0, hp1, hp2, hp3);
          } else {
            unsigned char* dst_s = abuf + ((size_t)rank * 32 + nl0 / 8) * 2048 + (size_t)row * 16;
            if (pooled) pair_act_stage<NV, true>(acc, bq, bias, lane, dst_s, t0 + nl0);
            else pair_act_stage<NV, false>(acc, bq, bias, lane, dst_s, t0 + nl0);
            // my part of K-block (half*2 + st) is in shared memory: order it before the MMAs (async proxy) and
            // before the pusher's bulk copy, then one arrival per warp
            const long long t__ = PROF_CLOCK();
            tc::fence_before_sync();
            tc::fence_proxy_async();            // generic smem writes -> async proxy (MMA reads, pusher's bulk store)
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(a_own + st);
            w_sig += PROF_CLOCK() - t__;
          }
        }
        w_kind[d.epi] += PROF_CLOCK() - tl0;
        if (d.epi == EPI_HEAD) {
          tc::fence_before_sync();
          // combine the two column halves of the CTA, then one float4 of partial sums per row
          if (half == 1) reinterpret_cast<float4*>(head_s)[row] = make_float4(hp0, hp1, hp2, hp3);
          bar_sync_epi();
          if (half == 0) {
            const float4 o4 = reinterpret_cast<const float4*>(head_s)[row];
            P.head_partial[((size_t)tile * 2 + rank) * 128 + row] = make_float4(hp0 + o4.x, hp1 + o4.y, hp2 + o4.z, hp3 + o4.w);
          }
          bar_sync_epi();
        } else {
          ++xops;
        }
      }
    }
    if (leader && P.prof) {
      long long* pr = P.prof + (size_t)blockIdx.x * 16;
      pr[4] = PROF_CLOCK() - t_epi0; pr[5] = w_acc; pr[6] = w_fence; pr[7] = w_sig; pr[12] = w_push;
      pr[8] = w_kind[0]; pr[9] = w_kind[1]; pr[10] = w_kind[2]; pr[11] = w_kind[3];
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::cluster_sync_all();               // nobody exits while its peer may still signal it
  if (warp == 1) { tc::fence_after_sync(); tc::tmem_dealloc(tmem_base, 512); }
}

// =====================================================================================================================
// Quad variant: a cluster of FOUR CTAs = 2 row tiles x 2 column halves.  CTA (h, t) -- cluster rank 2h + t -- owns output
// columns [256h, 256h+256) of tile t.  The two CTAs with the same h form a tcgen05 CTA pair: ONE thread of the even one
// issues tcgen05.mma.cta_group::2 (M = 256: 128 rows from each CTA, N = 256) for both, and each CTA streams only HALF of
// the weight K-block (its 128 of the 256 output rows, 16 KB) -- the L2 -> SM weight traffic per tile, which bounds the
// pair kernel above, is halved.  The A-operand exchange runs between the CTAs with the same t (rank ^ 2) exactly as in
// the pair kernel.  Weights are read in the PK128 layout ([N/128][K/8][128][8]): CTA rank r streams N-tile r.
// The odd CTA of a pair has no MMA thread; its warp 1 forwards "my operands for ring step s are in place" to the even
// CTA (one relaxed remote mbarrier arrival per ring slot).
constexpr int kQRing = 6;
constexpr int kQSlot = 16384;
constexpr size_t kQSmem = (size_t)kPAbuf + kQRing * kQSlot + 512;

// MERGE (NV > 1): after the view mean only 128/NV rows of a tile are distinct, so a CTA parks the pooled rows of NV
// consecutive tiles (fp32, in an L2-resident scratch), merges them into ONE full tile and runs blocks 3-4 + the head
// once per NV tiles instead of on NV-fold duplicated rows.  `last` = this iteration closes its group.
template <int NV, bool MERGE>
__device__ __forceinline__ bool quad_last(int it, int q, int n_clusters, int n_pairs) {
  return !MERGE || (it % NV) == NV - 1 || q + n_clusters >= n_pairs;
}

template <int NV, bool MERGE>
__global__ void __cluster_dims__(4, 1, 1) __launch_bounds__(kPThreads, 1) mlp_quad_fwd_kernel(const PairParams P, const __grid_constant__ QuadMaps M) {
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* abuf = smem;                                   // 128 KB resident A operand
  unsigned char* ring = smem + kPAbuf;                          // 6 x 16 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + kQRing * kQSlot);
  uint64_t* full = bars;                  // [6]  my ring slot landed
  uint64_t* empty = bars + 6;             // [6]  ring slot consumed in BOTH CTAs of the MMA pair (multicast commit)
  uint64_t* peer_full = bars + 12;        // [6]  (even CTA) the odd CTA's operands of this ring step are in place
  uint64_t* accf = bars + 18;             // [2]  accumulator of op (parity) complete (multicast commit)
  uint64_t* a_peer = bars + 36;           // [4]  exchange partner's K-block p landed (one 16 KB push each: the first remote
                                          //      K-block is needed four K-steps after the first own one)
  uint64_t* a_own = bars + 22;            // [4]  own K-block p written (one arrival per epilogue warp; pusher waits here)
  uint64_t* a_own2 = bars + 32;           // [4]  (even CTA) K-block p written in BOTH CTAs of the MMA pair (16 arrivals)
  uint64_t* pushed = bars + 26;           // [1]  partner received my whole half
  uint64_t* peer_free = bars + 27;        // [1]  partner's A buffer may receive my next operand
  uint64_t* odd_peer = bars + 28;         // [4]  (even CTA, tma2) the partner's K-block p landed in the ODD CTA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 40);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cr = tc::cluster_ctarank();
  const int h = (int)(cr >> 1), t = (int)(cr & 1);
  const uint32_t partner = cr ^ 2u;       // same tile, other column half
  const uint32_t mma_peer = cr ^ 1u;      // same column half, other tile
  const uint16_t pair_mask = (uint16_t)(3u << (2 * h));
  const int cluster = blockIdx.x >> 2, n_clusters = gridDim.x >> 2;
  const int n_pairs = (P.mv_tiles + 1) >> 1;
  auto tile_of = [&](int q) { const int x = 2 * q + t; return x < P.mv_tiles ? x : P.mv_tiles - 1; };

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kQRing; ++s) { tc::mbar_init(full + s, 1); tc::mbar_init(empty + s, 1); tc::mbar_init(peer_full + s, 1); }
    tc::mbar_init(accf + 0, 1); tc::mbar_init(accf + 1, 1);
    for (int i = 0; i < 4; ++i) { tc::mbar_init(a_own + i, 8); tc::mbar_init(a_own2 + i, 16); }
    for (int i = 0; i < 4; ++i) { tc::mbar_init(a_peer + i, 1); tc::mbar_init(odd_peer + i, 1); }
    tc::mbar_init(pushed, 1); tc::mbar_init(peer_free, 1);
    tc::fence_barrier_init();
  }
  __syncthreads();
  tc::cluster_sync_all();               // every CTA's barriers exist before anything can signal them
  if (warp == 1) tc::tmem_alloc_2cta(tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::cluster_sync_all();
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= ring producer (every CTA: its half of the weights, its tile's streamed operands) =============
    {     // the whole warp runs the loop (warp-uniform operands -> uniform registers); one elected lane issues the copies
      const uint64_t pol_keep = tc::policy_evict_last(), pol_stream = tc::policy_evict_first();
      const uint64_t pol_norm = tc::policy_evict_normal();
      int slot = 0; uint32_t phase = 0;
      // fill(): one 16 KB K-block into the next ring slot.  Plain mode: cp.async.bulk completing on MY full barrier (the
      // odd CTA then forwards "in place" to the even CTA).  tma2 mode: tensor-map TMA whose bytes are counted on the EVEN
      // CTA's full barrier, which its producer arms for both CTAs' 2 x 16 KB.
      const uint32_t even_rank = cr & ~1u;
      auto fill = [&](const unsigned char* src, uint64_t pol, const CUtensorMap* map, int row0) {
        tc::mbar_wait(empty + slot, phase ^ 1);
        if (tc::elect_one()) {
          if (P.tma2) {
            if (t == 0) tc::mbar_arrive_expect_tx(full + slot, 2 * kQSlot);
            tc::tma_load_2d_2cta(ring + (size_t)slot * kQSlot, map, 0, row0, tc::cluster_addr(full + slot, even_rank), pol);
          } else {
            tc::mbar_arrive_expect_tx(full + slot, kQSlot);
            tc::bulk_g2s_hint(ring + (size_t)slot * kQSlot, src, kQSlot, full + slot, pol);
          }
        }
        __syncwarp();
        if (++slot == kQRing) { slot = 0; phase ^= 1; }
      };
      auto w_map = [&](const unsigned char* w) -> const CUtensorMap* {      // the tensor map of a weight image
        if (w == P.w_in) return &M.w[0];
        if (w == P.w_zc[0]) return &M.wc[0];
        if (w == P.w_zc[1]) return &M.wc[1];
        for (int i = 0; i < 3; ++i) if (w == P.w_z[i]) return &M.w[1 + i];
        for (int i = 0; i < 5; ++i) if (w == P.w_0[i]) return &M.w[4 + i];
        for (int i = 0; i < 5; ++i) if (w == P.w_1[i]) return &M.w[9 + i];
        return &M.w[0];
      };
      POp d;
      for (int q = cluster, it = 0; q < n_pairs; q += n_clusters, ++it) {
        const int n_ops = quad_last<NV, MERGE>(it, q, n_clusters, n_pairs) ? kPOps : 7;
        const int tile = tile_of(q);
        const unsigned char* lat = P.lat_pk + (size_t)tile * 131072;
        if (q + n_clusters < n_pairs && lane == 0) {     // pull the next tile's inputs from HBM into L2 while this one computes
          const int nt = tile_of(q + n_clusters);
          tc::bulk_prefetch_l2(P.lat_pk + (size_t)nt * 131072, 131072);
          tc::bulk_prefetch_l2(P.code_pk + (size_t)nt * 16384, 16384);
        }
        for (int o = 0; o < n_ops; ++o) {
          pair_get_op(P, o, d);
          if (d.code) {
            fill(P.code_pk + (size_t)tile * 16384, (o == 4 || !P.w_zc[0]) ? pol_stream : pol_norm, &M.code, tile * 32);
            fill(d.code_w + (size_t)cr * 16384, pol_keep, P.tma2 ? w_map(d.code_w) : nullptr, (int)cr * 32);
          }
          if (d.lat_w) {
            const CUtensorMap* wm = P.tma2 ? w_map(d.lat_w) : nullptr;
            for (int kb = 0; kb < 8; ++kb) {
              fill(lat + (size_t)kb * 16384, o == 4 ? pol_stream : pol_norm, &M.lat, tile * 256 + kb * 32);   // last use: evict first
              fill(d.lat_w + (size_t)cr * 131072 + (size_t)kb * 16384, pol_keep, wm, (int)cr * 256 + kb * 32);
            }
          }
          if (d.abuf_w) {
            const CUtensorMap* wm = P.tma2 ? w_map(d.abuf_w) : nullptr;
            for (int i = 0; i < 8; ++i) {
              const int kb = pair_kb_order(h, i);
              fill(d.abuf_w + (size_t)cr * 131072 + (size_t)kb * 16384, pol_keep, wm, (int)cr * 256 + kb * 32);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (t == 0) {              // the whole warp runs the loop (warp-uniform operands); one elected lane issues
      // ================= MMA issuer for the CTA pair =================
      constexpr uint32_t idesc = tc::make_idesc_bf16(256, 256, 0, 0);
      int slot = 0; uint32_t phase = 0;
      int opc = 0, aops = 0, titer = 0;
      POp d;
      const uint32_t abuf0 = tc::smem_u32(abuf), ring0 = tc::smem_u32(ring);
      long long w_ring = 0, w_own = 0, w_peer = 0, w_pfull = 0, w_ring_lat = 0, w_pfull_lat = 0, t_start = PROF_CLOCK();
      for (int q = cluster; q < n_pairs; q += n_clusters, ++titer) {
        const uint32_t x_col = (titer & 1) ? 256u : 0u;           // X / T swap regions every tile
        const int n_ops = quad_last<NV, MERGE>(titer, q, n_clusters, n_pairs) ? kPOps : 7;
        for (int o = 0; o < n_ops; ++o, ++opc) {
          pair_get_op(P, o, d);
          const uint32_t d_tmem = tmem_base + (d.acc_x ? x_col : (256u - x_col));
          uint32_t accumulate = d.fresh ? 0u : 1u;
          auto step_ring_ab = [&]() {        // A and B both come from consecutive ring slots
            const int sa = slot;
            // (descriptors first: they depend on the slot numbers only and must not sit between "operands landed" and
            // the first MMA -- the issuing warp's latency per K-step is on the kernel's critical path)
            const int sb_ = sa + 1 == kQRing ? 0 : sa + 1;
            uint64_t ad0 = tc::make_smem_desc(ring0 + (uint32_t)sa * kQSlot, 2048, 128);
            uint64_t bd0 = tc::make_smem_desc(ring0 + (uint32_t)sb_ * kQSlot, 2048, 128);
            asm volatile("" : "+l"(ad0), "+l"(bd0));
            TIMED(w_ring_lat, tc::mbar_wait(full + sa, phase));
            if (!P.tma2) TIMED(w_pfull_lat, tc::mbar_wait(peer_full + sa, phase));
            if (++slot == kQRing) { slot = 0; phase ^= 1; }
            const int sb = slot;
            TIMED(w_ring_lat, tc::mbar_wait(full + sb, phase));
            if (!P.tma2) TIMED(w_pfull_lat, tc::mbar_wait(peer_full + sb, phase));
            if (++slot == kQRing) { slot = 0; phase ^= 1; }
            tc::fence_after_sync();
            if (tc::elect_one()) {
#pragma unroll
              for (int j = 0; j < 4; ++j)        // + 4 KB per K = 16: 256 in the descriptor's 16-byte address field
                tc::mma_bf16_2cta(d_tmem, ad0 + (uint64_t)(j * 256), bd0 + (uint64_t)(j * 256), idesc,
                                  j == 0 ? accumulate : 1u);
              tc::mma_commit_2cta_mc(empty + sa, pair_mask);
              tc::mma_commit_2cta_mc(empty + sb, pair_mask);
            }
            __syncwarp();
            accumulate = 1;
          };
          if (d.code) step_ring_ab();
          if (d.lat_w)
            for (int kb = 0; kb < 8; ++kb) step_ring_ab();
          if (d.abuf_w) {
            const uint32_t par = (uint32_t)(aops & 1);
            for (int i = 0; i < 8; ++i) {
              const int kb = pair_kb_order(h, i);
              uint64_t ad0 = tc::make_smem_desc(abuf0 + (uint32_t)kb * 16384u, 2048, 128);
              uint64_t bd0 = tc::make_smem_desc(ring0 + (uint32_t)slot * kQSlot, 2048, 128);
              asm volatile("" : "+l"(ad0), "+l"(bd0));                      // (before the waits, see step_ring_ab)
#if DNR_CLUSTER_PROF
              if (i < 4) TIMED(w_own, tc::mbar_wait(a_own2 + i, par));
              if (i >= 4) TIMED(w_peer, tc::mbar_wait(a_peer + (i - 4), par));
              if (P.tma2 && i >= 4) TIMED(w_peer, tc::mbar_wait(odd_peer + (i - 4), par));     // ... and in the odd CTA
              TIMED(w_ring, tc::mbar_wait(full + slot, phase));
#else
              // production build: the operand-block barrier and the ring slot are polled together
              if (i < 4) tc::mbar_wait2(a_own2 + i, par, full + slot, phase);
              else {
                tc::mbar_wait2(a_peer + (i - 4), par, full + slot, phase);
                if (P.tma2) tc::mbar_wait(odd_peer + (i - 4), par);                       // ... and in the odd CTA
              }
#endif
              if (!P.tma2) TIMED(w_pfull, tc::mbar_wait(peer_full + slot, phase));
              tc::fence_after_sync();
              if (tc::elect_one()) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  tc::mma_bf16_2cta(d_tmem, ad0 + (uint64_t)(j * 256), bd0 + (uint64_t)(j * 256), idesc,
                                    j == 0 ? accumulate : 1u);
                tc::mma_commit_2cta_mc(empty + slot, pair_mask);
              }
              __syncwarp();
              accumulate = 1;
              if (++slot == kQRing) { slot = 0; phase ^= 1; }
            }
            ++aops;
          }
          if (tc::elect_one()) tc::mma_commit_2cta_mc(accf + (opc & 1), pair_mask);
          __syncwarp();
        }
      }
      if (P.prof && lane == 0) {
        long long* pr = P.prof + (size_t)blockIdx.x * 16;
        pr[0] = PROF_CLOCK() - t_start; pr[1] = w_ring; pr[2] = w_own; pr[3] = w_peer; pr[13] = w_pfull;
        pr[14] = w_ring_lat; pr[15] = w_pfull_lat;
      }
    } else if (lane == 0) {
      // ================= odd CTA: forward "operands of ring step s in place" to the even CTA =================
      int slot = 0; uint32_t phase = 0;
      int aops = 0;
      POp d;
      auto forward = [&]() {
        tc::mbar_wait(full + slot, phase);
        tc::mbar_arrive_remote_relaxed(peer_full + slot, mma_peer);
        if (++slot == kQRing) { slot = 0; phase ^= 1; }
      };
      for (int q = cluster, it = 0; q < n_pairs; q += n_clusters, ++it) {
        const int n_ops = quad_last<NV, MERGE>(it, q, n_clusters, n_pairs) ? kPOps : 7;
        for (int o = 0; o < n_ops; ++o) {
          pair_get_op(P, o, d);
          if (P.tma2) {
            // the ring needs no forwarding (my loads complete on the even CTA's barrier); what the even CTA still has to
            // learn from me is that the partner's half of the A operand has landed in MY buffer
            if (d.abuf_w) {
              const uint32_t par = (uint32_t)(aops & 1);
              for (int e = 0; e < 4; ++e) {
                tc::mbar_wait(a_peer + e, par);
                tc::mbar_arrive_remote_relaxed(odd_peer + e, mma_peer);
              }
              ++aops;
            }
            continue;
          }
          if (d.code) { forward(); forward(); }
          if (d.lat_w)
            for (int kb = 0; kb < 8; ++kb) { forward(); forward(); }
          if (d.abuf_w) {
            const uint32_t par = (uint32_t)(aops & 1);
            for (int i = 0; i < 8; ++i) {
              if (i >= 4) tc::mbar_wait(a_peer + (i - 4), par);   // (own K-blocks: the epilogue warps signal the even CTA directly)
              forward();
            }
            ++aops;
          }
        }
      }
    }
  }
  else if (warp == 2) {
    // ================= pusher: my finished K-blocks -> the partner's resident A buffer (bulk copy over DSMEM) =========
    {     // whole warp, one elected lane issues the copies (tc::elect_one)
      int xops = 0;
      unsigned char* mine = abuf + (size_t)h * 65536;
      for (int q = cluster, it = 0; q < n_pairs; q += n_clusters, ++it) {
        // A operands produced per iteration: by the epilogues of ops 0..5, and - when the iteration closes its group -
        // by the merge after op 6 and the epilogues of ops 7..9
        const int n_events = quad_last<NV, MERGE>(it, q, n_clusters, n_pairs) ? kPOps - 1 : 6;
        for (int o = 0; o < n_events; ++o, ++xops) {
          for (int kb = 0; kb < 4; ++kb) {                    // one push per K-block, as soon as it is written
            tc::mbar_wait(a_own + kb, (uint32_t)(xops & 1));
            if (kb == 0) tc::mbar_wait(peer_free, (uint32_t)(xops & 1));
            if (tc::elect_one()) tc::bulk_s2s_peer(mine + (size_t)kb * 16384, 16384, a_peer + kb, partner);
            __syncwarp();
          }
        }
      }
    }
  } else if (warp == 3) {
    // ================= receiver: arm the landing barriers, grant the credit, acknowledge =================
    if (lane == 0) {
      int opc = 0, xops = 0;
      for (int q = cluster, it = 0; q < n_pairs; q += n_clusters, ++it) {
        const bool last = quad_last<NV, MERGE>(it, q, n_clusters, n_pairs);
        const int n_ops = last ? kPOps : 7;
        for (int o = 0; o < n_ops; ++o, ++opc) {
          if (!(o < 6 || (last && o < kPOps - 1))) continue;                  // this op's output is no A operand
          tc::mbar_wait(accf + (opc & 1), (uint32_t)((opc >> 1) & 1));     // the MMAs of this op no longer read my buffer
          for (int e = 0; e < 4; ++e) tc::mbar_arrive_expect_tx(a_peer + e, 16384);
          tc::mbar_arrive_remote_relaxed(peer_free, partner);
          for (int e = 3; e >= 0; --e) tc::mbar_wait(a_peer + e, (uint32_t)(xops & 1));
          tc::mbar_arrive_remote_relaxed(pushed, partner);
          ++xops;
        }
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue (the pair kernel's; column half = h; + park / merge when MERGE) =================
    const int quad = warp & 3;
    const int half = (warp - 4) >> 2;
    const int row = quad * 32 + lane;
    const bool leader = (warp == 4 && lane == 0);
    constexpr int SPWc = 32 / NV;                    // samples per 32-row slab (RowMap)
    constexpr int U = 4 * SPWc;                      // distinct (pooled) rows per view tile
    float* park = MERGE ? P.merge_scratch + (size_t)blockIdx.x * (128 * 256) : nullptr;     // [NV*U <= 128][256] fp32
    const uint64_t pol_keep = tc::policy_evict_last();
    int opc = 0, xops = 0, titer = 0;
    long long w_acc = 0, w_merge = 0, t_epi0 = PROF_CLOCK();
    long long w_kind[4] = {0, 0, 0, 0};
    POp d;
    auto signal_kblock = [&](int st) {               // my part of own K-block st is in shared memory
      tc::fence_before_sync();
      tc::fence_proxy_async();                       // generic smem writes -> async proxy (MMA reads, pusher's bulk copy)
      __syncwarp();
      if (lane == 0) {
        tc::mbar_arrive(a_own + st);
        tc::mbar_arrive_remote_relaxed(a_own2 + st, cr & ~1u);     // the MMA issuer of my pair
      }
    };
    for (int q = cluster; q < n_pairs; q += n_clusters, ++titer) {
      const int tile = tile_of(q);
      const bool valid = 2 * q + t < P.mv_tiles;
      const uint32_t x_col = (titer & 1) ? 256u : 0u;
      const int pos = MERGE ? titer % NV : 0;                     // position of this tile in its merge group
      const bool last = quad_last<NV, MERGE>(titer, q, n_clusters, n_pairs);
      const int n_ops = last ? kPOps : 7;
      for (int o = 0; o < n_ops; ++o, ++opc) {
        pair_get_op(P, o, d);
        TIMED(w_acc, tc::mbar_wait(accf + (opc & 1), (uint32_t)((opc >> 1) & 1)));
        tc::fence_after_sync();
        const uint32_t t0 = tmem_base + ((uint32_t)(quad * 32) << 16) + (d.acc_x ? x_col : (256u - x_col));
        const long long tl0 = PROF_CLOCK();
        if (MERGE && o == 6) {
          // ---- park: mean over the view rows (resnetfc.py:9-14,148-151); the first row of every sample stores it
          const bool first_view = (lane % NV == 0) && (lane / NV < SPWc);
          const int j_own = quad * SPWc + lane / NV;
#pragma unroll 1
          for (int st = 0; st < 4; ++st) {
            uint32_t acc[32];
            const int nl0 = st * 64 + half * 32;
            tc::tmem_ld32(t0 + nl0, acc);
            tc::tmem_ld_wait();
            if (NV == 3) {
              // only the first row of a sample is parked: (a0 + a1) + a2 from the two lanes below -- the bits view_mean<3>
              // gives that row, without its per-lane operand selection
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const float v = __uint_as_float(acc[j]);
                const float x = __shfl_down_sync(0xffffffffu, v, 1), y = __shfl_down_sync(0xffffffffu, v, 2);
                acc[j] = __float_as_uint(((v + x) + y) * (1.0f / 3.0f));
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) acc[j] = __float_as_uint(view_mean<NV>(__uint_as_float(acc[j]), lane));
            }
            if (first_view) {
              // scratch layout [64 column chunks of 4][128 merged rows] x 16 B: the lanes of a warp write (and the
              // merge reads) consecutive 16-byte words; evict_last keeps it in L2 against the streaming operands
              uint4* dst = reinterpret_cast<uint4*>(park) + (size_t)(nl0 / 4) * 128 + (pos * U + j_own);
#pragma unroll
              for (int g = 0; g < 8; ++g)
                tc::st_hint_v4(dst + (size_t)g * 128, acc[4 * g], acc[4 * g + 1], acc[4 * g + 2], acc[4 * g + 3], pol_keep);
            }
          }
          w_kind[EPI_MEAN] += PROF_CLOCK() - tl0;
          if (!last) continue;
          const long long tm0 = PROF_CLOCK();
          bar_sync_epi();                  // the parked rows of the whole group are visible to all epilogue threads
          // ---- merge: thread = row m of the merged tile = pooled row (m % U) of the group's tile (m / U)
          const int i_src = row / U, j_src = row % U;
          const bool have = i_src <= pos;
          const uint4* src_row = reinterpret_cast<const uint4*>(park) + (i_src * U + j_src);     // (+ chunk * 128)
          uint4 nxt[8];                    // the next piece's rows travel from L2 while this piece is processed
          auto fetch = [&](int st) {
#pragma unroll
            for (int g = 0; g < 8; ++g)
              nxt[g] = have ? tc::ld_cg_hint_v4(src_row + (size_t)((st * 64 + half * 32) / 4 + g) * 128, pol_keep)
                            : make_uint4(0, 0, 0, 0);
          };
          fetch(0);
#pragma unroll 1                   // (compact code: this path runs once per group and is cold in the instruction cache)
          for (int st = 0; st < 4; ++st) {
            uint32_t acc[32];
            const int nl0 = st * 64 + half * 32;
#pragma unroll
            for (int g = 0; g < 8; ++g) { acc[4 * g] = nxt[g].x; acc[4 * g + 1] = nxt[g].y; acc[4 * g + 2] = nxt[g].z; acc[4 * g + 3] = nxt[g].w; }
            if (st < 3) fetch(st + 1);
            float4 bq[8];
            const float* bias = d.bias + h * 256 + nl0;
#pragma unroll
            for (int g = 0; g < 8; ++g) bq[g] = __ldg(reinterpret_cast<const float4*>(bias) + g);
            if (st == 0 && xops > 0) tc::mbar_wait(pushed, (uint32_t)((xops - 1) & 1));
            tc::tmem_st32(t0 + nl0, acc);                           // the merged tile's residual stream
            unsigned char* dst_s = abuf + ((size_t)h * 32 + nl0 / 8) * 2048 + (size_t)row * 16;
            pair_act_stage<1, false>(acc, bq, bias, lane, dst_s, t0 + nl0);
            tc::tmem_st_wait();
            signal_kblock(st);
          }
          w_merge += PROF_CLOCK() - tm0;
          ++xops;
          continue;
        }
        const bool pooled = !MERGE && o == 6;
        float hp0 = 0.f, hp1 = 0.f, hp2 = 0.f, hp3 = 0.f;
#pragma unroll 1
        for (int st = 0; st < 4; ++st) {
          uint32_t acc[32];
          const int nl0 = st * 64 + half * 32;
          tc::tmem_ld32(t0 + nl0, acc);
          float4 bq[8];                                        // the piece's biases travel while the TMEM load does
          if (NV != 3) {                                       // (NV = 3: the three-way mean needs the registers)
#pragma unroll
            for (int g = 0; g < 8; ++g) bq[g] = __ldg(reinterpret_cast<const float4*>(d.bias + h * 256 + nl0) + g);
          }
          // the previous operand has left my half of the A buffer (the pusher's bulk copies have read it)
          if (st == 0 && d.epi != EPI_HEAD && xops > 0) tc::mbar_wait(pushed, (uint32_t)((xops - 1) & 1));
          tc::tmem_ld_wait();
          const float* bias = d.bias + h * 256 + nl0;
          if (NV == 3 && d.epi == EPI_HEAD) {
#pragma unroll
            for (int g = 0; g < 8; ++g) bq[g] = __ldg(reinterpret_cast<const float4*>(bias) + g);
          }
          if (d.epi == EPI_HEAD) {
            pair_head_stage(acc, bq, P.w_out + h * 256 + nl0, hp0, hp1, hp2, hp3);
          } else {
            unsigned char* dst_s = abuf + ((size_t)h * 32 + nl0 / 8) * 2048 + (size_t)row * 16;
            if (pooled) pair_act_stage<NV, true>(acc, bq, bias, lane, dst_s, t0 + nl0);
            else pair_act_stage<NV, false>(acc, bq, bias, lane, dst_s, t0 + nl0);
            signal_kblock(st);
          }
        }
        w_kind[d.epi] += PROF_CLOCK() - tl0;
        if (d.epi == EPI_HEAD) {
          tc::fence_before_sync();
          // the two column groups of a row are summed by head_finalize (4 partials per row and tile)
          if (!MERGE) {
            if (valid)
              P.head_partial[(((size_t)tile * 2 + h) * 2 + half) * 128 + row] = make_float4(hp0, hp1, hp2, hp3);
          } else {
            // merged row -> the first view row of its sample in the tile it came from (what head_finalize reads)
            const int i_src = row / U, j_src = row % U;
            const int x = 2 * (q - (pos - i_src) * n_clusters) + t;
            if (i_src <= pos && x < P.mv_tiles) {
              const int lrow = (j_src / SPWc) * 32 + (j_src % SPWc) * NV;
              P.head_partial[(((size_t)x * 2 + h) * 2 + half) * 128 + lrow] = make_float4(hp0, hp1, hp2, hp3);
            }
          }
        } else {
          ++xops;
        }
      }
    }
    if (leader && P.prof) {
      long long* pr = P.prof + (size_t)blockIdx.x * 16;
      pr[4] = PROF_CLOCK() - t_epi0; pr[5] = w_acc; pr[12] = w_merge;
      pr[8] = w_kind[0]; pr[9] = w_kind[1]; pr[10] = w_kind[2]; pr[11] = w_kind[3];
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::cluster_sync_all();               // nobody exits (or frees TMEM) while another CTA may still signal or read it
  if (warp == 1) { tc::fence_after_sync(); tc::tmem_dealloc_2cta(tmem_base, 512); }
}

// quad kernel: 4 partial sums per row (2 column halves x 2 column groups)
__global__ void head_finalize4_kernel(const float4* __restrict__ part, const float* __restrict__ bout, int NV, int SPW,
                                      long long n_samples, float4* __restrict__ out) {
  const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_samples) return;
  const long long per_tile = 4LL * SPW;
  const long long tile = s / per_tile;
  const int j = (int)(s % per_tile);
  const int row = (j / SPW) * 32 + (j % SPW) * NV;
  const float4 a0 = part[(tile * 4 + 0) * 128 + row], a1 = part[(tile * 4 + 1) * 128 + row];
  const float4 b0 = part[(tile * 4 + 2) * 128 + row], b1 = part[(tile * 4 + 3) * 128 + row];
  // same association as the pair kernel: (half 0 + half 1) per column half, then the two halves, then the bias
  const float o0 = (a0.x + a1.x) + (b0.x + b1.x) + bout[0], o1 = (a0.y + a1.y) + (b0.y + b1.y) + bout[1];
  const float o2 = (a0.z + a1.z) + (b0.z + b1.z) + bout[2], o3 = (a0.w + a1.w) + (b0.w + b1.w) + bout[3];
  out[s] = make_float4(1.f / (1.f + expf(-o0)), 1.f / (1.f + expf(-o1)), 1.f / (1.f + expf(-o2)), fmaxf(o3, 0.f));
}

// out[s] = [sigmoid, sigmoid, sigmoid, relu](partial_0 + partial_1 + b_out) read at the leader row of sample s
__global__ void head_finalize_kernel(const float4* __restrict__ part, const float* __restrict__ bout, int NV, int SPW,
                                     long long n_samples, float4* __restrict__ out) {
  const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_samples) return;
  const long long per_tile = 4LL * SPW;
  const long long tile = s / per_tile;
  const int j = (int)(s % per_tile);
  const int row = (j / SPW) * 32 + (j % SPW) * NV;
  const float4 a = part[(tile * 2 + 0) * 128 + row], b = part[(tile * 2 + 1) * 128 + row];
  const float o0 = a.x + b.x + bout[0], o1 = a.y + b.y + bout[1], o2 = a.z + b.z + bout[2], o3 = a.w + b.w + bout[3];
  out[s] = make_float4(1.f / (1.f + expf(-o0)), 1.f / (1.f + expf(-o1)), 1.f / (1.f + expf(-o2)), fmaxf(o3, 0.f));
}

}  // namespace dnr

// Whole ResnetFC forward + lin_out partial sums on CTA pairs (see file header).  weights: 14 PK256 pointers
// [lin_in, lin_z0..2, fc_0 x5, fc_1 x5]; biases: 11 fp32[512] pointers [cb0..cb5 (cumulative residual biases),
// fc_0 bias x5]; w_out fp32 [4 x 512]; head_partial: [mv_tiles*2*128] float4; max_clusters: 0 = all SM pairs, else
// the number of clusters to launch (to run beside another kernel on the SMs that one leaves idle).
static int pair_fwd_impl(const void* lat_pk, const void* code_pk, const void* const* weights,
                         const void* const* code_weights, const float* const* biases, const float* w_out, int NV,
                         int mv_tiles, int max_clusters, float* head_partial, long long* prof, cudaStream_t stream);
DNR_EXPORT int dnr_mlp_pair_fwd(const void* lat_pk, const void* code_pk, const void* const* weights,
                                const float* const* biases, const float* w_out, int NV, int mv_tiles, int max_clusters,
                                float* head_partial, long long* prof, cudaStream_t stream) {
  return pair_fwd_impl(lat_pk, code_pk, weights, nullptr, biases, w_out, NV, mv_tiles, max_clusters, head_partial, prof, stream);
}
// NOVEL_PE field (src/models/novel_pe/pe_novel_pixelnerf.py:263-275, d_latent = 518): fc_1 of blocks 0 and 1 gets a second
// K-step on the 64-wide code operand, whose columns 55..60 carry the six positional values -- code_weights[b]: PK256
// [512 x 64] image of [0 (55) | lin_z[b+1].weight[:, 512:518] | 0 (3)] (diner_b200.field.PackedPE); lat_pk is then the
// variant's final_latent.  Otherwise dnr_mlp_pair_fwd.
DNR_EXPORT int dnr_mlp_pair_fwd_pe(const void* lat_pk, const void* code_pk, const void* const* weights,
                                   const void* const* code_weights, const float* const* biases, const float* w_out, int NV,
                                   int mv_tiles, int max_clusters, float* head_partial, cudaStream_t stream) {
  DNR_REQUIRE(code_weights && code_weights[0] && code_weights[1], "dnr_mlp_pair_fwd_pe: null code weights");
  return pair_fwd_impl(lat_pk, code_pk, weights, code_weights, biases, w_out, NV, mv_tiles, max_clusters, head_partial,
                       nullptr, stream);
}
static int pair_fwd_impl(const void* lat_pk, const void* code_pk, const void* const* weights,
                         const void* const* code_weights, const float* const* biases, const float* w_out, int NV,
                         int mv_tiles, int max_clusters, float* head_partial, long long* prof, cudaStream_t stream) {
  DNR_REQUIRE(NV >= 1 && NV <= 4, "dnr_mlp_pair_fwd: NV must be in [1,4]");
  DNR_REQUIRE(max_clusters >= 0, "dnr_mlp_pair_fwd: max_clusters must be >= 0 (0 = one cluster per SM pair)");
  DNR_REQUIRE(lat_pk && code_pk && weights && biases && w_out && head_partial, "dnr_mlp_pair_fwd: null argument");
  if (mv_tiles <= 0) return 0;
  dnr::PairParams P;
  P.lat_pk = (const unsigned char*)lat_pk; P.code_pk = (const unsigned char*)code_pk;
  P.w_in = (const unsigned char*)weights[0];
  for (int i = 0; i < 2; ++i) P.w_zc[i] = code_weights ? (const unsigned char*)code_weights[i] : nullptr;
  for (int i = 0; i < 3; ++i) P.w_z[i] = (const unsigned char*)weights[1 + i];
  for (int i = 0; i < 5; ++i) { P.w_0[i] = (const unsigned char*)weights[4 + i]; P.w_1[i] = (const unsigned char*)weights[9 + i]; }
  for (int i = 0; i < 6; ++i) P.cb[i] = biases[i];
  for (int i = 0; i < 5; ++i) P.b_0[i] = biases[6 + i];
  P.w_out = w_out; P.head_partial = (float4*)head_partial;
  P.NV = NV; P.SPW = 32 / NV; P.mv_tiles = mv_tiles; P.prof = prof; P.tma2 = 0;
  void (*kern)(const dnr::PairParams) = NV == 1 ? dnr::mlp_pair_fwd_kernel<1>
                                         : NV == 2 ? dnr::mlp_pair_fwd_kernel<2>
                                         : NV == 3 ? dnr::mlp_pair_fwd_kernel<3> : dnr::mlp_pair_fwd_kernel<4>;
  dnr::DeviceInfo* di = dnr::device_info();
  DNR_REQUIRE(di != nullptr, "dnr_mlp_pair_fwd: no current CUDA device");
  if (!di->pair_configured[NV]) {
    DNR_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dnr::kPSmem),
              "dnr_mlp_pair_fwd/attr");
    di->pair_configured[NV] = true;
  }
  const int all_pairs = di->num_sms / 2;
  const int cap = max_clusters > 0 && max_clusters < all_pairs ? max_clusters : all_pairs;
  int clusters = mv_tiles < cap ? mv_tiles : cap;
  kern<<<2 * clusters, dnr::kPThreads, dnr::kPSmem, stream>>>(P);
  DNR_CHECK_LAUNCH("dnr_mlp_pair_fwd");
  return 0;
}

namespace dnr {
// cuTensorMapEncodeTiled through the runtime (no -lcuda): a 2-D bf16 view [bytes / 512][256] of a packed buffer with a box
// of 32 rows = one 16 KB K-block, no swizzle / interleave (the buffer already is the shared-memory image).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}
static bool encode_kblock_map(CUtensorMap* m, const void* base, size_t bytes) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn || bytes % 16384 != 0 || ((uintptr_t)base & 15)) return false;
  const cuuint64_t dims[2] = {256, (cuuint64_t)(bytes / 512)};
  const cuuint64_t strides[1] = {512};
  const cuuint32_t box[2] = {256, 32};
  const cuuint32_t estr[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
// Weight images rarely change between launches: keep their tensor maps (per thread; keyed by base pointer and size -- a
// tensor map describes addresses, not contents, so a refilled buffer at the same address needs no new map).
static bool cached_kblock_map(CUtensorMap* out, const void* base, size_t bytes) {
  struct Entry { const void* base; size_t bytes; CUtensorMap m; };
  static thread_local Entry cache[32];
  static thread_local int used = 0, next = 0;
  for (int i = 0; i < used; ++i)
    if (cache[i].base == base && cache[i].bytes == bytes) { *out = cache[i].m; return true; }
  if (!encode_kblock_map(out, base, bytes)) return false;
  Entry& e = cache[next];
  e.base = base; e.bytes = bytes; e.m = *out;
  next = (next + 1) % 32;
  if (used < 32) ++used;
  return true;
}
// how many 4-CTA clusters the device holds at once (GPCs whose SM count is not a multiple of 4 leave SMs idle); -1 on error
typedef void (*QuadKernel)(const PairParams, const QuadMaps);
static QuadKernel quad_kernel(int NV, bool merge) {
  if (merge) return NV == 2 ? mlp_quad_fwd_kernel<2, true> : NV == 3 ? mlp_quad_fwd_kernel<3, true> : mlp_quad_fwd_kernel<4, true>;
  return NV == 1 ? mlp_quad_fwd_kernel<1, false> : NV == 2 ? mlp_quad_fwd_kernel<2, false>
         : NV == 3 ? mlp_quad_fwd_kernel<3, false> : mlp_quad_fwd_kernel<4, false>;
}
constexpr int kQuadMaxClusters = 37;      // merge scratch: 4 * 37 CTAs x 128 KB (dnr_mlp_quad_scratch_bytes)
static int quad_max_clusters(int NV) {
  if (NV < 1 || NV > 4) return -1;
  DeviceInfo* di = device_info();
  if (!di) return -1;
  int* cached = di->quad_clusters;
  if (cached[NV]) return cached[NV];
  QuadKernel kern = quad_kernel(NV, false);
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kQSmem) != cudaSuccess) return -1;
  if (NV > 1 && cudaFuncSetAttribute(quad_kernel(NV, true), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kQSmem) !=
                    cudaSuccess)
    return -1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(4 * kQuadMaxClusters, 1, 1); cfg.blockDim = dim3(kPThreads, 1, 1); cfg.dynamicSmemBytes = kQSmem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 4; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, (const void*)kern, &cfg) != cudaSuccess) return -1;
  if (n > kQuadMaxClusters) n = kQuadMaxClusters;       // the parking scratch is sized for 4 * 37 CTAs
  cached[NV] = n;
  return n;
}
}  // namespace dnr

// Number of 4-CTA clusters dnr_mlp_quad_fwd keeps resident (<= 37); the SMs beyond 4x that number stay idle and can
// run dnr_mlp_pair_fwd clusters concurrently.  Negative on error.
DNR_EXPORT int dnr_mlp_quad_max_clusters(int NV) { return dnr::quad_max_clusters(NV); }

// Bytes of the scratch that enables the quad kernel's merge of pooled rows (see MERGE above).
DNR_EXPORT long long dnr_mlp_quad_scratch_bytes(void) { return 4LL * dnr::kQuadMaxClusters * 128 * 256 * 4; }

// The quad variant (clusters of 4 CTAs, tcgen05 cta_group::2; see above).  weights: the same 14 matrices as
// dnr_mlp_pair_fwd but in the PK128 layout [N/128][K/8][128][8]; head_partial: [mv_tiles*4*128] float4
// (dnr_head_finalize with n_partials = 4); merge_scratch: null, or dnr_mlp_quad_scratch_bytes() bytes -> blocks 3-4 and
// the head run once per NV tiles on merged pooled rows (NV > 1; the partial sums land where they would otherwise).
static int quad_fwd_impl(const void* lat_pk, const void* code_pk, const void* const* weights_pk128,
                         const void* const* code_weights, const float* const* biases, const float* w_out, int NV,
                         int mv_tiles, void* merge_scratch, float* head_partial, long long* prof, cudaStream_t stream);
DNR_EXPORT int dnr_mlp_quad_fwd(const void* lat_pk, const void* code_pk, const void* const* weights_pk128,
                                const float* const* biases, const float* w_out, int NV, int mv_tiles,
                                void* merge_scratch, float* head_partial, long long* prof, cudaStream_t stream) {
  return quad_fwd_impl(lat_pk, code_pk, weights_pk128, nullptr, biases, w_out, NV, mv_tiles, merge_scratch, head_partial,
                       prof, stream);
}
// NOVEL_PE counterpart of dnr_mlp_quad_fwd (see dnr_mlp_pair_fwd_pe): code_weights_pk128[b] = PK128 [512 x 64] images.
DNR_EXPORT int dnr_mlp_quad_fwd_pe(const void* lat_pk, const void* code_pk, const void* const* weights_pk128,
                                   const void* const* code_weights_pk128, const float* const* biases, const float* w_out,
                                   int NV, int mv_tiles, void* merge_scratch, float* head_partial, cudaStream_t stream) {
  DNR_REQUIRE(code_weights_pk128 && code_weights_pk128[0] && code_weights_pk128[1], "dnr_mlp_quad_fwd_pe: null code weights");
  return quad_fwd_impl(lat_pk, code_pk, weights_pk128, code_weights_pk128, biases, w_out, NV, mv_tiles, merge_scratch,
                       head_partial, nullptr, stream);
}
static int quad_fwd_impl(const void* lat_pk, const void* code_pk, const void* const* weights_pk128,
                         const void* const* code_weights, const float* const* biases, const float* w_out, int NV,
                         int mv_tiles, void* merge_scratch, float* head_partial, long long* prof, cudaStream_t stream) {
  DNR_REQUIRE(NV >= 1 && NV <= 4, "dnr_mlp_quad_fwd: NV must be in [1,4]");
  DNR_REQUIRE(lat_pk && code_pk && weights_pk128 && biases && w_out && head_partial, "dnr_mlp_quad_fwd: null argument");
  if (mv_tiles <= 0) return 0;
  dnr::PairParams P;
  P.lat_pk = (const unsigned char*)lat_pk; P.code_pk = (const unsigned char*)code_pk;
  P.w_in = (const unsigned char*)weights_pk128[0];
  for (int i = 0; i < 2; ++i) P.w_zc[i] = code_weights ? (const unsigned char*)code_weights[i] : nullptr;
  for (int i = 0; i < 3; ++i) P.w_z[i] = (const unsigned char*)weights_pk128[1 + i];
  for (int i = 0; i < 5; ++i) {
    P.w_0[i] = (const unsigned char*)weights_pk128[4 + i]; P.w_1[i] = (const unsigned char*)weights_pk128[9 + i];
  }
  for (int i = 0; i < 6; ++i) P.cb[i] = biases[i];
  for (int i = 0; i < 5; ++i) P.b_0[i] = biases[6 + i];
  P.w_out = w_out; P.head_partial = (float4*)head_partial;
  P.NV = NV; P.SPW = 32 / NV; P.mv_tiles = mv_tiles; P.prof = prof; P.tma2 = 0;
  P.merge_scratch = (float*)merge_scratch;
  dnr::QuadMaps M;
  memset(&M, 0, sizeof(M));
  {
    // operand delivery: tensor-map TMA completing on the even CTA's barrier (tma2) unless DNR_QUAD_TMA2=0
    static const char* e = getenv("DNR_QUAD_TMA2");
    P.tma2 = (e && e[0] == '0') ? 0 : 1;
    if (P.tma2) {
      bool ok = dnr::cached_kblock_map(&M.w[0], weights_pk128[0], (size_t)512 * 64 * 2);
      for (int i = 1; i < 14 && ok; ++i) ok = dnr::cached_kblock_map(&M.w[i], weights_pk128[i], (size_t)512 * 512 * 2);
      for (int i = 0; i < 2 && ok && code_weights; ++i) ok = dnr::cached_kblock_map(&M.wc[i], code_weights[i], (size_t)512 * 64 * 2);
      ok = ok && dnr::encode_kblock_map(&M.lat, lat_pk, (size_t)mv_tiles * 131072);
      ok = ok && dnr::encode_kblock_map(&M.code, code_pk, (size_t)mv_tiles * 16384);
      if (!ok) { P.tma2 = 0; memset(&M, 0, sizeof(M)); }     // driver entry point missing: the plain bulk-copy delivery
    }
  }
  dnr::QuadKernel kern = dnr::quad_kernel(NV, merge_scratch != nullptr && NV > 1);
  const int max_cl = dnr::quad_max_clusters(NV);
  DNR_REQUIRE(max_cl >= 1, "dnr_mlp_quad_fwd: the device cannot hold a cluster of 4 CTAs with 225 KB of shared memory each");
  const int n_pairs = (mv_tiles + 1) / 2;
  const int clusters = n_pairs < max_cl ? n_pairs : max_cl;
  kern<<<4 * clusters, dnr::kPThreads, dnr::kQSmem, stream>>>(P, M);
  DNR_CHECK_LAUNCH("dnr_mlp_quad_fwd");
  return 0;
}

// n_partials = 2 (dnr_mlp_pair_fwd) or 4 (dnr_mlp_quad_fwd) partial lin_out sums per row and tile.
DNR_EXPORT int dnr_head_finalize(const float* head_partial, const float* bout, int NV, long long n_samples,
                                 int n_partials, float* rgbsigma, cudaStream_t stream) {
  DNR_REQUIRE(NV >= 1 && NV <= 8, "dnr_head_finalize: NV must be in [1,8]");
  DNR_REQUIRE(n_partials == 2 || n_partials == 4, "dnr_head_finalize: n_partials must be 2 or 4");
  if (n_samples <= 0) return 0;
  const int blocks = (int)((n_samples + 255) / 256);
  if (n_partials == 2)
    dnr::head_finalize_kernel<<<blocks, 256, 0, stream>>>((const float4*)head_partial, bout, NV, 32 / NV, n_samples,
                                                         (float4*)rgbsigma);
  else
    dnr::head_finalize4_kernel<<<blocks, 256, 0, stream>>>((const float4*)head_partial, bout, NV, 32 / NV, n_samples,
                                                          (float4*)rgbsigma);
  DNR_CHECK_LAUNCH("dnr_head_finalize");
  return 0;
}

// =====================================================================================================================
// Allocation-free variant of dnr_mlp_quad_fwd.  NO COUNTERPART IN THE REFERENCE: added
// for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that
// forbid it).  The parking scratch of the MERGE path lives in a file-scope device array instead of a caller-allocated
// buffer: kStaticQuadClusters clusters x 4 CTAs x 128 KB = 4 MiB, held by every GPU that loads the library until exit.
// Fewer clusters than dnr_mlp_quad_fwd keeps resident (8 instead of 33), so it is a small-batch / latency entry point.
namespace dnr {
constexpr int kStaticQuadClusters = 8;                 // 32 CTAs
constexpr int kStaticQuadMaxTiles = 4096;              // 524 288 view rows per call (stated maximum of this entry point)
constexpr size_t kStaticQuadFloats = (size_t)4 * kStaticQuadClusters * 128 * 256;
static __device__ __align__(256) float g_quad_static_scratch[kStaticQuadFloats];
// In host code the array's name is only a host-side stand-in; the device address comes from cudaGetSymbolAddress, once
// per device, with that device current.
static float* g_quad_static_ptr[64] = {};
}  // namespace dnr

// dnr_mlp_quad_fwd with the merge scratch inside the library.  Allocates nothing (no cudaMalloc / cudaFree / library
// allocator); inputs and outputs are the caller's.  mv_tiles <= dnr_mlp_quad_static_max_tiles() (4096), else
// cudaErrorInvalidValue with a message.  The scratch is shared by every call on a device and keeps the last call's
// contents: nothing in it has to start at zero (a merge group reads only the rows it parked itself, missing rows are
// zero-filled in registers), but two calls that OVERLAP on the GPU (different streams, or graph branches) race on it --
// serialise them.
DNR_EXPORT int dnr_mlp_quad_static_max_tiles(void) { return dnr::kStaticQuadMaxTiles; }
DNR_EXPORT int dnr_mlp_quad_fwd_static(const void* lat_pk, const void* code_pk, const void* const* weights_pk128,
                                       const float* const* biases, const float* w_out, int NV, int mv_tiles,
                                       float* head_partial, cudaStream_t stream) {
  DNR_REQUIRE(NV >= 1 && NV <= 4, "dnr_mlp_quad_fwd_static: NV must be in [1,4]");
  DNR_REQUIRE(lat_pk && code_pk && weights_pk128 && biases && w_out && head_partial, "dnr_mlp_quad_fwd_static: null argument");
  DNR_REQUIRE(mv_tiles <= dnr::kStaticQuadMaxTiles,
              "dnr_mlp_quad_fwd_static: mv_tiles exceeds the maximum of 4096 tiles (524288 view rows) of the allocation-free "
              "entry point; split the batch or use dnr_mlp_quad_fwd");
  if (mv_tiles <= 0) return 0;
  int dev = -1;
  DNR_CHECK(cudaGetDevice(&dev), "dnr_mlp_quad_fwd_static/device");
  DNR_REQUIRE(dev >= 0 && dev < 64, "dnr_mlp_quad_fwd_static: device index out of range");
  if (!dnr::g_quad_static_ptr[dev]) {          // set-up, once per device, that device current
    float* ptr = nullptr;
    DNR_CHECK(cudaGetSymbolAddress((void**)&ptr, dnr::g_quad_static_scratch), "dnr_mlp_quad_fwd_static/symbol");
    DNR_REQUIRE(ptr != nullptr, "dnr_mlp_quad_fwd_static: cudaGetSymbolAddress returned null");
    dnr::g_quad_static_ptr[dev] = ptr;
  }
  dnr::PairParams P;
  P.lat_pk = (const unsigned char*)lat_pk; P.code_pk = (const unsigned char*)code_pk;
  P.w_in = (const unsigned char*)weights_pk128[0];
  for (int i = 0; i < 3; ++i) P.w_z[i] = (const unsigned char*)weights_pk128[1 + i];
  for (int i = 0; i < 5; ++i) {
    P.w_0[i] = (const unsigned char*)weights_pk128[4 + i]; P.w_1[i] = (const unsigned char*)weights_pk128[9 + i];
  }
  for (int i = 0; i < 6; ++i) P.cb[i] = biases[i];
  for (int i = 0; i < 5; ++i) P.b_0[i] = biases[6 + i];
  P.w_out = w_out; P.head_partial = (float4*)head_partial;
  P.NV = NV; P.SPW = 32 / NV; P.mv_tiles = mv_tiles; P.prof = nullptr; P.tma2 = 0;
  P.w_zc[0] = P.w_zc[1] = nullptr;
  P.merge_scratch = dnr::g_quad_static_ptr[dev];
  dnr::QuadKernel kern = dnr::quad_kernel(NV, NV > 1);
  const int max_cl = dnr::quad_max_clusters(NV);      // also sets the kernels' shared-memory attribute on this device
  DNR_REQUIRE(max_cl >= 1, "dnr_mlp_quad_fwd_static: the device cannot hold a cluster of 4 CTAs with 225 KB of shared memory each");
  const int n_pairs = (mv_tiles + 1) / 2;
  int clusters = n_pairs < dnr::kStaticQuadClusters ? n_pairs : dnr::kStaticQuadClusters;
  if (clusters > max_cl) clusters = max_cl;
  dnr::QuadMaps M;                                     // (plain bulk copies: this entry point keeps the first-version path)
  memset(&M, 0, sizeof(M));
  kern<<<4 * clusters, dnr::kPThreads, dnr::kQSmem, stream>>>(P, M);
  DNR_CHECK_LAUNCH("dnr_mlp_quad_fwd_static");
  return 0;
}
