// Persistent layer-chain kernel for the ResnetFC field MLP, forward / inference
// (reference resnetfc.py:129-159 + combine() :9-14, evaluated per PixelNeRF.forward pixelnerf.py:128-134).
//
// The per-layer GEMM kernel (gemm.cu) round-trips every activation through HBM: the fp32 residual stream
// alone is 2 KB per row and layer, which makes the residual layers HBM-bound (ncu: 4.2 TB/s, tensor pipe
// 17 %).  Here ONE persistent CTA per SM walks a group of NV view-row tiles (+ the pooled tile they reduce
// to) through all 13 layers.  Activations of the tile live in a 1 MB per-CTA scratch in global memory that is
// re-used for every tile and therefore stays L2-resident (148 MB allocated, ~80 MB hot); HBM only sees the
// gathered inputs (lat/code operands) once and the final activation once.  Weights stream from L2 as before.
//
//   warp 0      : bulk-copy producer (cp.async.bulk -> 4-stage smem ring)
//   warp 1      : tcgen05.mma issuer (M=128, N=256, K=16; two 256-column accumulators ping-pong in TMEM)
//   warps 2..9  : epilogue (two warps per TMEM lane quadrant, 128 columns each): bias, residual, view mean,
//                 ReLU, bf16 repack -> scratch
// Every layer ("op") is two jobs (the two 256-column halves of the 512 outputs).  The next op's K-blocks
// 0..3 only need the first job's columns, blocks 4..7 the second job's, so the producer waits on two
// per-job mbarriers and the MMAs of op l+1 start while the epilogue of op l's second half still runs.
#include "common.cuh"
#include "tc05.cuh"

namespace dnr {

constexpr int kFBM = 128, kFBN = 256, kFBK = 64;
constexpr int kFStages = 4;
constexpr int kFAStage = kFBM * kFBK * 2;     // 16 KB
constexpr int kFBStage = kFBN * kFBK * 2;     // 32 KB
constexpr int kFEpiWarps = 8;
constexpr int kFThreads = (2 + kFEpiWarps) * 32;   // 320
constexpr size_t kFSmem = (size_t)kFStages * (kFAStage + kFBStage) + 256;
constexpr int kHidden = 512;
constexpr size_t kTileXR = (size_t)128 * kHidden * 4;   // 256 KB
constexpr size_t kTilePK = (size_t)128 * kHidden * 2;   // 128 KB
constexpr size_t kScratchPerCta = 2 * kTileXR + 4 * kTilePK;   // x_v, x_p, xa_v, net_v, xa_p, net_p = 1 MB

struct FusedParams {
  const unsigned char* lat_pk;    // PK128 [mv_tiles x 512]
  const unsigned char* code_pk;   // PK128 [mv_tiles x 64]
  const unsigned char* w_in;      // PK256 [512 x 64]
  const unsigned char* w_z[3];    // PK256 [512 x 512]
  const unsigned char* w_0[5];
  const unsigned char* w_1[5];
  const float* b_first;           // lin_in.bias + lin_z[0].bias
  const float* b_0[5];
  const float* b_1[5];            // fc_1.bias (+ lin_z[b+1].bias for b < 2)
  unsigned char* scratch;         // gridDim.x * kScratchPerCta
  unsigned char* xa_out;          // PK128 [n_groups x 128 x 512]: relu(x) after the last block, pooled rows
  int NV, SPW;
  int mv_tiles;                   // view-row tiles
  int n_groups;                   // ceil(mv_tiles / NV)
  long long n_samples;
};

struct OpDesc {
  int nseg;
  const unsigned char* A[2];
  const unsigned char* B[2];
  int K[2];
  const float* bias;
  const float* res_in;     // XR tile or null
  float* res_out;          // XR tile or null
  unsigned char* act_out;  // PK tile
  int mean;                // view mean-pool epilogue (outputs indexed by pooled row)
  int pooled_base;         // first pooled row of this view tile inside the group's pooled tile
  long long sample_base;   // first sample of the group
  int a_stream[2];         // segment comes straight from HBM (lat / code): evict-first in L2
  int out_stream;          // act_out leaves the kernel (xa_out): evict-first
};

// Deterministic schedule shared by the three roles.  Returns false when the op is skipped (tail group).
__device__ __forceinline__ bool fused_get_op(const FusedParams& P, unsigned char* scr, long long group, int op,
                                             OpDesc& d) {
  float* x_v = reinterpret_cast<float*>(scr);
  float* x_p = reinterpret_cast<float*>(scr + kTileXR);
  unsigned char* xa_v = scr + 2 * kTileXR;
  unsigned char* net_v = xa_v + kTilePK;
  unsigned char* xa_p = net_v + kTilePK;
  unsigned char* net_p = xa_p + kTilePK;
  const int NV = P.NV;
  d.nseg = 1; d.A[1] = nullptr; d.B[1] = nullptr; d.K[0] = kHidden; d.K[1] = 0;
  d.res_in = nullptr; d.res_out = nullptr; d.mean = 0; d.pooled_base = 0;
  d.a_stream[0] = d.a_stream[1] = 0; d.out_stream = 0;
  d.sample_base = group * (long long)(NV * 4 * P.SPW);
  if (op < NV * 7) {
    const int t = op / 7, o = op % 7;
    const long long vt = group * NV + t;
    if (vt >= P.mv_tiles) return false;
    const unsigned char* lat = P.lat_pk + (size_t)vt * kTilePK;
    if (o == 0) {                       // x = lin_in(code) + lin_z[0](lat)
      d.nseg = 2;
      d.A[0] = P.code_pk + (size_t)vt * (128 * 64 * 2); d.B[0] = P.w_in; d.K[0] = 64;
      d.A[1] = lat; d.B[1] = P.w_z[0]; d.K[1] = kHidden;
      d.bias = P.b_first; d.res_out = x_v; d.act_out = xa_v; d.a_stream[0] = d.a_stream[1] = 1;
    } else if (o & 1) {                 // net = relu(fc_0(relu(x)))
      const int b = (o - 1) >> 1;
      d.A[0] = xa_v; d.B[0] = P.w_0[b]; d.bias = P.b_0[b]; d.act_out = net_v;
    } else {                            // x += fc_1(net) [+ lin_z[b+1](lat)]  /  view mean after block 2
      const int b = (o - 2) >> 1;
      d.A[0] = net_v; d.B[0] = P.w_1[b]; d.bias = P.b_1[b]; d.res_in = x_v;
      if (b < 2) {
        d.nseg = 2; d.A[1] = lat; d.B[1] = P.w_z[b + 1]; d.K[1] = kHidden; d.a_stream[1] = 1;
        d.res_out = x_v; d.act_out = xa_v;
      } else {
        d.mean = NV; d.pooled_base = t * 4 * P.SPW; d.res_out = x_p; d.act_out = xa_p;
      }
    }
    return true;
  }
  const int q = op - NV * 7;            // pooled blocks 3, 4
  const int b = 3 + (q >> 1);
  if ((q & 1) == 0) {
    d.A[0] = xa_p; d.B[0] = P.w_0[b]; d.bias = P.b_0[b]; d.act_out = net_p;
  } else {
    d.A[0] = net_p; d.B[0] = P.w_1[b]; d.bias = P.b_1[b]; d.res_in = x_p;
    if (b == 3) { d.res_out = x_p; d.act_out = xa_p; }
    else { d.res_out = nullptr; d.act_out = P.xa_out + (size_t)group * kTilePK; d.out_stream = 1; }
  }
  return true;
}

__device__ __forceinline__ float4 ldcg4(const float4* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

__device__ __forceinline__ void fused_epilogue_chunk(const OpDesc& d, const FusedParams& P, int n0, int row_in_tile,
                                                     int lane, const uint32_t (&acc)[32], const float4 (&r)[8],
                                                     uint64_t pol_keep, uint64_t pol_stream) {
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    const float4 b = __ldg(reinterpret_cast<const float4*>(d.bias + n0 + j));
    v[j] = __uint_as_float(acc[j]) + b.x; v[j + 1] = __uint_as_float(acc[j + 1]) + b.y;
    v[j + 2] = __uint_as_float(acc[j + 2]) + b.z; v[j + 3] = __uint_as_float(acc[j + 3]) + b.w;
  }
  if (d.res_in) {
#pragma unroll
    for (int c = 0; c < 8; ++c) { v[c * 4] += r[c].x; v[c * 4 + 1] += r[c].y; v[c * 4 + 2] += r[c].z; v[c * 4 + 3] += r[c].w; }
  }
  int out_row = row_in_tile;
  bool write = true;
  if (d.mean) {
    const int NV = d.mean, SPW = P.SPW;
    float s[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) s[j] = v[j];
    for (int dd = 1; dd < NV; ++dd) {
#pragma unroll
      for (int j = 0; j < 32; ++j) s[j] += __shfl_down_sync(0xffffffffu, v[j], dd);
    }
    const float nv = (float)NV;
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = s[j] / nv;               // torch.mean over the view rows
    out_row = d.pooled_base + (row_in_tile >> 5) * SPW + lane / NV;
    write = (lane % NV == 0) && (lane < SPW * NV) && (d.sample_base + out_row < P.n_samples);
  }
  if (!write) return;
  if (d.res_out) {
    float4* op = reinterpret_cast<float4*>(d.res_out) + (size_t)(n0 / 4) * 128 + out_row;
#pragma unroll
    for (int c = 0; c < 8; ++c)
      tc::st_hint_v4(op + (size_t)c * 128, __float_as_uint(v[c * 4]), __float_as_uint(v[c * 4 + 1]),
                     __float_as_uint(v[c * 4 + 2]), __float_as_uint(v[c * 4 + 3]), pol_keep);
  }
  unsigned char* ap = d.act_out + ((size_t)(n0 / 8) * 128 + out_row) * 16;
  const uint64_t pol_act = d.out_stream ? pol_stream : pol_keep;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint32_t w[4];
#pragma unroll
    for (int q = 0; q < 4; ++q)
      w[q] = pack_bf16x2(fmaxf(v[c * 8 + q * 2], 0.f), fmaxf(v[c * 8 + q * 2 + 1], 0.f));
    tc::st_hint_v4(ap + (size_t)c * 2048, w[0], w[1], w[2], w[3], pol_act);
  }
}

__global__ void __launch_bounds__(kFThreads, 1) mlp_fused_fwd_kernel(const FusedParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* sA = smem;
  unsigned char* sB = smem + (size_t)kFStages * kFAStage;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)kFStages * (kFAStage + kFBStage));
  uint64_t* full = bars;                  // [kFStages]
  uint64_t* empty = bars + kFStages;      // [kFStages]
  uint64_t* tfull = bars + 2 * kFStages;  // [2]
  uint64_t* tempty = tfull + 2;           // [2]
  uint64_t* done = tempty + 2;            // [2] epilogue of the op's first / second job has landed in scratch
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* scr = P.scratch + (size_t)blockIdx.x * kScratchPerCta;
  const int n_ops = P.NV * 7 + 4;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kFStages; ++s) { tc::mbar_init(full + s, 1); tc::mbar_init(empty + s, 1); }
    for (int b = 0; b < 2; ++b) {
      tc::mbar_init(tfull + b, 1); tc::mbar_init(tempty + b, kFEpiWarps * 32); tc::mbar_init(done + b, kFEpiWarps * 32);
    }
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= producer =================
    {     // the whole warp runs the loop (warp-uniform operands -> uniform registers); one elected lane issues (tc::elect_one)
      int stage = 0; uint32_t phase = 0;
      int c = 0;                                   // ops executed so far by this CTA
      const uint64_t pol_keep = tc::policy_evict_last(), pol_stream = tc::policy_evict_first();
      OpDesc d;
      for (long long group = blockIdx.x; group < P.n_groups; group += gridDim.x) {
        for (int op = 0; op < n_ops; ++op) {
          if (!fused_get_op(P, scr, group, op, d)) continue;
          const uint32_t dep_parity = (uint32_t)((c - 1) & 1);
          for (int n_tile = 0; n_tile < 2; ++n_tile) {
            for (int sg = 0; sg < d.nseg; ++sg) {
              const int K = d.K[sg];
              const unsigned char* a = d.A[sg];
              const unsigned char* b = d.B[sg] + (size_t)n_tile * (K / 8) * 4096;
              for (int kb = 0; kb < K / kFBK; ++kb) {
                if (c > 0 && n_tile == 0 && sg == 0) {
                  // K-blocks 0..3 of the input need the previous op's first job, 4..7 its second job
                  if (kb == 0) tc::mbar_wait(done + 0, dep_parity);
                  if ((d.K[0] >= kHidden && kb == 4) || (d.K[0] < kHidden && kb == 0)) tc::mbar_wait(done + 1, dep_parity);
                }
                tc::mbar_wait(empty + stage, phase ^ 1);
                if (tc::elect_one()) {
                  tc::mbar_arrive_expect_tx(full + stage, kFAStage + kFBStage);
                  tc::bulk_g2s_hint(sA + (size_t)stage * kFAStage, a + (size_t)kb * kFAStage, kFAStage, full + stage,
                                    d.a_stream[sg] ? pol_stream : pol_keep);
                  tc::bulk_g2s_hint(sB + (size_t)stage * kFBStage, b + (size_t)kb * kFBStage, kFBStage, full + stage, pol_keep);
                }
                __syncwarp();
                if (++stage == kFStages) { stage = 0; phase ^= 1; }
              }
            }
          }
          ++c;
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    {     // whole warp, one elected lane issues
      constexpr uint32_t idesc = tc::make_idesc_bf16(kFBM, kFBN, 0, 0);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      OpDesc d;
      for (long long group = blockIdx.x; group < P.n_groups; group += gridDim.x) {
        for (int op = 0; op < n_ops; ++op) {
          if (!fused_get_op(P, scr, group, op, d)) continue;
          for (int n_tile = 0; n_tile < 2; ++n_tile, ++it) {
            const int buf = it & 1;
            tc::mbar_wait(tempty + buf, ((it >> 1) & 1) ^ 1);
            tc::fence_after_sync();
            const uint32_t d_tmem = tmem_base + (uint32_t)buf * kFBN;
            uint32_t accumulate = 0;
            for (int sg = 0; sg < d.nseg; ++sg) {
              for (int kb = 0; kb < d.K[sg] / kFBK; ++kb) {
                tc::mbar_wait(full + stage, phase);
                tc::fence_after_sync();
                const uint32_t a0 = tc::smem_u32(sA + (size_t)stage * kFAStage);
                const uint32_t b0 = tc::smem_u32(sB + (size_t)stage * kFBStage);
                if (tc::elect_one()) {
#pragma unroll
                  for (int j = 0; j < kFBK / 16; ++j) {
                    const uint64_t ad = tc::make_smem_desc(a0 + j * 2 * 2048, 2048, 128);
                    const uint64_t bd = tc::make_smem_desc(b0 + j * 2 * 4096, 4096, 128);
                    tc::mma_bf16(d_tmem, ad, bd, idesc, j == 0 ? accumulate : 1u);
                  }
                  tc::mma_commit(empty + stage);
                }
                __syncwarp();
                accumulate = 1;
                if (++stage == kFStages) { stage = 0; phase ^= 1; }
              }
            }
            if (tc::elect_one()) tc::mma_commit(tfull + buf);
            __syncwarp();
          }
        }
      }
    }
  } else {
    // ================= epilogue =================
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row_in_tile = quad * 32 + lane;
    int it = 0;
    OpDesc d;
    const uint64_t pol_keep = tc::policy_evict_last(), pol_stream = tc::policy_evict_first();
    for (long long group = blockIdx.x; group < P.n_groups; group += gridDim.x) {
      for (int op = 0; op < n_ops; ++op) {
        if (!fused_get_op(P, scr, group, op, d)) continue;
        for (int n_tile = 0; n_tile < 2; ++n_tile, ++it) {
          const int buf = it & 1;
          const int nb = n_tile * kFBN + half * 128;
          const float4* rp = d.res_in ? reinterpret_cast<const float4*>(d.res_in) + (size_t)(nb / 4) * 128 + row_in_tile
                                      : nullptr;
          float4 r[8];
          if (rp) {
#pragma unroll
            for (int c = 0; c < 8; ++c) r[c] = ldcg4(rp + (size_t)c * 128);
          }
          tc::mbar_wait(tfull + buf, (it >> 1) & 1);
          tc::fence_after_sync();
          const uint32_t t0 = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)buf * kFBN + half * 128;
#pragma unroll 1
          for (int ch = 0; ch < 4; ++ch) {
            uint32_t acc[32];
            tc::tmem_ld32(t0 + ch * 32, acc);
            float4 rn[8];
            if (rp && ch < 3) {
#pragma unroll
              for (int c = 0; c < 8; ++c) rn[c] = ldcg4(rp + (size_t)((ch + 1) * 8 + c) * 128);
            }
            tc::tmem_ld_wait();
            fused_epilogue_chunk(d, P, nb + ch * 32, row_in_tile, lane, acc, r, pol_keep, pol_stream);
            if (rp && ch < 3) {
#pragma unroll
              for (int c = 0; c < 8; ++c) r[c] = rn[c];
            }
          }
          tc::fence_before_sync();
          tc::mbar_arrive(tempty + buf);
          // make this thread's scratch writes visible to the bulk-copy (async proxy) reads of the next op
          __threadfence();
          asm volatile("fence.proxy.async;" ::: "memory");
          tc::mbar_arrive(done + n_tile);
        }
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 1) { tc::fence_after_sync(); tc::tmem_dealloc(tmem_base, 512); }
}

}  // namespace dnr

// Whole ResnetFC forward (all 13 linear layers, view mean fused) for mv_tiles view-row tiles produced by
// dnr_encode_fwd; leaves relu(x) after the last block in xa_out (pooled tiles of 4*NV*floor(32/NV) samples each),
// ready for dnr_head_fwd_grouped.  weights: [w_in, w_z0..2, w_00..04, w_10..14] (PK256), biases:
// [b_first, b_00..04, b_10..14] (fp32 [512]).  scratch: dnr_mlp_fused_scratch_bytes() bytes.
DNR_EXPORT long long dnr_mlp_fused_scratch_bytes(int max_ctas) {
  int n = 148;
  if (max_ctas > 0 && max_ctas < n) n = max_ctas;
  return (long long)n * (long long)dnr::kScratchPerCta;
}

DNR_EXPORT int dnr_mlp_fused_fwd(const void* lat_pk, const void* code_pk, const void* const* weights,
                                 const float* const* biases, int NV, int mv_tiles, long long n_samples, void* scratch,
                                 void* xa_out, int max_ctas, cudaStream_t stream) {
  DNR_REQUIRE(NV >= 1 && NV <= 8, "dnr_mlp_fused_fwd: NV must be in [1,8]");
  DNR_REQUIRE(lat_pk && code_pk && weights && biases && scratch && xa_out, "dnr_mlp_fused_fwd: null argument");
  if (mv_tiles <= 0) return 0;
  dnr::FusedParams P;
  P.lat_pk = (const unsigned char*)lat_pk; P.code_pk = (const unsigned char*)code_pk;
  P.w_in = (const unsigned char*)weights[0];
  for (int i = 0; i < 3; ++i) P.w_z[i] = (const unsigned char*)weights[1 + i];
  for (int i = 0; i < 5; ++i) { P.w_0[i] = (const unsigned char*)weights[4 + i]; P.w_1[i] = (const unsigned char*)weights[9 + i]; }
  P.b_first = biases[0];
  for (int i = 0; i < 5; ++i) { P.b_0[i] = biases[1 + i]; P.b_1[i] = biases[6 + i]; }
  P.scratch = (unsigned char*)scratch; P.xa_out = (unsigned char*)xa_out;
  P.NV = NV; P.SPW = 32 / NV; P.mv_tiles = mv_tiles; P.n_groups = (mv_tiles + NV - 1) / NV; P.n_samples = n_samples;
  dnr::DeviceInfo* di = dnr::device_info();
  DNR_REQUIRE(di != nullptr, "dnr_mlp_fused_fwd: no current CUDA device");
  if (!di->fused_configured) {
    DNR_CHECK(cudaFuncSetAttribute(dnr::mlp_fused_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)dnr::kFSmem), "dnr_mlp_fused_fwd/attr");
    di->fused_configured = true;
  }
  const int num_sms = di->num_sms;
  int grid = P.n_groups;
  int cap = num_sms < 148 ? num_sms : 148;
  if (max_ctas > 0 && max_ctas < cap) cap = max_ctas;
  if (grid > cap) grid = cap;
  dnr::mlp_fused_fwd_kernel<<<grid, dnr::kFThreads, dnr::kFSmem, stream>>>(P);
  DNR_CHECK_LAUNCH("dnr_mlp_fused_fwd");
  return 0;
}
