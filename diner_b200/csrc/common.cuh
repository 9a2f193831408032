// Shared helpers for the diner_b200 CUDA kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#ifndef DNR_EXPORT
#define DNR_EXPORT extern "C" __attribute__((visibility("default")))
#endif

namespace dnr {

// ---- error reporting ---------------------------------------------------------------------
void set_last_error(const char* where, cudaError_t e);
void set_last_error_msg(const char* msg);

#define DNR_CHECK_LAUNCH(where)                                  \
  do {                                                           \
    cudaError_t e__ = cudaGetLastError();                        \
    if (e__ != cudaSuccess) {                                    \
      dnr::set_last_error(where, e__);                           \
      return (int)e__;                                           \
    }                                                            \
  } while (0)

#define DNR_CHECK(call, where)                                   \
  do {                                                           \
    cudaError_t e__ = (call);                                    \
    if (e__ != cudaSuccess) {                                    \
      dnr::set_last_error(where, e__);                           \
      return (int)e__;                                           \
    }                                                            \
  } while (0)

#define DNR_REQUIRE(cond, msg)                                   \
  do {                                                           \
    if (!(cond)) {                                               \
      dnr::set_last_error_msg(msg);                              \
      return (int)cudaErrorInvalidValue;                         \
    }                                                            \
  } while (0)

// ---- per-device launch state ---------------------------------------------------------------------
// cudaFuncSetAttribute and the SM count are per DEVICE: one process may render on several GPUs in turn, so nothing of
// this lives in function-local statics.  device_info() returns the slot of the calling thread's current device
// (nullptr when there is none); the flags are idempotent, so unsynchronised first use from two threads is harmless.
struct DeviceInfo {
  int num_sms;
  bool gemm_configured, fused_configured, wgrad2_configured, pair_configured[5];
  size_t sampler_smem, wgrad_smem;
  int quad_clusters[5];
};
DeviceInfo* device_info();

// ---- row tiling of the MLP batch ------------------------------------------------------------
// Per-view rows are ordered sample-major / view-minor and grouped so that the NV rows of one
// sample never straddle a warp of the GEMM epilogue (32 TMEM lanes): each warp-slab of 32 rows
// holds SPW = 32 / NV samples (NV*SPW valid rows, the rest zero padding).
struct RowMap {
  int NV;        // source views
  int SPW;       // samples per 32-row slab
  __host__ __device__ static RowMap make(int nv) {
    RowMap m; m.NV = nv; m.SPW = 32 / nv; return m;
  }
  // number of 128-row tiles needed for n samples
  __host__ __device__ int tiles_for(long long n_samples) const {
    long long per_tile = 4LL * SPW;
    return (int)((n_samples + per_tile - 1) / per_tile);
  }
  // row -> (sample, view); returns false for padding rows
  __host__ __device__ bool decode(long long row, long long n_samples, long long& sample, int& view) const {
    long long slab = row >> 5;
    int l = (int)(row & 31);
    if (l >= SPW * NV) return false;
    sample = slab * SPW + l / NV;
    view = l % NV;
    return sample < n_samples;
  }
};

// ---- packed operand layouts -------------------------------------------------------------------
// PK<TH>: bf16 matrix [R x F] stored as [R/TH][F/8][TH][8]: every 8-feature chunk of a TH-row tile
// is TH*16 contiguous bytes.  This is the tcgen05 canonical no-swizzle ("interleaved") smem image
// for both a K-major operand (rows = M/N, features = K: SBO = 128 B, LBO = TH*16 B) and an
// MN-major operand (features = M/N, rows = K: SBO = TH*16 B, LBO = 128 B), so a K-block of any
// width is one contiguous range that a single cp.async.bulk moves into shared memory.
template <int TH>
__host__ __device__ inline size_t pk_offset_bytes(long long r, int f, int F) {
  long long tile = r / TH;
  int rr = (int)(r % TH);
  return ((size_t)(tile * (F / 8) + f / 8) * TH + rr) * 16 + (size_t)(f % 8) * 2;
}
// XR: fp32 matrix [R x N] stored as [R/128][N/4][128] float4 (the order the GEMM epilogue
// reads TMEM in: one row per lane, 4 columns per register quad) -> 512 B contiguous per warp.
__host__ __device__ inline size_t xr_offset_floats(long long r, int n, int N) {
  long long tile = r / 128;
  int rr = (int)(r % 128);
  return ((size_t)(tile * (N / 4) + n / 4) * 128 + rr) * 4 + (n % 4);
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float bf16lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- counter-based RNG (Philox4x32-10) for the performance-mode sampler ------------------------
struct Philox {
  uint32_t key0, key1;
  __device__ __forceinline__ Philox(uint64_t seed) : key0((uint32_t)seed), key1((uint32_t)(seed >> 32)) {}
  __device__ __forceinline__ uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const {
    uint32_t k0 = key0, k1 = key1;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
      c0 = n0; c1 = n1; c2 = n2; c3 = n3;
      k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x & 0xffffffu) * (1.0f / 16777216.0f); }

}  // namespace dnr
