// Small SIMT kernels around the tcgen05 GEMMs of the field MLP: weight packing, the 512->4 output head
// with its activations (reference resnetfc.py:158, pixelnerf.py:137-145) and the backward of the view
// mean-pool (resnetfc.py:9-14).
#include "common.cuh"

namespace dnr {

// W (row-major fp32, [rows_w x cols_w], leading dimension ld) -> PK<256> bf16 [N_pad x K_pad] with
// out(n,k) = transpose ? W[k][n] : W[n][k]; zero outside the source.
__global__ void pack_weight_kernel(const float* __restrict__ W, int rows_w, int cols_w, int ld, int transpose,
                                   int N_pad, int K_pad, __nv_bfloat16* __restrict__ out) {
  const size_t total = (size_t)N_pad * K_pad;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    // idx enumerates the packed image: [N_pad/256][K_pad/8][256][8]
    const int e = (int)(idx & 7);
    const int r = (int)((idx >> 3) & 255);
    const size_t chunk = idx >> 11;
    const int kc = (int)(chunk % (K_pad / 8));
    const int nt = (int)(chunk / (K_pad / 8));
    const int n = nt * 256 + r, k = kc * 8 + e;
    float v = 0.f;
    if (!transpose) { if (n < rows_w && k < cols_w) v = W[(size_t)n * ld + k]; }
    else            { if (k < rows_w && n < cols_w) v = W[(size_t)k * ld + n]; }
    out[idx] = __float2bfloat16_rn(v);
  }
}

constexpr int kHid = 512;

// out[s, 0:3] = sigmoid(lin_out(xa)[0:3]),  out[s, 3] = relu(lin_out(xa)[3]); xa = relu(x) is PK<128> bf16.
__global__ void __launch_bounds__(128) head_fwd_kernel(const unsigned char* __restrict__ xa_pk,
                                                       const unsigned char* __restrict__ xa_lo_pk,   // or null
                                                       const float* __restrict__ Wout,  // [4,512]
                                                       const float* __restrict__ bout,  // [4]
                                                       long long n_samples, int samples_per_tile,
                                                       float4* __restrict__ out) {
  __shared__ float sw[4 * kHid];
  for (int i = threadIdx.x; i < 4 * kHid; i += blockDim.x) sw[i] = Wout[i];
  __syncthreads();
  // tiles hold samples_per_tile (<= 128) samples each: 128 for the per-layer path, 4*NV*floor(32/NV) for the
  // fused kernel's pooled tiles
  const long long row = (int)threadIdx.x < samples_per_tile ? (long long)blockIdx.x * samples_per_tile + threadIdx.x
                                                            : n_samples;
  const unsigned char* base = xa_pk + (size_t)blockIdx.x * (kHid / 8) * 2048 + (size_t)threadIdx.x * 16;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll 4
  for (int c = 0; c < kHid / 8; ++c) {
    const uint4 p = __ldg(reinterpret_cast<const uint4*>(base + (size_t)c * 2048));
    const uint32_t pv[4] = {p.x, p.y, p.z, p.w};
    uint4 pl = make_uint4(0u, 0u, 0u, 0u);
    if (xa_lo_pk) pl = __ldg(reinterpret_cast<const uint4*>(xa_lo_pk + (base - xa_pk) + (size_t)c * 2048));
    const uint32_t plv[4] = {pl.x, pl.y, pl.z, pl.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float x0 = bf16lo(pv[q]) + bf16lo(plv[q]), x1 = bf16hi(pv[q]) + bf16hi(plv[q]);     // (+ 0 without a lo image)
      const int k = c * 8 + q * 2;
      a0 = fmaf(x0, sw[k], a0); a0 = fmaf(x1, sw[k + 1], a0);
      a1 = fmaf(x0, sw[kHid + k], a1); a1 = fmaf(x1, sw[kHid + k + 1], a1);
      a2 = fmaf(x0, sw[2 * kHid + k], a2); a2 = fmaf(x1, sw[2 * kHid + k + 1], a2);
      a3 = fmaf(x0, sw[3 * kHid + k], a3); a3 = fmaf(x1, sw[3 * kHid + k + 1], a3);
    }
  }
  if (row < n_samples) {
    a0 += bout[0]; a1 += bout[1]; a2 += bout[2]; a3 += bout[3];
    out[row] = make_float4(1.f / (1.f + expf(-a0)), 1.f / (1.f + expf(-a1)), 1.f / (1.f + expf(-a2)), fmaxf(a3, 0.f));
  }
}

// Backward of the head: d_raw = d_out * act'(out); d_x = (d_raw @ Wout) * [xa > 0]  (fp32 XR + bf16 PK);
// d_raw is also written as a PK<128> [rows x 128] bf16 operand (columns 0..3) for the wgrad GEMM.
__global__ void __launch_bounds__(128) head_bwd_kernel(const unsigned char* __restrict__ xa_pk,
                                                       const float* __restrict__ Wout, const float4* __restrict__ out,
                                                       const float4* __restrict__ d_out, long long n_samples,
                                                       float* __restrict__ dx_xr, unsigned char* __restrict__ dx_pk,
                                                       unsigned char* __restrict__ draw_pk) {
  __shared__ float sw[4 * kHid];
  for (int i = threadIdx.x; i < 4 * kHid; i += blockDim.x) sw[i] = Wout[i];
  __syncthreads();
  const long long row = (long long)blockIdx.x * 128 + threadIdx.x;
  float g0 = 0.f, g1 = 0.f, g2 = 0.f, g3 = 0.f;
  if (row < n_samples) {
    const float4 y = out[row], d = d_out[row];
    g0 = d.x * y.x * (1.f - y.x); g1 = d.y * y.y * (1.f - y.y); g2 = d.z * y.z * (1.f - y.z);
    g3 = y.w > 0.f ? d.w : 0.f;
  }
  *reinterpret_cast<uint4*>(draw_pk + (size_t)blockIdx.x * (128 / 8) * 2048 + (size_t)threadIdx.x * 16) =
      make_uint4(pack_bf16x2(g0, g1), pack_bf16x2(g2, g3), 0u, 0u);
  const unsigned char* xb = xa_pk + (size_t)blockIdx.x * (kHid / 8) * 2048 + (size_t)threadIdx.x * 16;
  unsigned char* pb = dx_pk + (size_t)blockIdx.x * (kHid / 8) * 2048 + (size_t)threadIdx.x * 16;
  float4* xr = reinterpret_cast<float4*>(dx_xr) + (size_t)blockIdx.x * (kHid / 4) * 128 + threadIdx.x;
#pragma unroll 2
  for (int c = 0; c < kHid / 8; ++c) {
    const uint4 p = __ldg(reinterpret_cast<const uint4*>(xb + (size_t)c * 2048));
    const uint32_t pv[4] = {p.x, p.y, p.z, p.w};
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int k = c * 8 + e;
      const float xa = (e & 1) ? bf16hi(pv[e >> 1]) : bf16lo(pv[e >> 1]);
      const float g = g0 * sw[k] + g1 * sw[kHid + k] + g2 * sw[2 * kHid + k] + g3 * sw[3 * kHid + k];
      v[e] = xa > 0.f ? g : 0.f;
    }
    if (dx_xr) {
      xr[(size_t)(c * 2) * 128] = make_float4(v[0], v[1], v[2], v[3]);
      xr[(size_t)(c * 2 + 1) * 128] = make_float4(v[4], v[5], v[6], v[7]);
    }
    *reinterpret_cast<uint4*>(pb + (size_t)c * 2048) =
        make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
  }
}

// Backward of the view mean: d_x[(sample, view)] = d_mean[sample] / NV, for every view row (zeros on padding).
__global__ void __launch_bounds__(128) mean_bwd_kernel(const float* __restrict__ dmean_xr, int NV, long long n_samples,
                                                       float* __restrict__ dx_xr, unsigned char* __restrict__ dx_pk) {
  const RowMap rm = RowMap::make(NV);
  const long long row = (long long)blockIdx.x * 128 + threadIdx.x;
  long long sample; int view;
  const bool valid = rm.decode(row, n_samples, sample, view);
  const float inv = 1.f / (float)NV;
  const float4* src = nullptr;
  if (valid) src = reinterpret_cast<const float4*>(dmean_xr) + (size_t)(sample >> 7) * (kHid / 4) * 128 + (sample & 127);
  float4* xr = reinterpret_cast<float4*>(dx_xr) + (size_t)blockIdx.x * (kHid / 4) * 128 + threadIdx.x;
  unsigned char* pb = dx_pk + (size_t)blockIdx.x * (kHid / 8) * 2048 + (size_t)threadIdx.x * 16;
#pragma unroll 2
  for (int c = 0; c < kHid / 8; ++c) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    if (valid) { a = __ldg(src + (size_t)(c * 2) * 128); b = __ldg(src + (size_t)(c * 2 + 1) * 128); }
    a.x *= inv; a.y *= inv; a.z *= inv; a.w *= inv; b.x *= inv; b.y *= inv; b.z *= inv; b.w *= inv;
    if (dx_xr) { xr[(size_t)(c * 2) * 128] = a; xr[(size_t)(c * 2 + 1) * 128] = b; }
    *reinterpret_cast<uint4*>(pb + (size_t)c * 2048) =
        make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
  }
}

// The same for the bf16 gradient stream: the pooled gradient arrives as its PK<128> bf16 image (indexed by sample).
__global__ void __launch_bounds__(128) mean_bwd_pk_kernel(const unsigned char* __restrict__ dmean_pk, int NV,
                                                          long long n_samples, unsigned char* __restrict__ dx_pk) {
  const RowMap rm = RowMap::make(NV);
  const long long row = (long long)blockIdx.x * 128 + threadIdx.x;
  long long sample; int view;
  const bool valid = rm.decode(row, n_samples, sample, view);
  const float inv = 1.f / (float)NV;
  const unsigned char* src = nullptr;
  if (valid) src = dmean_pk + ((size_t)(sample >> 7) * (kHid / 8) * 128 + (size_t)(sample & 127)) * 16;
  unsigned char* pb = dx_pk + (size_t)blockIdx.x * (kHid / 8) * 2048 + (size_t)threadIdx.x * 16;
#pragma unroll 4
  for (int c = 0; c < kHid / 8; ++c) {
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (valid) {
      const uint4 a = __ldg(reinterpret_cast<const uint4*>(src + (size_t)c * 2048));
      o = make_uint4(pack_bf16x2(bf16lo(a.x) * inv, bf16hi(a.x) * inv), pack_bf16x2(bf16lo(a.y) * inv, bf16hi(a.y) * inv),
                     pack_bf16x2(bf16lo(a.z) * inv, bf16hi(a.z) * inv), pack_bf16x2(bf16lo(a.w) * inv, bf16hi(a.w) * inv));
    }
    *reinterpret_cast<uint4*>(pb + (size_t)c * 2048) = o;
  }
}

// debugging / tests: row-major fp32 <-> packed layouts
__global__ void pack_rows_kernel(const float* __restrict__ src, long long rows, int F, long long rows_pad,
                                 __nv_bfloat16* __restrict__ out) {
  const size_t total = (size_t)rows_pad * F;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int e = (int)(idx & 7);
    const int r = (int)((idx >> 3) & 127);
    const size_t chunk = idx >> 10;
    const int fc = (int)(chunk % (F / 8));
    const long long t = (long long)(chunk / (F / 8));
    const long long row = t * 128 + r;
    const int f = fc * 8 + e;
    out[idx] = __float2bfloat16_rn(row < rows ? src[(size_t)row * F + f] : 0.f);
  }
}
__global__ void unpack_rows_kernel(const __nv_bfloat16* __restrict__ pk, long long rows, int F, float* __restrict__ out) {
  const size_t total = (size_t)rows * F;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const long long row = (long long)(idx / F);
    const int f = (int)(idx % F);
    out[idx] = __bfloat162float(pk[pk_offset_bytes<128>(row, f, F) / 2]);
  }
}
__global__ void xr_to_rows_kernel(const float* __restrict__ xr, long long rows, int N, float* __restrict__ out) {
  const size_t total = (size_t)rows * N;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const long long row = (long long)(idx / N);
    const int n = (int)(idx % N);
    out[idx] = xr[xr_offset_floats(row, n, N)];
  }
}
__global__ void rows_to_xr_kernel(const float* __restrict__ src, long long rows, int N, long long rows_pad,
                                  float* __restrict__ xr) {
  const size_t total = (size_t)rows_pad * N;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const long long row = (long long)(idx / N);
    const int n = (int)(idx % N);
    xr[xr_offset_floats(row, n, N)] = row < rows ? src[(size_t)row * N + n] : 0.f;
  }
}

static inline int grid_for(size_t total, int block = 256) {
  size_t b = (total + block - 1) / block;
  if (b > 148 * 16) b = 148 * 16;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace dnr

// nn.Linear weight (fp32 row-major [out x in]) -> packed bf16 B operand; transpose=1 packs W^T for dgrad.
DNR_EXPORT int dnr_pack_weight(const float* W, int rows_w, int cols_w, int ld, int transpose, int N_pad, int K_pad,
                               void* out_pk256, cudaStream_t stream) {
  DNR_REQUIRE(N_pad % 256 == 0 && K_pad % 8 == 0, "dnr_pack_weight: N_pad %% 256 and K_pad %% 8 must be 0");
  dnr::pack_weight_kernel<<<dnr::grid_for((size_t)N_pad * K_pad), 256, 0, stream>>>(W, rows_w, cols_w, ld, transpose,
                                                                                   N_pad, K_pad, (__nv_bfloat16*)out_pk256);
  DNR_CHECK_LAUNCH("dnr_pack_weight");
  return 0;
}

// lin_out + sigmoid/relu (resnetfc.py:158, pixelnerf.py:137-145).
DNR_EXPORT int dnr_head_fwd(const void* xa_pk, const void* xa_lo_pk, const float* Wout, const float* bout,
                            long long n_samples, int samples_per_tile, float* rgbsigma, cudaStream_t stream) {
  DNR_REQUIRE(samples_per_tile >= 1 && samples_per_tile <= 128, "dnr_head_fwd: samples_per_tile must be in [1,128]");
  if (n_samples <= 0) return 0;
  const int tiles = (int)((n_samples + samples_per_tile - 1) / samples_per_tile);
  dnr::head_fwd_kernel<<<tiles, 128, 0, stream>>>((const unsigned char*)xa_pk, (const unsigned char*)xa_lo_pk, Wout, bout,
                                                  n_samples, samples_per_tile,
                                                  (float4*)rgbsigma);
  DNR_CHECK_LAUNCH("dnr_head_fwd");
  return 0;
}

DNR_EXPORT int dnr_head_bwd(const void* xa_pk, const float* Wout, const float* rgbsigma, const float* d_rgbsigma,
                            long long n_samples, float* dx_xr, void* dx_pk, void* draw_pk, cudaStream_t stream) {
  if (n_samples <= 0) return 0;
  const int tiles = (int)((n_samples + 127) / 128);
  dnr::head_bwd_kernel<<<tiles, 128, 0, stream>>>((const unsigned char*)xa_pk, Wout, (const float4*)rgbsigma,
                                                  (const float4*)d_rgbsigma, n_samples, dx_xr, (unsigned char*)dx_pk,
                                                  (unsigned char*)draw_pk);
  DNR_CHECK_LAUNCH("dnr_head_bwd");
  return 0;
}

DNR_EXPORT int dnr_mean_bwd(const float* dmean_xr, int NV, long long n_samples, int view_tiles, float* dx_xr,
                            void* dx_pk, cudaStream_t stream) {
  if (view_tiles <= 0) return 0;
  dnr::mean_bwd_kernel<<<view_tiles, 128, 0, stream>>>(dmean_xr, NV, n_samples, dx_xr, (unsigned char*)dx_pk);
  DNR_CHECK_LAUNCH("dnr_mean_bwd");
  return 0;
}

DNR_EXPORT int dnr_mean_bwd_pk(const void* dmean_pk, int NV, long long n_samples, int view_tiles, void* dx_pk,
                               cudaStream_t stream) {
  DNR_REQUIRE(NV >= 1 && NV <= 8, "dnr_mean_bwd_pk: NV must be in [1,8]");
  if (view_tiles <= 0) return 0;
  dnr::mean_bwd_pk_kernel<<<view_tiles, 128, 0, stream>>>((const unsigned char*)dmean_pk, NV, n_samples, (unsigned char*)dx_pk);
  DNR_CHECK_LAUNCH("dnr_mean_bwd_pk");
  return 0;
}

// Layout converters (used by tests and by the host wrapper for small tensors).
DNR_EXPORT int dnr_pack_rows(const float* src, long long rows, int F, long long rows_pad, void* out_pk128,
                             cudaStream_t stream) {
  DNR_REQUIRE(F % 8 == 0 && rows_pad % 128 == 0, "dnr_pack_rows: F %% 8, rows_pad %% 128");
  dnr::pack_rows_kernel<<<dnr::grid_for((size_t)rows_pad * F), 256, 0, stream>>>(src, rows, F, rows_pad,
                                                                                (__nv_bfloat16*)out_pk128);
  DNR_CHECK_LAUNCH("dnr_pack_rows");
  return 0;
}
DNR_EXPORT int dnr_unpack_rows(const void* pk128, long long rows, int F, float* out, cudaStream_t stream) {
  dnr::unpack_rows_kernel<<<dnr::grid_for((size_t)rows * F), 256, 0, stream>>>((const __nv_bfloat16*)pk128, rows, F, out);
  DNR_CHECK_LAUNCH("dnr_unpack_rows");
  return 0;
}
DNR_EXPORT int dnr_xr_to_rows(const float* xr, long long rows, int N, float* out, cudaStream_t stream) {
  dnr::xr_to_rows_kernel<<<dnr::grid_for((size_t)rows * N), 256, 0, stream>>>(xr, rows, N, out);
  DNR_CHECK_LAUNCH("dnr_xr_to_rows");
  return 0;
}
DNR_EXPORT int dnr_rows_to_xr(const float* src, long long rows, int N, long long rows_pad, float* xr,
                              cudaStream_t stream) {
  dnr::rows_to_xr_kernel<<<dnr::grid_for((size_t)rows_pad * N), 256, 0, stream>>>(src, rows, N, rows_pad, xr);
  DNR_CHECK_LAUNCH("dnr_rows_to_xr");
  return 0;
}
