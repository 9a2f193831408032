// One C-ABI call per rendered ray batch (evaluation): the sequence that diner_b200/renderer.py + field.py issue call by
// call -- depth-guided sampler, then per balanced chunk of rays {projection / gather, the ResnetFC on 4-CTA clusters with
// 2-CTA clusters on the SMs those cannot occupy (second stream), head}, then ONE compositing launch -- as a host routine of
// the library.  Replaces NeRFRendererDGS.forward (reference src/models/nerf_renderer.py:399-424 -> :286-365) for a host
// that is not Python, and takes ~0.3 ms of interpreter time out of every call of a host that is (the reference's callers
// hand over 4 096 rays per call, src/models/diner.py:57,86-92).
//
// Nothing here computes: every kernel is launched through the entry points of the other translation units, in the order
// and with the arguments of the Python sequencing, so the results are the same bits (tests/test_gpu_frame.py).
// Device memory comes from the caller (dnr_render_workspace_bytes); the library keeps, per device, one non-blocking side
// stream and a few events for the fork / join and for timing the two MLP kernels (no device memory).
#include "common.cuh"
#include "../../include/diner_b200.h"
#include <stdlib.h>

namespace dnr {

constexpr int kRenderDevices = 64;
constexpr int kPending = 128;     // timing records in flight (a host that runs many frames ahead of the GPU fills them)

// throughput of one 2-CTA cluster relative to one 4-CTA cluster by view count (measured on the BASELINE.json shapes,
// diner_b200/field.py PAIR_RATE); the split adapts from there
static const float kPairRate[5] = {0.37f, 0.37f, 0.34f, 0.32f, 0.30f};

struct SplitTiming {
  cudaEvent_t q0, q1, p0, p1;
  int tiles_q, tiles_p;
  bool used;
};

struct RenderState {
  cudaStream_t side;
  cudaEvent_t fork, join;
  bool ready;
  float rate[5];              // per view count
  SplitTiming pending[kPending];
  double quad_ms, pair_ms;    // cumulative device time of the timed 4-CTA / 2-CTA cluster launches (dnr_render_timings)
  long long n_timed;
};
static RenderState g_render[kRenderDevices];      // zero-initialised

static RenderState* render_state() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kRenderDevices) return nullptr;
  RenderState* s = &g_render[dev];
  if (!s->ready) {
    if (cudaStreamCreateWithFlags(&s->side, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&s->fork, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&s->join, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    for (int i = 0; i < kPending; ++i) {
      SplitTiming& t = s->pending[i];
      if (cudaEventCreate(&t.q0) != cudaSuccess || cudaEventCreate(&t.q1) != cudaSuccess ||
          cudaEventCreate(&t.p0) != cudaSuccess || cudaEventCreate(&t.p1) != cudaSuccess) return nullptr;
      t.used = false;
    }
    for (int nv = 0; nv < 5; ++nv) s->rate[nv] = kPairRate[nv];
    s->ready = true;
  }
  return s;
}

// fold completed (quad, pair) timings into the running totals and, with `adapt`, into the pair rate
// (field.py::_update_split)
static void update_split(RenderState* s, int NV, int spare, int n_quad, bool adapt) {
  for (int i = 0; i < kPending; ++i) {
    SplitTiming& t = s->pending[i];
    if (!t.used) continue;
    if (cudaEventQuery(t.q1) != cudaSuccess || cudaEventQuery(t.p1) != cudaSuccess) { (void)cudaGetLastError(); continue; }
    float tq = 0.f, tp = 0.f;
    if (cudaEventElapsedTime(&tq, t.q0, t.q1) == cudaSuccess && cudaEventElapsedTime(&tp, t.p0, t.p1) == cudaSuccess) {
      s->quad_ms += tq; s->pair_ms += tp; ++s->n_timed;
      if (adapt && tq > 0.05f && tp > 0.05f && spare > 0 && n_quad > 0) {
        float measured = ((float)t.tiles_p / tp / (float)spare) / ((float)t.tiles_q / tq / (float)n_quad);
        measured = fminf(fmaxf(measured, 0.15f), 0.6f);
        s->rate[NV] += (0.97f * measured - s->rate[NV]) / 3.0f;    // a third of the way, aim 3 % low
      }
    }
    (void)cudaGetLastError();
    t.used = false;
  }
}

static inline long long align_up(long long x) { return (x + 255) & ~255LL; }

struct Workspace {
  long long z, rgbsigma, weights, lat, code, part, scratch, total;
};

static Workspace plan(int NV, long long n_rays, int K, long long rows_per_chunk, bool need_z, bool need_w) {
  Workspace w;
  const long long per_chunk = rows_per_chunk / ((long long)K * NV) > 0 ? rows_per_chunk / ((long long)K * NV) : 1;
  const long long n_chunks = n_rays > 0 ? (n_rays + per_chunk - 1) / per_chunk : 1;
  const long long max_rays = n_rays > 0 ? (n_rays + n_chunks - 1) / n_chunks : 0;
  const long long mv = RowMap::make(NV).tiles_for(max_rays * K);
  long long off = 0;
  w.z = off; off += need_z ? align_up(n_rays * K * 4) : 0;
  w.rgbsigma = off; off += align_up(n_rays * K * 16);
  w.weights = off; off += need_w ? align_up(n_rays * K * 4) : 0;
  w.lat = off; off += align_up(mv * 128 * 512 * 2);
  w.code = off; off += align_up(mv * 128 * 64 * 2);
  w.part = off; off += align_up(mv * 4 * 128 * 16);
  w.scratch = off; off += align_up(dnr_mlp_quad_scratch_bytes());
  w.total = off;
  return w;
}

}  // namespace dnr

// Bytes of device workspace dnr_render_fwd needs for n_rays rays (all scenes), K samples per ray, NV views, and chunks of
// at most rows_per_chunk per-view MLP rows (<= 0: 2^22); the worst case over the optional outputs.
DNR_EXPORT long long dnr_render_workspace_bytes(int NV, long long n_rays, int K, long long rows_per_chunk) {
  if (NV < 1 || NV > 4 || n_rays < 0 || K < 1) return -1;
  if (rows_per_chunk <= 0) rows_per_chunk = 1LL << 22;
  return dnr::plan(NV, n_rays, K, rows_per_chunk, true, true).total;
}

DNR_EXPORT int dnr_render_fwd(const dnr_scene_desc* sc, const dnr_mlp_desc* mlp, const float* rays, long long NR, int K,
                              int C, int G, unsigned long long seed, const float* z_in, int white_bkgd,
                              long long rows_per_chunk, void* workspace, long long workspace_bytes, float* rgb,
                              float* depth, float* weights_out, float* z_out, int* info, dnr_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  DNR_REQUIRE(sc && mlp && rays && rgb && depth, "dnr_render_fwd: null argument");
  DNR_REQUIRE(sc->NV >= 1 && sc->NV <= 4, "dnr_render_fwd: NV must be in [1,4] (the cluster kernels)");
  DNR_REQUIRE(sc->SB >= 1 && NR >= 0 && K >= 1 && G >= 0 && G <= K && C >= 1, "dnr_render_fwd: bad sizes");
  if (rows_per_chunk <= 0) rows_per_chunk = 1LL << 22;
  const int NV = sc->NV;
  const long long n_rays = (long long)sc->SB * NR;
  if (info) for (int i = 0; i < 8; ++i) info[i] = 0;
  if (n_rays == 0) return 0;
  DNR_REQUIRE(n_rays * K < (1LL << 31), "dnr_render_fwd: too many samples for one call");
  const dnr::Workspace ws = dnr::plan(NV, n_rays, K, rows_per_chunk, z_out == nullptr, weights_out == nullptr);
  DNR_REQUIRE(workspace != nullptr && workspace_bytes >= ws.total, "dnr_render_fwd: workspace too small (dnr_render_workspace_bytes)");
  DNR_REQUIRE(((uintptr_t)workspace & 255) == 0, "dnr_render_fwd: workspace must be 256-byte aligned");
  dnr::DeviceInfo* di = dnr::device_info();
  dnr::RenderState* st = dnr::render_state();
  DNR_REQUIRE(di != nullptr && st != nullptr, "dnr_render_fwd: no current CUDA device / cannot create the side stream");
  unsigned char* base = (unsigned char*)workspace;
  float* z = z_out ? z_out : (float*)(base + ws.z);
  float* rgbsigma = (float*)(base + ws.rgbsigma);
  float* weights = weights_out ? weights_out : (float*)(base + ws.weights);
  unsigned char* lat = base + ws.lat;
  unsigned char* code = base + ws.code;
  float* part = (float*)(base + ws.part);
  void* scratch = base + ws.scratch;
  int launches = 0, rc = 0;

  // ---- samples: depth-guided sampler, or the caller's (sorted) depths ----
  if (z_in) {
    if (z != z_in)
      DNR_CHECK(cudaMemcpyAsync(z, z_in, (size_t)n_rays * K * 4, cudaMemcpyDeviceToDevice, stream), "dnr_render_fwd/z");
  } else {
    rc = dnr_sampler_fwd(rays, sc->poses, sc->focal, sc->cxy, sc->depth, sc->depth_std_padded, sc->normal, sc->lin_tab,
                         nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, seed, sc->SB, NV, (int)NR, sc->H, sc->W, C, K,
                         G, sc->pad, sc->depth_diff_max, sc->img_w, sc->img_h, sc->std_sx, sc->std_sy, 0, 0, z, nullptr,
                         nullptr, nullptr, nullptr, stream_);
    if (rc) return rc;
    ++launches;
  }

  // ---- field: balanced chunks of rays (renderer.py::balanced_chunks) ----
  const long long per_chunk = rows_per_chunk / ((long long)K * NV) > 0 ? rows_per_chunk / ((long long)K * NV) : 1;
  const long long n_chunks = (n_rays + per_chunk - 1) / per_chunk;
  const long long cbase = n_rays / n_chunks, cextra = n_rays % n_chunks;
  const int per_tile = 4 * (32 / NV);                       // samples per 128-row view tile
  const int n_quad = dnr_mlp_quad_max_clusters(NV);
  DNR_REQUIRE(n_quad > 0, "dnr_render_fwd: the 4-CTA-cluster kernel cannot run on this device");
  const int spare = (di->num_sms - 4 * n_quad) / 2 > 0 ? (di->num_sms - 4 * n_quad) / 2 : 0;
  const char* adapt_env = getenv("DNR_ADAPT_SPLIT");
  const bool adapt = !(adapt_env && adapt_env[0] == '0');
  long long r0 = 0;
  int last_q = 0, last_p = 0;
  for (long long c = 0; c < n_chunks; ++c) {
    const long long n = cbase + (c < cextra ? 1 : 0);
    const long long n_samples = n * K;
    const int mv = dnr::RowMap::make(NV).tiles_for(n_samples);
    rc = dnr_encode_fwd(rays + r0 * 8, z + r0 * K, sc->poses, sc->focal, sc->cxy, sc->depth, sc->latent_nhwc_bf16, sc->freqs,
                        sc->phases, sc->SB, NV, (int)NR, K, sc->H, sc->W, sc->Hl, sc->Wl, sc->L, sc->NF, r0, n, sc->img_w,
                        sc->img_h, sc->lat_sx, sc->lat_sy, lat, code, stream_);
    if (rc) return rc;
    ++launches;
    dnr::update_split(st, NV, spare, n_quad, adapt);
    if (!adapt) st->rate[NV] = dnr::kPairRate[NV];
    const float r = st->rate[NV];
    const float share = spare > 0 ? r * spare / (r * spare + n_quad) : 0.f;
    int mv_q = mv;
    if (share > 0.f && spare > 0 && mv >= 64 * spare) {
      int side_tiles = (int)((double)mv * (double)share);
      if (side_tiles < 2) side_tiles = 2;
      side_tiles = side_tiles / 2 * 2;
      mv_q = mv - side_tiles > 0 ? mv - side_tiles : 0;
    }
    const int n_p = mv - mv_q;
    float* out = rgbsigma + r0 * K * 4;
    dnr::SplitTiming* tm = nullptr;
    if (n_p > 0 && mv_q > 0) {                                // fork: the pair clusters run beside the quad clusters
      DNR_CHECK(cudaEventRecord(st->fork, stream), "dnr_render_fwd/fork");
      DNR_CHECK(cudaStreamWaitEvent(st->side, st->fork, 0), "dnr_render_fwd/fork");
      for (int i = 0; i < dnr::kPending; ++i)                   // (a free slot: the two kernels of this chunk are timed)
        if (!st->pending[i].used) { tm = &st->pending[i]; break; }
    }
    if (mv_q > 0) {
      if (tm) DNR_CHECK(cudaEventRecord(tm->q0, stream), "dnr_render_fwd/timing");
      rc = dnr_mlp_quad_fwd(lat, code, mlp->w128, mlp->pair_biases, mlp->w_out, NV, mv_q, NV > 1 ? scratch : nullptr, part,
                            nullptr, stream_);
      if (rc) return rc;
      if (tm) DNR_CHECK(cudaEventRecord(tm->q1, stream), "dnr_render_fwd/timing");
      const long long nq = n_samples < (long long)mv_q * per_tile ? n_samples : (long long)mv_q * per_tile;
      rc = dnr_head_finalize(part, mlp->b_out, NV, nq, 4, out, stream_);
      if (rc) return rc;
      launches += 2;
    }
    if (n_p > 0) {
      const long long s0 = (long long)mv_q * per_tile;         // first sample of the pair clusters' tiles
      cudaStream_t ps = mv_q > 0 ? st->side : stream;
      float* part_p = part + (size_t)mv_q * 4 * 128 * 4;
      if (tm) DNR_CHECK(cudaEventRecord(tm->p0, ps), "dnr_render_fwd/timing");
      rc = dnr_mlp_pair_fwd(lat + (size_t)mv_q * 128 * 512 * 2, code + (size_t)mv_q * 128 * 64 * 2, mlp->w256,
                            mlp->pair_biases, mlp->w_out, NV, n_p, mv_q > 0 ? spare : 0, part_p, nullptr, (dnr_stream_t)ps);
      if (rc) return rc;
      if (tm) {
        DNR_CHECK(cudaEventRecord(tm->p1, ps), "dnr_render_fwd/timing");
        tm->tiles_q = mv_q; tm->tiles_p = n_p; tm->used = true;
      }
      rc = dnr_head_finalize(part_p, mlp->b_out, NV, n_samples - s0, 2, out + s0 * 4, (dnr_stream_t)ps);
      if (rc) return rc;
      launches += 2;
      if (mv_q > 0) {                                          // join
        DNR_CHECK(cudaEventRecord(st->join, st->side), "dnr_render_fwd/join");
        DNR_CHECK(cudaStreamWaitEvent(stream, st->join, 0), "dnr_render_fwd/join");
      }
    }
    last_q = mv_q > 0 ? mv_q : mv; last_p = mv_q > 0 ? n_p : 0;
    r0 += n;
  }

  // ---- one compositing launch per call (nerf_renderer.py:344-365) ----
  rc = dnr_composite_fwd(rgbsigma, z, rays, (int)n_rays, K, white_bkgd, weights, rgb, depth, stream_);
  if (rc) return rc;
  ++launches;
  if (info) { info[0] = launches; info[1] = (int)n_chunks; info[2] = last_q; info[3] = last_p; info[4] = last_q + last_p; }
  return 0;
}

// Cumulative device time (ms, CUDA events on the launching streams) of the 4-CTA / 2-CTA cluster MLP launches that
// dnr_render_fwd has issued and timed on the current device since the library was loaded, and their number (pairs of
// launches).  Records still in flight are not included: synchronise first.  For benchmarks: the kernel's duration inside
// a timed region without a second, instrumented pass.
DNR_EXPORT int dnr_render_timings(double* quad_ms, double* pair_ms, long long* n_timed) {
  dnr::RenderState* st = dnr::render_state();
  DNR_REQUIRE(st != nullptr, "dnr_render_timings: no current CUDA device");
  dnr::update_split(st, 0, 0, 0, false);
  if (quad_ms) *quad_ms = st->quad_ms;
  if (pair_ms) *pair_ms = st->pair_ms;
  if (n_timed) *n_timed = st->n_timed;
  return 0;
}
