// On-disk depth format of the reference's datasets -> the renderer's depth / depth-std maps.
//
// The datasets store depth as uint16 PNG planes in units of 1e-4 m.  FaceScape keeps one "triptych" image
// [gt | TransMVSNet prediction | confidence] per view plus a mesh-rendered depth image, and picks or merges them by
// `depth_type` (src/data/facescape.py:69-106); the confidence becomes a standard deviation through a per-dataset
// linear law (facescape.py:54-56,266; dtu.py:68-70,223; multiface.py:306-311).  DTU stores depth and confidence as
// separate images and rescales both (dtu.py:104-107,119).  PNG inflation stays with the host image library; this
// kernel consumes the decoded uint16 planes and produces both fp32 maps in one pass (HBM-bound: 4-6 B read, 8 B
// written per pixel), with the reference's fp32 operation order (separate multiply / add roundings, no FMA).
#include "common.cuh"

namespace dnr {

struct DepthDecodeParams {
  const uint16_t* pred;     // MVS depth plane (nullable)
  const uint16_t* conf;     // MVS confidence plane (nullable)
  const uint16_t* mesh;     // mesh depth plane, dense [N,H,W] (nullable)
  long long img_stride;     // elements between images of pred / conf
  int row_stride;           // elements between rows of pred / conf
  int N, H, W;
  int depth_type;           // 0 original (MVS), 1 mesh, 2 merge
  float scale;              // 1e-4
  float div0, mul1;         // depth / div0 * mul1, each one fp32 rounding (DTU: 0.7/872, scale_factor); 1 = none
  float std_a, std_b;       // std = a * conf + b
  float std_const;          // used with kStdConst
  int flags;
  float* depth;             // [N,H,W]
  float* std;               // [N,H,W]
};

enum { kStdClip0 = 1, kStdZeroNoDepth = 2, kConfScaled = 4, kStdConst = 8 };

__global__ void __launch_bounds__(256) depth_decode_kernel(const DepthDecodeParams P) {
  const long long per = (long long)P.H * P.W;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= per * P.N) return;
  const int n = (int)(idx / per);
  const int rem = (int)(idx % per);
  const int y = rem / P.W, x = rem % P.W;
  const size_t src = (size_t)n * P.img_stride + (size_t)y * P.row_stride + x;
  float d_mvs = 0.f, c_mvs = 0.f;
  if (P.pred) d_mvs = __fmul_rn((float)P.pred[src], P.scale);                  // facescape.py:88
  if (P.conf) c_mvs = __fmul_rn((float)P.conf[src], P.scale);                  // facescape.py:89
  float d = d_mvs, c = c_mvs;
  if (P.depth_type != 0) {
    const float d_mesh = __fmul_rn((float)P.mesh[idx], P.scale);               // facescape.py:76
    const float c_mesh = d_mesh == 0.f ? 0.f : 0.8f;                           // facescape.py:77-79
    d = d_mesh; c = c_mesh;
    if (P.depth_type == 2) {                                                   // merge, facescape.py:96-101
      if (d_mesh == 0.f && d_mvs != 0.f) d = d_mvs;
      if (c_mesh == 0.f && c_mvs != 0.f) c = c_mvs;
    }
  }
  d = __fmul_rn(__fdiv_rn(d, P.div0), P.mul1);                                  // dtu.py:107,119 (x / 1, x * 1 exact)
  if (P.flags & kConfScaled) c = __fmul_rn(__fdiv_rn(c, P.div0), P.mul1);       // dtu.py:221 reads conf like depth
  float s = (P.flags & kStdConst) ? P.std_const : __fadd_rn(__fmul_rn(P.std_a, c), P.std_b);
  if (P.flags & kStdClip0) s = fmaxf(s, 0.f);                                   // multiface.py:310
  if ((P.flags & kStdZeroNoDepth) && d == 0.f) s = 0.f;                         // multiface.py:311
  P.depth[idx] = d;
  P.std[idx] = s;
}

}  // namespace dnr

// Decoded uint16 depth planes -> fp32 depth and depth-std maps.  pred / conf are planes inside a larger image
// (FaceScape triptych: pass the pointer to the plane's first pixel, row_stride = 3 W) or dense images; mesh is dense.
// depth_type: 0 = MVS planes, 1 = mesh (confidence 0.8 where there is depth), 2 = mesh with holes filled from MVS.
// flags: 1 clip std at 0, 2 std = 0 where depth == 0, 4 confidence is rescaled like depth, 8 constant std.
// div0 must not be 0.
// Replaces FacescapeDataSet.read_depth + conf2std (src/data/facescape.py:54-56,69-106,266), the PNG branch of
// DTUDataSet.read_depth + conf2std (src/data/dtu.py:68-70,104-107,119,221-223) and multiface.py:103-109,305-311.
DNR_EXPORT int dnr_decode_depth_u16(const uint16_t* pred, const uint16_t* conf, long long img_stride, int row_stride,
                                    const uint16_t* mesh, int N, int H, int W, int depth_type, float scale, float div0,
                                    float mul1, float std_a, float std_b, float std_const, int flags, float* depth_out,
                                    float* std_out, cudaStream_t stream) {
  DNR_REQUIRE(N >= 0 && H >= 1 && W >= 1, "dnr_decode_depth_u16: bad sizes");
  DNR_REQUIRE(depth_type >= 0 && depth_type <= 2, "dnr_decode_depth_u16: depth_type must be 0 (original), 1 (mesh) or 2 (merge)");
  DNR_REQUIRE(flags >= 0 && flags < 16, "dnr_decode_depth_u16: unknown flag");
  const long long n = (long long)N * H * W;
  if (n == 0) return 0;
  DNR_REQUIRE(depth_out && std_out, "dnr_decode_depth_u16: null output");
  DNR_REQUIRE(depth_type == 0 || mesh, "dnr_decode_depth_u16: depth_type mesh / merge needs the mesh plane");
  DNR_REQUIRE(depth_type == 1 || pred, "dnr_decode_depth_u16: depth_type original / merge needs the MVS depth plane");
  DNR_REQUIRE(row_stride >= W || !(pred || conf), "dnr_decode_depth_u16: row_stride smaller than W");
  dnr::DepthDecodeParams P;
  P.pred = depth_type == 1 ? nullptr : pred; P.conf = depth_type == 1 ? nullptr : conf; P.mesh = mesh;
  P.img_stride = img_stride; P.row_stride = row_stride; P.N = N; P.H = H; P.W = W; P.depth_type = depth_type;
  P.scale = scale; P.div0 = div0; P.mul1 = mul1; P.std_a = std_a; P.std_b = std_b; P.std_const = std_const;
  P.flags = flags; P.depth = depth_out; P.std = std_out;
  dnr::depth_decode_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(P);
  DNR_CHECK_LAUNCH("dnr_decode_depth_u16");
  return 0;
}
