/* diner_b200 -- C ABI of the B200 (sm_100a) implementation of DINER's per-ray rendering hot path.
 *
 * This header is the drop-in boundary below the Python plugin class `diner_b200.NeRFRendererDGS`
 * (which itself replaces the reference's `src.models.nerf_renderer.NeRFRendererDGS`, selected by the
 * `renderer.module` string of the reference's YAML configs, reference src/models/diner.py:48).
 * The reference is pure Python/PyTorch, so there is no reference FFI to copy; every entry point
 * below cites the reference Python function (file:line, relative to the reference repo) it replaces.
 *
 * Conventions
 *   - plain C: raw DEVICE pointers, explicit sizes, a cudaStream_t passed as `void*`; no torch types
 *   - every function returns 0 on success, otherwise a cudaError_t value (cudaErrorInvalidValue for
 *     argument violations); `dnr_last_error()` returns a thread-local description
 *   - all work is enqueued on `stream`; nothing synchronises, nothing allocates device memory
 *   - there is no CPU implementation behind this ABI
 *
 * Packed operand layouts (see DESIGN.md "Data layout in HBM")
 *   PK128 : bf16 matrix [rows x F] stored as [rows/128][F/8][128][8]  (tcgen05 canonical no-swizzle image)
 *   PK256 : bf16 weights [N x K] stored as [N/256][K/8][256][8]
 *   XR    : fp32 matrix [rows x N] stored as [rows/128][N/4][128] float4
 *   view rows: the per-view MLP batch is ordered sample-major / view-minor in 32-row slabs holding
 *   floor(32/NV) samples each (`dnr_view_row_tiles` gives the number of 128-row tiles).
 */
#ifndef DINER_B200_H
#define DINER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* dnr_stream_t; /* cudaStream_t */

/* ---- library ------------------------------------------------------------------------------------------- */
int dnr_abi_version(void);
const char* dnr_last_error(void);

/* ---- depth-guided sampler ------------------------------------------------------------------------------- */
/* Exponentially padded depth-std map, built once per scene.  Replaces the per-call rebuild inside
 * SpatialEncoder.index_depth_std (src/models/image_encoder.py:153-180) -> exponential_padding
 * (src/util/torch_helpers.py:100-122).  gain_tab[e] = exp(e/12*ln2), e in [0,pad).
 * depth_std [n_maps,H,W] -> out [n_maps,H+2*pad,W+2*pad]. */
int dnr_pad_depth_std(const float* depth_std, const float* gain_tab, int n_maps, int H, int W, int pad, float* out,
                      dnr_stream_t stream);

/* Replaces NeRFRendererDGS.sample_coarse + sample_depthguided + fill_up_uniform_samples
 * (src/models/nerf_renderer.py:39-63, 65-284, 367-397) incl. the three nearest taps
 * (image_encoder.py:129-204) and weighted_mean_n_std (torch_helpers.py:294-302).
 *   rays [SB*NR,8] = [o3,d3,near,far]; poses [SB,NV,4,4] world->cam; focal,cxy [SB,NV,2];
 *   depth [SB,NV,H,W]; depth_std_padded [SB,NV,H+2pad,W+2pad]; normal [SB,NV,3,H,W];
 *   lin_tab [C] = linspace(0,1-1/C,C).
 *   Jitter: u_coarse [SB*NR,C] | NULL (-> Philox(seed)); n_gauss / u_fill: recorded draws of the reference,
 *   either packed in the reference's order (gauss_off/fill_off [SB*NR] = exclusive prefix sums of ray_mask /
 *   n_missing, obtained from a count_only=1 pass) or, with unpacked_noise=1, [SB*NR,G] / [SB*NR,K].
 *   z_out [SB*NR,K] ascending.  Debug outputs (nullable): ray_mask_out, n_missing_out [SB*NR],
 *   sel_idx_out [SB*NR,K-G] (candidate index per shortlist rank, -1 = empty), p_out [SB*NR,C] likelihood.
 *   cand_offset [SB*NR,C,3] | NULL: NOVEL variant (src/models/novel/nerf_novel_renderer.py:104-107) -- every candidate
 *   position is moved by the offset of its nearest target-mesh vertex (from dnr_novel_deform) before it is projected. */
int dnr_sampler_fwd(const float* rays, const float* poses, const float* focal, const float* cxy, const float* depth,
                    const float* depth_std_padded, const float* normal, const float* lin_tab, const float* u_coarse,
                    const float* n_gauss, const float* u_fill, const int* gauss_off, const int* fill_off,
                    const float* cand_offset, unsigned long long seed, int SB, int NV, int NR, int H, int W, int C, int K, int G, int pad,
                    float depth_diff_max, float img_w, float img_h, float std_sx, float std_sy, int count_only,
                    int unpacked_noise, float* z_out, int* ray_mask_out, int* n_missing_out, int* sel_idx_out,
                    float* p_out, dnr_stream_t stream);

/* ---- projection + gather + encoding --------------------------------------------------------------------- */
/* Per-scene repack of encoder.latent (SB*NV,C,Hl,Wl) fp32 NCHW -> (SB*NV,Hl,Wl,C) bf16 channels-last, and the
 * inverse for the feature gradient. */
int dnr_pack_latent_nhwc(const float* latent_nchw, int n_maps, int C, int Hl, int Wl, void* out_bf16,
                         dnr_stream_t stream);
int dnr_unpack_dlatent_nchw(const float* d_latent_nhwc, int n_maps, int C, int Hl, int Wl, float* out_nchw,
                            dnr_stream_t stream);
/* Number of 128-row tiles of the per-view MLP batch (-1 on bad NV). */
int dnr_view_row_tiles(long long n_samples, int NV);

/* Replaces PixelNeRF.forward's projection / gather / encoding (src/models/pixelnerf.py:91-128) with
 * SpatialEncoder.index + index_depth (image_encoder.py:97-151) and PositionalEncoding.forward
 * (positional_encoding.py:33-53).  For the chunk of n_rays rays starting at ray_offset of the (SB*NR) batch
 * (rays/z point at the chunk's first row) it writes the MLP A operands lat_pk PK128[rows x 512] and
 * in_pk PK128[rows x 64] (55 valid columns). */
int dnr_encode_fwd(const float* rays, const float* z, const float* poses, const float* focal, const float* cxy,
                   const float* depth, const void* latent_nhwc_bf16, const float* freqs, const float* phases, int SB,
                   int NV, int NR, int K, int H, int W, int Hl, int Wl, int L, int NF, long long ray_offset,
                   long long n_rays, float img_w, float img_h, float lat_sx, float lat_sy, void* lat_pk, void* in_pk,
                   dnr_stream_t stream);
/* Verification-mode variant (NeRFRendererDGS(precision="split"), DESIGN.md "Precision contract"): the latent is a
 * bf16 hi + lo pair (x ~ hi + lo to 2^-17 relative) and both operands leave as hi + lo pairs of PK128 images;
 * otherwise identical to dnr_encode_fwd. */
int dnr_encode_fwd_split(const float* rays, const float* z, const float* poses, const float* focal, const float* cxy,
                         const float* depth, const void* latent_hi, const void* latent_lo, const float* freqs,
                         const float* phases, int SB, int NV, int NR, int K, int H, int W, int Hl, int Wl, int L, int NF,
                         long long ray_offset, long long n_rays, float img_w, float img_h, float lat_sx, float lat_sy,
                         void* lat_pk_hi, void* in_pk_hi, void* lat_pk_lo, void* in_pk_lo, dnr_stream_t stream);
/* NOVEL variant of dnr_encode_fwd (src/models/novel/novel_pixelnerf.py:143-241): pts_in / pts_gen [n_rays*K,3] are the
 * chunk's sample points deformed with tgt_in_offsets / tgt_gen_offsets (dnr_novel_deform, mode samples); pts_in replaces
 * o + z d in the projection chain, pts_gen is projected by the ONE generic camera (gen_poses [SB,1,4,4], gen_focal /
 * gen_cxy [SB,1,2], gen_img_w/h; :73-80) into the learned map gen_latent [Hg,Wg,512] bf16 channels-last (:46,108-141) and the
 * feature row is gen_latent[gen_uv] + latent[uv] (:196).  Backward: dnr_encode_novel_bwd. */
int dnr_encode_novel_fwd(const float* rays, const float* z, const float* pts_in, const float* pts_gen, const float* poses,
                         const float* focal, const float* cxy, const float* depth, const void* latent_nhwc_bf16,
                         const float* gen_poses, const float* gen_focal, const float* gen_cxy,
                         const void* gen_latent_nhwc_bf16, int Hg, int Wg, float gen_img_w, float gen_img_h, float gen_sx,
                         float gen_sy, const float* freqs, const float* phases, int SB, int NV, int NR, int K, int H, int W,
                         int Hl, int Wl, int L, int NF, long long ray_offset, long long n_rays, float img_w, float img_h,
                         float lat_sx, float lat_sy, void* lat_pk, void* in_pk, dnr_stream_t stream);
/* Backward of dnr_encode_novel_fwd w.r.t. the two feature maps (novel_pixelnerf.py:196 under autograd; the NOVEL training
 * step optimises the field's parameters only, novel.py:280-285,361-363, so the deformed points carry no gradient):
 * d_lat_xr [rows_pad x 512] fp32 XR is scattered into d_latent_nhwc [SB,NV,Hl,Wl,512] fp32 at the footprint of pts_in in
 * each source view and into d_gen_latent_nhwc [Hg,Wg,512] fp32 at the footprint of pts_gen under the generic camera (one
 * learned map for all scenes and views, :110).  Both are accumulated into; either may be null.  d_gen_xr: gradient rows
 * of the gen_latent taps when they are not those of the encoder feature (NOVEL_PE), null = d_lat_xr. */
int dnr_encode_novel_bwd(const float* rays, const float* z, const float* pts_in, const float* pts_gen, const float* poses,
                         const float* focal, const float* cxy, const float* depth, const float* gen_poses,
                         const float* gen_focal, const float* gen_cxy, int Hg, int Wg, float gen_img_w, float gen_img_h,
                         float gen_sx, float gen_sy, const float* freqs, const float* phases, int SB, int NV, int NR, int K,
                         int H, int W, int Hl, int Wl, int L, int NF, long long ray_offset, long long n_rays, float img_w,
                         float img_h, float lat_sx, float lat_sy, const float* d_lat_xr, const float* d_gen_xr,
                         float* d_latent_nhwc, float* d_gen_latent_nhwc, dnr_stream_t stream);
/* NOVEL_PE variant of dnr_encode_novel_fwd (src/models/novel_pe/pe_novel_pixelnerf.py:200-275): lat_pk receives the
 * encoder feature alone and gen_pk [rows_pad x 512] bf16 PK128 the gen_latent taps (the variant's deformation layer, a
 * 518 -> 512 nn.Linear on [feature | 6 positional values], is applied before the two are added, :265-268: one
 * dnr_mlp_gemm with gen_pk as res_in_pk); columns 55..60 of in_pk carry [src_pos_encodings 3 | tgt_pos_encodings 3]:
 * pe_src [SB,NV,3,Hp,Wp] fp32 tapped bilinearly (border) at the row's source uv rescaled by pe_sx / pe_sy (:133-166),
 * pe_tgt [SB,1,3,Ht,Wt] fp32 at the projection of the UNDEFORMED point o + z d by the target camera (tgt_poses [SB,1,4,4],
 * tgt_focal / tgt_cxy [SB,1,2], tgt_img_w/h) rescaled by tgt_sx / tgt_sy (:168-198,223-226,248-251). */
int dnr_encode_novel_pe_fwd(const float* rays, const float* z, const float* pts_in, const float* pts_gen,
                            const float* poses, const float* focal, const float* cxy, const float* depth,
                            const void* latent_nhwc_bf16, const float* gen_poses, const float* gen_focal,
                            const float* gen_cxy, const void* gen_latent_nhwc_bf16, int Hg, int Wg, float gen_img_w,
                            float gen_img_h, float gen_sx, float gen_sy, const float* pe_src, int Hp, int Wp, float pe_sx,
                            float pe_sy, const float* tgt_poses, const float* tgt_focal, const float* tgt_cxy,
                            float tgt_img_w, float tgt_img_h, const float* pe_tgt, int Ht, int Wt, float tgt_sx,
                            float tgt_sy, const float* freqs, const float* phases, int SB, int NV, int NR, int K, int H,
                            int W, int Hl, int Wl, int L, int NF, long long ray_offset, long long n_rays, float img_w,
                            float img_h, float lat_sx, float lat_sy, void* lat_pk, void* gen_pk, void* in_pk,
                            dnr_stream_t stream);
/* Nearest-vertex deformation of the NOVEL variant at frame scale (src/models/novel/nerf_novel_renderer.py:40-50 applied at
 * :104-107 to every sampler candidate and at :441-442 to every sample point): K = 1 nearest neighbour (squared L2, first
 * minimum) against a target mesh given Morton-sorted in groups of 32 with one bounding sphere per group
 * (verts_sorted [SB][G*32] float4 = x, y, z, original index bits; spheres [SB][G] float4 = centre, radius; supers
 * [SB][ceil(G/8)][2] float4 = box lo / hi of 8 consecutive groups; subs [SB][G][4] float4 = spheres of the four runs of 8
 * vertices inside a group, nullable) -- the search is pruned by boxes and spheres and equals the brute force
 * (dnr_deform_points) bit for bit.  candidates = 1: per_ray = C, the
 * candidate depths are recomputed exactly as dnr_sampler_fwd draws them (lin_tab, u_coarse | Philox(seed)) and out_a
 * receives the OFFSETS [SB*NR,C,3] for dnr_sampler_fwd's cand_offset.  candidates = 0: per_ray = K, z given, out_a / out_b
 * (nullable) receive the points o + z d moved by offsets_a / offsets_b ([SB][V][3], indexed by original vertex). */
int dnr_novel_deform(const float* rays, const float* z, const float* lin_tab, const float* u_coarse,
                     unsigned long long seed, const void* verts_sorted, const void* spheres, const void* supers,
                     const void* subs, const float* offsets_a, const float* offsets_b, int SB, int NR, int per_ray, int V,
                     int G, int candidates, float* out_a, float* out_b, int* idx_out, dnr_stream_t stream);
/* Backward of SpatialEncoder.index w.r.t. the latent map (grid_sample bilinear backward,
 * image_encoder.py:117-123): d_latent_nhwc [SB,NV,Hl,Wl,512] fp32 += scatter(d_lat_xr XR[rows x 512]). */
int dnr_encode_bwd(const float* rays, const float* z, const float* poses, const float* focal, const float* cxy,
                   const float* depth, const float* freqs, const float* phases, int SB, int NV, int NR, int K, int H,
                   int W, int Hl, int Wl, int L, int NF, long long ray_offset, long long n_rays, float img_w,
                   float img_h, float lat_sx, float lat_sy, const float* d_lat_xr, float* d_latent_nhwc,
                   dnr_stream_t stream);

/* ---- field MLP (ResnetFC) on tcgen05 -------------------------------------------------------------------- */
/* nn.Linear weight (fp32 row-major [rows_w x cols_w], leading dimension ld) -> PK256 bf16 [N_pad x K_pad];
 * transpose=1 packs W^T (dgrad operand).  Zero outside the source. */
int dnr_pack_weight(const float* W, int rows_w, int cols_w, int ld, int transpose, int N_pad, int K_pad,
                    void* out_pk256, dnr_stream_t stream);

/* One fused linear layer of ResnetFC (src/models/resnetfc.py:61-69,129-159), forward or dgrad:
 *   D[rows x N] = sum_{seg<3} A_seg[rows x K_seg] * B_seg[N x K_seg]^T   (bf16 in, fp32 accumulate in TMEM)
 * epilogue: + bias[N]; * [mask_pk > 0] (ReLU backward); + res_in (XR fp32) or + res_in_pk (PK128 bf16: the residual
 * GRADIENT stream of the backward pass carried in bf16); optional mean over the NV view rows of
 * each sample (combine(), resnetfc.py:9-14; outputs then indexed by sample); outputs res_out (XR fp32) and/or
 * act_out (PK128 bf16, relu_act=1 applies ReLU; act_lo_out, nullable, receives the bf16 image of the rounding residue
 * v - bf16(v) for the verification mode).  A: PK128, B: PK256, K multiple of 64, N multiple of 256.
 * max_ctas > 0 caps the persistent grid.  mean_nv < 0 is the backward of that mean: res_in is a POOLED stream indexed by
 * sample, and row (sample, view) receives res_in[sample] / |mean_nv| (n_samples bounds the valid rows). */
int dnr_mlp_gemm(const void* A0, const void* B0, int K0, const void* A1, const void* B1, int K1, const void* A2,
                 const void* B2, int K2, int m_tiles, int N, const float* bias, const void* mask_pk,
                 const float* res_in, const void* res_in_pk, float* res_out, void* act_out, void* act_lo_out, int relu_act,
                 int mean_nv, long long n_samples, int max_ctas, dnr_stream_t stream);

/* Whole ResnetFC forward in ONE persistent kernel (inference): all 13 linear layers of resnetfc.py:129-159 with
 * the view mean (combine(), :9-14) fused, activations kept in an L2-resident per-CTA scratch instead of HBM.
 * Inputs are the operands written by dnr_encode_fwd (mv_tiles view-row tiles).  weights = 14 PK256 pointers
 * [lin_in, lin_z0..2, fc_0 of blocks 0..4, fc_1 of blocks 0..4]; biases = 11 fp32[512] pointers
 * [lin_in+lin_z0, fc_0 x5, fc_1 (+lin_z[b+1]) x5] (host arrays of device pointers).  xa_out: PK128
 * [ceil(mv_tiles/NV)*128 x 512] = relu(x) after the last block, pooled tile g holding samples
 * [g*S, (g+1)*S), S = 4*NV*floor(32/NV) -> dnr_head_fwd(..., samples_per_tile = S). */
long long dnr_mlp_fused_scratch_bytes(int max_ctas);
int dnr_mlp_fused_fwd(const void* lat_pk, const void* code_pk, const void* const* weights, const float* const* biases,
                      int NV, int mv_tiles, long long n_samples, void* scratch, void* xa_out, int max_ctas,
                      dnr_stream_t stream);

/* The same forward on CLUSTERS OF TWO CTAs with the fp32 residual stream resident in TMEM (csrc/mlp_pair.cu): CTA r
 * of a pair owns output columns [256r, 256r+256) of every layer and the two exchange only the bf16 A operand
 * (64 KB per layer) by bulk copies over distributed shared memory, so nothing of the layer chain touches HBM or
 * L2 except the weights.  lin_out is fused as
 * partial sums.  weights as for dnr_mlp_fused_fwd; biases = 11 fp32[512] pointers [cb0..cb5, fc_0 bias x5] where
 * cb0 = lin_in+lin_z0, cb1..5 = cb0 plus the cumulative fc_1 (+lin_z[b+1]) biases of blocks 0..4;
 * head_partial: [mv_tiles*2*128] float4, consumed by dnr_head_finalize (sigmoid / relu, pixelnerf.py:137-145). */
int dnr_mlp_pair_fwd(const void* lat_pk, const void* code_pk, const void* const* weights, const float* const* biases,
                     const float* w_out, int NV, int mv_tiles, int max_clusters /* 0 = all SM pairs */,
                     float* head_partial, long long* prof /* nullable: [grid][16] cycle counters */,
                     dnr_stream_t stream);
/* The same on clusters of FOUR CTAs (2 row tiles x 2 column halves) with tcgen05.mma.cta_group::2: the two CTAs of a
 * column half share every weight K-block, each streaming half of it, which halves the L2 -> SM traffic that bounds the
 * pair kernel.  weights_pk128: the same 14 matrices in the PK128 layout [N/128][K/8][128][8] bf16;
 * head_partial: [mv_tiles*4*128] float4.
 * merge_scratch: null, or dnr_mlp_quad_scratch_bytes() bytes of device memory: for NV > 1 every CTA then parks the
 * pooled rows (128/NV distinct rows per tile after the view mean) of NV consecutive tiles, merges them into one full tile
 * and runs blocks 3-4 and lin_out once per NV tiles instead of on NV-fold duplicated rows. */
long long dnr_mlp_quad_scratch_bytes(void);
int dnr_mlp_quad_fwd(const void* lat_pk, const void* code_pk, const void* const* weights_pk128,
                     const float* const* biases, const float* w_out, int NV, int mv_tiles, void* merge_scratch,
                     float* head_partial, long long* prof /* nullable */, dnr_stream_t stream);
/* NOVEL_PE field (src/models/novel_pe/pe_novel_pixelnerf.py:263-275: ResnetFC with d_latent = 518): the cluster kernels
 * with a second K-step on the 64-wide code operand in fc_1 of blocks 0 and 1 -- columns 55..60 of that operand carry the six
 * positional values (dnr_encode_novel_pe_fwd), code_weights[b] is the [512 x 64] image of [0 (55) | lin_z[b+1].weight[:,
 * 512:518] | 0 (3)] in the kernel's weight layout (PK256 for pair, PK128 for quad), weights[0] merges lin_in with
 * lin_z[0]'s positional columns, and lat_pk is the variant's final_latent.  Otherwise as dnr_mlp_pair_fwd / _quad_fwd. */
int dnr_mlp_pair_fwd_pe(const void* lat_pk, const void* code_pk, const void* const* weights,
                        const void* const* code_weights, const float* const* biases, const float* w_out, int NV,
                        int mv_tiles, int max_clusters, float* head_partial, dnr_stream_t stream);
int dnr_mlp_quad_fwd_pe(const void* lat_pk, const void* code_pk, const void* const* weights_pk128,
                        const void* const* code_weights_pk128, const float* const* biases, const float* w_out, int NV,
                        int mv_tiles, void* merge_scratch, float* head_partial, dnr_stream_t stream);
/* Allocation-free variant of dnr_mlp_quad_fwd.  It has NO COUNTERPART IN THE REFERENCE: it was added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid it).  The merge scratch lives in a file-scope device array of the library (4 MiB = 8 clusters x
 * 4 CTAs x 128 KB, held by every GPU that loads the library until exit); the call allocates nothing and the caller still
 * passes inputs and outputs.  mv_tiles must not exceed dnr_mlp_quad_static_max_tiles() (4096 tiles = 524 288 view rows),
 * else cudaErrorInvalidValue with a message in dnr_last_error().  All calls on a device share the array, which keeps the
 * last call's contents (nothing in it needs to start at zero): calls that overlap on the GPU race -- serialise them.
 * Results are bit-identical to dnr_mlp_quad_fwd with a merge scratch. */
int dnr_mlp_quad_static_max_tiles(void);
int dnr_mlp_quad_fwd_static(const void* lat_pk, const void* code_pk, const void* const* weights_pk128,
                            const float* const* biases, const float* w_out, int NV, int mv_tiles, float* head_partial,
                            dnr_stream_t stream);
/* Number of 4-CTA clusters the quad kernel keeps resident; SMs beyond 4x that stay idle and can run
 * dnr_mlp_pair_fwd clusters (max_clusters) on another stream at the same time. */
int dnr_mlp_quad_max_clusters(int NV);
/* n_partials = 2 after dnr_mlp_pair_fwd, 4 after dnr_mlp_quad_fwd. */
int dnr_head_finalize(const float* head_partial, const float* bout, int NV, long long n_samples, int n_partials,
                      float* rgbsigma, dnr_stream_t stream);

/* Weight / bias gradient of one nn.Linear (autograd of F.linear in resnetfc.py):
 *   dW[n,k] += sum_rows dY[row,n] * X[row,k],  db[n] += sum_rows dY[row,n]
 * dY PK128[rows x Fy], X PK128[rows x Fx]; only n < n_valid, k < k_valid are written. */
int dnr_mlp_wgrad(const void* dY_pk, int Fy, const void* X_pk, int Fx, int row_tiles, float* dW, int ldw, int n_valid,
                  int k_valid, float* db, dnr_stream_t stream);

/* lin_out + activations: rgbsigma[s] = [sigmoid(o0..2), relu(o3)], o = lin_out(xa)
 * (resnetfc.py:158, src/models/pixelnerf.py:137-145); xa = relu(x) PK128[tiles*128 x 512], tile t holding samples
 * [t*samples_per_tile, (t+1)*samples_per_tile) in its first rows (128 for the per-layer path); xa_lo_pk (nullable):
 * the residue image of the verification mode, xa = hi + lo. */
int dnr_head_fwd(const void* xa_pk, const void* xa_lo_pk, const float* Wout, const float* bout, long long n_samples,
                 int samples_per_tile, float* rgbsigma, dnr_stream_t stream);
/* Backward of the head: dx (XR fp32, nullable + PK128 bf16) and d_raw as PK128[rows x 128] (columns 0..3) for wgrad. */
int dnr_head_bwd(const void* xa_pk, const float* Wout, const float* rgbsigma, const float* d_rgbsigma,
                 long long n_samples, float* dx_xr, void* dx_pk, void* draw_pk, dnr_stream_t stream);
/* Backward of the view mean (resnetfc.py:9-14): d_x[(sample,view)] = d_mean[sample] / NV. */
/* dx_xr may be null (only the bf16 operand image is written; the fp32 stream is then folded into the next dgrad GEMM
 * through dnr_mlp_gemm's mean_nv < 0). */
int dnr_mean_bwd(const float* dmean_xr, int NV, long long n_samples, int view_tiles, float* dx_xr, void* dx_pk,
                 dnr_stream_t stream);
/* The same for the bf16 gradient stream (diner_b200.field.GRAD_STREAM = "bf16"): the pooled gradient is its PK128 bf16 image
 * indexed by sample; d_x[(sample,view)] = bf16(d_mean[sample] / NV) as the PK128 image of the view rows. */
int dnr_mean_bwd_pk(const void* dmean_pk, int NV, long long n_samples, int view_tiles, void* dx_pk, dnr_stream_t stream);

/* Layout converters (tests, small tensors). */
int dnr_pack_rows(const float* src, long long rows, int F, long long rows_pad, void* out_pk128, dnr_stream_t stream);
int dnr_unpack_rows(const void* pk128, long long rows, int F, float* out, dnr_stream_t stream);
int dnr_xr_to_rows(const float* xr, long long rows, int N, float* out, dnr_stream_t stream);
int dnr_rows_to_xr(const float* src, long long rows, int N, long long rows_pad, float* xr, dnr_stream_t stream);

/* ---- alpha compositing ----------------------------------------------------------------------------------- */
/* Replaces the blending tail of NeRFRendererDGS.composite (src/models/nerf_renderer.py:299-301,344-365).
 * rgbsigma [n_rays,K,4] (activated), z [n_rays,K] ascending, rays [n_rays,8] -> weights [n_rays,K] (nullable),
 * rgb [n_rays,3], depth [n_rays]. */
int dnr_composite_fwd(const float* rgbsigma, const float* z, const float* rays, int n_rays, int K, int white_bkgd,
                      float* weights, float* rgb, float* depth, dnr_stream_t stream);
/* Its backward w.r.t. rgbsigma (d_depth, d_weights nullable). */
int dnr_composite_bwd(const float* rgbsigma, const float* z, const float* rays, const float* d_rgb,
                      const float* d_depth, const float* d_weights, int n_rays, int K, int white_bkgd,
                      float* d_rgbsigma, dnr_stream_t stream);

/* ---- the whole evaluation render in one call --------------------------------------------------------------------- */
/* Device-side description of the encoded scenes (what PixelNeRF.encode() leaves behind, src/models/pixelnerf.py:34-53,
 * after the per-scene repacks above): all pointers are device pointers that stay valid for the call. */
typedef struct dnr_scene_desc {
  const float* poses;            /* [SB,NV,4,4] world -> camera */
  const float* focal;            /* [SB,NV,2] */
  const float* cxy;              /* [SB,NV,2] */
  const float* depth;            /* [SB,NV,H,W] */
  const float* depth_std_padded; /* [SB,NV,H+2pad,W+2pad] (dnr_pad_depth_std) */
  const float* normal;           /* [SB,NV,3,H,W] */
  const void* latent_nhwc_bf16;  /* [SB,NV,Hl,Wl,L] (dnr_pack_latent_nhwc) */
  const float* freqs;            /* [2*NF] */
  const float* phases;           /* [2*NF] */
  const float* lin_tab;          /* [C] = linspace(0, 1 - 1/C, C) */
  int SB, NV, H, W, Hl, Wl, L, NF, pad;
  float img_w, img_h, std_sx, std_sy, lat_sx, lat_sy, depth_diff_max;
} dnr_scene_desc;
/* The packed ResnetFC (dnr_pack_weight): the 14 matrices lin_in, lin_z x3, fc_0 x5, fc_1 x5 in both cluster-kernel layouts
 * (HOST arrays of device pointers, as dnr_mlp_quad_fwd / dnr_mlp_pair_fwd take them), the 11 bias vectors of
 * dnr_mlp_pair_fwd, and lin_out in fp32. */
typedef struct dnr_mlp_desc {
  const void* const* w128;
  const void* const* w256;
  const float* const* pair_biases;
  const float* w_out;            /* [4,512] */
  const float* b_out;            /* [4] */
} dnr_mlp_desc;
/* Bytes of device workspace for dnr_render_fwd: n_rays = SB*NR rays, K samples per ray, NV views, chunks of at most
 * rows_per_chunk per-view MLP rows (<= 0: 2^22). */
long long dnr_render_workspace_bytes(int NV, long long n_rays, int K, long long rows_per_chunk);
/* NeRFRendererDGS.forward for evaluation (src/models/nerf_renderer.py:399-424 -> sample_depthguided :65-284,
 * fill_up_uniform_samples :367-397, composite :286-365) as ONE host call: dnr_sampler_fwd (in-kernel Philox jitter from
 * `seed`; or z_in [SB*NR,K], sorted sample depths supplied by the caller), then per balanced chunk of rays dnr_encode_fwd,
 * dnr_mlp_quad_fwd with dnr_mlp_pair_fwd on the SMs the 4-CTA clusters cannot occupy (a side stream the library keeps per
 * device; the split adapts to the two kernels' measured times unless DNR_ADAPT_SPLIT=0) and dnr_head_finalize, then one
 * dnr_composite_fwd.  rays [SB*NR,8]; outputs rgb [SB*NR,3], depth [SB*NR], nullable weights_out / z_out [SB*NR,K].
 * workspace: 256-byte aligned device memory of dnr_render_workspace_bytes(...) bytes, owned by the caller; the call
 * allocates no device memory.  info (nullable, HOST int[8]): kernel launches, chunks, tiles of the last chunk on the 4-CTA
 * / 2-CTA clusters, their sum.  Same bits as the call-by-call sequence.  One call at a time per device and process (the
 * side stream is shared).  NV <= 4. */
int dnr_render_fwd(const dnr_scene_desc* scene, const dnr_mlp_desc* mlp, const float* rays, long long NR, int K, int C,
                   int G, unsigned long long seed, const float* z_in, int white_bkgd, long long rows_per_chunk,
                   void* workspace, long long workspace_bytes, float* rgb, float* depth, float* weights_out, float* z_out,
                   int* info, dnr_stream_t stream);

/* Cumulative device time (ms) of the 4-CTA / 2-CTA cluster MLP launches dnr_render_fwd has issued and timed (CUDA events
 * on the launching streams) on the current device since the library was loaded, and the number of timed launch pairs;
 * launches still in flight are not included (synchronise first).  HOST pointers, nullable. */
int dnr_render_timings(double* quad_ms, double* pair_ms, long long* n_timed);

/* ---- input producers next to the path ---------------------------------------------------- */

/* Pixel-centre rays of B target cameras: rays [B,H,W,8] = [origin 3 | unit direction 3 | near | far] from world->cam
 * extrinsics [B,4,4], intrinsics [B,3,3], z_near / z_far [B].  Replaces gen_rays (src/util/cam_geometry.py:36-79). */
int dnr_gen_rays(const float* extrinsics, const float* intrinsics, int B, int W, int H, const float* z_near,
                 const float* z_far, float* rays, dnr_stream_t stream);

/* Rays of n selected pixels per camera: pix_idx int64 [B,n] holds flat pixel indices y*W+x (negative indices count from
 * the end, as in torch); rays [B,n,8] equal dnr_gen_rays(...)[b, pix_idx[b,i]] bit for bit.  Replaces gen_rays followed by
 * the advanced-index gather of DINER.calc_losses (src/models/diner.py:222-258).  bad_flag (device int, nullable) is set to
 * 1 when an index is out of range (that ray is then generated for pixel 0). */
int dnr_gen_rays_at(const float* extrinsics, const float* intrinsics, int B, int W, int H, const float* z_near,
                    const float* z_far, const long long* pix_idx, int n, float* rays, int* bad_flag, dnr_stream_t stream);

/* Camera-space normal maps [N,3,H,W] from depth maps [N,1,H,W] and intrinsics [N,3,3] by central differences of the
 * back-projected points, with the reference's hole-cleaning rule and NaN behaviour.  Replaces depth2normal
 * (src/util/depth2normal.py:6-87) including the axis its torch.cross (no dim) takes when N == 3. */
int dnr_depth2normal(const float* dmap, const float* K, int N, int H, int W, float* normal, dnr_stream_t stream);

/* On-disk depth format: decoded uint16 PNG planes in units of `scale` (1e-4 m) -> fp32 depth
 * [N,H,W] and depth-std [N,H,W] maps in one pass.  `pred` / `conf` point at the first pixel of the MVS depth /
 * confidence plane of image 0 (FaceScape triptych [gt|pred|conf]: tri + W and tri + 2 W, row_stride = 3 W,
 * img_stride = 3 W H; dense images: row_stride = W, img_stride = W H); `mesh` is a dense [N,H,W] plane.
 * depth_type 0 = MVS planes ("original"), 1 = mesh depth with confidence 0.8 where it is non-zero, 2 = mesh with its
 * holes filled from the MVS planes ("merge").  depth = u16 * scale [/ div0 * mul1]; std = std_a * conf + std_b.
 * flags: 1 clip std at 0; 2 std = 0 where depth == 0; 4 conf is rescaled (/ div0 * mul1) like depth; 8 std = std_const.
 * Every operation is a separate fp32 rounding in the reference's order.  Replaces FacescapeDataSet.read_depth +
 * conf2std (src/data/facescape.py:54-56,69-106,266), the PNG branch of DTUDataSet.read_depth + conf2std
 * (src/data/dtu.py:68-70,104-107,119,221-223) and MultiFaceDataset.read_depth + std law (src/data/multiface.py:103-109,
 * 305-311).  PNG inflation stays in the host image library. */
int dnr_decode_depth_u16(const uint16_t* pred, const uint16_t* conf, long long img_stride, int row_stride,
                         const uint16_t* mesh, int N, int H, int W, int depth_type, float scale, float div0, float mul1,
                         float std_a, float std_b, float std_const, int flags, float* depth_out, float* std_out,
                         dnr_stream_t stream);

/* Nearest-vertex deformation (first piece): out[sb,i] = points[sb,i] + offsets[sb, argmin_v
 * |points[sb,i] - vertices[sb,v]|^2] (brute-force K = 1 search, squared distance accumulated x, y, z in fp32, the lowest index
 * wins a tie).  points / out [SB,B,3], vertices / offsets [SB,V,3]; idx_out [SB,B] int32 and dist_out [SB,B] (squared) are
 * optional.  Replaces NeRFRendererDGS.deform_points of the NOVEL variants (src/models/novel/nerf_novel_renderer.py:40-50,
 * pytorch3d.ops.knn_points + gather + add).  Parity with pytorch3d is unpinned: it is not installed in the build image. */
int dnr_deform_points(const float* points, const float* vertices, const float* offsets, int SB, long long B, int V,
                      float* out, int* idx_out, float* dist_out, dnr_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DINER_B200_H */
