"""Host-side sequencing of the field evaluation (reference PixelNeRF.forward, src/models/pixelnerf.py:55-145,
and ResnetFC.forward, src/models/resnetfc.py:129-159) on the CUDA kernels, forward and backward.

    encode (project + gather + codes)            dnr_encode_fwd
    x0   = lin_in(code) + lin_z[0](lat)          one GEMM, two K segments
    for b in 0..2:  net = relu(fc_0(relu(x)))    GEMM, ReLU epilogue
                    x  += fc_1(net) [+ lin_z[b+1](lat)]    GEMM, residual epilogue (b == 2: + view mean)
    for b in 3..4:  same on the pooled rows
    rgb, sigma = sigmoid / relu(lin_out(relu(x)))          dnr_head_fwd

Activations never exist in row-major form: every GEMM epilogue writes the next GEMM's A operand directly
in the packed tensor-core layout (csrc/common.cuh).  The backward pass runs the same kernels on the
transposed weights (dgrad) and dnr_mlp_wgrad on the saved operands.
"""
from __future__ import annotations

from dataclasses import dataclass, field as dc_field

import contextlib
import os

import torch

from . import _lib
from .scene import SceneBuffers

HID = 512
IN_PAD = 64
USE_FUSED_FORWARD = True     # tests flip this to compare the fused kernel with the per-layer path
PAIR_PROF = None             # debug: int64 tensor [148*8] receiving cycle counters of the pair kernel
# inference MLP: "quad" / "pair" (mlp_pair.cu: 4- / 2-CTA clusters) | "chain" (mlp_fused.cu) | "layers" (gemm.cu)
FORWARD_KERNEL = "quad"
# With "quad": share of a chunk's tiles rendered by 2-CTA clusters on the SMs that the 4-CTA clusters cannot use
# (the quad kernel keeps 33 clusters = 132 of 148 SMs resident on a B200), on a second stream.  0 disables; None =
# spare * r / (spare * r + quad_clusters), r = throughput of one pair cluster relative to one quad cluster, measured
# on the BASELINE.json shapes (profiles/r1_quad_kernel.md) and rounded down: a late pair kernel is the tail of the
# step while an early one only idles 16 SMs.  The quad kernel merges pooled rows (MERGE_POOLED), the pair kernel does
# not, so r falls with the view count.
PAIR_RATE = {1: 0.37, 2: 0.34, 3: 0.32, 4: 0.30}
PAIR_SHARE = None
# The split adapts itself: both kernels of a chunk are bracketed by CUDA events, and once a pair of timings has
# completed (polled without synchronising) the rate of that (device, NV) moves towards the measured one.  The split
# never changes the result (both kernels produce the same bits, tests/test_gpu_field.py), only the balance.
ADAPT_SPLIT = os.environ.get("DNR_ADAPT_SPLIT", "1") != "0"     # 0: fixed split from PAIR_RATE (reproducible captures)
_split_state = {}            # (device, NV) -> dict(rate=float, pending=[(e_q0, e_q1, e_p0, e_p1, tiles_q, tiles_p)])
# Backward pass: precision of the residual GRADIENT stream d x_b between blocks.  "bf16": the stream is the bf16 operand
# image the next dgrad GEMM reads anyway (each block adds its dgrad in fp32 and rounds once) -- 3 KB instead of 7 KB per
# row in the HBM-bound fc_0 dgrad layers.  "fp32": a separate fp32 stream is read and written by every block (round 1).
GRAD_STREAM = os.environ.get("DNR_GRAD_STREAM", "bf16")
PE_FORWARD_LAYERS = False     # NOVEL_PE evaluation on the per-layer kernels instead of the cluster kernels (tests, A/B)
MERGE_POOLED = True          # quad kernel: run blocks 3-4 + head once per NV tiles on merged pooled rows (NV > 1)
_side_stream = {}
LAST_SPLIT = {}              # bench instrumentation: tiles of the most recent inference MLP launch (main / side kernel)


def _mlp_signature(mlp):
    return tuple((p.data_ptr(), p._version) for p in mlp.parameters())


def _pack_weight(w: torch.Tensor, transpose: bool = False) -> torch.Tensor:
    """nn.Linear weight (out, in) -> bf16 tensor-core image [N/256][K/8][256][8] of W (or of W^T), zero-padded."""
    w = w.detach().to(torch.float32).contiguous()
    rows, cols = w.shape
    n, k = (cols, rows) if transpose else (rows, cols)
    n_pad, k_pad = (n + 255) // 256 * 256, (k + 63) // 64 * 64
    out = torch.empty(n_pad * k_pad, device=w.device, dtype=torch.bfloat16)
    _lib.call("dnr_pack_weight", _lib.ptr(w), rows, cols, cols, 1 if transpose else 0, n_pad, k_pad, _lib.ptr(out),
              _lib.stream_ptr())
    return out


class PackedMLP:
    """bf16 tensor-core images of the ResnetFC weights (W for forward, W^T for dgrad) + fp32 biases."""

    def __init__(self, mlp, need_transposed: bool, split: bool = False):
        self.sig = _mlp_signature(mlp)
        self.has_t = need_transposed
        self.has_split = split
        pack = _pack_weight

        f = lambda t: t.detach().to(torch.float32).contiguous()
        self.w_in = pack(mlp.lin_in.weight)                      # [512 x 64]
        self.w_z = [pack(l.weight) for l in mlp.lin_z]
        self.w_0 = [pack(b.fc_0.weight) for b in mlp.blocks]
        self.w_1 = [pack(b.fc_1.weight) for b in mlp.blocks]
        self.b_first = f(mlp.lin_in.bias + mlp.lin_z[0].bias)
        self.b_0 = [f(b.fc_0.bias) for b in mlp.blocks]
        # fc_1 bias of block b merged with the lin_z bias that is added right after it (b = 0, 1)
        self.b_1 = [f(mlp.blocks[b].fc_1.bias + (mlp.lin_z[b + 1].bias if b + 1 < len(mlp.lin_z) else 0))
                    for b in range(len(mlp.blocks))]
        self.w_out = f(mlp.lin_out.weight)
        self.b_out = f(mlp.lin_out.bias)
        # host arrays of device pointers for the fused forward kernel (dnr_mlp_fused_fwd)
        import ctypes as C
        ws = [self.w_in, *self.w_z, *self.w_0, *self.w_1]
        bs = [self.b_first, *self.b_0, *self.b_1]
        self.w_ptrs = (C.c_void_p * len(ws))(*[t.data_ptr() for t in ws])
        self.b_ptrs = (C.c_void_p * len(bs))(*[t.data_ptr() for t in bs])
        # cumulative residual-stream biases for the CTA-pair kernel (the stream lives in TMEM without biases)
        cb = [self.b_first]
        for b in range(3):
            cb.append(cb[-1] + self.b_1[b])
        cb.append(cb[-1] + self.b_1[3])
        cb.append(cb[-1] + self.b_1[4])
        self.cb = [t.contiguous() for t in cb]
        pb = [*self.cb, *self.b_0]
        self.pair_b_ptrs = (C.c_void_p * len(pb))(*[t.data_ptr() for t in pb])
        # PK128 images of the forward weights for the quad kernel: [N/256][K/8][256][8] -> [N/128][K/8][128][8]
        def to128(w, k_pad):
            return w.view(-1, k_pad // 8, 2, 128, 8).permute(0, 2, 1, 3, 4).contiguous().view(-1)
        self.w128 = [to128(self.w_in, IN_PAD)] + [to128(w, HID) for w in ws[1:]]
        self.w128_ptrs = (C.c_void_p * len(self.w128))(*[t.data_ptr() for t in self.w128])
        if need_transposed:
            self.wt_z = [pack(l.weight, True) for l in mlp.lin_z]
            self.wt_0 = [pack(b.fc_0.weight, True) for b in mlp.blocks]
            self.wt_1 = [pack(b.fc_1.weight, True) for b in mlp.blocks]
        if split:
            # verification mode: the rounding residue W - bf16(W) as a second bf16 image of every forward weight
            lo = lambda w: pack(w.detach().float() - w.detach().float().bfloat16().float())
            self.lo_in = lo(mlp.lin_in.weight)
            self.lo_z = [lo(l.weight) for l in mlp.lin_z]
            self.lo_0 = [lo(b.fc_0.weight) for b in mlp.blocks]
            self.lo_1 = [lo(b.fc_1.weight) for b in mlp.blocks]


class _Lin:
    def __init__(self, weight, bias):
        self.weight, self.bias = weight, bias


class PackedPE(PackedMLP):
    """NOVEL_PE field (src/models/novel_pe/pe_novel_pixelnerf.py): ResnetFC with d_latent = 518 behind a 518 -> 512
    `deformation_layer`.  The six positional values of a row travel in columns 55..60 of the 64-wide geometric-code
    operand (dnr_encode_novel_pe_fwd), so every 518-wide nn.Linear becomes two K segments of one GEMM:

        W[:, :512] on the 512-wide operand   +   [0 (55) | W[:, 512:518] | 0 (3)] on the code operand

    and lin_in (55 columns of the same operand) merges with lin_z[0]'s positional columns into ONE 512 x 64 image.  With
    that, the attributes of PackedMLP keep their meaning (w_in, w_z, ...) and the per-layer sequence needs only the extra
    segment `w_zc[b]` in blocks 1-2 and the deformation GEMM (`wd`, `wdc`, `b_d`) in front."""

    def __init__(self, model, need_transposed: bool):
        mlp, dl = model.mlp_fine, model.deformation_layer
        d_in = mlp.lin_in.weight.shape[1]
        if mlp.lin_z[0].weight.shape[1] != HID + 6 or dl.weight.shape != (HID, HID + 6) or d_in + 6 > IN_PAD:
            raise NotImplementedError("diner_b200: the NOVEL_PE field needs d_latent = 518 and a 518 -> 512 deformation layer")
        self.d_in = d_in

        def code_side(w6, w_code=None):
            """(512, 6) positional columns [+ (512, d_in) code columns] -> dense (512, d_in + 6) over operand columns 0..60"""
            left = w_code.detach().float() if w_code is not None else torch.zeros(HID, d_in, device=w6.device)
            return torch.cat([left, w6.detach().float()], 1)

        wz = [l.weight.detach().float() for l in mlp.lin_z]
        shim = _Lin(None, None)
        shim.lin_in = _Lin(code_side(wz[0][:, HID:], mlp.lin_in.weight), mlp.lin_in.bias)
        shim.lin_z = [_Lin(w[:, :HID], l.bias) for w, l in zip(wz, mlp.lin_z)]
        shim.blocks, shim.lin_out = mlp.blocks, mlp.lin_out
        shim.parameters = lambda: [*mlp.parameters(), *dl.parameters()]
        super().__init__(shim, need_transposed)
        wd = dl.weight.detach().float()
        self.w_zc = [None] + [_pack_weight(code_side(w[:, HID:])) for w in wz[1:]]      # blocks 1, 2
        self.wd, self.wdc = _pack_weight(wd[:, :HID]), _pack_weight(code_side(wd[:, HID:]))
        self.b_d = dl.bias.detach().float().contiguous()
        if need_transposed:
            self.wt_d = _pack_weight(wd[:, :HID], True)
        # the cluster kernels' second code K-step (dnr_mlp_pair_fwd_pe / dnr_mlp_quad_fwd_pe): PK256 and PK128 images
        import ctypes as C
        self.w_zc128 = [w.view(-1, IN_PAD // 8, 2, 128, 8).permute(0, 2, 1, 3, 4).contiguous().view(-1) for w in self.w_zc[1:]]
        self.zc_ptrs = (C.c_void_p * 2)(*[t.data_ptr() for t in self.w_zc[1:]])
        self.zc128_ptrs = (C.c_void_p * 2)(*[t.data_ptr() for t in self.w_zc128])


_packed_cache: dict = {}
_packed_pe_cache: dict = {}


def get_packed_pe(model, need_transposed: bool) -> PackedPE:
    """PackedPE of a NOVEL_PE model, cached like get_packed (same invalidation rules)."""
    sig = tuple((q.data_ptr(), q._version) for q in (*model.mlp_fine.parameters(), *model.deformation_layer.parameters()))
    hit = _packed_pe_cache.get(id(model))
    if hit is not None and hit.sig == sig and (hit.has_t or not need_transposed):
        return hit
    pk = PackedPE(model, need_transposed)
    _packed_pe_cache.clear()
    _packed_pe_cache[id(model)] = pk
    return pk


def get_packed(mlp, need_transposed: bool, split: bool = False) -> PackedMLP:
    """Packed weights of `mlp`, cached on the parameters' (data_ptr, _version).  Writers that bypass the version counter
    (`p.data.copy_`, c10d broadcasts, fused optimizers working on .data) must call `invalidate_caches()`."""
    hit = _packed_cache.get(id(mlp))
    if (hit is not None and hit.sig == _mlp_signature(mlp) and (hit.has_t or not need_transposed)
            and (hit.has_split or not split)):
        return hit
    pk = PackedMLP(mlp, need_transposed or (hit is not None and hit.sig == _mlp_signature(mlp) and hit.has_t),
                   split or (hit is not None and hit.sig == _mlp_signature(mlp) and hit.has_split))
    _packed_cache.clear()
    _packed_cache[id(mlp)] = pk
    return pk


def invalidate_caches() -> None:
    """Forget the packed weights (and, via scene.invalidate, the per-scene repacks)."""
    _packed_cache.clear()
    _packed_pe_cache.clear()


_scratch: dict = {}


def _update_split(state: dict, spare: int, n_quad: int) -> None:
    """Fold completed (quad, pair) timings into the pair rate: rate = (tiles_p / t_p / spare) / (tiles_q / t_q / n_quad)."""
    keep = []
    for rec in state["pending"]:
        e_q0, e_q1, e_p0, e_p1, tiles_q, tiles_p = rec
        if not (e_q1.query() and e_p1.query()):
            keep.append(rec)
            continue
        t_q, t_p = e_q0.elapsed_time(e_q1), e_p0.elapsed_time(e_p1)
        if t_q > 0.05 and t_p > 0.05 and spare > 0:            # ignore launches too small to time
            measured = (tiles_p / t_p / spare) / (tiles_q / t_q / n_quad)
            measured = min(max(measured, 0.15), 0.6)
            # move a third of the way, aim 3 % low: a late pair kernel is the tail, an early one only idles 16 SMs
            state["rate"] += (0.97 * measured - state["rate"]) / 3.0
    state["pending"] = keep


def _quad_scratch(dev) -> torch.Tensor:
    """Parking space of the quad kernel's pooled rows (128 KB per CTA, L2-resident)."""
    k = ("quad", str(dev), torch.cuda.current_stream(dev).cuda_stream)      # one per stream: launches on different
    if k not in _scratch:                                                   # streams may overlap
        _scratch[k] = torch.zeros(_lib.load().dnr_mlp_quad_scratch_bytes(), device=dev, dtype=torch.uint8)
    return _scratch[k]


def _fused_scratch(dev) -> torch.Tensor:
    """Per-CTA activation scratch of the fused forward kernel (1 MB per SM; re-used by every tile, L2-resident)."""
    k = ("chain", str(dev), torch.cuda.current_stream(dev).cuda_stream)
    if k not in _scratch:
        n = _lib.load().dnr_mlp_fused_scratch_bytes(0)
        _scratch[k] = torch.zeros(n, device=dev, dtype=torch.uint8)
    return _scratch[k]


def _gemm(A0, B0, K0, m_tiles, N=HID, A1=None, B1=None, K1=0, A2=None, B2=None, K2=0, bias=None, mask=None,
          res_in=None, res_out=None, act_out=None, relu=0, mean_nv=0, n_samples=0, max_ctas=0, act_lo=None, res_in_pk=None):
    p = _lib.ptr
    _lib.call("dnr_mlp_gemm", p(A0), p(B0), K0, p(A1), p(B1), K1, p(A2), p(B2), K2, m_tiles, N, p(bias), p(mask),
              p(res_in), p(res_in_pk), p(res_out), p(act_out), p(act_lo), relu, mean_nv, n_samples, max_ctas,
              _lib.stream_ptr())


def _gemm3(A, W, K, m_tiles, **kw):
    """Verification mode: A = (hi, lo), W = (hi, lo) bf16 pairs; A W^T ~ hi hi^T + hi lo^T + lo hi^T as the three K
    segments of ONE launch (the lo lo^T term is 2^-16 of the product and dropped)."""
    (ah, al), (wh, wl) = A, W
    _gemm(ah, wh, K, m_tiles, A1=ah, B1=wl, K1=K, A2=al, B2=wh, K2=K, **kw)


@dataclass
class FieldSaved:
    """Operands kept for the backward pass (all in packed layouts)."""
    n_samples: int = 0
    mv: int = 0
    ms: int = 0
    lat: torch.Tensor = None
    code: torch.Tensor = None
    xa: list = dc_field(default_factory=list)      # relu(x_b) entering block b (b = 0..4); xa[5] = relu(x_5)
    net: list = dc_field(default_factory=list)     # relu(fc_0(.)) of block b
    out: torch.Tensor = None
    lat_src: torch.Tensor = None                   # NOVEL_PE: the encoder feature alone (`lat` is then final_latent)


def _encode_args(sc: SceneBuffers, NR: int, K: int, ray0: int, n_rays: int):
    return (sc.SB, sc.NV, NR, K, sc.H, sc.W, sc.Hl, sc.Wl, sc.L, sc.num_freqs, ray0, n_rays,
            sc.img_w, sc.img_h, sc.lat_sx, sc.lat_sy)


def field_forward(sc: SceneBuffers, pk: PackedMLP, rays: torch.Tensor, z: torch.Tensor, NR: int, ray0: int,
                  save: bool, out: torch.Tensor = None, precision: str = "bf16", novel: dict = None):
    """Evaluate the field on one chunk of rays.

    rays (n,8) / z (n,K) are the chunk's rows of the flattened (SB*NR) batch, `ray0` the index of its first
    ray in that batch (the scene of a ray is ray // NR).  Returns (rgbsigma (n*K,4) fp32, FieldSaved | None);
    `out` (contiguous (n*K,4) fp32) receives the result in place when given.  `novel` (diner_b200.novel):
    the chunk's deformed sample points and the generic camera / learned map of the NOVEL variant.
    """
    p, st = _lib.ptr, _lib.stream_ptr()
    dev = rays.device
    n, K = z.shape
    n_samples = n * K
    NV = sc.NV
    mv = _lib.load().dnr_view_row_tiles(n_samples, NV)
    ms = (n_samples + 127) // 128
    bf = dict(device=dev, dtype=torch.bfloat16)
    f32 = dict(device=dev, dtype=torch.float32)
    if out is None:
        out = torch.empty(n_samples, 4, **f32)
    assert out.is_contiguous() and out.shape == (n_samples, 4) and out.dtype == torch.float32
    if precision == "split":
        return _field_forward_split(sc, pk, rays, z, NR, ray0, save, out)

    lat = torch.empty(mv * 128 * HID, **bf)
    code = torch.empty(mv * 128 * IN_PAD, **bf)
    pe = novel is not None and novel.get("pe") is not None
    lat_src = None
    if pe:
        # NOVEL_PE (pe_novel_pixelnerf.py:260-275): encoder feature, gen_latent taps and [code | pe 6] leave as three
        # operands; final_latent = gen + deformation_layer([feature | pe 6]) is one GEMM with the taps as residual operand
        g, e = novel, novel["pe"]
        gen_pk = torch.empty(mv * 128 * HID, **bf)
        _lib.call("dnr_encode_novel_pe_fwd", p(rays), p(z), p(g["pts_in"]), p(g["pts_gen"]), p(sc.poses), p(sc.focal),
                  p(sc.cxy), p(sc.depth), p(sc.latent_nhwc), p(g["poses"]), p(g["focal"]), p(g["cxy"]), p(g["latent"]),
                  g["Hg"], g["Wg"], g["img_w"], g["img_h"], g["sx"], g["sy"], p(e["src"]), e["Hp"], e["Wp"], e["sx"],
                  e["sy"], p(e["tgt_poses"]), p(e["tgt_focal"]), p(e["tgt_cxy"]), e["tgt_img_w"], e["tgt_img_h"],
                  p(e["tgt"]), e["Ht"], e["Wt"], e["tgt_sx"], e["tgt_sy"], p(sc.freqs), p(sc.phases),
                  *_encode_args(sc, NR, K, ray0, n), p(lat), p(gen_pk), p(code), st)
        lat_src, lat = lat, torch.empty(mv * 128 * HID, **bf)
        _gemm(lat_src, pk.wd, HID, mv, A1=code, B1=pk.wdc, K1=IN_PAD, bias=pk.b_d, res_in_pk=gen_pk, act_out=lat)
        del gen_pk
    elif novel is not None:
        g = novel
        _lib.call("dnr_encode_novel_fwd", p(rays), p(z), p(g["pts_in"]), p(g["pts_gen"]), p(sc.poses), p(sc.focal),
                  p(sc.cxy), p(sc.depth), p(sc.latent_nhwc), p(g["poses"]), p(g["focal"]), p(g["cxy"]), p(g["latent"]),
                  g["Hg"], g["Wg"], g["img_w"], g["img_h"], g["sx"], g["sy"], p(sc.freqs), p(sc.phases),
                  *_encode_args(sc, NR, K, ray0, n), p(lat), p(code), st)
    else:
        _lib.call("dnr_encode_fwd", p(rays), p(z), p(sc.poses), p(sc.focal), p(sc.cxy), p(sc.depth), p(sc.latent_nhwc),
                  p(sc.freqs), p(sc.phases), *_encode_args(sc, NR, K, ray0, n), p(lat), p(code), st)

    if not save and FORWARD_KERNEL in ("quad", "pair") and NV <= 4 and not (pe and PE_FORWARD_LAYERS):
        # inference: CTA-cluster kernels, residual stream in TMEM, lin_out fused (csrc/mlp_pair.cu)
        per_tile = 4 * (32 // NV)                                # samples per 128-row view tile (RowMap)
        mv_q, spare = mv, 0
        if FORWARD_KERNEL == "quad":
            n_quad = _lib.load().dnr_mlp_quad_max_clusters(NV)
            n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
            spare = max(0, (n_sm - 4 * n_quad) // 2)             # SM pairs the quad kernel leaves idle
            state = _split_state.setdefault((str(dev), NV, MERGE_POOLED),
                                            dict(rate=PAIR_RATE[NV] if MERGE_POOLED else 0.37, pending=[]))
            _update_split(state, spare, n_quad)
            r = state["rate"]
            share = PAIR_SHARE if PAIR_SHARE is not None else r * spare / (r * spare + n_quad)
            if share > 0 and spare > 0 and (PAIR_SHARE is not None or mv >= 64 * spare):
                mv_q = max(0, mv - max(2, int(mv * share)) // 2 * 2)
        else:
            mv_q = 0
        LAST_SPLIT.update(tiles_main=mv_q if mv_q > 0 else mv, tiles_side=mv - mv_q if mv_q > 0 else 0, tiles=mv,
                          kernel=FORWARD_KERNEL)
        cur = torch.cuda.current_stream(dev)
        side = None
        if 0 < mv_q < mv:                                        # fork: the pair clusters run beside the quad clusters
            side = _side_stream.setdefault(str(dev), torch.cuda.Stream(dev))
            side.wait_stream(cur)
        timing = side is not None and ADAPT_SPLIT and PAIR_SHARE is None and len(state["pending"]) < 8
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if timing else None
        if mv_q > 0:
            part_q = torch.empty(mv_q * 4 * 128, 4, **f32)
            if timing:
                ev[0].record(cur)
            if pe:
                _lib.call("dnr_mlp_quad_fwd_pe", p(lat), p(code), pk.w128_ptrs, pk.zc128_ptrs, pk.pair_b_ptrs, p(pk.w_out),
                          NV, mv_q, p(_quad_scratch(dev)) if MERGE_POOLED and NV > 1 else None, p(part_q), st)
            else:
                _lib.call("dnr_mlp_quad_fwd", p(lat), p(code), pk.w128_ptrs, pk.pair_b_ptrs, p(pk.w_out), NV, mv_q,
                          p(_quad_scratch(dev)) if MERGE_POOLED and NV > 1 else None, p(part_q), p(PAIR_PROF), st)
            if timing:
                ev[1].record(cur)
            _lib.call("dnr_head_finalize", p(part_q), p(pk.b_out), NV, min(n_samples, mv_q * per_tile), 4, p(out), st)
        if mv_q < mv:
            n_p = mv - mv_q
            part_p = torch.empty(n_p * 2 * 128, 4, **f32)
            s0 = mv_q * per_tile                                 # first sample of the pair clusters' tiles
            lat_p, code_p, out_p = lat[mv_q * 128 * HID:], code[mv_q * 128 * IN_PAD:], out[s0:]
            with torch.cuda.stream(side) if side is not None else contextlib.nullcontext():
                st2 = _lib.stream_ptr()
                if timing:
                    ev[2].record(side)
                if pe:
                    _lib.call("dnr_mlp_pair_fwd_pe", p(lat_p), p(code_p), pk.w_ptrs, pk.zc_ptrs, pk.pair_b_ptrs, p(pk.w_out),
                              NV, n_p, spare if side is not None else 0, p(part_p), st2,
                              tag="overlapped" if side is not None else None)
                else:
                    _lib.call("dnr_mlp_pair_fwd", p(lat_p), p(code_p), pk.w_ptrs, pk.pair_b_ptrs, p(pk.w_out), NV, n_p,
                              spare if side is not None else 0, p(part_p), p(PAIR_PROF) if mv_q == 0 else None, st2,
                              tag="overlapped" if side is not None else None)
                if timing:
                    ev[3].record(side)
                    state["pending"].append((*ev, mv_q, n_p))
                _lib.call("dnr_head_finalize", p(part_p), p(pk.b_out), NV, n_samples - s0, 2, p(out_p), st2)
            if side is not None:
                cur.wait_stream(side)                            # join
        return out, None
    if not save and FORWARD_KERNEL != "layers" and not pe:
        # inference: the whole ResnetFC in one persistent kernel, activations stay in L2 (csrc/mlp_fused.cu)
        S = 4 * NV * (32 // NV)
        groups = (mv + NV - 1) // NV
        xa = torch.empty(groups * 128 * HID, **bf)
        _lib.call("dnr_mlp_fused_fwd", p(lat), p(code), pk.w_ptrs, pk.b_ptrs, NV, mv, n_samples,
                  p(_fused_scratch(dev)), p(xa), 0, st)
        _lib.call("dnr_head_fwd", p(xa), None, p(pk.w_out), p(pk.b_out), n_samples, S, p(out), st)
        return out, None
    sv = FieldSaved(n_samples=n_samples, mv=mv, ms=ms, lat=lat, code=code, lat_src=lat_src) if save else None
    zc = (lambda b: dict(A2=code, B2=pk.w_zc[b], K2=IN_PAD)) if pe else (lambda b: {})     # NOVEL_PE: lin_z's pe columns
    x = torch.empty(mv * 128 * HID, **f32)                      # residual stream, XR layout (updated in place)
    new_act = lambda tiles: torch.empty(tiles * 128 * HID, **bf)
    xa = new_act(mv)
    net = new_act(mv)
    # x0 = lin_in(code) + lin_z[0](lat)                           resnetfc.py:139,152-154
    _gemm(code, pk.w_in, IN_PAD, mv, A1=lat, B1=pk.w_z[0], K1=HID, bias=pk.b_first, res_out=x, act_out=xa, relu=1)
    xm = xam = None
    for b in range(3):
        if save:
            sv.xa.append(xa)
        # net = relu(fc_0(relu(x)))                               resnetfc.py:62
        _gemm(xa, pk.w_0[b], HID, mv, bias=pk.b_0[b], act_out=net, relu=1)
        if save:
            sv.net.append(net)
        if b < 2:
            xa_next = new_act(mv) if save else xa
            # x = x + fc_1(net) + lin_z[b+1](lat)                 resnetfc.py:63-69,152-154
            _gemm(net, pk.w_1[b], HID, mv, A1=lat, B1=pk.w_z[b + 1], K1=HID, bias=pk.b_1[b], res_in=x, res_out=x,
                  act_out=xa_next, relu=1, **zc(b + 1))
            xa = xa_next
            if save:
                net = new_act(mv)
        else:
            # last per-view block: the epilogue averages the NV rows of each sample (resnetfc.py:147-151)
            # the mean epilogue writes the rows of valid samples only: the padding rows of the LAST pooled tile (both
            # layouts keep a tile contiguous) must be zero -- they are read again as wgrad operands (0 x NaN = NaN)
            xm = torch.empty(ms * 128 * HID, **f32)
            xam = torch.empty(ms * 128 * HID, **bf)
            xm[(ms - 1) * 128 * HID:].zero_()
            xam[(ms - 1) * 128 * HID:].zero_()
            _gemm(net, pk.w_1[b], HID, mv, bias=pk.b_1[b], res_in=x, res_out=xm, act_out=xam, relu=1, mean_nv=NV,
                  n_samples=n_samples)
    del x
    xa, net = xam, torch.empty(ms * 128 * HID, **bf)      # fc_0 writes every row of every tile
    for b in (3, 4):
        if save:
            sv.xa.append(xa)
        _gemm(xa, pk.w_0[b], HID, ms, bias=pk.b_0[b], act_out=net, relu=1)
        if save:
            sv.net.append(net)
        xa_next = new_act(ms) if save else xa
        _gemm(net, pk.w_1[b], HID, ms, bias=pk.b_1[b], res_in=xm, res_out=xm, act_out=xa_next, relu=1)
        xa = xa_next
        if save and b == 3:
            net = new_act(ms)
    _lib.call("dnr_head_fwd", p(xa), None, p(pk.w_out), p(pk.b_out), n_samples, 128, p(out), st)
    if save:
        sv.xa.append(xa)
        sv.out = out
    return out, sv


def _field_forward_split(sc: SceneBuffers, pk: PackedMLP, rays, z, NR: int, ray0: int, save: bool, out):
    """Verification mode of field_forward (NeRFRendererDGS(precision="split")): the same kernels, operand layouts,
    row maps and epilogues as the per-layer path, but every bf16 operand x is carried as a pair (hi = bf16(x),
    lo = bf16(x - hi)) and every product as hi*hi + hi*lo + lo*hi -- 2^-16 relative per product instead of 2^-8, so the
    result can be held to ~1e-4 against the fp32 reference and an indexing error cannot hide in the bf16 budget.
    The backward pass (when `save`) runs on the hi images, i.e. it is the ordinary bf16 backward."""
    p, st = _lib.ptr, _lib.stream_ptr()
    dev = rays.device
    n, K = z.shape
    n_samples = n * K
    NV = sc.NV
    mv = _lib.load().dnr_view_row_tiles(n_samples, NV)
    ms = (n_samples + 127) // 128
    bf = dict(device=dev, dtype=torch.bfloat16)
    f32 = dict(device=dev, dtype=torch.float32)
    pair = lambda tiles, width=HID: (torch.empty(tiles * 128 * width, **bf), torch.empty(tiles * 128 * width, **bf))
    lat, code = pair(mv), pair(mv, IN_PAD)
    _lib.call("dnr_encode_fwd_split", p(rays), p(z), p(sc.poses), p(sc.focal), p(sc.cxy), p(sc.depth),
              p(sc.latent_nhwc), p(sc.latent_lo()), p(sc.freqs), p(sc.phases), *_encode_args(sc, NR, K, ray0, n),
              p(lat[0]), p(code[0]), p(lat[1]), p(code[1]), st)
    sv = FieldSaved(n_samples=n_samples, mv=mv, ms=ms, lat=lat[0], code=code[0]) if save else None
    x = torch.empty(mv * 128 * HID, **f32)
    xa = pair(mv)
    # x0 = lin_in(code) + lin_z[0](lat)
    _gemm3(code, (pk.w_in, pk.lo_in), IN_PAD, mv, bias=pk.b_first, res_out=x)
    _gemm3(lat, (pk.w_z[0], pk.lo_z[0]), HID, mv, res_in=x, res_out=x, act_out=xa[0], act_lo=xa[1], relu=1)
    xm = xam = None
    for b in range(3):
        if save:
            sv.xa.append(xa[0])
        net = pair(mv)
        _gemm3(xa, (pk.w_0[b], pk.lo_0[b]), HID, mv, bias=pk.b_0[b], act_out=net[0], act_lo=net[1], relu=1)
        if save:
            sv.net.append(net[0])
        if b < 2:
            xa = pair(mv)
            _gemm3(net, (pk.w_1[b], pk.lo_1[b]), HID, mv, bias=pk.b_1[b], res_in=x, res_out=x)
            _gemm3(lat, (pk.w_z[b + 1], pk.lo_z[b + 1]), HID, mv, res_in=x, res_out=x, act_out=xa[0], act_lo=xa[1],
                   relu=1)
        else:
            xm = torch.empty(ms * 128 * HID, **f32)
            xam = pair(ms)
            xm[(ms - 1) * 128 * HID:].zero_()
            xam[0][(ms - 1) * 128 * HID:].zero_()
            xam[1][(ms - 1) * 128 * HID:].zero_()
            _gemm3(net, (pk.w_1[b], pk.lo_1[b]), HID, mv, bias=pk.b_1[b], res_in=x, res_out=xm, act_out=xam[0],
                   act_lo=xam[1], relu=1, mean_nv=NV, n_samples=n_samples)
    del x
    xa = xam
    for b in (3, 4):
        if save:
            sv.xa.append(xa[0])
        net = pair(ms)
        _gemm3(xa, (pk.w_0[b], pk.lo_0[b]), HID, ms, bias=pk.b_0[b], act_out=net[0], act_lo=net[1], relu=1)
        if save:
            sv.net.append(net[0])
        xa = pair(ms)
        _gemm3(net, (pk.w_1[b], pk.lo_1[b]), HID, ms, bias=pk.b_1[b], res_in=xm, res_out=xm, act_out=xa[0],
               act_lo=xa[1], relu=1)
    _lib.call("dnr_head_fwd", p(xa[0]), p(xa[1]), p(pk.w_out), p(pk.b_out), n_samples, 128, p(out), st)
    if save:
        sv.xa.append(xa[0])
        sv.out = out
    return out, sv


def _wgrad(dY, Fy, X, Fx, tiles, dW, n_valid, k_valid, db):
    p = _lib.ptr
    _lib.call("dnr_mlp_wgrad", p(dY), Fy, p(X), Fx, tiles, p(dW), dW.shape[1], n_valid, k_valid, p(db),
              _lib.stream_ptr())


def field_backward(sc: SceneBuffers, pk: PackedMLP, mlp, sv: FieldSaved, rays, z, NR: int, ray0: int,
                   d_out: torch.Tensor, grads: dict, d_latent_nhwc, novel=None, d_gen_nhwc=None):
    """Backward of field_forward.  Accumulates parameter gradients into `grads` (name -> fp32 tensor shaped like
    the parameter, zero-initialised by the caller) and, if not None, the feature gradient into
    d_latent_nhwc (SB,NV,Hl,Wl,512) fp32.  NOVEL variant (`novel` = the dict field_forward was given): the same feature
    gradient also goes to the learned map, d_gen_nhwc (Hg,Wg,512) fp32, if not None."""
    p, st = _lib.ptr, _lib.stream_ptr()
    dev = d_out.device
    n, K = z.shape
    mv, ms, n_samples, NV = sv.mv, sv.ms, sv.n_samples, sc.NV
    bf = dict(device=dev, dtype=torch.bfloat16)
    f32 = dict(device=dev, dtype=torch.float32)
    G = lambda name: grads[name]

    # ---- head: lin_out + activations ----
    bf16_stream = GRAD_STREAM == "bf16"
    dx = None if bf16_stream else torch.empty(ms * 128 * HID, **f32)           # d(block output), XR (fp32 stream only)
    dx_pk = torch.empty(ms * 128 * HID, **bf)
    draw = torch.zeros(ms * 128 * 128, **bf)
    _lib.call("dnr_head_bwd", p(sv.xa[5]), p(pk.w_out), p(sv.out), p(d_out.contiguous()), n_samples, p(dx), p(dx_pk),
              p(draw), st)
    _wgrad(draw, 128, sv.xa[5], HID, ms, G("lin_out.weight"), 4, HID, G("lin_out.bias"))

    def block_bwd(b, tiles, dx, dx_pk, pooled_dx=None):
        """dx = d x_{b+1} -> returns d x_b' (gradient at the block's input, after its lin_z was added).  With
        `pooled_dx` the fp32 stream entering the block is the POOLED gradient of the view mean: the dgrad epilogue
        reads row (sample, view) as pooled_dx[sample] / NV and writes the per-view stream for the first time.
        With the bf16 gradient stream `dx` is unused: the stream IS dx_pk."""
        _wgrad(dx_pk, HID, sv.net[b], HID, tiles, G(f"blocks.{b}.fc_1.weight"), HID, HID, G(f"blocks.{b}.fc_1.bias"))
        dnet = torch.empty(tiles * 128 * HID, **bf)
        _gemm(dx_pk, pk.wt_1[b], HID, tiles, mask=sv.net[b], act_out=dnet)
        _wgrad(dnet, HID, sv.xa[b], HID, tiles, G(f"blocks.{b}.fc_0.weight"), HID, HID, G(f"blocks.{b}.fc_0.bias"))
        dx_pk2 = torch.empty(tiles * 128 * HID, **bf)
        if bf16_stream:
            _gemm(dnet, pk.wt_0[b], HID, tiles, mask=sv.xa[b], res_in_pk=dx_pk, act_out=dx_pk2)
        elif pooled_dx is not None:
            dx = torch.empty(tiles * 128 * HID, **f32)
            _gemm(dnet, pk.wt_0[b], HID, tiles, mask=sv.xa[b], res_in=pooled_dx, res_out=dx, act_out=dx_pk2,
                  mean_nv=-NV, n_samples=n_samples)
        else:
            _gemm(dnet, pk.wt_0[b], HID, tiles, mask=sv.xa[b], res_in=dx, res_out=dx, act_out=dx_pk2)
        return dx, dx_pk2

    for b in (4, 3):
        dx, dx_pk = block_bwd(b, ms, dx, dx_pk)
    # ---- view mean ----
    # only the bf16 operand image of d x / NV is materialised per view row; with the fp32 stream the per-view fp32 values
    # are read from the pooled gradient by block 2's dgrad epilogue (2 KB per view row less to write and to read back)
    dxv_pk = torch.empty(mv * 128 * HID, **bf)
    if bf16_stream:
        _lib.call("dnr_mean_bwd_pk", p(dx_pk), NV, n_samples, mv, p(dxv_pk), st)
    else:
        _lib.call("dnr_mean_bwd", p(dx), NV, n_samples, mv, None, p(dxv_pk), st)
    pooled, dxv = dx, None
    del dx_pk
    pe = sv.lat_src is not None
    dz_pk = [None, None, None]
    for b in (2, 1, 0):
        dxv, dxv_pk = block_bwd(b, mv, dxv, dxv_pk, pooled_dx=pooled if b == 2 else None)
        dz_pk[b] = dxv_pk
        # (NOVEL_PE: the weight is (512, 518); its first 512 columns are written here, leading dimension 518)
        _wgrad(dxv_pk, HID, sv.lat, HID, mv, G(f"lin_z.{b}.weight"), HID, HID, G(f"lin_z.{b}.bias"))
    if pe:
        _field_backward_pe(sc, pk, sv, rays, z, NR, ray0, dz_pk, grads, d_latent_nhwc, novel, d_gen_nhwc)
        return
    _wgrad(dz_pk[0], HID, sv.code, IN_PAD, mv, G("lin_in.weight"), HID, mlp.d_in, G("lin_in.bias"))
    # ---- feature map ----
    if d_latent_nhwc is not None or d_gen_nhwc is not None:
        dlat = torch.empty(mv * 128 * HID, **f32)
        _gemm(dz_pk[0], pk.wt_z[0], HID, mv, A1=dz_pk[1], B1=pk.wt_z[1], K1=HID, A2=dz_pk[2], B2=pk.wt_z[2], K2=HID,
              res_out=dlat)
        if novel is not None:
            g = novel
            _lib.call("dnr_encode_novel_bwd", p(rays), p(z), p(g["pts_in"]), p(g["pts_gen"]), p(sc.poses), p(sc.focal),
                      p(sc.cxy), p(sc.depth), p(g["poses"]), p(g["focal"]), p(g["cxy"]), g["Hg"], g["Wg"], g["img_w"],
                      g["img_h"], g["sx"], g["sy"], p(sc.freqs), p(sc.phases), *_encode_args(sc, NR, K, ray0, n), p(dlat),
                      None, p(d_latent_nhwc), p(d_gen_nhwc), st)
            return
        _lib.call("dnr_encode_bwd", p(rays), p(z), p(sc.poses), p(sc.focal), p(sc.cxy), p(sc.depth), p(sc.freqs),
                  p(sc.phases), *_encode_args(sc, NR, K, ray0, n), p(dlat), p(d_latent_nhwc), st)


def _field_backward_pe(sc, pk, sv, rays, z, NR, ray0, dz_pk, grads, d_latent_nhwc, novel, d_gen_nhwc):
    """Tail of field_backward for the NOVEL_PE field (pe_novel_pixelnerf.py:260-275 under autograd): the code-side columns
    of lin_in / lin_z / deformation_layer, the deformation layer itself, and the two feature-map gradients."""
    p, st = _lib.ptr, _lib.stream_ptr()
    dev = dz_pk[0].device
    n, K = z.shape
    mv, d_in = sv.mv, pk.d_in
    f32 = dict(device=dev, dtype=torch.float32)
    bf = dict(device=dev, dtype=torch.bfloat16)
    G = lambda name: grads[name]
    # code operand columns: [code d_in | pe 6 | 0]; one 512 x 64 weight gradient per consumer
    tmp = torch.zeros(4, HID, IN_PAD, **f32)
    for b in range(3):
        _wgrad(dz_pk[b], HID, sv.code, IN_PAD, mv, tmp[b], HID, d_in + 6, G("lin_in.bias") if b == 0 else None)
    # d final_latent = sum_b lin_z[b][:, :512]^T dz_b: fp32 rows for the gen_latent scatter, bf16 operand for what follows
    dfin = torch.empty(mv * 128 * HID, **f32) if d_gen_nhwc is not None else None
    dfin_pk = torch.empty(mv * 128 * HID, **bf)
    _gemm(dz_pk[0], pk.wt_z[0], HID, mv, A1=dz_pk[1], B1=pk.wt_z[1], K1=HID, A2=dz_pk[2], B2=pk.wt_z[2], K2=HID,
          res_out=dfin, act_out=dfin_pk)
    # deformation_layer: conditioned = W_d [feature | pe 6] + b_d, final_latent = gen + conditioned   (:265-268)
    _wgrad(dfin_pk, HID, sv.lat_src, HID, mv, G("deformation_layer.weight"), HID, HID, G("deformation_layer.bias"))
    _wgrad(dfin_pk, HID, sv.code, IN_PAD, mv, tmp[3], HID, d_in + 6, None)
    G("lin_in.weight").add_(tmp[0, :, :d_in])
    for b in range(3):
        G(f"lin_z.{b}.weight")[:, HID:].add_(tmp[b, :, d_in:d_in + 6])
    G("deformation_layer.weight")[:, HID:].add_(tmp[3, :, d_in:d_in + 6])
    if d_latent_nhwc is None and d_gen_nhwc is None:
        return
    dlat = None
    if d_latent_nhwc is not None:
        dlat = torch.empty(mv * 128 * HID, **f32)
        _gemm(dfin_pk, pk.wt_d, HID, mv, res_out=dlat)
    g = novel
    _lib.call("dnr_encode_novel_bwd", p(rays), p(z), p(g["pts_in"]), p(g["pts_gen"]), p(sc.poses), p(sc.focal), p(sc.cxy),
              p(sc.depth), p(g["poses"]), p(g["focal"]), p(g["cxy"]), g["Hg"], g["Wg"], g["img_w"], g["img_h"], g["sx"],
              g["sy"], p(sc.freqs), p(sc.phases), *_encode_args(sc, NR, K, ray0, n), p(dlat), p(dfin), p(d_latent_nhwc),
              p(d_gen_nhwc), st)
