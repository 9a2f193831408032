"""ctypes loader for the C-ABI CUDA library (include/diner_b200.h).

There is no CPU fallback: if the library is missing it is built in-tree with nvcc; if that fails, or a
kernel returns a CUDA error, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os

from . import _build

_lib = None
ABI_VERSION = 19

# instrumentation used by bench.py: number of kernel launches issued through the ABI, and (when `profile` is a
# list) one (name, start_event, end_event, tag) record per call for per-kernel CUDA-event timing
launch_count = 0
profile = None

c_p = C.c_void_p
c_i = C.c_int
c_ll = C.c_longlong
c_f = C.c_float
c_ull = C.c_ulonglong



class SceneDesc(C.Structure):
    """dnr_scene_desc of include/diner_b200.h."""
    _fields_ = ([(n, c_p) for n in ("poses", "focal", "cxy", "depth", "depth_std_padded", "normal", "latent_nhwc_bf16",
                                   "freqs", "phases", "lin_tab")]
                + [(n, c_i) for n in ("SB", "NV", "H", "W", "Hl", "Wl", "L", "NF", "pad")]
                + [(n, c_f) for n in ("img_w", "img_h", "std_sx", "std_sy", "lat_sx", "lat_sy", "depth_diff_max")])


class MlpDesc(C.Structure):
    """dnr_mlp_desc of include/diner_b200.h."""
    _fields_ = [(n, c_p) for n in ("w128", "w256", "pair_biases", "w_out", "b_out")]


_SIGNATURES = {
    "dnr_abi_version": (c_i, []),
    "dnr_last_error": (C.c_char_p, []),
    "dnr_composite_fwd": (c_i, [c_p, c_p, c_p, c_i, c_i, c_i, c_p, c_p, c_p, c_p]),
    "dnr_composite_bwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_p, c_p]),
    "dnr_gen_rays": (c_i, [c_p, c_p, c_i, c_i, c_i, c_p, c_p, c_p, c_p]),
    "dnr_gen_rays_at": (c_i, [c_p, c_p, c_i, c_i, c_i, c_p, c_p, c_p, c_i, c_p, c_p, c_p]),
    "dnr_depth2normal": (c_i, [c_p, c_p, c_i, c_i, c_i, c_p, c_p]),
    "dnr_decode_depth_u16": (c_i, [c_p, c_p, c_ll, c_i, c_p, c_i, c_i, c_i, c_i] + [c_f] * 6 + [c_i, c_p, c_p, c_p]),
    "dnr_deform_points": (c_i, [c_p, c_p, c_p, c_i, c_ll, c_i, c_p, c_p, c_p, c_p]),
    "dnr_pad_depth_std": (c_i, [c_p, c_p, c_i, c_i, c_i, c_i, c_p, c_p]),
    "dnr_sampler_fwd": (c_i, [c_p] * 14 + [c_ull] + [c_i] * 9 + [c_f] * 5 + [c_i, c_i] + [c_p] * 6),
    "dnr_view_row_tiles": (c_i, [c_ll, c_i]),
    "dnr_encode_fwd": (c_i, [c_p] * 9 + [c_i] * 10 + [c_ll, c_ll] + [c_f] * 4 + [c_p, c_p, c_p]),
    "dnr_encode_fwd_split": (c_i, [c_p] * 10 + [c_i] * 10 + [c_ll, c_ll] + [c_f] * 4 + [c_p] * 5),
    "dnr_encode_novel_fwd": (c_i, [c_p] * 13 + [c_i, c_i] + [c_f] * 4 + [c_p, c_p] + [c_i] * 10 + [c_ll, c_ll] + [c_f] * 4
                             + [c_p] * 3),
    "dnr_encode_novel_bwd": (c_i, [c_p] * 11 + [c_i, c_i] + [c_f] * 4 + [c_p, c_p] + [c_i] * 10 + [c_ll, c_ll] + [c_f] * 4
                             + [c_p] * 5),
    "dnr_encode_novel_pe_fwd": (c_i, [c_p] * 13 + [c_i, c_i] + [c_f] * 4 + [c_p, c_i, c_i, c_f, c_f] + [c_p] * 3
                                + [c_f, c_f] + [c_p, c_i, c_i, c_f, c_f] + [c_p, c_p] + [c_i] * 10 + [c_ll, c_ll]
                                + [c_f] * 4 + [c_p] * 4),
    "dnr_novel_deform": (c_i, [c_p] * 4 + [c_ull] + [c_p] * 6 + [c_i] * 6 + [c_p] * 4),
    "dnr_encode_bwd": (c_i, [c_p] * 8 + [c_i] * 10 + [c_ll, c_ll] + [c_f] * 4 + [c_p, c_p, c_p]),
    "dnr_pack_latent_nhwc": (c_i, [c_p, c_i, c_i, c_i, c_i, c_p, c_p]),
    "dnr_unpack_dlatent_nchw": (c_i, [c_p, c_i, c_i, c_i, c_i, c_p, c_p]),
    "dnr_mlp_gemm": (c_i, [c_p, c_p, c_i, c_p, c_p, c_i, c_p, c_p, c_i, c_i, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i,
                           c_ll, c_i, c_p]),
    "dnr_mlp_wgrad": (c_i, [c_p, c_i, c_p, c_i, c_i, c_p, c_i, c_i, c_i, c_p, c_p]),
    "dnr_pack_weight": (c_i, [c_p, c_i, c_i, c_i, c_i, c_i, c_i, c_p, c_p]),
    "dnr_head_fwd": (c_i, [c_p, c_p, c_p, c_p, c_ll, c_i, c_p, c_p]),
    "dnr_mlp_fused_scratch_bytes": (c_ll, [c_i]),
    "dnr_mlp_pair_fwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_p, c_p, c_p]),
    "dnr_mlp_quad_max_clusters": (c_i, [c_i]),
    "dnr_mlp_quad_scratch_bytes": (c_ll, []),
    "dnr_mlp_quad_fwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_p, c_p, c_p, c_p]),
    "dnr_mlp_pair_fwd_pe": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_p, c_p]),
    "dnr_mlp_quad_fwd_pe": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_p, c_p, c_p]),
    "dnr_mlp_quad_static_max_tiles": (c_i, []),
    "dnr_mlp_quad_fwd_static": (c_i, [c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_p, c_p]),
    "dnr_head_finalize": (c_i, [c_p, c_p, c_i, c_ll, c_i, c_p, c_p]),
    "dnr_mlp_fused_fwd": (c_i, [c_p, c_p, c_p, c_p, c_i, c_i, c_ll, c_p, c_p, c_i, c_p]),
    "dnr_head_bwd": (c_i, [c_p, c_p, c_p, c_p, c_ll, c_p, c_p, c_p, c_p]),
    "dnr_mean_bwd": (c_i, [c_p, c_i, c_ll, c_i, c_p, c_p, c_p]),
    "dnr_mean_bwd_pk": (c_i, [c_p, c_i, c_ll, c_i, c_p, c_p]),
    "dnr_render_workspace_bytes": (c_ll, [c_i, c_ll, c_i, c_ll]),
    "dnr_render_fwd": (c_i, [c_p, c_p, c_p, c_ll, c_i, c_i, c_i, c_ull, c_p, c_i, c_ll, c_p, c_ll, c_p, c_p, c_p, c_p, c_p,
                             c_p]),
    "dnr_render_timings": (c_i, [c_p, c_p, c_p]),
    "dnr_pack_rows": (c_i, [c_p, c_ll, c_i, c_ll, c_p, c_p]),
    "dnr_unpack_rows": (c_i, [c_p, c_ll, c_i, c_p, c_p]),
    "dnr_xr_to_rows": (c_i, [c_p, c_ll, c_i, c_p, c_p]),
    "dnr_rows_to_xr": (c_i, [c_p, c_ll, c_i, c_ll, c_p, c_p]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)


def lib_path() -> str:
    return _build.LIB_PATH


def load(build_if_missing: bool = True):
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    if build_if_missing:
        _build.build()          # no-op when the library is newer than every source; file-locked against other ranks
    elif not os.path.exists(path):
        raise RuntimeError(f"{path} is missing and building was disabled; diner_b200 has no CPU fallback")
    lib = C.CDLL(path)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.dnr_abi_version() != ABI_VERSION:
        raise RuntimeError(f"{path} has ABI version {lib.dnr_abi_version()}, expected {ABI_VERSION}: rebuild it "
                           "(python -m diner_b200._build --force)")
    _lib = lib
    return lib


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def call(name: str, *args, tag=None):
    global launch_count
    lib = load()
    launch_count += 1
    if profile is not None:
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = getattr(lib, name)(*args)
        e1.record()
        profile.append((name, e0, e1, tag))
    else:
        rc = getattr(lib, name)(*args)
    if rc != 0:
        msg = lib.dnr_last_error()
        raise RuntimeError(f"{name} failed (code {rc}): {msg.decode() if msg else '?'}")


def stream_ptr():
    """Current stream of the current device (NeRFRendererDGS.forward and the autograd backwards run inside
    `torch.cuda.device(rays.device)`, so this is the stream of the device the tensors live on)."""
    import torch
    return torch.cuda.current_stream().cuda_stream
