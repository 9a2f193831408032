"""Allocation-free entry point `dnr_mlp_quad_fwd_static` (no counterpart in
the reference) against the existing `dnr_mlp_quad_fwd`, through the C ABI, on calls in a row -- run with `-m gpu`.

Collected last on purpose (file name): a failure here hides no other test.  The test
is self-contained: it passes alone in a fresh process and after the other tests (the library's static scratch may then
already hold another call's contents -- nothing in it has to start at zero).
"""
import dataclasses

import pytest
import torch

pytestmark = pytest.mark.gpu


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from diner_b200 import _lib
    _lib.load()
    return torch.device("cuda", 0)


def _operands(NV, n_rays, dev, seed):
    """The MLP's packed operands for n_rays rays of a small synthetic scene with NV source views (dnr_encode_fwd)."""
    from diner_b200 import _lib, field as F_
    from diner_b200.scene import get_scene
    from tools.synthetic import WORKLOADS, make_scene
    wl = dataclasses.replace(WORKLOADS["tiny"], name=f"noalloc_nv{NV}", NV=NV, tgt_H=24, tgt_W=24)
    scene = make_scene(wl, seed=seed)
    K = wl.n_samples
    m = scene.model.to_device(dev)
    rays = scene.rays[:, :n_rays].contiguous().to(dev)
    gen = torch.Generator().manual_seed(seed)
    u = torch.rand(1, n_rays, K, generator=gen).sort(-1).values.to(dev)
    z = (rays[..., 6:7] + (rays[..., 7:8] - rays[..., 6:7]) * u).view(-1, K).contiguous()
    sc = get_scene(m)
    pk = F_.get_packed(m.mlp_fine, need_transposed=False)
    n_samples = n_rays * K
    mv = _lib.load().dnr_view_row_tiles(n_samples, NV)
    lat = torch.empty(mv * 128 * 512, device=dev, dtype=torch.bfloat16)
    code = torch.empty(mv * 128 * 64, device=dev, dtype=torch.bfloat16)
    p, st = _lib.ptr, _lib.stream_ptr()
    _lib.call("dnr_encode_fwd", p(rays.view(-1, 8)), p(z), p(sc.poses), p(sc.focal), p(sc.cxy), p(sc.depth),
              p(sc.latent_nhwc), p(sc.freqs), p(sc.phases), *F_._encode_args(sc, n_rays, K, 0, n_rays), p(lat), p(code), st)
    return dict(m=m, pk=pk, lat=lat, code=code, mv=mv, n_samples=n_samples, NV=NV)


def _finalize(part, op, dev):
    from diner_b200 import _lib
    out = torch.empty(op["n_samples"], 4, device=dev, dtype=torch.float32)
    _lib.call("dnr_head_finalize", _lib.ptr(part), _lib.ptr(op["pk"].b_out), op["NV"], op["n_samples"], 4, _lib.ptr(out),
              _lib.stream_ptr())
    return out


def test_static_scratch_entry_matches_existing_entry_on_calls_in_a_row():
    dev = _cuda()
    from diner_b200 import _lib, field as F_
    p = _lib.ptr
    lib = _lib.load()
    assert lib.dnr_mlp_quad_static_max_tiles() == 4096
    # calls in a row with different view counts and batch sizes: the parked-row layout depends on NV, so every call
    # finds the previous call's rows in the shared scratch (333 rays x NV=2 = 209 tiles: many merge groups per cluster;
    # 130 rays x NV=3: groups of three, a last group that ends early; 37 rays x NV=2: fewer tile pairs than clusters)
    cases = [(2, 333, 1), (3, 130, 2), (2, 37, 3), (4, 96, 4), (2, 333, 1)]
    outs = []
    for NV, n_rays, seed in cases:
        op = _operands(NV, n_rays, dev, seed)
        pk, mv = op["pk"], op["mv"]
        # existing entry point, merge scratch allocated by the caller
        want_part = torch.empty(mv * 4 * 128, 4, device=dev, dtype=torch.float32)
        _lib.call("dnr_mlp_quad_fwd", p(op["lat"]), p(op["code"]), pk.w128_ptrs, pk.pair_b_ptrs, p(pk.w_out), NV, mv,
                  p(F_._quad_scratch(dev)), p(want_part), None, _lib.stream_ptr())
        want = _finalize(want_part, op, dev)
        # new entry point: outputs are the caller's, the scratch is the library's; nothing may be allocated by the call
        got_part = torch.empty(mv * 4 * 128, 4, device=dev, dtype=torch.float32)
        torch.cuda.synchronize()
        alloc0, free0 = torch.cuda.memory_allocated(dev), torch.cuda.mem_get_info(dev)[0]
        _lib.call("dnr_mlp_quad_fwd_static", p(op["lat"]), p(op["code"]), pk.w128_ptrs, pk.pair_b_ptrs, p(pk.w_out), NV, mv,
                  p(got_part), _lib.stream_ptr())
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == alloc0
        if outs:                                   # (the very first launch of a kernel may load its code lazily)
            assert torch.cuda.mem_get_info(dev)[0] >= free0
        got = _finalize(got_part, op, dev)
        torch.cuda.synchronize()
        assert torch.isfinite(got).all() and got.shape == (op["n_samples"], 4)
        # the tolerance the tests use between the cluster kernels (tests/test_gpu_field.py: torch.equal(quad, pair)):
        # every row goes through the same MMAs in the same order whatever the number of clusters
        assert torch.equal(got, want), (NV, n_rays, float((got - want).abs().max()))
        outs.append(got.clone())
        op["m"].to_device("cpu")
    assert torch.equal(outs[0], outs[-1])          # same inputs, other calls in between: same bits
    assert float(outs[0][:, 3].max()) > 0          # live outputs, not zeros


def test_static_scratch_entry_rejects_sizes_beyond_its_maximum():
    _cuda()
    from diner_b200 import _lib
    lib = _lib.load()
    big = lib.dnr_mlp_quad_static_max_tiles() + 1
    dummy = torch.zeros(16, device="cuda")
    import ctypes as C
    ptrs = (C.c_void_p * 14)(*[dummy.data_ptr()] * 14)
    with pytest.raises(RuntimeError, match="exceeds the maximum of 4096 tiles"):
        _lib.call("dnr_mlp_quad_fwd_static", _lib.ptr(dummy), _lib.ptr(dummy), ptrs, ptrs, _lib.ptr(dummy), 2, big,
                  _lib.ptr(dummy), _lib.stream_ptr())
    with pytest.raises(RuntimeError, match="NV must be"):
        _lib.call("dnr_mlp_quad_fwd_static", _lib.ptr(dummy), _lib.ptr(dummy), ptrs, ptrs, _lib.ptr(dummy), 5, 4,
                  _lib.ptr(dummy), _lib.stream_ptr())
