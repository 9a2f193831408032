"""CPU tests: the C-ABI library loads and exports what include/diner_b200.h declares, host-side validation
and error behaviour, ray sharding over a 2-rank gloo group.  No kernels are launched here."""
import ctypes
import os
import re
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_decls():
    src = open(os.path.join(ROOT, "include", "diner_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    decls = {}
    for m in re.finditer(r"(?:int|long long|const char\*)\s+(dnr_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.S):
        args = m.group(2).strip()
        n = 0 if args in ("", "void") else len([a for a in args.split(",") if a.strip()])
        decls[m.group(1)] = n
    return decls


def test_library_builds_and_exports_every_declared_symbol():
    from diner_b200 import _build, _lib
    path = _build.build()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)
    decls = _header_decls()
    assert len(decls) >= 20
    for name in decls:
        assert hasattr(lib, name), f"{name} declared in include/diner_b200.h but not exported"
    # the ctypes table of the host layer binds exactly the declared entry points, with the declared arity
    assert set(_lib.EXPORTED_SYMBOLS) == set(decls)
    for name, (_, argtypes) in _lib._SIGNATURES.items():
        assert len(argtypes) == decls[name], f"{name}: ctypes arity {len(argtypes)} != header {decls[name]}"
    lib.dnr_abi_version.restype = ctypes.c_int
    assert lib.dnr_abi_version() == _lib.ABI_VERSION


def test_argument_validation_needs_no_gpu():
    """Entry points reject bad arguments before touching the device (cudaErrorInvalidValue = 1)."""
    from diner_b200 import _lib
    lib = _lib.load()
    assert lib.dnr_view_row_tiles(1000, 2) == (1000 + 63) // 64
    assert lib.dnr_view_row_tiles(1000, 3) == (1000 + 39) // 40
    assert lib.dnr_view_row_tiles(10, 9) == -1
    rc = lib.dnr_composite_fwd(None, None, None, 10, 0, 1, None, None, None, None)
    assert rc == 1 and b"K must be" in lib.dnr_last_error()
    rc = lib.dnr_mlp_gemm(None, None, 64, None, None, 0, None, None, 0, 1, 100, None, None, None, None, None, None, None, 0,
                          0, 0, 0, None)
    assert rc == 1 and b"multiple of 256" in lib.dnr_last_error()
    with pytest.raises(RuntimeError, match="dnr_mlp_wgrad failed"):
        _lib.call("dnr_mlp_wgrad", None, 100, None, 64, 1, None, 64, 1, 1, None, None)
    # the single-call render: workspace planning is host arithmetic, argument checks come before any device call
    ws = lib.dnr_render_workspace_bytes(2, 65536, 40, 0)
    rows = 32768 * 40 * 2                                   # two balanced chunks of 32 768 rays (2^22-row limit)
    assert ws >= rows * (1024 + 128) + 65536 * 40 * 16 and ws % 256 == 0
    assert lib.dnr_render_workspace_bytes(2, 65536, 40, 1 << 20) < ws          # smaller chunks, smaller operand buffers
    assert lib.dnr_render_workspace_bytes(5, 65536, 40, 0) == -1 and lib.dnr_render_workspace_bytes(2, -1, 40, 0) == -1
    rc = lib.dnr_render_fwd(None, None, None, 10, 40, 1000, 15, 0, None, 1, 0, None, 0, None, None, None, None, None, None)
    assert rc == 1 and b"null argument" in lib.dnr_last_error()
    sd, md = _lib.SceneDesc(SB=1, NV=7), _lib.MlpDesc()
    dummy = ctypes.c_float(0)
    rc = lib.dnr_render_fwd(ctypes.byref(sd), ctypes.byref(md), ctypes.byref(dummy), 10, 40, 1000, 15, 0, None, 1, 0, None, 0,
                            ctypes.byref(dummy), ctypes.byref(dummy), None, None, None, None)
    assert rc == 1 and b"NV must be in [1,4]" in lib.dnr_last_error()
    sd.NV = 2
    rc = lib.dnr_render_fwd(ctypes.byref(sd), ctypes.byref(md), ctypes.byref(dummy), 10, 40, 1000, 15, 0, None, 1, 0, None, 0,
                            ctypes.byref(dummy), ctypes.byref(dummy), None, None, None, None)
    assert rc == 1 and b"workspace too small" in lib.dnr_last_error()


def test_renderer_signature_and_errors():
    import inspect
    from diner_b200 import NeRFRendererDGS, RenderResult
    from diner_b200.scene import validate_model
    from tools.synthetic import make_scene
    sig = inspect.signature(NeRFRendererDGS.__init__)
    # reference constructor (src/models/nerf_renderer.py:23-30): same names, same defaults
    want = dict(n_samples=40, n_depth_candidates=1000, n_gaussian=15, eval_batch_size=100000, white_bkgd=True)
    for k, v in want.items():
        assert sig.parameters[k].default == v
    fsig = inspect.signature(NeRFRendererDGS.forward)
    assert list(fsig.parameters)[:4] == ["self", "model", "rays", "want_weights"]
    r = NeRFRendererDGS()
    assert isinstance(r, torch.nn.Module) and len(list(r.parameters())) == 0 and len(list(r.buffers())) == 0
    scene = make_scene("tiny", calibrate=False)
    with pytest.raises(AssertionError):
        r(scene.model, scene.rays[0])                        # rays must be 3-D (nerf_renderer.py:412)
    with pytest.raises(RuntimeError, match="no CPU path"):
        r(scene.model, scene.rays)                           # CPU tensors: fail loudly, never fall back
    validate_model(scene.model)
    scene.model.mlp_fine.combine_type = "max"
    with pytest.raises(NotImplementedError):
        validate_model(scene.model)
    scene.model.mlp_fine.combine_type = "average"
    scene.model.mlp_fine.activation = torch.nn.Softplus()
    with pytest.raises(NotImplementedError):
        validate_model(scene.model)
    out = RenderResult(fine=RenderResult(rgb=1))
    assert out.fine.rgb == 1


def test_shard_range_is_a_balanced_partition():
    from diner_b200.sharding import shard_range
    for n in (0, 1, 7, 4096, 65536, 120000):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


class _StubRenderer:
    """Stands in for the CUDA renderer in the gloo test: a pure function of each ray, so sharding is checkable."""

    def __call__(self, model, rays, **kw):
        from diner_b200 import RenderResult
        rgb = torch.stack([rays[..., 0] + rays[..., 3], rays[..., 1] * 2, rays[..., 7] - rays[..., 6]], -1)
        return RenderResult(fine=RenderResult(rgb=rgb, depth=rays[..., :3].sum(-1)))


def _worker(rank, world, port, q):
    import types
    from diner_b200.sharding import allreduce_gradients, broadcast_scene, render_sharded
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    torch.manual_seed(0)
    rays = torch.randn(2, 1001, 8)                            # not divisible by the world size
    rgb, depth = render_sharded(_StubRenderer(), None, rays)
    full = _StubRenderer()(None, rays)
    ok = torch.equal(rgb, full.fine.rgb) and torch.equal(depth, full.fine.depth)
    # scene broadcast: rank 1 starts with garbage and must end up with rank 0's tensors
    torch.manual_seed(rank)
    enc = types.SimpleNamespace(latent=torch.randn(1, 2, 8, 4, 4), depths=torch.randn(1, 2, 1, 4, 4),
                                depths_std=torch.randn(1, 2, 1, 4, 4), normals=torch.randn(1, 2, 3, 4, 4))
    model = types.SimpleNamespace(encoder=enc, poses=torch.randn(1, 2, 4, 4), focal=torch.randn(1, 2, 2),
                                  c=torch.randn(1, 2, 2), image_shape=torch.randn(2), mlp_fine=torch.nn.Linear(4, 4))
    from diner_b200 import field as F_, scene as S_
    S_._cache[123] = "stale scene repack"                      # c10d broadcasts / .data writes do not bump _version:
    F_._packed_cache[456] = "stale packed weights"             # broadcast_scene must drop the renderer's caches itself
    w_before = model.mlp_fine.weight.detach().clone()
    broadcast_scene(model, src=0)
    ok = ok and len(S_._cache) == 0 and len(F_._packed_cache) == 0
    torch.manual_seed(0)
    want = dict(latent=torch.randn(1, 2, 8, 4, 4), depths=torch.randn(1, 2, 1, 4, 4), depths_std=torch.randn(1, 2, 1, 4, 4),
                normals=torch.randn(1, 2, 3, 4, 4))
    for k, v in want.items():                                   # every tensor of the flat bucket lands in its slot
        ok = ok and torch.equal(getattr(model.encoder, k), v)
    want_m = dict(poses=torch.randn(1, 2, 4, 4), focal=torch.randn(1, 2, 2), c=torch.randn(1, 2, 2), image_shape=torch.randn(2))
    for k, v in want_m.items():
        ok = ok and torch.equal(getattr(model, k), v)
    w0 = torch.nn.Linear(4, 4).weight.detach()                  # rank 0's Linear, drawn right after the tensors above
    ok = ok and torch.equal(model.mlp_fine.weight.detach(), w0) and (rank == 0 or not torch.equal(w_before, w0))
    # gradient all-reduce (mean)
    p = torch.nn.Parameter(torch.zeros(3)); p.grad = torch.full((3,), float(rank + 1))
    allreduce_gradients([p])
    ok = ok and torch.allclose(p.grad, torch.full((3,), (1 + world) / 2))
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_sharded_render_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]


def test_patch_selection_host_logic_matches_loop_restatement():
    """callers.patch_pixels is torch index arithmetic (no kernel): check it on the CPU against the loop form of
    diner.py:231-247 in oracle/producers_oracle.py."""
    import numpy as np
    from diner_b200.callers import patch_pixels, patch_pixels_at, random_pixels
    from oracle import producers_oracle as P
    SB, H, W = 4, 40, 56
    torch.manual_seed(0)
    alpha = (torch.rand(SB, 1, H, W) > 0.5).float()
    for patch in (1, 4, 9, 16):
        idx, centres = patch_pixels(alpha, patch, generator=torch.Generator().manual_seed(patch))
        flat = (centres[:, 0] + centres[:, 1] * W).numpy()
        assert np.array_equal(idx.numpy(), P.patch_pixel_indices(flat, W, patch))
        assert np.array_equal(patch_pixels_at(torch.from_numpy(flat), W, patch).numpy(), idx.numpy())
        elig = P.eligible_patch_centres(alpha.numpy(), patch)
        assert all(elig[s, int(centres[s, 1]), int(centres[s, 0])] > 0 for s in range(SB))
        assert int(idx.min()) >= 0 and int(idx.max()) < H * W
    pix = random_pixels(SB, H, W, 128, generator=torch.Generator().manual_seed(1))
    assert pix.shape == (SB, 128) and int(pix.min()) >= 0 and int(pix.max()) < H * W


def test_numpy_philox_known_answers():
    """Random123's kat_vectors for philox4x32-10: all-zero and all-one counter/key."""
    from oracle.philox import philox4x32_10
    assert [int(x) for x in philox4x32_10(0, 0, 0, 0, 0, 0)] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    f = 0xffffffff
    assert [int(x) for x in philox4x32_10(f, f, f, f, f, f)] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
