"""CPU restatement of the two per-scene / per-frame INPUT PRODUCERS next to the hot path:

  * gen_rays      -- reference src/util/cam_geometry.py:36-79  (pixel-centre rays of a target camera)
  * depth2normal  -- reference src/util/depth2normal.py:6-87   (camera-space normals by central differences,
                     with the reference's hole "cleaning" rule)

Plain numpy, one rounding per reference op, loops only where the reference uses advanced indexing.  TEST
INFRASTRUCTURE: imported by tests/ only; the product path (diner_b200/producers.py) never touches it.

Parity pinning: tests/golden/producers.npz holds the outputs of the UNMODIFIED reference functions on seeded inputs
(oracle/make_golden_producers.py); tests/test_oracle.py requires this restatement to reproduce them (rays: bit-exact
except the 3x3 rotation, whose accumulation order belongs to the BLAS -> 1e-6; normals: 1e-6, NaN pattern equal).
"""
from __future__ import annotations

import numpy as np

f32 = np.float32


def gen_rays(extrinsics, intrinsics, W, H, z_near, z_far):
    """(B,4,4) world->cam, (B,3,3) -> (B,H,W,8) = [origin 3 | unit direction 3 | near | far]; cam_geometry.py:36-79."""
    E = np.asarray(extrinsics, f32)
    K = np.asarray(intrinsics, f32)
    B = E.shape[0]
    fx, fy = K[:, 0, 0], K[:, 1, 1]                      # :54
    cx, cy = K[:, 0, 2], K[:, 1, 2]                      # :55
    xs = np.arange(0.5, W, 1.0, dtype=f32)               # pixel centres, x fastest (:57-59)
    ys = np.arange(0.5, H, 1.0, dtype=f32)
    out = np.empty((B, H, W, 8), f32)
    for b in range(B):
        px = ((xs - cx[b]) / fx[b]).astype(f32)[None, :].repeat(H, 0)          # :61
        py = ((ys - cy[b]) / fy[b]).astype(f32)[:, None].repeat(W, 1)
        one = np.ones_like(px)
        nrm = np.sqrt(((px * px + py * py).astype(f32) + one * one).astype(f32))   # pow(2).sum(-1).sqrt()  :63
        d_cam = np.stack((px / nrm, py / nrm, one / nrm), -1).astype(f32)
        R_c2w = E[b, :3, :3].T                                                  # :66
        d_w = (d_cam.reshape(-1, 3).astype(np.float64) @ R_c2w.T.astype(np.float64)).astype(f32)   # :67 (BLAS order free)
        center = (-(R_c2w.astype(np.float64) @ E[b, :3, 3].astype(np.float64))).astype(f32)        # :70
        out[b, ..., 0:3] = center
        out[b, ..., 3:6] = d_w.reshape(H, W, 3)
        out[b, ..., 6] = f32(z_near[b])
        out[b, ..., 7] = f32(z_far[b])
    return out


def depth2normal(dmap, K):
    """(N,1,H,W) depth, (N,3,3) -> (N,3,H,W) normals; depth2normal.py:6-87."""
    d = np.asarray(dmap, f32)
    K = np.asarray(K, f32)
    N, _, H, W = d.shape
    xs = np.arange(0.5, W, 1.0, dtype=f32)
    ys = np.arange(0.5, H, 1.0, dtype=f32)
    out = np.empty((N, H, W, 3), f32)
    if H == 3 or W == 3:
        raise NotImplementedError("depth2normal: H == 3 or W == 3 changes the axis of the reference's torch.cross")
    with np.errstate(invalid="ignore", divide="ignore"):
        vs, hs, downs, ups, rights, lefts = [], [], [], [], [], []
        for n in range(N):
            rx = ((xs - K[n, 0, 2]) / K[n, 0, 0]).astype(f32)[None, :].repeat(H, 0)      # :27-28
            ry = ((ys - K[n, 1, 2]) / K[n, 1, 1]).astype(f32)[:, None].repeat(W, 1)
            pts = np.stack((rx * d[n, 0], ry * d[n, 0], d[n, 0]), 0).astype(f32)          # :30  (3,H,W)
            p = np.pad(pts, ((0, 0), (1, 1), (1, 1)), mode="edge")                         # :32
            down, up = p[:, 2:, 1:-1], p[:, :-2, 1:-1]                                     # :46-49
            right, left = p[:, 1:-1, 2:], p[:, 1:-1, :-2]
            vs.append((down - up).astype(f32).transpose(1, 2, 0))                          # (H,W,3)
            hs.append((right - left).astype(f32).transpose(1, 2, 0))
            downs.append(down[0]); ups.append(up[0]); rights.append(right[0]); lefts.append(left[0])
        v, h = np.stack(vs), np.stack(hs)                                                   # (N,H,W,3)
        # torch.cross WITHOUT dim (:53) takes the FIRST axis of size 3.  For N == 3 (e.g. SB=1, NV=3) that is the
        # batch axis: the reference crosses the views with each other, element by element.  Restated as is.
        ax = 0 if N == 3 else 3
        a1, a2 = np.roll(v, -1, ax), np.roll(v, -2, ax)
        b1, b2 = np.roll(h, -1, ax), np.roll(h, -2, ax)
        cr = ((a1 * b2).astype(f32) - (a2 * b1).astype(f32)).astype(f32)
        nrm = np.sqrt((((cr[..., 0] * cr[..., 0]).astype(f32) + (cr[..., 1] * cr[..., 1]).astype(f32)).astype(f32)
                       + (cr[..., 2] * cr[..., 2]).astype(f32)).astype(f32))
        raw_all = (cr / nrm[..., None]).astype(f32)                                         # :54
        yy, xx = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
        for n in range(N):
            raw = raw_all[n]
            # cleaning (:57-77): a pixel with a hole (x == 0) below / above / right / left takes the raw normal of the
            # pixel one step AWAY from the hole (offsets accumulate, then clip)
            oy = (-(downs[n] == 0).astype(np.int64)) + (ups[n] == 0).astype(np.int64)
            ox = (-(rights[n] == 0).astype(np.int64)) + (lefts[n] == 0).astype(np.int64)
            sy = np.clip(yy + oy, 0, H - 1)
            sx = np.clip(xx + ox, 0, W - 1)
            move = (oy != 0) | (ox != 0)
            res = raw.copy()
            res[move] = raw[sy[move], sx[move]]                                             # :77 (gathers the raw map)
            res[d[n, 0] == 0] = 0                                                           # :78
            out[n] = res
    return out.transpose(0, 3, 1, 2).copy()                                                 # :84


# ---- caller-side ray handling (reference src/models/diner.py; not importable here: pytorch_lightning is absent, so
# ---- these two are restated from the cited lines and "parity unpinned" by a run of the reference) -----------------

def rays_at(extrinsics, intrinsics, W, H, z_near, z_far, pix_idcs):
    """rays.view(SB, H*W, 8)[arange(SB)[:, None], pix_idcs] -- diner.py:256-257."""
    rays = gen_rays(extrinsics, intrinsics, W, H, z_near, z_far).reshape(len(extrinsics), H * W, 8)
    pix = np.asarray(pix_idcs, np.int64)
    return rays[np.arange(rays.shape[0])[:, None], pix]


def patch_pixel_indices(centers, W, patch):
    """Flat pixel indices of the patch around each centre -- diner.py:234,238-247, loop form.
    centers: (SB,) flat indices as torch.multinomial returns them."""
    pad = (patch + 1) // 2                                   # :234
    out = np.empty((len(centers), patch * patch), np.int64)
    for n, c in enumerate(np.asarray(centers, np.int64).reshape(-1)):
        cx, cy = c % W, c // W                               # :239
        i = 0
        for h in range(patch):                               # pix_coords[h, w] = (x=w, y=h) - pad  (:240-243)
            for w in range(patch):
                x, y = cx + (w - pad), cy + (h - pad)        # :244
                out[n, i] = x + y * W                        # :245
                i += 1
    return out


def eligible_patch_centres(alpha, patch):
    """The mask torch.multinomial draws from (:232-237): foreground minus a border of pad pixels."""
    fg = np.array(alpha[:, 0], np.float32, copy=True)
    pad = (patch + 1) // 2
    fg[..., :pad] = 0
    fg[..., :pad, :] = 0
    fg[..., -pad:] = 0
    fg[..., -pad:, :] = 0
    return fg
