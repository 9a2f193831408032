"""Synthetic scenes for the DINER rendering hot path (input producer, stock PyTorch).

The reference gets its per-scene inputs from ``PixelNeRF.encode`` (reference
``src/models/pixelnerf.py:35-53``): a ResNet34 feature map, per-view depth / depth-std /
normal maps and the source cameras.  None of that is on the accelerated path, and neither the
datasets nor the reference sources exist on the GPU box, so tests and ``bench.py`` build a
*duck-typed* stand-in here that exposes exactly the attributes the renderer reads
(SURVEY.md section 8b):

    model.poses (SB,NV,4,4)  model.focal (SB,NV,2)  model.c (SB,NV,2)  model.image_shape (2,)=[W,H]
    model.encoder.{latent, depths, depths_std, normals, feature_padding, index_interp,
                   index_padding, nviews, nobjects}
    model.poscode / model.depthcode  (num_freqs, _freqs, _phases, include_input, d_in, d_out)
    model.mlp_fine  (lin_in, lin_z[0..2], blocks[0..4].fc_0/fc_1, lin_out, combine_layer, ...)

The geometry is an analytic ellipsoid rendered into every source camera, so the depth maps are
multi-view consistent and the depth-guided sampler has real work to do.
Nothing in this file is imported by the CUDA path; it only *produces inputs* for it.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from types import SimpleNamespace

import torch
import torch.nn as nn
import torch.nn.functional as F

# ----------------------------------------------------------------------------------------
# workload shapes (BASELINE.json configs; constants from the reference data loaders)
# ----------------------------------------------------------------------------------------


@dataclass(frozen=True)
class Workload:
    name: str
    SB: int
    NV: int
    H: int            # source image height
    W: int            # source image width
    tgt_H: int
    tgt_W: int
    near: float
    far: float
    scale: float      # ellipsoid scale
    white_bkgd: bool
    std_law: str      # "facescape" | "dtu" | "multiface"
    n_samples: int = 40
    n_candidates: int = 1000
    n_gaussian: int = 15


WORKLOADS = {
    # config 0: CPU plumbing case (reference-runnable)
    "plumbing64": Workload("plumbing64", 1, 2, 256, 256, 64, 64, 1.0, 2.5, 1.0, True, "facescape"),
    # config 1: DTU-shape eval, 3 views 300x400, black background (train_dtu.yaml:52-58)
    "dtu": Workload("dtu", 1, 3, 300, 400, 300, 400, 0.3211, 1.2041, 0.45, False, "dtu"),
    # config 2: FaceScape eval, 2 views 256^2 (facescape.py:20-21,45)
    "facescape": Workload("facescape", 1, 2, 256, 256, 256, 256, 1.0, 2.5, 1.0, True, "facescape"),
    # config 3: FaceScape training step, SB=2, one 64x64 patch per scene
    "facescape_train": Workload("facescape_train", 2, 2, 256, 256, 64, 64, 1.0, 2.5, 1.0, True, "facescape"),
    # config 4: Multiface stress, 4 views 1024x672, 1M rays (multiface.py:23-24,47,306-310)
    "multiface": Workload("multiface", 1, 4, 1024, 672, 1024, 1024, 0.5, 1.5, 0.6, True, "multiface"),
    # small cases for unit tests
    "tiny": Workload("tiny", 1, 2, 64, 80, 16, 16, 1.0, 2.5, 1.0, True, "facescape"),
    "tiny3": Workload("tiny3", 2, 3, 48, 64, 12, 12, 0.3211, 1.2041, 0.45, False, "dtu"),
}

_SEMI_AXES = (0.11, 0.15, 0.12)


# ----------------------------------------------------------------------------------------
# cameras
# ----------------------------------------------------------------------------------------

def look_at_extrinsics(center: torch.Tensor) -> torch.Tensor:
    """World->camera 4x4 (OpenCV: x right, y down, z forward) for a camera at `center` looking
    at the origin.  Same convention as the reference rays (cam_geometry.py:56-78)."""
    z = -center / center.norm()
    down = torch.tensor([0.0, -1.0, 0.0], dtype=center.dtype)
    x = torch.linalg.cross(down, z)
    x = x / x.norm()
    y = torch.linalg.cross(z, x)
    R = torch.stack([x, y, z], dim=0)
    E = torch.eye(4, dtype=center.dtype)
    E[:3, :3] = R
    E[:3, 3] = -R @ center
    return E


def _cam_center(az_deg: float, el_deg: float, dist: float) -> torch.Tensor:
    az, el = math.radians(az_deg), math.radians(el_deg)
    return torch.tensor([dist * math.sin(az) * math.cos(el),
                         dist * math.sin(el),
                         -dist * math.cos(az) * math.cos(el)], dtype=torch.float64)


def _intrinsics(H: int, W: int, dist: float, scale: float) -> torch.Tensor:
    # the ellipsoid (height 0.30*scale at distance dist) projects to ~90 % of min(H,W)
    f = 3.0 * min(H, W) * dist / scale
    K = torch.eye(3, dtype=torch.float64)
    K[0, 0] = f
    K[1, 1] = f
    K[0, 2] = W / 2.0
    K[1, 2] = H / 2.0
    return K


def render_ellipsoid_depth(E: torch.Tensor, K: torch.Tensor, H: int, W: int, scale: float) -> torch.Tensor:
    """z-depth of the ellipsoid |x/a|=1 seen from camera (E,K); 0 where the ray misses."""
    a = torch.tensor(_SEMI_AXES, dtype=torch.float64) * scale
    R, t = E[:3, :3], E[:3, 3]
    C = -R.T @ t
    v, u = torch.meshgrid(torch.arange(0.5, H, 1.0, dtype=torch.float64),
                          torch.arange(0.5, W, 1.0, dtype=torch.float64), indexing="ij")
    d_cam = torch.stack([(u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], torch.ones_like(u)], -1)
    d = d_cam @ R          # rows: R^T d_cam
    o = C / a
    dd = d / a
    qa = (dd * dd).sum(-1)
    qb = 2.0 * (dd * o).sum(-1)
    qc = (o * o).sum() - 1.0
    disc = qb * qb - 4 * qa * qc
    hit = disc > 0
    tt = (-qb - torch.sqrt(disc.clamp(min=0))) / (2 * qa)
    depth = torch.where(hit & (tt > 0), tt, torch.zeros_like(tt))
    return depth


def depth_to_normals(depth: torch.Tensor, K: torch.Tensor) -> torch.Tensor:
    """(N,1,H,W) depth -> (N,3,H,W) camera-space normals by central differences
    (cross(vertical, horizontal), i.e. pointing at the camera like depth2normal.py:47-55).
    Pixels without depth, or touching a hole, get the zero normal (which passes the sampler's
    `dot <= 0` test just as the reference's cleaned / zeroed normals do)."""
    N, _, H, W = depth.shape
    v, u = torch.meshgrid(torch.arange(0.5, H, 1.0), torch.arange(0.5, W, 1.0), indexing="ij")
    rays = torch.stack([(u[None] - K[:, 0, 2].view(N, 1, 1)) / K[:, 0, 0].view(N, 1, 1),
                        (v[None] - K[:, 1, 2].view(N, 1, 1)) / K[:, 1, 1].view(N, 1, 1),
                        torch.ones(N, H, W)], dim=1)                       # N,3,H,W
    pts = rays * depth
    pts = F.pad(pts, [1, 1, 1, 1], mode="replicate")
    vdiff = pts[:, :, 2:, 1:-1] - pts[:, :, :-2, 1:-1]
    hdiff = pts[:, :, 1:-1, 2:] - pts[:, :, 1:-1, :-2]
    n = torch.linalg.cross(vdiff, hdiff, dim=1)
    n = n / n.norm(dim=1, keepdim=True).clamp(min=1e-20)
    valid = F.pad((depth != 0).float(), [1, 1, 1, 1], mode="replicate")
    ok = (valid[:, :, 2:, 1:-1] * valid[:, :, :-2, 1:-1] * valid[:, :, 1:-1, 2:] * valid[:, :, 1:-1, :-2]
          * valid[:, :, 1:-1, 1:-1])
    return n * ok


def conf_to_std(conf: torch.Tensor, law: str) -> torch.Tensor:
    if law == "facescape":        # facescape.py:54-56
        return -1.582e-2 * conf + 1.649e-2
    if law == "dtu":              # dtu.py:68-70
        return -2.5679e-2 * conf + 3.2818e-2
    if law == "multiface":        # multiface.py:306-310
        return torch.full_like(conf, 1e-3)
    raise ValueError(law)


# ----------------------------------------------------------------------------------------
# duck-typed field model (parameter names match the reference's ResnetFC state dict,
# resnetfc.py:74-127, so real checkpoints / the real PixelNeRF can be passed instead)
# ----------------------------------------------------------------------------------------

class _Block(nn.Module):
    def __init__(self, d: int):
        super().__init__()
        self.fc_0 = nn.Linear(d, d)
        self.fc_1 = nn.Linear(d, d)
        self.activation = nn.ReLU()
        self.shortcut = None


class FieldMLP(nn.Module):
    """Parameter container with the layout of the reference ResnetFC (no forward: the renderer
    reads the tensors; the torch evaluation used for calibration lives in `field_torch`)."""

    def __init__(self, d_in=55, d_latent=512, d_hidden=512, d_out=4, n_blocks=5, combine_layer=3):
        super().__init__()
        self.d_in, self.d_latent, self.d_hidden, self.d_out = d_in, d_latent, d_hidden, d_out
        self.n_blocks, self.combine_layer, self.combine_type = n_blocks, combine_layer, "average"
        self.lin_in = nn.Linear(d_in, d_hidden)
        self.lin_out = nn.Linear(d_hidden, d_out)
        self.blocks = nn.ModuleList([_Block(d_hidden) for _ in range(n_blocks)])
        self.lin_z = nn.ModuleList([nn.Linear(d_latent, d_hidden) for _ in range(min(combine_layer, n_blocks))])
        self.activation = nn.ReLU()
        for lin in [self.lin_in, self.lin_out, *self.lin_z, *[b.fc_0 for b in self.blocks]]:
            nn.init.constant_(lin.bias, 0.0)
            nn.init.kaiming_normal_(lin.weight, a=0, mode="fan_in")
        for b in self.blocks:
            # the stock init zeroes fc_1 (resnetfc.py:47) which makes every block the identity and
            # the whole render degenerate; use a live init instead
            nn.init.constant_(b.fc_1.bias, 0.0)
            nn.init.kaiming_normal_(b.fc_1.weight, a=0, mode="fan_in")
            with torch.no_grad():
                b.fc_1.weight.mul_(0.3)


class PosCode(nn.Module):
    """Buffers of the reference PositionalEncoding (positional_encoding.py:15-31)."""

    def __init__(self, num_freqs=6, d_in=3, freq_factor=6.28, include_input=True):
        super().__init__()
        self.num_freqs, self.d_in, self.include_input = num_freqs, d_in, include_input
        self.freqs = freq_factor * 2.0 ** torch.arange(0, num_freqs)
        self.d_out = num_freqs * 2 * d_in + (d_in if include_input else 0)
        self.register_buffer("_freqs", torch.repeat_interleave(self.freqs, 2).view(1, -1, 1))
        ph = torch.zeros(2 * num_freqs)
        ph[1::2] = math.pi * 0.5
        self.register_buffer("_phases", ph.view(1, -1, 1))

    def forward(self, x):
        shp = x.shape
        x = x.reshape(-1, shp[-1])
        e = x.unsqueeze(1).repeat(1, self.num_freqs * 2, 1)
        e = torch.sin(torch.addcmul(self._phases, e, self._freqs)).view(x.shape[0], -1)
        if self.include_input:
            e = torch.cat((x, e), dim=-1)
        return e.reshape(*shp[:-1], self.d_out)


class SynthField(nn.Module):
    """Stand-in for PixelNeRF after encode(): holds cameras, maps and the MLP."""

    def __init__(self):
        super().__init__()
        self.poscode = PosCode(d_in=3)
        self.depthcode = PosCode(d_in=1)
        self.mlp_fine = FieldMLP()
        self.encoder = SimpleNamespace(latent=None, depths=None, depths_std=None, normals=None,
                                       feature_padding=32.0, index_interp="bilinear", index_padding="border",
                                       nviews=None, nobjects=None)
        self.poses = None
        self.focal = None
        self.c = None
        self.image_shape = None

    def to_device(self, device):
        self.to(device)
        for k in ("latent", "depths", "depths_std", "normals"):
            setattr(self.encoder, k, getattr(self.encoder, k).to(device))
        for k in ("poses", "focal", "c", "image_shape"):
            setattr(self, k, getattr(self, k).to(device))
        return self


def gen_rays(extrinsics, intrinsics, W, H, z_near, z_far):
    """Pixel-centre rays [o(3), d(3), near, far] for (B,4,4) world->cam extrinsics; same
    convention as the reference caller (cam_geometry.py:36-79).  Input producer."""
    B = extrinsics.shape[0]
    dev = extrinsics.device
    v, u = torch.meshgrid(torch.arange(0.5, H, 1.0, device=dev), torch.arange(0.5, W, 1.0, device=dev),
                          indexing="ij")
    pix = torch.stack((u, v), -1).unsqueeze(0).expand(B, -1, -1, -1)
    focal = torch.stack((intrinsics[:, 0, 0], intrinsics[:, 1, 1]), -1).view(B, 1, 1, 2)
    c = intrinsics[:, :2, 2].view(B, 1, 1, 2)
    pc = (pix - c) / focal
    pc = torch.cat((pc, torch.ones_like(pc[..., :1])), -1)
    dirs_cam = pc / pc.pow(2).sum(-1, keepdim=True).sqrt()
    R_c2w = extrinsics[:, :3, :3].transpose(1, 2)
    dirs = (R_c2w @ dirs_cam.view(B, -1, 3).transpose(1, 2)).transpose(1, 2).view(B, H, W, 3)
    centers = (-R_c2w @ extrinsics[:, :3, 3:]).view(B, 1, 1, 3).expand(-1, H, W, -1)
    near = z_near.view(B, 1, 1, 1).expand(-1, H, W, -1)
    far = z_far.view(B, 1, 1, 1).expand(-1, H, W, -1)
    return torch.cat((centers, dirs, near, far), -1)


def source_azimuths(NV: int):
    if NV == 1:
        return [20.0]
    lo, hi = -35.0, 35.0
    return [lo + (hi - lo) * i / (NV - 1) for i in range(NV)]


def make_scene(workload: Workload | str, seed: int = 0, latent_dtype=torch.float32, perturb_depth=True,
               calibrate=True) -> SimpleNamespace:
    """Build model + target rays for a workload on the CPU (move with scene.model.to_device)."""
    wl = WORKLOADS[workload] if isinstance(workload, str) else workload
    g = torch.Generator().manual_seed(1000 + seed)
    dist = 0.5 * (wl.near + wl.far)
    model = SynthField()
    torch.manual_seed(seed + 1)
    model.mlp_fine = FieldMLP()

    poses, Ks, depths, stds = [], [], [], []
    for sb in range(wl.SB):
        p_s, k_s, d_s, s_s = [], [], [], []
        for i, az in enumerate(source_azimuths(wl.NV)):
            E = look_at_extrinsics(_cam_center(az + 3.0 * sb, 5.0 * ((i % 2) * 2 - 1), dist))
            K = _intrinsics(wl.H, wl.W, dist, wl.scale)
            d = render_ellipsoid_depth(E, K, wl.H, wl.W, wl.scale).float()
            conf = torch.rand(wl.H, wl.W, generator=g) * 0.8 + 0.2
            std = conf_to_std(conf, wl.std_law) * (d != 0)
            if perturb_depth:
                d = d + torch.randn(wl.H, wl.W, generator=g) * std * 0.5
                d = d * (std != 0)
            p_s.append(E.float()); k_s.append(K.float()); d_s.append(d); s_s.append(std)
        poses.append(torch.stack(p_s)); Ks.append(torch.stack(k_s))
        depths.append(torch.stack(d_s)); stds.append(torch.stack(s_s))
    poses = torch.stack(poses)                       # SB,NV,4,4
    Ks = torch.stack(Ks)                             # SB,NV,3,3
    depths = torch.stack(depths).unsqueeze(2)        # SB,NV,1,H,W
    stds = torch.stack(stds).unsqueeze(2)
    normals = depth_to_normals(depths.flatten(0, 1), Ks.flatten(0, 1)).view(wl.SB, wl.NV, 3, wl.H, wl.W)

    Hl, Wl = (wl.H + 128) // 2, (wl.W + 128) // 2
    # post-ReLU-like feature statistics (latent ~1.3 +- 1.1)
    latent = (torch.randn(wl.SB, wl.NV, 512, Hl, Wl, generator=g) * 0.9 + 0.6).clamp_(min=0).to(latent_dtype)

    model.encoder.latent = latent
    model.encoder.depths = depths
    model.encoder.depths_std = stds
    model.encoder.normals = normals
    model.encoder.nviews, model.encoder.nobjects = wl.NV, wl.SB
    model.poses = poses
    model.focal = torch.stack((Ks[..., 0, 0], Ks[..., 1, 1]), -1)
    model.c = Ks[..., :2, 2].clone()
    model.image_shape = torch.tensor([float(wl.W), float(wl.H)])

    # target cameras
    tE = torch.stack([look_at_extrinsics(_cam_center(2.0 * sb, 8.0, dist)).float() for sb in range(wl.SB)])
    tK = torch.stack([_intrinsics(wl.tgt_H, wl.tgt_W, dist, wl.scale).float() for _ in range(wl.SB)])
    rays = gen_rays(tE, tK, wl.tgt_W, wl.tgt_H, torch.full((wl.SB,), wl.near), torch.full((wl.SB,), wl.far))
    rays = rays.view(wl.SB, wl.tgt_H * wl.tgt_W, 8).contiguous()

    scene = SimpleNamespace(workload=wl, model=model, rays=rays, target_extrinsics=tE, target_intrinsics=tK,
                            src_intrinsics=Ks)
    if calibrate:
        calibrate_sigma(scene, seed=seed)
    return scene


def target_rays(scene, azimuth_deg: float, elevation_deg: float = 8.0) -> torch.Tensor:
    """Rays (SB, tgt_H*tgt_W, 8) of a target camera on the sweep ring around the scene (one frame of the
    reference's create_cam_sweep loop, diner.py:178-199)."""
    wl = scene.workload
    dist = 0.5 * (wl.near + wl.far)
    tE = torch.stack([look_at_extrinsics(_cam_center(azimuth_deg + 2.0 * sb, elevation_deg, dist)).float()
                      for sb in range(wl.SB)])
    rays = gen_rays(tE, scene.target_intrinsics, wl.tgt_W, wl.tgt_H, torch.full((wl.SB,), wl.near),
                    torch.full((wl.SB,), wl.far))
    return rays.view(wl.SB, wl.tgt_H * wl.tgt_W, 8).contiguous()


# ----------------------------------------------------------------------------------------
# stock-torch field evaluation, used ONLY to calibrate the random-init density head
# ----------------------------------------------------------------------------------------

@torch.no_grad()
def field_torch(model, xyz, viewdirs):
    """Raw MLP output (SB,B,4) before sigmoid/relu.  Plain torch ops; input-producer side."""
    SB, B, _ = xyz.shape
    NV = model.encoder.nviews
    R, t = model.poses[:, :, :3, :3], model.poses[:, :, :3, 3]
    x = torch.einsum("svij,sbj->svbi", R, xyz) + t.unsqueeze(2)
    d = torch.einsum("svij,sbj->svbi", R, viewdirs)
    uv = x[..., :2] / x[..., 2:] * model.focal.unsqueeze(2) + model.c.unsqueeze(2)
    uv = uv / model.image_shape * 2 - 1
    lat = model.encoder.latent.flatten(0, 1).float()
    size = torch.tensor([lat.shape[-1], lat.shape[-2]], dtype=torch.float32)
    uv_l = uv * ((size - 2 * model.encoder.feature_padding) / size)
    feat = F.grid_sample(lat, uv_l.flatten(0, 1).unsqueeze(2), mode="bilinear", padding_mode="border",
                         align_corners=False)[..., 0].transpose(1, 2)           # SBNV,B,512
    dep = F.grid_sample(model.encoder.depths.flatten(0, 1), uv.flatten(0, 1).unsqueeze(2), mode="nearest",
                        padding_mode="border", align_corners=False)[:, 0, :, 0]  # SBNV,B
    dd = dep.view(SB, NV, B) - x[..., 2]
    zin = torch.cat((model.poscode(x), d, model.depthcode(dd.unsqueeze(-1))), -1).flatten(0, 1)
    m = model.mlp_fine
    h = m.lin_in(zin)
    for b in range(m.n_blocks):
        if b == m.combine_layer:
            h = h.view(SB, NV, B, -1).mean(1)
        if b < m.combine_layer:
            h = h + m.lin_z[b](feat)
        blk = m.blocks[b]
        h = h + blk.fc_1(torch.relu(blk.fc_0(torch.relu(h))))
    return m.lin_out(torch.relu(h)).view(SB, B, -1)


@torch.no_grad()
def calibrate_sigma(scene, seed=0, n_rays=256, target_std=150.0):
    """Rescale lin_out row 3 so that the raw density logit has median 0 / std `target_std` on points
    near the surface -- otherwise relu(sigma)==0 everywhere and the render is empty."""
    model, rays, wl = scene.model, scene.rays, scene.workload
    g = torch.Generator().manual_seed(77 + seed)
    idx = torch.randint(0, rays.shape[1], (n_rays,), generator=g)
    r = rays[:, idx]
    z = wl.near + (wl.far - wl.near) * (0.35 + 0.3 * torch.rand(wl.SB, n_rays, 16, generator=g))
    pts = (r[:, :, None, :3] + z.unsqueeze(-1) * r[:, :, None, 3:6]).reshape(wl.SB, -1, 3)
    dirs = r[:, :, None, 3:6].expand(-1, -1, 16, -1).reshape(wl.SB, -1, 3)
    raw = field_torch(model, pts, dirs)[..., 3]
    med, std = raw.median(), raw.std().clamp(min=1e-6)
    s = target_std / std
    w, b = model.mlp_fine.lin_out.weight, model.mlp_fine.lin_out.bias
    w[3] *= s
    b[3] = b[3] * s - med * s


# ----------------------------------------------------------------------------------------
# NOVEL variant inputs: a target mesh with per-vertex offsets, the generic camera and the learned
# second feature map of novel_pixelnerf.py:34-46,73-80.  Input producers only.
# ----------------------------------------------------------------------------------------

def add_novel_inputs(scene, n_vertices: int = 400, seed: int = 0, offset_std: float = 0.004, gen_hw=None):
    """Adds scene.target_vertices / tgt_in_offsets / tgt_gen_offsets (SB,V,3) and model.gen_poses (SB,1,4,4),
    gen_focal / gen_c (SB,1,2), gen_image_shape (2,), gen_latent (512,Hg,Wg) to a scene from make_scene."""
    wl, m = scene.workload, scene.model
    g = torch.Generator().manual_seed(9000 + seed)
    a = torch.tensor(_SEMI_AXES) * wl.scale
    # vertices: quasi-uniform directions on the ellipsoid (the "mesh" the offsets live on)
    u = torch.randn(wl.SB, n_vertices, 3, generator=g)
    u = u / u.norm(dim=-1, keepdim=True)
    scene.target_vertices = (u * a).contiguous()
    scene.tgt_in_offsets = (torch.randn(wl.SB, n_vertices, 3, generator=g) * offset_std * wl.scale).contiguous()
    scene.tgt_gen_offsets = (torch.randn(wl.SB, n_vertices, 3, generator=g) * offset_std * wl.scale).contiguous()
    dist = 0.5 * (wl.near + wl.far)
    E = torch.stack([look_at_extrinsics(_cam_center(0.0, 0.0, dist)).float() for _ in range(wl.SB)])
    K = _intrinsics(wl.H, wl.W, dist, wl.scale).float()
    m.gen_poses = E.unsqueeze(1).contiguous()                                        # SB,1,4,4
    m.gen_focal = torch.stack((K[0, 0], K[1, 1])).view(1, 1, 2).expand(wl.SB, 1, 2).contiguous()
    m.gen_c = K[:2, 2].view(1, 1, 2).expand(wl.SB, 1, 2).contiguous()
    m.gen_image_shape = torch.tensor([float(wl.W), float(wl.H)])
    Hg, Wg = gen_hw or ((wl.H + 128) // 2 - 8, (wl.W + 128) // 2 + 8)                  # its own size on purpose
    m.gen_latent = (torch.randn(512, Hg, Wg, generator=g) * 0.5).contiguous()
    return scene


def add_novel_pe_inputs(scene, seed: int = 0):
    """NOVEL_PE inputs on top of add_novel_inputs (src/models/novel_pe/pe_novel_pixelnerf.py:21,27,35,45-49,57-99):
    model.pos_encodings (SB,NV,3,Hp,Wp), model.tgt_pos_encoding (SB,1,3,Ht,Wt), the target camera model.tgt_poses
    (SB,1,4,4) / tgt_focal / tgt_c (SB,1,2) / tgt_image_shape (2,), model.deformation_layer = nn.Linear(518, 512), and an
    MLP with d_latent = 518 (the scene's MLP with six more lin_z columns, so the calibrated density head stays alive)."""
    wl, m = scene.workload, scene.model
    g = torch.Generator().manual_seed(9100 + seed)
    m.pos_encodings = (torch.randn(wl.SB, wl.NV, 3, wl.H, wl.W, generator=g) * 0.5).contiguous()
    m.tgt_pos_encoding = (torch.randn(wl.SB, 1, 3, wl.H + 16, wl.W + 24, generator=g) * 0.5).contiguous()   # its own size
    dist = 0.5 * (wl.near + wl.far)
    E = torch.stack([look_at_extrinsics(_cam_center(20.0, 5.0, dist)).float() for _ in range(wl.SB)])
    K = _intrinsics(wl.H, wl.W, dist, wl.scale).float()
    m.tgt_poses = E.unsqueeze(1).contiguous()
    m.tgt_focal = torch.stack((K[0, 0], K[1, 1])).view(1, 1, 2).expand(wl.SB, 1, 2).contiguous()
    m.tgt_c = K[:2, 2].view(1, 1, 2).expand(wl.SB, 1, 2).contiguous()
    m.tgt_image_shape = torch.tensor([float(wl.W), float(wl.H)])
    old = m.mlp_fine
    new = FieldMLP(d_in=old.d_in, d_latent=old.d_latent + 6)
    with torch.no_grad():
        sd = old.state_dict()
        for k, v in new.state_dict().items():
            if k.startswith("lin_z.") and k.endswith(".weight"):
                v[:, :old.d_latent] = sd[k]
                v[:, old.d_latent:] = torch.randn(v.shape[0], 6, generator=g) * 0.5
            else:
                v.copy_(sd[k])
    m.mlp_fine = new
    dl = nn.Linear(old.d_latent + 6, old.d_latent)
    with torch.no_grad():
        dl.weight[:, :old.d_latent] = 0.8 * torch.eye(old.d_latent) + 0.02 * torch.randn(old.d_latent, old.d_latent, generator=g)
        dl.weight[:, old.d_latent:] = 0.3 * torch.randn(old.d_latent, 6, generator=g)
        dl.bias.copy_(0.05 * torch.randn(old.d_latent, generator=g))
    m.deformation_layer = dl
    return scene


def novel_to_device(scene, device):
    m = scene.model
    for k in ("gen_poses", "gen_focal", "gen_c", "gen_image_shape", "gen_latent", "pos_encodings", "tgt_pos_encoding",
              "tgt_poses", "tgt_focal", "tgt_c", "tgt_image_shape"):
        if getattr(m, k, None) is None:
            continue
        setattr(m, k, getattr(m, k).to(device))
    for k in ("target_vertices", "tgt_in_offsets", "tgt_gen_offsets"):
        setattr(scene, k, getattr(scene, k).to(device))
    m.to_device(device)
    return scene
