"""Seeded synthetic inputs shaped like GUAVA's rasterizer calls.

Recipe: SURVEY.md section 8(d).  Camera convention is the reference's
(`utils/graphics_utils.py:23-50`): `viewmatrix` is the transposed world-to-camera
matrix, `projmatrix` is `viewmatrix @ proj^T` (row-vector convention, flat index
j*4+i, read as m[0],m[4],m[8],m[12] by `auxiliary.h:70-89`), tan(fov/2) = 1/24,
z_near = 0.01, z_far = 100, camera 22 units from the cloud
(`dataset/data_loader.py:377-394`), orbit of `utils/camera_utils.py:72-97`
(yaw +-0.35 rad, pitch +-0.3 rad).

Everything is generated on the CPU with an explicit torch.Generator so the
same tensors exist on the GPU box, in this container and for the oracle.
"""
from __future__ import annotations

import math
from typing import Dict, List

import torch

NUM_CHANNELS = 32


def proj_matrix(tanfovx: float, tanfovy: float, z_near: float = 0.01, z_far: float = 100.0) -> torch.Tensor:
    """graphics_utils.py:23-42 generalised to tanfovx != tanfovy (config C5)."""
    top = tanfovy * z_near
    right = tanfovx * z_near
    m = torch.zeros(4, 4, dtype=torch.float32)
    m[0, 0] = 2.0 * z_near / (2.0 * right)
    m[1, 1] = 2.0 * z_near / (2.0 * top)
    m[3, 2] = 1.0
    m[2, 2] = z_far / (z_far - z_near)
    m[2, 3] = -(z_far * z_near) / (z_far - z_near)
    return m


def look_at_w2c(campos: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """World-to-camera with +z forward (z_sign = 1.0 in the reference)."""
    f = target - campos
    f = f / f.norm()
    up = torch.tensor([0.0, 1.0, 0.0])
    r = torch.linalg.cross(up, f)
    r = r / r.norm()
    u = torch.linalg.cross(f, r)
    R = torch.stack([r, u, f], 0)
    w2c = torch.eye(4, dtype=torch.float32)
    w2c[:3, :3] = R
    w2c[:3, 3] = -R @ campos
    return w2c


def camera(view: int, num_views: int, tanfovx: float = 1.0 / 24, tanfovy: float = 1.0 / 24,
           radius: float = 22.0, yaw_range: float = 0.35, pitch_range: float = 0.3) -> Dict[str, torch.Tensor]:
    """View `view` of `num_views` on the reference's novel-view orbit."""
    ph = 2.0 * math.pi * view / max(num_views, 1)
    yaw = yaw_range * math.sin(ph)
    pitch = pitch_range * math.cos(ph) - 0.05
    if num_views <= 1:
        yaw, pitch = 0.0, 0.0
    target = torch.zeros(3)
    campos = torch.tensor([
        radius * math.sin(yaw) * math.cos(pitch),
        radius * math.sin(pitch),
        -radius * math.cos(yaw) * math.cos(pitch),
    ], dtype=torch.float32)
    w2c = look_at_w2c(campos, target)
    view_t = w2c.transpose(0, 1).contiguous()
    proj_t = proj_matrix(tanfovx, tanfovy).transpose(0, 1).contiguous()
    full = (view_t.unsqueeze(0).bmm(proj_t.unsqueeze(0))).squeeze(0).contiguous()
    return dict(viewmatrix=view_t, projmatrix=full, campos=campos.clone(),
                tanfovx=float(tanfovx), tanfovy=float(tanfovy))


def guava_cloud(P: int, seed: int, scale_mult: float = 1.0) -> Dict[str, torch.Tensor]:
    """Two-sided ellipsoidal shell ~ an upper-body avatar."""
    g = torch.Generator().manual_seed(int(seed))
    d = torch.randn(P, 3, generator=g)
    d = d / d.norm(dim=1, keepdim=True).clamp_min(1e-12)
    radii = torch.tensor([0.55, 0.75, 0.25])
    xyz = d * radii * (1.0 + 0.02 * torch.randn(P, 1, generator=g))
    s0 = 0.0054 * math.sqrt(1e5 / P) * scale_mult
    scales = s0 * torch.exp(0.4 * torch.randn(P, 3, generator=g))
    rot = torch.randn(P, 4, generator=g)
    rot = rot / rot.norm(dim=1, keepdim=True).clamp_min(1e-12)
    opacity = torch.sigmoid(1.0 + 2.0 * torch.randn(P, 1, generator=g)).clamp(0.002, 0.999)
    feats = torch.randn(P, NUM_CHANNELS, generator=g)
    feats[:, :3] = torch.sigmoid(feats[:, :3])
    return dict(xyz=xyz.float().contiguous(), scaling=scales.float().contiguous(),
                rotation=rot.float().contiguous(), opacity=opacity.float().contiguous(),
                features_color=feats.float().contiguous())


def upstream_grads(H: int, W: int, seed: int):
    """dL/dcolor = N(0,1)/(32 H W); dL/dinvdepth = 0 (GUAVA's loss never reads
    depth, so autograd feeds zeros: SURVEY 3.2)."""
    g = torch.Generator().manual_seed(int(seed) + 7919)
    dc = torch.randn(NUM_CHANNELS, H, W, generator=g) / (NUM_CHANNELS * H * W)
    return dc.float().contiguous(), torch.zeros(1, H, W, dtype=torch.float32)


CONFIGS = {
    # name: (P, H, W, views, scale_mult, fwd_bwd)
    "C1": dict(P=10_000, H=512, W=512, views=1, scale_mult=1.0),
    "C2": dict(P=100_000, H=512, W=512, views=8, scale_mult=1.0),
    "C3": dict(P=100_000, H=512, W=512, views=16, scale_mult=1.0),
    "C3-202k": dict(P=202_000, H=512, W=512, views=16, scale_mult=1.0),
    "C4": dict(P=300_000, H=1024, W=1024, views=240, scale_mult=1.0),
    "C5": dict(P=1_000_000, H=1080, W=1920, views=1, scale_mult=4.0),
}


def make_views(config_id: int, P: int, H: int, W: int, views: List[int], num_views: int,
               scale_mult: float = 1.0, shared_cloud: bool = False) -> Dict[str, torch.Tensor]:
    """Stacked [V, ...] tensors for a set of view indices; seed = 1000*config + view."""
    tanfovy = 1.0 / 24
    tanfovx = tanfovy * W / H
    clouds, cams = [], []
    for v in views:
        seed = 1000 * config_id + (0 if shared_cloud else v)
        clouds.append(guava_cloud(P, seed, scale_mult))
        cams.append(camera(v, num_views, tanfovx, tanfovy))
    out = {k: torch.stack([c[k] for c in clouds], 0) for k in clouds[0]}
    out["viewmatrix"] = torch.stack([c["viewmatrix"] for c in cams], 0)
    out["projmatrix"] = torch.stack([c["projmatrix"] for c in cams], 0)
    out["campos"] = torch.stack([c["campos"] for c in cams], 0)
    out["tanfovx"] = tanfovx
    out["tanfovy"] = tanfovy
    out["H"], out["W"] = H, W
    return out


# ----------------------------------------------------------------------------- pose-driven avatar
def pose_avatar(P: int, V: int, seed: int, rows: int = 83, cols: int = 126) -> Dict[str, torch.Tensor]:
    """A GUAVA-shaped avatar for the pose-driven workload (what models/UbodyAvatar/ubody_gaussian.py:245-289 takes):
    a lat-long triangle mesh of the cloud's ellipsoid shell -- rows*cols = 10 458 vertices and 2*(rows-1)*cols =
    20 664 faces with the defaults, SMPL-X has 10 475 / 20 908 -- one Gaussian per vertex and N = P - Nv Gaussians bound
    to random faces with barycentric coordinates and a local offset, per-view attribute sets [V, ...] (the reference's
    batch dimension).  World-space sizes follow guava_cloud (s0 = 0.0054 sqrt(1e5 / P))."""
    g = torch.Generator().manual_seed(int(seed))
    Nv = rows * cols
    N = P - Nv
    if N < 0:
        raise ValueError("P must be at least rows * cols")
    th = torch.linspace(0.04, math.pi - 0.04, rows)[:, None].expand(rows, cols)
    ph = (torch.arange(cols, dtype=torch.float32) * (2.0 * math.pi / cols))[None, :].expand(rows, cols)
    radii = torch.tensor([0.55, 0.75, 0.25])
    verts = torch.stack([torch.sin(th) * torch.cos(ph), torch.cos(th), torch.sin(th) * torch.sin(ph)], -1).reshape(Nv, 3) * radii
    r = torch.arange(rows - 1)[:, None]
    c = torch.arange(cols)[None, :]
    v00, v10 = r * cols + c, (r + 1) * cols + c
    v01, v11 = r * cols + (c + 1) % cols, (r + 1) * cols + (c + 1) % cols
    faces = torch.cat([torch.stack([v00, v10, v01], -1).reshape(-1, 3), torch.stack([v01, v10, v11], -1).reshape(-1, 3)], 0)
    F = faces.shape[0]
    face_size = 0.5 * (2.0 * math.pi * 0.55 / cols + math.pi * 0.75 / rows)      # typical (edge + height) / 2 of a face
    s0 = 0.0054 * math.sqrt(1e5 / P)
    rn = lambda *s: torch.randn(*s, generator=g)
    unit = lambda x: x / x.norm(dim=-1, keepdim=True).clamp_min(1e-12)
    bary = torch.rand(N, 3, generator=g)
    bary = bary / bary.sum(-1, keepdim=True)
    # faces shrink towards the poles: bind in proportion to face area (uniform over the surface) and size the bound
    # Gaussians by their face's own scale (utils/graphics_utils.py:61-80: s = (|e0| + |a2 . e1|) / 2), so that world
    # sizes and tile-list lengths are those of guava_cloud
    e0, e1 = verts[faces[:, 1]] - verts[faces[:, 0]], verts[faces[:, 2]] - verts[faces[:, 0]]
    a0 = unit(e0)
    a1 = unit(torch.linalg.cross(a0, e1))
    a2 = -unit(torch.linalg.cross(a1, a0))
    s_face = 0.5 * (e0.norm(dim=-1) + (a2 * e1).sum(-1).abs())
    area = 0.5 * torch.linalg.cross(e0, e1).norm(dim=-1)
    binding = torch.multinomial(area / area.sum(), N, replacement=True, generator=g) if N else torch.zeros(0, dtype=torch.long)
    inv_s = (1.0 / s_face[binding].clamp_min(1e-4))[None, :, None]

    def feats(n):
        f = rn(V, n, NUM_CHANNELS)
        f[..., :3] = torch.sigmoid(f[..., :3])
        return f

    opac = lambda n: torch.sigmoid(1.0 + 2.0 * rn(V, n, 1)).clamp(0.002, 0.999)
    return dict(
        vertices=verts.float().contiguous(), faces=faces.int().contiguous(),
        uv_binding_face=binding.int().contiguous(), uv_face_bary=bary.float().contiguous(),
        smplx_rotation=unit(rn(V, Nv, 4)), smplx_scaling=s0 * torch.exp(0.4 * rn(V, Nv, 3)), smplx_opacity=opac(Nv),
        smplx_features_color=feats(Nv), uv_local_xyz=(0.3 * face_size) * rn(V, N, 3) * inv_s, uv_rotation=unit(rn(V, N, 4)),
        uv_scaling=s0 * torch.exp(0.4 * rn(V, N, 3)) * inv_s, uv_opacity=opac(N), uv_features_color=feats(N))


def pose_of(vertices: torch.Tensor, step: int, V: int):
    """Posed mesh of `step` for V views: a rigid sway of the rest mesh (rotation about y, small translation) -- the
    per-step data GUAVA moves to the device: vertices [V,Nv,3] and per-vertex transforms [V,Nv,4,4]."""
    Nv = vertices.shape[0]
    outv, outt = [], []
    for v in range(V):
        a = 0.2 * math.sin(0.37 * step + 1.3 * v)
        R = torch.tensor([[math.cos(a), 0.0, math.sin(a)], [0.0, 1.0, 0.0], [-math.sin(a), 0.0, math.cos(a)]])
        t = torch.tensor([0.02 * math.sin(0.11 * step + v), 0.015 * math.cos(0.23 * step), 0.0])
        T = torch.eye(4)
        T[:3, :3] = R
        T[:3, 3] = t
        outv.append(vertices @ R.T + t)
        outt.append(T[None].expand(Nv, 4, 4))
    return torch.stack(outv, 0).float().contiguous(), torch.stack(outt, 0).float().contiguous()
