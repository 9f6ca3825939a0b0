"""Pure-PyTorch float64 restatement of the splatting equations with autograd.

Independent of oracle/splat_oracle.c (different language, dense evaluation, no
hand-written backward): it exists to cross-check the oracle's forward values and,
through torch.autograd, the oracle's *hand-restated* backward formulas
(DGR32/cuda_rasterizer/backward.cu).  Small inputs only (dense [pixels x P]).

Equations: forward.cu:74-109 (cov2D), :114-148 (cov3D), :179-268 (preprocess),
:290-396 (compositing).  `tile_members[t]` (Gaussian ids in tile t, depth order)
is taken from the caller because a Gaussian only contributes to pixels of tiles
inside its 3-sigma rectangle (rasterizer_impl.cu:70-111).
"""
from __future__ import annotations

import torch


def quat_to_R(q):
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1).reshape(-1, 3, 3)
    return R   # row-major, as typed in forward.cu:130-134


def render(means3D, scales, rotations, opacity, feats, viewmatrix, projmatrix, tanfovx, tanfovy, H, W, bg,
           tile_members, ndc_offset=None, scale_modifier=1.0, antialiasing=False):
    """All tensors float64.  Returns color [C,H,W], invdepth [H,W]."""
    P = means3D.shape[0]
    C = feats.shape[1]
    V = viewmatrix.reshape(4, 4)      # transposed convention: p_view = [p,1] @ V
    PR = projmatrix.reshape(4, 4)
    ones = torch.ones(P, 1, dtype=means3D.dtype)
    ph = torch.cat([means3D, ones], 1)
    t = ph @ V                         # [P,4]
    hom = ph @ PR
    p_w = 1.0 / (hom[:, 3] + 1e-7)
    proj = hom[:, :2] * p_w[:, None]
    if ndc_offset is not None:
        proj = proj + ndc_offset[:, :2]
    focal_y = H / (2.0 * tanfovy)
    focal_x = W / (2.0 * tanfovx)

    # cov3D: GLM reads the typed rows as columns, so transpose(M)*M with M = S*R_glm is
    # Sigma = R S^2 R^T with R the ordinary (row-major, as typed) rotation matrix.
    Rt = quat_to_R(rotations)
    S = torch.diag_embed(scales * scale_modifier)
    Mm = Rt @ S
    Sigma = Mm @ Mm.transpose(1, 2)

    tz = t[:, 2]
    limx, limy = 1.3 * tanfovx, 1.3 * tanfovy
    txc = torch.clamp(t[:, 0] / tz, -limx, limx) * tz
    tyc = torch.clamp(t[:, 1] / tz, -limy, limy) * tz
    zeros = torch.zeros_like(tz)
    J = torch.stack([
        focal_x / tz, zeros, -(focal_x * txc) / (tz * tz),
        zeros, focal_y / tz, -(focal_y * tyc) / (tz * tz)], dim=1).reshape(-1, 2, 3)
    Wm = V[:3, :3].T                   # world->view rotation (rows)
    T = J @ Wm                         # [P,2,3]
    cov = T @ Sigma @ T.transpose(1, 2)
    a = cov[:, 0, 0]
    b = cov[:, 0, 1]
    c = cov[:, 1, 1]
    det_cov = a * c - b * b
    a = a + 0.3
    c = c + 0.3
    det = a * c - b * b
    h = torch.ones_like(det)
    if antialiasing:
        h = torch.sqrt(torch.clamp(det_cov / det, min=0.000025))
    conx, cony, conz = c / det, -b / det, a / det
    op = opacity.reshape(-1) * h
    px = ((proj[:, 0] + 1.0) * W - 1.0) * 0.5
    py = ((proj[:, 1] + 1.0) * H - 1.0) * 0.5
    invd = 1.0 / tz

    color = torch.zeros(C, H, W, dtype=means3D.dtype)
    invdepth = torch.zeros(H, W, dtype=means3D.dtype)
    gx = (W + 15) // 16
    for tile, members in enumerate(tile_members):
        if len(members) == 0 and bg is None:
            continue
        ty0, tx0 = (tile // gx) * 16, (tile % gx) * 16
        ys = torch.arange(ty0, min(ty0 + 16, H))
        xs = torch.arange(tx0, min(tx0 + 16, W))
        yy, xx = torch.meshgrid(ys.to(means3D.dtype), xs.to(means3D.dtype), indexing="ij")
        Tacc = torch.ones_like(yy)
        done = torch.zeros_like(yy, dtype=torch.bool)
        Cacc = torch.zeros(C, *yy.shape, dtype=means3D.dtype)
        Dacc = torch.zeros_like(yy)
        for gid in members:
            dx = px[gid] - xx
            dy = py[gid] - yy
            power = -0.5 * (conx[gid] * dx * dx + conz[gid] * dy * dy) - cony[gid] * dx * dy
            alpha = torch.clamp(op[gid] * torch.exp(power), max=0.99)
            ok = (power <= 0) & (alpha >= 1.0 / 255.0) & ~done
            test_T = Tacc * (1 - alpha)
            stop = ok & (test_T < 0.0001)
            done = done | stop
            ok = ok & ~stop
            w = torch.where(ok, alpha * Tacc, torch.zeros_like(alpha))
            Cacc = Cacc + feats[gid][:, None, None] * w[None]
            Dacc = Dacc + invd[gid] * w
            Tacc = torch.where(ok, test_T, Tacc)
        color[:, ty0:ty0 + len(ys), tx0:tx0 + len(xs)] = Cacc + Tacc[None] * bg[:, None, None]
        invdepth[ty0:ty0 + len(ys), tx0:tx0 + len(xs)] = Dacc
    return color, invdepth
