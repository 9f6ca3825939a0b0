"""Batched replacement for the rasterization part of GUAVA's renderer module.

Reference: models/UbodyAvatar/gaussian_render.py:19-85
(`GaussianRenderer.forward_nueral_refine`).  The reference loops over the batch
in Python, builds one `GaussianRasterizationSettings` and one
`GaussianRasterizer_32` per view, and `torch.stack`s the results (:37-76).
`render_views` takes the same `gaussian_assets` / `cam_params` dictionaries and
produces the same output dictionary from ONE launch chain over all views,
writing straight into the stacked [B,32,H,W] tensor.  The neural refiner
(StyleUNet, :73) is outside the hot path; pass it as `refiner` to get
"renders", otherwise that key is omitted.  `raw_renders` / `extra_renders` are views of the stacked
image (no copies, as in :71,:83); `refiner_input=True` adds "refiner_input": the same image as a
channels-last bf16 tensor written by the compositing kernel itself.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch

from .rasterizer import BatchRasterizationSettings, rasterize_gaussians_batch


def _as_float_list(x, B):
    if torch.is_tensor(x):
        return x.reshape(-1).to(torch.float32)
    return torch.tensor([float(v) for v in x], dtype=torch.float32) if hasattr(x, "__len__") \
        else torch.full((B,), float(x), dtype=torch.float32)


def render_views(gaussian_assets: Dict[str, torch.Tensor], cam_params: Dict[str, torch.Tensor], bg=None,
                 scaling_modifier: float = 1.0, antialiasing: bool = False,
                 refiner: Optional[Callable[[torch.Tensor], torch.Tensor]] = None,
                 refiner_input: bool = False) -> Dict[str, torch.Tensor]:
    mean_3d = gaussian_assets['xyz']
    opacity = gaussian_assets['opacity']
    scales = gaussian_assets['scaling']
    rotations = gaussian_assets['rotation']
    features_color = gaussian_assets['features_color']          # no .clone(): inputs are never mutated
    device = mean_3d.device
    B = int(cam_params['world_view_transform'].shape[0])

    P = mean_3d.shape[-2]
    mean_2d = torch.zeros((B, P, 3), dtype=torch.float32, requires_grad=True, device=device)   # :30-31
    if bg is None:
        bg = 0.0
    bg_t = torch.ones((B, features_color.shape[-1]), dtype=torch.float32, device=device) * bg   # :35

    H = int(cam_params['image_height'].reshape(-1)[0]) if torch.is_tensor(cam_params['image_height']) \
        else int(cam_params['image_height'])
    W = int(cam_params['image_width'].reshape(-1)[0]) if torch.is_tensor(cam_params['image_width']) \
        else int(cam_params['image_width'])
    tfx = _as_float_list(cam_params['tanfovx'], B).to(device)
    tfy = _as_float_list(cam_params['tanfovy'], B).to(device)
    if tfx.numel() == 1:
        tfx = tfx.expand(B)
    if tfy.numel() == 1:
        tfy = tfy.expand(B)
    tanfov = torch.stack([tfx, tfy], dim=1).contiguous()

    settings = BatchRasterizationSettings(
        image_height=H, image_width=W, tanfov=tanfov, bg=bg_t, scale_modifier=float(scaling_modifier),
        viewmatrix=cam_params['world_view_transform'], projmatrix=cam_params['full_proj_transform'],
        debug=False, antialiasing=bool(antialiasing))
    aux = None
    if refiner_input:
        # refiner hand-off: the compositing kernel's epilogue also writes the image as
        # channels-last bf16, the layout/precision a tensor-core first convolution of the refiner wants
        aux = torch.empty(B, H, W, features_color.shape[-1], dtype=torch.bfloat16, device=device)
    rendered_images, radiis, depth_images = rasterize_gaussians_batch(
        mean_3d, mean_2d, features_color, opacity, scales, rotations, settings, refiner_output=aux)

    out = {
        "raw_renders": rendered_images[:, :3],
        "viewspace_points": mean_2d,
        "radiis": radiis,
        "depths": depth_images,
        "extra_renders": rendered_images[:, 3:4],
        "rendered_images": rendered_images,
    }
    if aux is not None:
        # [B,32,H,W] view with channels_last strides: no copy; detached (inference hand-off)
        out["refiner_input"] = aux.permute(0, 3, 1, 2)
    if refiner is not None:
        out["renders"] = refiner(rendered_images)
    return out
