"""Fused deformation of GUAVA's Gaussians into rasterizer inputs.

Reference: models/UbodyAvatar/ubody_gaussian.py:245-289 (`Ubody_Gaussian.forward`).  There the
deformed cloud is produced by ~25 torch ops (face frames, [B,N,3,3] gathers, two einsums, roma
quaternion helpers, five `torch.cat`).  `deform_gaussians` takes the same ingredients and returns the
same `deformed_assets` dictionary from two kernels of libguava_rasterizer_b200.so
(csrc/deform.cu), differentiable w.r.t. the per-Gaussian attributes (uv_local_xyz, uv_rotation,
uv_scaling, smplx_rotation, and -- being plain slices of the output -- smplx_scaling, both opacities and
both feature tensors).  The mesh (deformed vertices, per-vertex transforms) is an input without
gradient, as in GUAVA where it comes from tracked SMPLX/FLAME coefficients (ubody_gaussian.py:250).
There is no CPU path: host tensors raise.
"""
from __future__ import annotations

import ctypes
from typing import Dict

import torch

from . import _lib

_ATTRS = ("smplx_rotation", "smplx_scaling", "smplx_opacity", "smplx_features_color",
          "uv_local_xyz", "uv_rotation", "uv_scaling", "uv_opacity", "uv_features_color")


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr() if t is not None and t.numel() else 0)


def _f32(t, shape, name):
    if not t.is_cuda:
        raise _lib.GuavaLibraryError(f"{name}: deform_gaussians has no CPU path (got a host tensor)")
    if t.dtype != torch.float32 or tuple(t.shape) != tuple(shape):
        raise ValueError(f"{name}: expected float32 {tuple(shape)}, got {t.dtype} {tuple(t.shape)}")
    return t.contiguous()


class _DeformGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, vertices, ver_transform, faces, binding, bary, smplx_rotation, smplx_scaling, smplx_opacity,
                smplx_features, uv_local_xyz, uv_rotation, uv_scaling, uv_opacity, uv_features):
        B, Nv = int(vertices.shape[0]), int(vertices.shape[1])
        N, F = int(binding.shape[0]), int(faces.shape[0])
        dev = vertices.device
        vertices = _f32(vertices, (B, Nv, 3), "vertices")
        ver_transform = _f32(ver_transform, (B, Nv, 4, 4), "ver_transform_mat")
        faces = faces.to(device=dev, dtype=torch.int32).contiguous()
        binding = binding.to(device=dev, dtype=torch.int32).contiguous()
        bary = _f32(bary, (N, 3), "uv_face_bary")
        smplx_rotation = _f32(smplx_rotation, (B, Nv, 4), "smplx_rotation")
        smplx_scaling = _f32(smplx_scaling, (B, Nv, 3), "smplx_scaling")
        smplx_opacity = _f32(smplx_opacity, (B, Nv, 1), "smplx_opacity")
        C = int(smplx_features.shape[-1])
        if C != 32:
            raise ValueError("features_color must have 32 channels")
        smplx_features = _f32(smplx_features, (B, Nv, C), "smplx_features_color")
        uv_local_xyz = _f32(uv_local_xyz, (B, N, 3), "uv_local_xyz")
        uv_rotation = _f32(uv_rotation, (B, N, 4), "uv_rotation")
        uv_scaling = _f32(uv_scaling, (B, N, 3), "uv_scaling")
        uv_opacity = _f32(uv_opacity, (B, N, 1), "uv_opacity")
        uv_features = _f32(uv_features, (B, N, C), "uv_features_color")
        P = Nv + N
        frames = torch.empty(B, max(F, 1), 16, device=dev)
        xyz = torch.empty(B, P, 3, device=dev)
        rot = torch.empty(B, P, 4, device=dev)
        sc = torch.empty(B, P, 3, device=dev)
        op = torch.empty(B, P, 1, device=dev)
        feat = torch.empty(B, P, C, device=dev)
        stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        with torch.cuda.device(dev):
            rc = _lib.load().guava_deform_gaussians_forward(
                B, Nv, N, F, _ptr(vertices), _ptr(ver_transform), _ptr(faces), _ptr(binding), _ptr(bary),
                _ptr(smplx_rotation), _ptr(smplx_scaling), _ptr(smplx_opacity), _ptr(smplx_features),
                _ptr(uv_local_xyz), _ptr(uv_rotation), _ptr(uv_scaling), _ptr(uv_opacity), _ptr(uv_features),
                _ptr(frames), _ptr(xyz), _ptr(rot), _ptr(sc), _ptr(op), _ptr(feat), stream)
        if rc != _lib.GUAVA_OK:
            raise _lib.GuavaLibraryError(f"guava_deform_gaussians_forward failed ({rc}): {_lib.last_error()}")
        ctx.save_for_backward(ver_transform, binding, smplx_rotation, frames)
        ctx.dims = (B, Nv, N, F)
        return xyz, rot, sc, op, feat

    @staticmethod
    def backward(ctx, g_xyz, g_rot, g_sc, g_op, g_feat):
        ver_transform, binding, smplx_rotation, frames = ctx.saved_tensors
        B, Nv, N, F = ctx.dims
        dev = frames.device
        z = lambda g, w: (torch.zeros(B, Nv + N, w, device=dev) if g is None else g.contiguous())
        g_xyz, g_rot, g_sc = z(g_xyz, 3), z(g_rot, 4), z(g_sc, 3)
        d_srot = torch.empty(B, Nv, 4, device=dev)
        d_loc = torch.empty(B, N, 3, device=dev)
        d_urot = torch.empty(B, N, 4, device=dev)
        d_usc = torch.empty(B, N, 3, device=dev)
        stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        with torch.cuda.device(dev):
            rc = _lib.load().guava_deform_gaussians_backward(
                B, Nv, N, F, _ptr(ver_transform), _ptr(binding), _ptr(smplx_rotation), _ptr(frames),
                _ptr(g_xyz), _ptr(g_rot), _ptr(g_sc), _ptr(d_srot), _ptr(d_loc), _ptr(d_urot), _ptr(d_usc), stream)
        if rc != _lib.GUAVA_OK:
            raise _lib.GuavaLibraryError(f"guava_deform_gaussians_backward failed ({rc}): {_lib.last_error()}")
        # the pass-through gradients as contiguous tensors of their own, one kernel (as slices of [B,Nv+N,w] each would be
        # cloned by autograd's accumulation into the leaf)
        want = ctx.needs_input_grad
        g_op = None if g_op is None else g_op.contiguous()
        g_feat = None if g_feat is None else g_feat.contiguous()
        new = lambda cond, n, w: torch.empty(B, n, w, device=dev) if cond else None
        d_ssc = new(want[6], Nv, 3)
        d_sop, d_uop = new(want[7] and g_op is not None, Nv, 1), new(want[12] and g_op is not None, N, 1)
        d_sft, d_uft = new(want[8] and g_feat is not None, Nv, 32), new(want[13] and g_feat is not None, N, 32)
        if any(t is not None for t in (d_ssc, d_sop, d_uop, d_sft, d_uft)):
            with torch.cuda.device(dev):
                rc = _lib.load().guava_deform_split_gradients(
                    B, Nv, N, _ptr(g_sc), _ptr(g_op), _ptr(g_feat), _ptr(d_ssc), _ptr(d_sop), _ptr(d_sft), _ptr(d_uop),
                    _ptr(d_uft), stream)
            if rc != _lib.GUAVA_OK:
                raise _lib.GuavaLibraryError(f"guava_deform_split_gradients failed ({rc}): {_lib.last_error()}")
        return (None, None, None, None, None, d_srot, d_ssc, d_sop, d_sft, d_loc, d_urot, d_usc, d_uop, d_uft)


def deform_gaussians(vertices: torch.Tensor, ver_transform_mat: torch.Tensor, faces: torch.Tensor,
                     uv_binding_face: torch.Tensor, uv_face_bary: torch.Tensor, *, smplx_rotation, smplx_scaling,
                     smplx_opacity, smplx_features_color, uv_local_xyz, uv_rotation, uv_scaling, uv_opacity,
                     uv_features_color, sh_degree: int = 0, validate: bool = False) -> Dict[str, torch.Tensor]:
    """`Ubody_Gaussian.forward` (ubody_gaussian.py:245-289) after the EHM call: `vertices` and
    `ver_transform_mat` are `smplx_deform_res["vertices"]` / `["ver_transform_mat"]`, `faces` is
    `smplx.faces_tensor`; the keyword tensors are the module's `_smplx_*` / `_uv_*` attributes.
    Returns the reference's `deformed_assets` dictionary."""
    if validate and uv_binding_face.numel():   # costs a device->host sync; the static binding is checked once by callers
        if int(uv_binding_face.max()) >= int(faces.shape[0]) or int(uv_binding_face.min()) < 0:
            raise IndexError("uv_binding_face out of range")
        if int(faces.max()) >= int(vertices.shape[1]) or int(faces.min()) < 0:
            raise IndexError("faces out of range")
    if vertices.requires_grad or ver_transform_mat.requires_grad:
        raise NotImplementedError("deform_gaussians: the mesh is an input without gradient")
    xyz, rot, sc, op, feat = _DeformGaussians.apply(
        vertices, ver_transform_mat, faces, uv_binding_face, uv_face_bary, smplx_rotation, smplx_scaling,
        smplx_opacity, smplx_features_color, uv_local_xyz, uv_rotation, uv_scaling, uv_opacity, uv_features_color)
    return {"features_color": feat, "xyz": xyz, "rotation": rot, "scaling": sc, "opacity": op,
            "sh_degree": sh_degree, "smplx_xyz_deform": vertices}
