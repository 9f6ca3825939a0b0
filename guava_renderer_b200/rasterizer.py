"""Host-side mirror of the reference's rasterizer interface, on top of the C ABI.

Mirrors, name for name and argument for argument:
  * the extension module `_C` of the reference
    (DGR32/ext.cpp:15-18, DGR32/rasterize_points.cu:35-244):
    `rasterize_gaussians`, `rasterize_gaussians_backward`, `mark_visible`;
  * the Python surface DGR32/diff_gaussian_rasterization_32/__init__.py:
    `GaussianRasterizationSettings` (:143-156), `_RasterizeGaussians` (:44-141),
    `GaussianRasterizer_32` (:158-207), `rasterize_gaussians` (:21-42);
and adds the batched multi-view entry.

DGR32 = /root/reference/submodules/diff-gaussian-rasterization-32.
PyTorch is used for device memory, streams and autograd plumbing only; all
arithmetic happens in libguava_rasterizer_b200.so.
"""
from __future__ import annotations

import ctypes
import importlib.util
import os
import threading
from typing import NamedTuple, Optional

import torch
import torch.nn as nn

from . import _lib

NUM_CHANNELS = 32

_tls = threading.local()


def _ptr(t: Optional[torch.Tensor]) -> int:
    """Device pointer, or NULL for None / empty tensors (the reference passes
    `torch.Tensor([])` for absent optional inputs, __init__.py:184-194)."""
    if t is None or t.numel() == 0:
        return 0
    return t.data_ptr()


def _f32c(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.float32:
        raise RuntimeError("expected a float32 tensor")
    return t if t.is_contiguous() else t.contiguous()


def _aligned16(t: torch.Tensor) -> torch.Tensor:
    t = _f32c(t)
    if t.numel() and (t.data_ptr() & 15):
        t = t.clone()   # fresh allocations are 512-byte aligned
    return t


class _CallBuffers:
    """The three resizable byte buffers of one forward call
    (reference: resizeFunctional, rasterize_points.cu:27-33)."""
    __slots__ = ("device", "bufs")

    def __init__(self, device):
        self.device = device
        self.bufs = [None, None, None]

    def alloc(self, which: int, nbytes: int) -> int:
        t = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
        self.bufs[which] = t
        return t.data_ptr()


def _alloc_trampoline(user, nbytes):
    try:
        return _tls.call.alloc(int(user or 0), nbytes)
    except Exception:   # out of memory etc.: report NULL to the C side
        _tls.alloc_error = True
        return 0


_ALLOC = _lib.ALLOC_FN(_alloc_trampoline)
_USER = [ctypes.c_void_p(0), ctypes.c_void_p(1), ctypes.c_void_p(2)]


def _check(rc: int) -> None:
    if rc == _lib.GUAVA_OK:
        return
    msg = _lib.last_error()
    if rc == _lib.GUAVA_ERR_ALLOC:
        raise torch.cuda.OutOfMemoryError(msg) if hasattr(torch.cuda, "OutOfMemoryError") else RuntimeError(msg)
    raise RuntimeError(msg)


def _stream_ptr(device) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class _device_of:
    """Make `t.device` current for the duration of a call (the reference has no
    device guard and silently uses the current device)."""
    __slots__ = ("ctx",)

    def __init__(self, t: torch.Tensor):
        if not t.is_cuda:
            raise RuntimeError("guava_renderer_b200: tensors must live on a CUDA device (there is no CPU path)")
        self.ctx = None
        if t.device.index != torch.cuda.current_device():
            self.ctx = torch.cuda.device(t.device)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *a):
        if self.ctx is not None:
            self.ctx.__exit__(*a)


# --------------------------------------------------------------------------
# `_C`-level functions (same names / positional arguments as the reference's
# pybind module)
# --------------------------------------------------------------------------
def rasterize_gaussians_C(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                          cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                          image_width, sh, degree, campos, prefiltered, antialiasing, debug):
    """rasterize_points.cu:35-124 RasterizeGaussiansCUDA.
    Returns (num_rendered, out_color[32,H,W], radii[P], geomBuffer, binningBuffer, imgBuffer, out_invdepth[1,H,W])."""
    lib = _lib.load()
    if means3D.ndimension() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    P = int(means3D.size(0))
    H, W = int(image_height), int(image_width)
    dev = means3D.device
    with _device_of(means3D):
        fopts = dict(dtype=torch.float32, device=dev)
        M = int(sh.size(1)) if (sh is not None and sh.numel() != 0 and sh.dim() > 1) else 0
        if P == 0:
            z = torch.empty(0, dtype=torch.uint8, device=dev)
            return (0, torch.zeros(NUM_CHANNELS, H, W, **fopts), torch.zeros(0, dtype=torch.int32, device=dev),
                    z, z.clone(), z.clone(), torch.zeros(1, H, W, **fopts))
        if colors is None or colors.numel() == 0:
            # rasterizer_impl.cu:244-247
            raise RuntimeError("For non-RGB, provide precomputed Gaussian colors!")
        # every pixel / every radius is written by the kernels: no fills needed
        # (the reference memsets 33.5 MB per 512x512 view, rasterize_points.cu:69-76)
        out_color = torch.empty(NUM_CHANNELS, H, W, **fopts)
        out_invdepth = torch.empty(1, H, W, **fopts)
        radii = torch.empty(P, dtype=torch.int32, device=dev)
        bufs = _CallBuffers(dev)
        _tls.call = bufs
        keep = [_f32c(background), _f32c(means3D), _aligned16(colors), _f32c(opacity),
                None if scales is None or scales.numel() == 0 else _f32c(scales),
                None if rotations is None or rotations.numel() == 0 else _f32c(rotations),
                None if cov3D_precomp is None or cov3D_precomp.numel() == 0 else _f32c(cov3D_precomp),
                _f32c(viewmatrix), _f32c(projmatrix), _f32c(campos)]
        rendered = ctypes.c_int(0)
        rc = lib.guava_rasterizer_forward(
            _ALLOC, _USER[0], _ALLOC, _USER[1], _ALLOC, _USER[2],
            P, int(degree), M, _ptr(keep[0]), W, H,
            _ptr(keep[1]), _ptr(sh) if (sh is not None and sh.is_cuda) else 0, _ptr(keep[2]), _ptr(keep[3]),
            _ptr(keep[4]), float(scale_modifier), _ptr(keep[5]), _ptr(keep[6]),
            _ptr(keep[7]), _ptr(keep[8]), _ptr(keep[9]),
            float(tan_fovx), float(tan_fovy), int(bool(prefiltered)),
            _ptr(out_color), _ptr(out_invdepth), int(bool(antialiasing)), _ptr(radii), int(bool(debug)),
            _stream_ptr(dev), ctypes.byref(rendered))
        _tls.call = None
        _check(rc)
        return (int(rendered.value), out_color, radii, bufs.bufs[0], bufs.bufs[1], bufs.bufs[2], out_invdepth)


def rasterize_gaussians_backward_C(background, means3D, radii, colors, opacities, scales, rotations,
                                   scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                                   dL_dout_color, dL_dout_invdepth, sh, degree, campos, geomBuffer, R,
                                   binningBuffer, imageBuffer, antialiasing, debug):
    """rasterize_points.cu:126-223 RasterizeGaussiansBackwardCUDA.
    Returns (dL_dmeans2D[P,3], dL_dcolors[P,32], dL_dopacity[P,1], dL_dmeans3D[P,3], dL_dcov3D[P,6],
             dL_dsh[P,M,3], dL_dscales[P,3], dL_drotations[P,4])."""
    lib = _lib.load()
    P = int(means3D.size(0))
    H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    dev = means3D.device
    M = int(sh.size(1)) if (sh is not None and sh.numel() != 0 and sh.dim() > 1) else 0
    with _device_of(means3D):
        has_inv = dL_dout_invdepth is not None and dL_dout_invdepth.numel() != 0 and dL_dout_invdepth.size(0) != 0
        # one zero fill for all gradient tensors (the reference issues ten)
        # dL_dcolors first: its rows must stay 16-byte aligned for the vector reductions
        widths = [32, 3, 1, 3, 6, 3, 4, 4] + ([1] if has_inv else [])
        flat = torch.zeros(P * sum(widths), dtype=torch.float32, device=dev)
        parts, o = [], 0
        for w in widths:
            parts.append(flat[o:o + P * w].view(P, w))
            o += P * w
        dL_dcolors, dL_dmeans2D, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dscales, dL_drotations, dL_dconic = parts[:8]
        dL_dinvdepths = parts[8] if has_inv else None
        dL_dsh = torch.zeros(P, M, 3, dtype=torch.float32, device=dev)
        if P != 0:
            keep = [_f32c(background), _f32c(means3D), _aligned16(colors), _f32c(opacities),
                    None if scales is None or scales.numel() == 0 else _f32c(scales),
                    None if rotations is None or rotations.numel() == 0 else _f32c(rotations),
                    None if cov3D_precomp is None or cov3D_precomp.numel() == 0 else _f32c(cov3D_precomp),
                    _f32c(viewmatrix), _f32c(projmatrix), _f32c(campos), _f32c(dL_dout_color),
                    _f32c(dL_dout_invdepth) if has_inv else None, radii.contiguous()]
            rc = lib.guava_rasterizer_backward(
                P, int(degree), M, int(R), _ptr(keep[0]), W, H,
                _ptr(keep[1]), 0, _ptr(keep[2]), _ptr(keep[3]), _ptr(keep[4]), float(scale_modifier),
                _ptr(keep[5]), _ptr(keep[6]), _ptr(keep[7]), _ptr(keep[8]), _ptr(keep[9]),
                float(tan_fovx), float(tan_fovy), _ptr(keep[12]),
                _ptr(geomBuffer), _ptr(binningBuffer), _ptr(imageBuffer),
                _ptr(keep[10]), _ptr(keep[11]),
                _ptr(dL_dmeans2D), _ptr(dL_dconic), _ptr(dL_dopacity), _ptr(dL_dcolors), _ptr(dL_dinvdepths),
                _ptr(dL_dmeans3D), _ptr(dL_dcov3D), 0, _ptr(dL_dscales), _ptr(dL_drotations),
                int(bool(antialiasing)), int(bool(debug)), _stream_ptr(dev))
            _check(rc)
        return (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)


def mark_visible_C(means3D, viewmatrix, projmatrix):
    """rasterize_points.cu:225-244 markVisible."""
    lib = _lib.load()
    P = int(means3D.size(0))
    with _device_of(means3D):
        present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
        if P != 0:
            m, v, p = _f32c(means3D), _f32c(viewmatrix), _f32c(projmatrix)
            rc = lib.guava_rasterizer_mark_visible(P, _ptr(m), _ptr(v), _ptr(p), _ptr(present),
                                                   _stream_ptr(means3D.device))
            _check(rc)
        return present


class _CModule:
    """The reference's `_C` module (ext.cpp:15-18) over ctypes: the same three functions, marshalled in Python.
    Kept as the binding of last resort for hosts without a C++ toolchain and for the ABI tests; the default `_C`
    is the compiled module below."""
    rasterize_gaussians = staticmethod(rasterize_gaussians_C)
    rasterize_gaussians_backward = staticmethod(rasterize_gaussians_backward_C)
    mark_visible = staticmethod(mark_visible_C)


TORCH_SO_PATH = os.path.join(os.path.dirname(_lib.SO_PATH), "_guava_torch.so")


def _load_compiled_binding():
    """lib/_guava_torch.so (csrc/torch_binding.cpp, built by build.py with g++): the pybind module that replaces the
    reference's `_C` -- rasterize_gaussians / rasterize_gaussians_backward / mark_visible with the reference's
    positional arguments, plus the batched pair.  Per-call host cost is microseconds, as for the reference's own
    binding; the ctypes marshalling cost ~0.4 ms per forward+backward."""
    _lib.load()          # the C ABI library first: fails loudly when the CUDA library itself is missing
    if not os.path.isfile(TORCH_SO_PATH):
        raise _lib.GuavaLibraryError(
            f"{TORCH_SO_PATH} is missing: the torch binding has not been built (run `python guava_renderer_b200/build.py`); "
            "set GUAVA_BINDING=ctypes to use the ctypes binding of the same C ABI instead")
    spec = importlib.util.spec_from_file_location("_guava_torch", TORCH_SO_PATH)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


BINDING = os.environ.get("GUAVA_BINDING", "compiled")
if BINDING not in ("compiled", "ctypes"):
    raise ValueError("GUAVA_BINDING must be 'compiled' or 'ctypes'")
_C = _load_compiled_binding() if BINDING == "compiled" else _CModule()


# --------------------------------------------------------------------------
# Python surface (same as DGR32/diff_gaussian_rasterization_32/__init__.py)
# --------------------------------------------------------------------------
def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        rs = raster_settings
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepths = _C.rasterize_gaussians(
            rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
            rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
            rs.sh_degree, rs.campos, rs.prefiltered, rs.antialiasing, rs.debug)
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities,
                              geomBuffer, binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        # an output the loss does not use arrives as None instead of a zero-filled tensor (the reference
        # lets autograd materialise both, two fills per view); a None invdepth gradient takes the
        # "no invdepth gradients" branch of rasterize_points.cu:176-182, which yields the same result
        ctx.set_materialize_grads(False)
        return color, radii, invdepths

    @staticmethod
    def backward(ctx, grad_out_color, _, grad_out_depth):
        rs = ctx.raster_settings
        if grad_out_color is None:
            grad_out_color = torch.zeros(NUM_CHANNELS, int(rs.image_height), int(rs.image_width),
                                         dtype=torch.float32, device=ctx.saved_tensors[1].device)
        if grad_out_depth is None:      # "no invdepth gradients": size(0) == 0, rasterize_points.cu:176-182
            grad_out_depth = torch.empty(0, dtype=torch.float32)
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities,
         geomBuffer, binningBuffer, imgBuffer) = ctx.saved_tensors
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh,
         grad_scales, grad_rotations) = _C.rasterize_gaussians_backward(
            rs.bg, means3D, radii, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
            cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color,
            grad_out_depth, sh, rs.sh_degree, rs.campos, geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer,
            rs.antialiasing, rs.debug)
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales,
                grad_rotations, grad_cov3Ds_precomp, None)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    antialiasing: bool


class GaussianRasterizer_32(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        rs = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, rs)


GaussianRasterizer = GaussianRasterizer_32   # un-suffixed alias named by BASELINE.json


# --------------------------------------------------------------------------
# Batched multi-view entry (replaces the Python loop of gaussian_render.py:37-67)
# --------------------------------------------------------------------------
class BatchRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfov: torch.Tensor        # [V,2] device: (tanfovx, tanfovy) per view
    bg: torch.Tensor            # [V,32]
    scale_modifier: float
    viewmatrix: torch.Tensor    # [V,4,4]
    projmatrix: torch.Tensor    # [V,4,4]
    debug: bool = False
    antialiasing: bool = False


def _vstride(t: torch.Tensor, V: int, per_view_numel: int) -> int:
    """0 when one tensor is shared by all views, else elements per view."""
    if t.numel() == per_view_numel:
        return 0
    if t.numel() != V * per_view_numel:
        raise RuntimeError("tensor must be [P,...] (shared) or [V,P,...] (per view)")
    return per_view_numel


# per Gaussian and view: dL_dcolors first (its rows must stay 16-byte aligned for the vector reductions),
# mean2D, opacity, mean3D, cov3D, scale, rotation, conic, invdepth
_GRAD_WIDTHS = [32, 3, 1, 3, 6, 3, 4, 4, 1]


class _RasterizeGaussiansBatch(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, colors_precomp, opacities, scales, rotations, settings, call=None):
        st = settings
        V = int(st.viewmatrix.shape[0])
        P = int(means3D.shape[-2])
        H, W = int(st.image_height), int(st.image_width)
        dev = means3D.device
        # per-call options (guava_call_options): refiner hand-off buffer, the caller's pinned status words,
        # no-host-wait mode -- given by render_views(refiner_input=True) / GraphedRenderStep
        call = call or {}
        aux = call.get("refiner_output")
        status = call.get("status_words")
        no_wait = bool(call.get("no_host_wait", False))
        # the training step's packed [P,43] buffer: the backward then sums over the views inside its kernels
        # (guava_rasterizer_backward_batch_summed) and hands autograd no per-view gradients
        ctx.summed_out = call.get("summed_out")
        fwd_only = not any(ctx.needs_input_grad)      # no gradient can be asked for: inference
        # the backward's [V,P,.] gradients must start at zero: the block is allocated here and zero-filled by the forward
        # under its tile scan (guava_call_options::zero_fill) instead of by a 9 us fill kernel in front of the backward
        ctx.grad_block = None
        zero_fill = ctx.summed_out
        if zero_fill is None and not fwd_only and V * P:
            ctx.grad_block = zero_fill = torch.empty(V * P * sum(_GRAD_WIDTHS), dtype=torch.float32, device=dev)
        m, c, o = _f32c(means3D), _f32c(colors_precomp), _f32c(opacities)
        s, r = _f32c(scales), _f32c(rotations)
        bg, vm, pm, tf = _f32c(st.bg), _f32c(st.viewmatrix), _f32c(st.projmatrix), _f32c(st.tanfov)
        strides = (_vstride(m, V, P * 3), _vstride(c, V, P * NUM_CHANNELS), _vstride(o, V, P),
                   _vstride(s, V, P * 3), _vstride(r, V, P * 4))
        if BINDING == "compiled":
            rendered, out_color, radii, geom, binning, img, out_invdepth = _C.rasterize_gaussians_batch(
                bg, m, c, o, s, r, float(st.scale_modifier), vm, pm, tf, H, W, bool(st.antialiasing), bool(st.debug),
                aux, status.data_ptr() if status is not None else 0, no_wait, fwd_only, zero_fill)
            bufs = [geom, binning, img]
        else:
            lib = _lib.load()
            with _device_of(means3D):
                fopts = dict(dtype=torch.float32, device=dev)
                out_color = torch.empty(V, NUM_CHANNELS, H, W, **fopts)
                out_invdepth = torch.empty(V, 1, H, W, **fopts)
                radii = torch.empty(V, P, dtype=torch.int32, device=dev)
                if V * P == 0:
                    out_color.zero_(); out_invdepth.zero_()
                c = _aligned16(c)
                cb = _CallBuffers(dev)
                _tls.call = cb
                rendered = ctypes.c_int(0)
                if aux is not None:
                    if aux.dtype != torch.bfloat16 or tuple(aux.shape) != (V, H, W, NUM_CHANNELS) or not aux.is_contiguous():
                        raise ValueError("refiner output must be a contiguous bfloat16 [V,H,W,32] tensor")
                    if V * P == 0:
                        aux.zero_()
                zf = zero_fill
                if zf is not None and (not zf.is_contiguous() or zf.device != dev):
                    raise ValueError("summed_out must be a contiguous tensor on the inputs' device")
                if zf is not None and V * P == 0:
                    zf.zero_()
                copts = _lib.CallOptions(refiner_output=_ptr(aux) if aux is not None else 0,
                                         status_words=status.data_ptr() if status is not None else 0, no_host_wait=no_wait,
                                         forward_only=fwd_only, zero_fill=zf.data_ptr() if zf is not None else 0,
                                         zero_fill_bytes=zf.numel() * zf.element_size() if zf is not None else 0)
                rc = lib.guava_rasterizer_forward_batch(
                    _ALLOC, _USER[0], _ALLOC, _USER[1], _ALLOC, _USER[2],
                    V, P, _ptr(bg), W, H,
                    _ptr(m), strides[0], _ptr(c), strides[1], _ptr(o), strides[2], _ptr(s), strides[3],
                    float(st.scale_modifier), _ptr(r), strides[4],
                    _ptr(vm), _ptr(pm), _ptr(tf), _ptr(out_color), _ptr(out_invdepth),
                    int(bool(st.antialiasing)), _ptr(radii), int(bool(st.debug)), _stream_ptr(dev),
                    ctypes.byref(rendered), ctypes.byref(copts))
                _tls.call = None
                _check(rc)
                rendered = int(rendered.value)
                bufs = cb.bufs
        _tls.last_buffers = bufs      # (geometry, binning, image) of the most recent batched forward
        # unused outputs' upstream gradients arrive as None, not as zero-filled tensors
        ctx.set_materialize_grads(False)
        ctx.settings = st
        ctx.num_rendered = int(rendered)
        ctx.strides = strides
        ctx.dims = (V, P, H, W)
        ctx.save_for_backward(m, c, o, s, r, bg, vm, pm, tf, radii, bufs[0], bufs[1], bufs[2])
        ctx.mark_non_differentiable(radii)
        return out_color, radii, out_invdepth

    @staticmethod
    def backward(ctx, grad_color, _, grad_invdepth):
        st = ctx.settings
        V, P, H, W = ctx.dims
        m, c, o, s, r, bg, vm, pm, tf, radii, geom, binning, img = ctx.saved_tensors
        dev = m.device
        has_inv = grad_invdepth is not None
        if grad_color is None:
            grad_color = torch.zeros(V, NUM_CHANNELS, H, W, dtype=torch.float32, device=dev)
        if ctx.summed_out is not None:
            pk = ctx.summed_out
            if pk.dtype != torch.float32 or tuple(pk.shape) != (P, 43) or not pk.is_contiguous() or pk.device != dev:
                raise ValueError("summed_out must be a contiguous float32 [P,43] tensor on the inputs' device")
            if BINDING == "compiled":
                _C.rasterize_gaussians_batch_backward_summed(
                    bg, m, radii, c, o, s, r, float(st.scale_modifier), vm, pm, tf, grad_color,
                    grad_invdepth if has_inv else None, geom, int(ctx.num_rendered), binning, img,
                    bool(st.antialiasing), bool(st.debug), pk)
            elif V * P:
                with _device_of(m):
                    gc = _f32c(grad_color)
                    gi = _f32c(grad_invdepth) if has_inv else None
                    b0 = pk.data_ptr()
                    _check(_lib.load().guava_rasterizer_backward_batch_summed(
                        V, P, int(ctx.num_rendered), _ptr(bg), W, H,
                        _ptr(m), ctx.strides[0], _ptr(c), ctx.strides[1], _ptr(o), ctx.strides[2],
                        _ptr(s), ctx.strides[3], float(st.scale_modifier), _ptr(r), ctx.strides[4],
                        _ptr(vm), _ptr(pm), _ptr(tf), _ptr(radii), _ptr(geom), _ptr(binning), _ptr(img),
                        _ptr(gc), _ptr(gi), b0, b0 + 12, b0 + 24, b0 + 40, b0 + 44, 43,
                        int(bool(st.antialiasing)), int(bool(st.debug)), _stream_ptr(dev)))
            return (None,) * 8
        block, ctx.grad_block = ctx.grad_block, None      # zero-filled by the forward; a repeated backward gets a fresh one
        if BINDING == "compiled":
            g_mean2D, g_color, g_opacity, g_mean3D, g_cov3D, g_scale, g_rot = _C.rasterize_gaussians_batch_backward(
                bg, m, radii, c, o, s, r, float(st.scale_modifier), vm, pm, tf, grad_color,
                grad_invdepth if has_inv else None, geom, int(ctx.num_rendered), binning, img,
                bool(st.antialiasing), bool(st.debug), block)
        else:
            lib = _lib.load()
            with _device_of(m):
                widths = _GRAD_WIDTHS if has_inv else _GRAD_WIDTHS[:8]
                n = V * P
                # one fill for all gradient tensors.  (Issuing it during the forward on a side stream was
                # measured slower inside the captured step: the fork/join costs more than the 9 us fill.)
                flat = block if block is not None else torch.zeros(n * sum(widths), dtype=torch.float32, device=dev)
                parts, off = [], 0
                for w in widths:
                    parts.append(flat[off:off + n * w].view(V, P, w))
                    off += n * w
                g_color, g_mean2D, g_opacity, g_mean3D, g_cov3D, g_scale, g_rot, g_conic = parts[:8]
                g_invd = parts[8] if has_inv else None
                if n:
                    gc = _f32c(grad_color)
                    gi = _f32c(grad_invdepth) if has_inv else None
                    rc = lib.guava_rasterizer_backward_batch(
                        V, P, int(ctx.num_rendered), _ptr(bg), W, H,
                        _ptr(m), ctx.strides[0], _ptr(c), ctx.strides[1], _ptr(o), ctx.strides[2],
                        _ptr(s), ctx.strides[3], float(st.scale_modifier), _ptr(r), ctx.strides[4],
                        _ptr(vm), _ptr(pm), _ptr(tf), _ptr(radii), _ptr(geom), _ptr(binning), _ptr(img),
                        _ptr(gc), _ptr(gi),
                        _ptr(g_mean2D), _ptr(g_conic), _ptr(g_opacity), _ptr(g_color), _ptr(g_invd),
                        _ptr(g_mean3D), _ptr(g_cov3D), _ptr(g_scale), _ptr(g_rot),
                        int(bool(st.antialiasing)), int(bool(st.debug)), _stream_ptr(dev))
                    _check(rc)

        def shape_like(g, stride, ref):
            # a tensor shared across views receives the sum of its per-view gradients
            g = g.sum(0) if stride == 0 and V != 1 else (g[0] if stride == 0 else g)
            return g.reshape(ref.shape)

        return (shape_like(g_mean3D, ctx.strides[0], m), g_mean2D, shape_like(g_color, ctx.strides[1], c),
                shape_like(g_opacity, ctx.strides[2], o), shape_like(g_scale, ctx.strides[3], s),
                shape_like(g_rot, ctx.strides[4], r), None, None)


def rasterize_gaussians_batch(means3D, means2D, colors_precomp, opacities, scales, rotations,
                              settings: BatchRasterizationSettings, refiner_output: Optional[torch.Tensor] = None,
                              status_words: Optional[torch.Tensor] = None, no_host_wait: bool = False,
                              summed_out: Optional[torch.Tensor] = None):
    """means3D [V,P,3] or [P,3]; means2D [V,P,3] (gradient sink only); colors_precomp [V,P,32] or [P,32];
    opacities [V,P,1] or [P,1]; scales [V,P,3] or [P,3]; rotations [V,P,4] or [P,4].
    Per-call options (guava_call_options): `refiner_output` bf16 [V,H,W,32] written by the compositing
    kernel's epilogue; `status_words` pinned int32[2] receiving (num_rendered, longest list | overflow << 31);
    `no_host_wait` enqueues the chain without waiting for num_rendered (CUDA-graph capture).
    `summed_out`: a float32 [P,43] tensor, zero-filled BY THE FORWARD (guava_call_options::zero_fill: on the SMs the
    one-block tile scan leaves idle); the backward then adds into it, inside its kernels, the sum over
    the views of the parameter gradients (means3D 3, scales 3, rotations 4, opacity 1, features 32 -- the training
    step's all-reduce payload, sharding.py) and the input tensors receive NO `.grad`.
    Returns (color [V,32,H,W], radii [V,P] int32, invdepth [V,1,H,W])."""
    call = None
    if refiner_output is not None or status_words is not None or no_host_wait or summed_out is not None:
        call = dict(refiner_output=refiner_output, status_words=status_words, no_host_wait=no_host_wait,
                    summed_out=summed_out)
    return _RasterizeGaussiansBatch.apply(means3D, means2D, colors_precomp, opacities, scales, rotations, settings, call)


class GaussianBatchRasterizer(nn.Module):
    def __init__(self, settings: BatchRasterizationSettings):
        super().__init__()
        self.settings = settings

    def forward(self, means3D, means2D, opacities, colors_precomp, scales, rotations):
        return rasterize_gaussians_batch(means3D, means2D, colors_precomp, opacities, scales, rotations,
                                         self.settings)
