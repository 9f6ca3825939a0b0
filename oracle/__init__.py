"""CPU oracle for the 32-channel Gaussian-splatting rasterizer.

TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / reference legs
may import this package.  ``guava_renderer_b200`` and the drop-in package
``diff_gaussian_rasterization_32`` never do, and fail loudly without their
CUDA library instead of falling back to anything in here.

``splat_oracle.c`` restates the reference algorithm
(/root/reference/submodules/diff-gaussian-rasterization-32/cuda_rasterizer/
forward.cu, backward.cu, rasterizer_impl.cu, auxiliary.h; every function cites
file:line).  This module compiles it with gcc (``build()``) and exposes numpy
entry points that mirror the reference call chain
(rasterizer_impl.cu:198-340 forward, :345-450 backward).

Parity status: see the header of splat_oracle.c and tests/golden/README.md.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "splat_oracle.c")
_BUILD = os.path.join(_HERE, "_build")
_SO = os.path.join(_BUILD, "libsplat_oracle.so")
_lib = None

NUM_CHANNELS = 32
BLOCK = 16


def build(force: bool = False) -> str:
    """gcc -O2 -ffp-contract=off (no compiler-introduced FMA: every fused op in
    the C file is an explicit fmaf) -mfma (hardware fmaf) -fopenmp."""
    os.makedirs(_BUILD, exist_ok=True)
    if (not force and os.path.isfile(_SO)
            and os.path.getmtime(_SO) >= os.path.getmtime(_SRC)):
        return _SO
    cmd = ["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-fopenmp", "-mfma",
           "-ffp-contract=off", "-fno-fast-math", "-fno-unsafe-math-optimizations",
           _SRC, "-o", _SO, "-lm"]
    subprocess.run(cmd, check=True)
    return _SO


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.isfile(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
            build()
        _lib = ctypes.CDLL(_SO)
        _lib.oracle_inclusive_sum.restype = ctypes.c_uint32
        _lib.oracle_higher_msb.restype = ctypes.c_int
        _lib.oracle_num_threads.restype = ctypes.c_int
    return _lib


def _p(a):
    if a is None:
        return ctypes.c_void_p(0)
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def set_num_threads(n: int) -> None:
    lib().oracle_set_num_threads(ctypes.c_int(int(n)))


def higher_msb(n: int) -> int:
    return int(lib().oracle_higher_msb(ctypes.c_uint32(n)))


def mark_visible(means3D, viewmatrix, projmatrix):
    means3D = _f32(means3D)
    P = means3D.shape[0]
    out = np.zeros(P, dtype=np.uint8)
    lib().oracle_mark_visible(ctypes.c_int(P), _p(means3D), _p(_f32(viewmatrix).reshape(-1)),
                              _p(_f32(projmatrix).reshape(-1)), _p(out))
    return out.astype(bool)


def forward(means3D, colors, opacities, scales, rotations, viewmatrix, projmatrix,
            tanfovx, tanfovy, H, W, bg, scale_modifier=1.0, cov3D_precomp=None,
            antialiasing=False, want_invdepth=True):
    """Full forward chain.  Returns a dict holding every intermediate the
    reference keeps in its geometry / binning / image buffers
    (rasterizer_impl.h:33-66) plus the outputs."""
    L = lib()
    means3D = _f32(means3D)
    P = int(means3D.shape[0])
    colors = _f32(colors).reshape(P, NUM_CHANNELS)
    opacities = _f32(opacities).reshape(P)
    scales = _f32(scales)
    rotations = _f32(rotations)
    cov3D_precomp = _f32(cov3D_precomp)
    V = _f32(viewmatrix).reshape(16)
    PR = _f32(projmatrix).reshape(16)
    bg = _f32(bg).reshape(NUM_CHANNELS)
    gx, gy = (W + BLOCK - 1) // BLOCK, (H + BLOCK - 1) // BLOCK

    radii = np.zeros(P, np.int32)
    means2D = np.zeros((P, 2), np.float32)
    depths = np.zeros(P, np.float32)
    cov3D = np.zeros((P, 6), np.float32)
    conic_opacity = np.zeros((P, 4), np.float32)
    tiles_touched = np.zeros(P, np.uint32)
    L.oracle_preprocess(ctypes.c_int(P), _p(means3D), _p(scales), ctypes.c_float(scale_modifier),
                        _p(rotations), _p(opacities), _p(cov3D_precomp), _p(V), _p(PR),
                        ctypes.c_int(W), ctypes.c_int(H), ctypes.c_float(tanfovx), ctypes.c_float(tanfovy),
                        ctypes.c_int(int(bool(antialiasing))),
                        _p(radii), _p(means2D), _p(depths), _p(cov3D), _p(conic_opacity), _p(tiles_touched))
    point_offsets = np.zeros(P, np.uint32)
    R = int(L.oracle_inclusive_sum(ctypes.c_int(P), _p(tiles_touched), _p(point_offsets))) if P else 0
    keys_unsorted = np.zeros(R, np.uint64)
    vals_unsorted = np.zeros(R, np.uint32)
    L.oracle_duplicate_with_keys(ctypes.c_int(P), _p(means2D), _p(depths), _p(point_offsets), _p(radii),
                                 ctypes.c_int(W), ctypes.c_int(H), _p(keys_unsorted), _p(vals_unsorted))
    bit = higher_msb(gx * gy)
    keys = np.zeros(R, np.uint64)
    point_list = np.zeros(R, np.uint32)
    L.oracle_sort_pairs(ctypes.c_uint32(R), _p(keys_unsorted), _p(vals_unsorted), _p(keys), _p(point_list),
                        ctypes.c_int(32 + bit))
    ranges = np.zeros((gx * gy, 2), np.uint32)
    L.oracle_identify_tile_ranges(ctypes.c_uint32(R), _p(keys), ctypes.c_int(gx * gy), _p(ranges))

    final_T = np.zeros(H * W, np.float32)
    n_contrib = np.zeros(H * W, np.uint32)
    out_color = np.zeros((NUM_CHANNELS, H, W), np.float32)
    out_invdepth = np.zeros((1, H, W), np.float32) if want_invdepth else None
    L.oracle_render_forward(ctypes.c_int(W), ctypes.c_int(H), _p(ranges), _p(point_list), _p(means2D),
                            _p(colors), _p(conic_opacity), _p(depths), _p(bg), _p(final_T), _p(n_contrib),
                            _p(out_color), _p(out_invdepth))
    return dict(P=P, H=H, W=W, num_rendered=R, radii=radii, means2D=means2D, depths=depths, cov3D=cov3D,
                conic_opacity=conic_opacity, tiles_touched=tiles_touched, point_offsets=point_offsets,
                keys_unsorted=keys_unsorted, point_list_unsorted=vals_unsorted, keys=keys,
                point_list=point_list, ranges=ranges, final_T=final_T, n_contrib=n_contrib,
                color=out_color, invdepth=out_invdepth,
                inputs=dict(means3D=means3D, colors=colors, opacities=opacities, scales=scales,
                            rotations=rotations, viewmatrix=V, projmatrix=PR, tanfovx=tanfovx,
                            tanfovy=tanfovy, bg=bg, scale_modifier=scale_modifier,
                            cov3D_precomp=cov3D_precomp, antialiasing=bool(antialiasing)))


def backward(fwd, dL_dcolor, dL_dinvdepth=None):
    """Backward chain (rasterizer_impl.cu:345-450).  ``fwd`` is the dict
    returned by forward().  Returns the 8 gradients of
    rasterize_points.cu:222 plus the two internal ones."""
    L = lib()
    P, H, W = fwd["P"], fwd["H"], fwd["W"]
    i = fwd["inputs"]
    dL_dcolor = _f32(dL_dcolor).reshape(NUM_CHANNELS, H, W)
    dL_dinvdepth = None if dL_dinvdepth is None else _f32(dL_dinvdepth).reshape(H * W)
    dL_dmean2D = np.zeros((P, 3), np.float32)
    dL_dconic = np.zeros((P, 4), np.float32)
    dL_dopacity = np.zeros((P, 1), np.float32)
    dL_dcolors = np.zeros((P, NUM_CHANNELS), np.float32)
    dL_dinvdepths = np.zeros((P, 1), np.float32) if dL_dinvdepth is not None else None
    L.oracle_render_backward(ctypes.c_int(P), ctypes.c_int(W), ctypes.c_int(H), _p(fwd["ranges"]),
                             _p(fwd["point_list"]), _p(i["bg"]), _p(fwd["means2D"]), _p(fwd["conic_opacity"]),
                             _p(i["colors"]), _p(fwd["depths"]), _p(fwd["final_T"]), _p(fwd["n_contrib"]),
                             _p(dL_dcolor), _p(dL_dinvdepth), _p(dL_dmean2D), _p(dL_dconic), _p(dL_dopacity),
                             _p(dL_dcolors), _p(dL_dinvdepths))
    dL_dmeans3D = np.zeros((P, 3), np.float32)
    dL_dcov3D = np.zeros((P, 6), np.float32)
    dL_dscales = np.zeros((P, 3), np.float32)
    dL_drots = np.zeros((P, 4), np.float32)
    use_precomp = i["cov3D_precomp"] is not None
    cov3Ds = i["cov3D_precomp"] if use_precomp else fwd["cov3D"]
    L.oracle_preprocess_backward(ctypes.c_int(P), _p(i["means3D"]), _p(fwd["radii"]),
                                 _p(None if use_precomp else i["scales"]),
                                 _p(None if use_precomp else i["rotations"]),
                                 ctypes.c_float(i["scale_modifier"]), _p(cov3Ds), _p(i["opacities"]),
                                 _p(i["viewmatrix"]), _p(i["projmatrix"]), ctypes.c_int(W), ctypes.c_int(H),
                                 ctypes.c_float(i["tanfovx"]), ctypes.c_float(i["tanfovy"]),
                                 ctypes.c_int(int(i["antialiasing"])), _p(dL_dmean2D), _p(dL_dconic),
                                 _p(dL_dinvdepths), _p(dL_dopacity), _p(dL_dmeans3D), _p(dL_dcov3D),
                                 _p(dL_dscales), _p(dL_drots))
    return dict(dL_dmeans2D=dL_dmean2D, dL_dcolors=dL_dcolors, dL_dopacity=dL_dopacity,
                dL_dmeans3D=dL_dmeans3D, dL_dcov3D=dL_dcov3D, dL_dscales=dL_dscales,
                dL_drotations=dL_drots, dL_dconic=dL_dconic, dL_dinvdepths=dL_dinvdepths)
