"""Shared helpers for the parity tests (test infrastructure; may import oracle/)."""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")

from guava_renderer_b200 import synthetic  # noqa: E402
import guava_renderer_b200 as grb  # noqa: E402


def inputs_for(P, H, W, seed, view=0, num_views=1, scale_mult=1.0, bg=None):
    c = synthetic.guava_cloud(P, seed, scale_mult)
    tanfovy = 1.0 / 24
    tanfovx = tanfovy * W / H
    cam = synthetic.camera(view, num_views, tanfovx, tanfovy)
    if bg is None:
        bg = torch.zeros(32)
    return dict(xyz=c["xyz"], features=c["features_color"], opacity=c["opacity"], scaling=c["scaling"],
                rotation=c["rotation"], viewmatrix=cam["viewmatrix"], projmatrix=cam["projmatrix"],
                campos=cam["campos"], tanfovx=tanfovx, tanfovy=tanfovy, H=H, W=W, bg=bg.float())


def oracle_forward(inp, antialiasing=False, scale_modifier=1.0):
    import oracle
    return oracle.forward(inp["xyz"].numpy(), inp["features"].numpy(), inp["opacity"].numpy(),
                          inp["scaling"].numpy(), inp["rotation"].numpy(), inp["viewmatrix"].numpy(),
                          inp["projmatrix"].numpy(), inp["tanfovx"], inp["tanfovy"], inp["H"], inp["W"],
                          inp["bg"].numpy(), scale_modifier=scale_modifier, antialiasing=antialiasing)


def settings_for(mod, inp, device, antialiasing=False, debug=False, scale_modifier=1.0):
    return mod.GaussianRasterizationSettings(
        image_height=inp["H"], image_width=inp["W"], tanfovx=inp["tanfovx"], tanfovy=inp["tanfovy"],
        bg=inp["bg"].to(device), scale_modifier=scale_modifier, viewmatrix=inp["viewmatrix"].to(device),
        projmatrix=inp["projmatrix"].to(device), sh_degree=0, campos=inp["campos"].to(device),
        prefiltered=False, debug=debug, antialiasing=antialiasing)


def c_forward(mod, inp, device, antialiasing=False, scale_modifier=1.0):
    """Call `_C.rasterize_gaussians` of `mod` (ours or the reference build) with the 20 positional
    arguments of DGR32/diff_gaussian_rasterization_32/__init__.py:61-82."""
    e = torch.Tensor([])
    d = lambda k: inp[k].to(device)
    return mod._C.rasterize_gaussians(
        d("bg"), d("xyz"), d("features"), d("opacity"), d("scaling"), d("rotation"), scale_modifier, e,
        d("viewmatrix"), d("projmatrix"), inp["tanfovx"], inp["tanfovy"], inp["H"], inp["W"], e, 0, d("campos"),
        False, antialiasing, False)


def decode_ours(geom, binning, img, V, P, W, H, capacity=None):
    """numpy views of the arrays inside our three opaque buffers.  The layout is the one THE CALL used
    (guava_buffer_layout_of_call reads the sort mode, the block-mask flag and the binning capacity the call
    resolved from the image buffer's header), not what the process-wide options say by now; `capacity` is
    accepted for older callers and ignored."""
    L, mode, masks, cap = grb._lib_mod.buffer_layout_of_call(img.data_ptr(), V, P, W, H)
    g = geom.cpu().numpy()
    i = img.cpu().numpy()
    b = binning.cpu().numpy()
    n = V * P
    tiles = V * ((W + 15) // 16) * ((H + 15) // 16)
    rec = g[L.splat_rec:L.splat_rec + n * 32].view(np.float32).reshape(n, 8)
    out = dict(
        means2D=rec[:, 0:2].copy(), conic_opacity=np.concatenate([rec[:, 2:5], rec[:, 5:6]], 1),
        rdepth=rec[:, 6].copy(),
        depths=g[L.depths:L.depths + n * 4].view(np.float32).copy(),
        rects=g[L.rects:L.rects + n * 8].view(np.uint16).reshape(n, 4).copy(),
        tiles_touched=g[L.tiles_touched:L.tiles_touched + n * 4].view(np.uint32).copy(),
        final_T=i[L.final_T:L.final_T + V * H * W * 4].view(np.float32).copy(),
        n_contrib=i[L.n_contrib:L.n_contrib + V * H * W * 4].view(np.uint32).copy(),
        ranges=i[L.ranges:L.ranges + tiles * 8].view(np.uint32).reshape(tiles, 2).copy(),
        num_rendered=int(i[L.num_rendered:L.num_rendered + 4].view(np.uint32)[0]),
    )
    R = out["num_rendered"]
    hdr = i[L.num_rendered:L.num_rendered + 32].view(np.uint32)
    out["sort_mode"], out["block_masks"], out["capacity"] = mode, masks, cap
    out["point_list"] = b[L.point_list:L.point_list + R * 4].view(np.uint32).copy()
    if masks:   # one byte per (list entry, 8x4 block of its tile): != 0 where the block got nothing from the entry
        plane = (cap + 255) // 256 * 256
        out["taken_marks"] = b[L.taken_marks:L.taken_marks + 8 * plane].reshape(8, plane)[:, :R].copy()
    if mode == 1:
        out["point_offsets"] = g[L.point_offsets:L.point_offsets + n * 4].view(np.uint32).copy()
    if mode == 1 or ((hdr[5] >> 16) & 1):
        out["keys"] = b[L.keys_sorted:L.keys_sorted + R * 8].view(np.uint64).copy()
    return out


def _align(x, a=128):
    return (x + a - 1) // a * a


def decode_reference(geomBuffer, binningBuffer, imgBuffer, P, H, W, R):
    """numpy views of GeometryState / BinningState / ImageState
    (DGR32/cuda_rasterizer/rasterizer_impl.cu:155-194, rasterizer_impl.h:21-28: every array
    starts at the next 128-byte boundary; torch allocations are 512-byte aligned)."""
    g = geomBuffer.cpu().numpy()
    b = binningBuffer.cpu().numpy()
    i = imgBuffer.cpu().numpy()
    o = 0
    out = {}

    def take(buf, o, count, dtype, width=1):
        o = _align(o)
        nbytes = count * width * np.dtype(dtype).itemsize
        arr = buf[o:o + nbytes].view(dtype).copy()
        return arr, o + nbytes

    out["depths"], o = take(g, o, P, np.float32)
    _, o = take(g, o, 3 * P, np.uint8)
    out["internal_radii"], o = take(g, o, P, np.int32)
    m2, o = take(g, o, P, np.float32, 2)
    out["means2D"] = m2.reshape(P, 2)
    c3, o = take(g, o, 6 * P, np.float32)
    out["cov3D"] = c3.reshape(P, 6)
    co, o = take(g, o, P, np.float32, 4)
    out["conic_opacity"] = co.reshape(P, 4)
    _, o = take(g, o, 3 * P, np.float32)
    out["tiles_touched"], o = take(g, o, P, np.uint32)
    N = H * W
    o = 0
    out["final_T"], o = take(i, o, N, np.float32)
    out["n_contrib"], o = take(i, o, N, np.uint32)
    rg, o = take(i, o, N, np.uint32, 2)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out["ranges"] = rg.reshape(N, 2)[:tiles]
    o = 0
    out["point_list"], o = take(b, o, R, np.uint32)
    out["point_list_unsorted"], o = take(b, o, R, np.uint32)
    out["keys"], o = take(b, o, R, np.uint64)
    out["keys_unsorted"], o = take(b, o, R, np.uint64)
    return out


def import_reference():
    """The unmodified reference build (oracle/_ref), or None when it has not been built."""
    if not os.path.isfile(os.path.join(REF_DIR, "diff_gaussian_rasterization_32", "_C.so")):
        return None
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "ref_diff_gaussian_rasterization_32",
        os.path.join(REF_DIR, "diff_gaussian_rasterization_32", "__init__.py"),
        submodule_search_locations=[os.path.join(REF_DIR, "diff_gaussian_rasterization_32")])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_diff_gaussian_rasterization_32"] = mod
    spec.loader.exec_module(mod)
    return mod


def rel_err(a, b):
    """max |a-b| / max|b| (tensor-scale relative error; a secondary statistic -- the gate is grad_gate)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    den = np.abs(b).max()
    return float(np.abs(a - b).max() / den) if den > 0 else float(np.abs(a).max())


def grad_stats(a, b, rtol=1e-4, atol_scale=1e-6):
    """Per-element comparison of a gradient tensor against the reference's:
        |a - b| <= rtol * |b| + atol_scale * max|b|
    Returns dict(viol = fraction of elements outside that bound, out3 = fraction outside 1e-3 relative
    (+ the same absolute floor), tensor = the tensor-scale error rel_err)."""
    a = np.asarray(a, np.float64).reshape(-1)
    b = np.asarray(b, np.float64).reshape(-1)
    m = np.abs(b).max() if b.size else 0.0
    d = np.abs(a - b)
    n = max(a.size, 1)
    return dict(viol=float((d > rtol * np.abs(b) + atol_scale * m).sum() / n),
                out3=float((d > 1e-3 * np.abs(b) + atol_scale * m).sum() / n),
                tensor=rel_err(a, b), max_abs=float(m))


# The reference's own backward is not bit-reproducible: 39 fp32 atomicAdd per contributing (pixel, Gaussian)
# pair land in arbitrary order (backward.cu:593-635), so two runs of the UNMODIFIED reference on identical
# inputs differ per element.  `test_reference_gradient_jitter_yardstick` measures that on the box; the gate
# below allows the same kind of deviation: every element within 1e-4 relative + 1e-6 of the tensor's
# maximum, except a small fraction (accumulation-order noise on cancelling sums), none of which may be
# outside 1e-3 relative by more than JITTER_OUT3.
GRAD_VIOL_MAX = 2e-3
GRAD_OUT3_MAX = 1e-4


def grad_gate(name, a, b, viol_max=GRAD_VIOL_MAX, out3_max=GRAD_OUT3_MAX):
    st = grad_stats(a, b)
    assert st["tensor"] <= 1e-4, (name, st)
    assert st["viol"] <= viol_max and st["out3"] <= out3_max, (name, st)
    return st
