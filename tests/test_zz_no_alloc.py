"""The no-allocation entry point `guava_deform_gaussians_forward_static` (include/guava_rasterizer.h,
no counterpart in the reference) against the existing
`guava_deform_gaussians_forward`, both called through the C ABI (ctypes), on two calls in a row.

Collected last on purpose (file name).  Self-contained: passes alone in a fresh
process and after the other tests (it touches no process-wide option).
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
MESH = ("vertices", "ver_transform_mat", "faces", "uv_binding_face", "uv_face_bary")
ATTRS = ("smplx_rotation", "smplx_scaling", "smplx_opacity", "smplx_features_color",
         "uv_local_xyz", "uv_rotation", "uv_scaling", "uv_opacity", "uv_features_color")
OUTS = (("xyz", 3), ("rotation", 4), ("scaling", 3), ("opacity", 1), ("features_color", 32))


def _p(t):
    return ctypes.c_void_p(t.data_ptr() if t is not None and t.numel() else 0)


def _inputs(B, Nv, F, N, seed):
    from oracle import deform_oracle as do      # seeded synthetic avatar (test infrastructure)
    a = do.synthetic_avatar(B, Nv, F, N, seed=seed)
    g = {k: v.to(DEV).contiguous() for k, v in a.items()}
    g["faces"] = g["faces"].to(torch.int32).contiguous()
    g["uv_binding_face"] = g["uv_binding_face"].to(torch.int32).contiguous()
    return g


def _outputs(B, P, fill):
    return {k: torch.full((B, P, w), fill, device=DEV) for k, w in OUTS}


def _call(lib, static, g, B, Nv, N, F, out, frames=None):
    ins = [_p(g[k]) for k in MESH] + [_p(g[k]) for k in ATTRS]
    outs = [_p(out[k]) for k, _ in OUTS]
    stream = ctypes.c_void_p(torch.cuda.current_stream(DEV).cuda_stream)
    if static:
        return lib.guava_deform_gaussians_forward_static(B, Nv, N, F, *ins, *outs, stream)
    return lib.guava_deform_gaussians_forward(B, Nv, N, F, *ins, _p(frames), *outs, stream)


def test_static_scratch_entry_point_matches_the_existing_one_on_two_calls_in_a_row():
    import guava_renderer_b200 as grb
    lib = grb._lib_mod.load()
    # two different shapes, the larger first: the second call finds the first call's frames in the shared array
    # (the first one sits exactly on the limit: B * F = 65536 = GUAVA_DEFORM_STATIC_MAX_FACE_FRAMES)
    shapes = [(2, 2000, 32768, 9000, 31), (1, 700, 900, 4000, 32)]
    torch.cuda.synchronize()
    for B, Nv, F, N, seed in shapes:
        g = _inputs(B, Nv, F, N, seed)
        P = Nv + N
        want = _outputs(B, P, 0.0)
        frames = torch.empty(B, F, 16, device=DEV)
        assert _call(lib, False, g, B, Nv, N, F, want, frames) == 0, grb._lib_mod.last_error()
        got = _outputs(B, P, float("nan"))          # every element must be written by the call
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(DEV)
        rc = _call(lib, True, g, B, Nv, N, F, got)
        assert rc == 0, grb._lib_mod.last_error()
        assert torch.cuda.memory_allocated(DEV) == before
        torch.cuda.synchronize()
        for k, _ in OUTS:
            w, o = want[k], got[k]
            assert torch.isfinite(o).all(), k
            # the tolerance of tests/test_deform_gpu.py (same kernels, same arithmetic: in fact bit-identical)
            assert float((o - w).abs().max()) <= 1e-5 * max(1.0, float(w.abs().max())), k
            assert torch.equal(o, w), k


def test_static_scratch_entry_point_rejects_sizes_beyond_its_maximum():
    import guava_renderer_b200 as grb
    lib = grb._lib_mod.load()
    B, Nv, F, N = 1, 8, 65537, 4             # B * F = 65537 > GUAVA_DEFORM_STATIC_MAX_FACE_FRAMES
    g = {k: torch.zeros(4, device=DEV) for k in MESH + ATTRS}     # never dereferenced: the size check comes first
    out = _outputs(B, Nv + N, 7.0)
    rc = _call(lib, True, g, B, Nv, N, F, out)
    assert rc == grb._lib_mod.GUAVA_ERR_INVALID_ARGUMENT
    msg = grb._lib_mod.last_error()
    assert "65536" in msg and "65537" in msg and "maximum" in msg
    torch.cuda.synchronize()
    assert all(float(out[k].min()) == 7.0 and float(out[k].max()) == 7.0 for k, _ in OUTS)   # nothing was launched
