"""CPU suite, part 2: host logic and the C-ABI boundary (no compute calls: there is no GPU here)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            src = open(os.path.join(ROOT, "include", fn)).read()
            src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
            names |= set(re.findall(r"\b(guava_[a-z_0-9]+)\s*\(", src))
    return sorted(names)


def test_library_loads_and_exports_every_declared_symbol():
    from guava_renderer_b200 import _lib
    lib = _lib.load()
    decl = _declared_symbols()
    assert len(decl) >= 11
    for name in decl:
        assert hasattr(lib, name), f"{name} declared in include/*.h but not exported"
    assert set(_lib.EXPORTS) <= set(decl)
    assert b"sm_100a" in lib.guava_version()


def test_shared_object_contains_sm100a_specific_code():
    from guava_renderer_b200 import _lib
    out = subprocess.run(["cuobjdump", "-lelf", _lib.SO_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    sass = subprocess.run(["cuobjdump", "-sass", _lib.SO_PATH], capture_output=True, text=True).stdout
    assert "UBLKCP" in sass and "SYNCS.ARRIVE.TRANS64" in sass      # cp.async.bulk + mbarrier tx-count (onesweep partitions)
    assert "FFMA2" in sass and "FMUL2" in sass                      # packed f32x2 arithmetic (compositors), sm_100 only
    assert "LDGSTS" in sass                                         # asynchronous feature-row staging


def test_options_and_layout():
    import guava_renderer_b200 as grb
    grb.set_option("capacity_hint", 123)
    assert grb.get_option("capacity_hint") == 123
    grb.set_option("capacity_hint", 0)
    with pytest.raises(ValueError):
        grb.set_option("no_such_option", 1)
    L = grb.buffer_layout(2, 1000, 100, 60, 5000)
    assert L.splat_rec == 0 and L.depths >= 2 * 1000 * 32
    assert L.geom_bytes % 256 == 0 and L.image_bytes % 256 == 0 and L.binning_bytes >= 5000 * 12
    assert L.point_list == 0            # backward relies on point_list being first for every capacity
    tiles = 2 * 7 * 4
    assert L.tile_cursor - L.ranges >= tiles * 8
    assert grb.launch_count() == 0 or grb.launch_count() > 0


def test_no_cpu_fallback_and_argument_validation():
    import diff_gaussian_rasterization_32 as dgr
    import diff_gaussian_rasterization as alias
    assert alias.GaussianRasterizer is dgr.GaussianRasterizer_32
    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug", "antialiasing")       # __init__.py:144-156
    st = dgr.GaussianRasterizationSettings(16, 16, 0.1, 0.1, torch.zeros(32), 1.0, torch.eye(4), torch.eye(4), 0,
                                           torch.zeros(3), False, False, False)
    r = dgr.GaussianRasterizer_32(st)
    x = torch.zeros(4, 3)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(means3D=x, means2D=x, opacities=torch.zeros(4, 1), scales=x, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(means3D=x, means2D=x, opacities=torch.zeros(4, 1), colors_precomp=torch.zeros(4, 32), scales=x,
          rotations=torch.zeros(4, 4), cov3D_precomp=torch.zeros(4, 6))
    # CPU tensors must be rejected loudly, never computed on the host
    with pytest.raises(RuntimeError, match="no CPU path|CUDA"):
        r(means3D=x, means2D=x, opacities=torch.zeros(4, 1), colors_precomp=torch.zeros(4, 32), scales=x,
          rotations=torch.zeros(4, 4))


def test_graphed_step_rejects_a_summed_backward_for_a_producer_and_host_tensors():
    """Argument errors of GraphedRenderStep that need no GPU: the view-summed backward hands a producer no per-view
    gradients to continue with, and host tensors are refused (there is no CPU path to capture)."""
    import guava_renderer_b200 as grb
    st = grb.BatchRasterizationSettings(16, 16, torch.zeros(1, 2), torch.zeros(1, 32), 1.0, torch.eye(4)[None], torch.eye(4)[None])
    with pytest.raises(ValueError, match="summed"):
        grb.GraphedRenderStep(None, None, None, None, None, st, dL_dcolor=torch.zeros(1, 32, 16, 16),
                              producer=lambda: {}, summed=True)
    x = torch.zeros(1, 4, 3)
    with pytest.raises(RuntimeError, match="no CPU path|CUDA"):
        grb.GraphedRenderStep(x, torch.zeros(1, 4, 32), torch.zeros(1, 4, 1), x, torch.zeros(1, 4, 4), st)


def test_c_abi_rejects_bad_arguments_before_touching_the_device():
    """Argument errors of the entry points added in round 2, through ctypes; each returns before any CUDA call (the
    pointers are never dereferenced), so this runs without a GPU."""
    import ctypes
    import guava_renderer_b200 as grb
    lib = grb._lib_mod.load()
    L = grb._lib_mod
    p = ctypes.c_void_p
    fake = p(0x1000)
    # guava_call_options::zero_fill must be 4-byte aligned and a multiple of 4 bytes long
    for ptr, nbytes in ((0x1002, 16), (0x1000, 18)):
        co = L.CallOptions(zero_fill=ptr, zero_fill_bytes=nbytes)
        n = ctypes.c_int(7)
        rc = lib.guava_rasterizer_forward_batch(L.ALLOC_FN(lambda u, s: 0), None, L.ALLOC_FN(lambda u, s: 0), None,
                                                L.ALLOC_FN(lambda u, s: 0), None, 0, 0, fake, 16, 16, fake, 0, fake, 0, fake, 0,
                                                fake, 0, 1.0, fake, 0, fake, fake, fake, fake, fake, 0, fake, 0, None,
                                                ctypes.byref(n), ctypes.byref(co))
        assert rc == L.GUAVA_ERR_INVALID_ARGUMENT and "zero_fill" in L.last_error()
    # a struct of another size (another version of the header) is refused
    co = L.CallOptions()
    co.struct_size = 8
    rc = lib.guava_rasterizer_forward_batch(L.ALLOC_FN(lambda u, s: 0), None, L.ALLOC_FN(lambda u, s: 0), None,
                                            L.ALLOC_FN(lambda u, s: 0), None, 0, 0, fake, 16, 16, fake, 0, fake, 0, fake, 0,
                                            fake, 0, 1.0, fake, 0, fake, fake, fake, fake, fake, 0, fake, 0, None, None,
                                            ctypes.byref(co))
    assert rc == L.GUAVA_ERR_INVALID_ARGUMENT and "struct_size" in L.last_error()
    # the summed backward: out_row_stride is 0 or at least 43 floats
    rc = lib.guava_rasterizer_backward_batch_summed(1, 4, 0, fake, 16, 16, fake, 0, fake, 0, fake, 0, fake, 0, 1.0, fake, 0,
                                                    fake, fake, fake, fake, fake, fake, fake, fake, None,
                                                    fake, fake, fake, fake, fake, 17, 0, 0, None)
    assert rc == L.GUAVA_ERR_INVALID_ARGUMENT and "out_row_stride" in L.last_error()
    # the deformation's gradient split: sizes, and 16-byte alignment of the feature tensors
    assert lib.guava_deform_split_gradients(-1, 4, 4, None, None, None, None, None, None, None, None, None) == \
        L.GUAVA_ERR_INVALID_ARGUMENT
    assert lib.guava_deform_split_gradients(1, 0, 0, None, None, None, None, None, None, None, None, None) == L.GUAVA_OK
    rc = lib.guava_deform_split_gradients(1, 4, 4, None, None, p(0x1004), None, None, fake, None, fake, None)
    assert rc == L.GUAVA_ERR_INVALID_ARGUMENT and "16-byte" in L.last_error()


def test_product_never_imports_the_oracle():
    for d in ("guava_renderer_b200", "diff_gaussian_rasterization_32", "diff_gaussian_rasterization"):
        for root, _dirs, files in os.walk(os.path.join(ROOT, d)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h")):
                    src = open(os.path.join(root, f)).read()
                    assert not re.search(r"^\s*(import|from)\s+oracle\b", src, flags=re.M), os.path.join(root, f)
                    assert "splat_oracle" not in src or f.endswith((".cu", ".cuh")), os.path.join(root, f)


def test_shard_views_partitions():
    from guava_renderer_b200.sharding import shard_views
    for n, w in [(16, 8), (240, 8), (240, 7), (5, 8), (0, 2)]:
        got = [shard_views(n, r, w) for r in range(w)]
        flat = [v for g in got for v in g]
        assert flat == list(range(n))
        assert max(len(g) for g in got) - min(len(g) for g in got) <= 1
    with pytest.raises(ValueError):
        shard_views(4, 4, 4)


_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from guava_renderer_b200 import sharding
dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
P, V = 50, 3
g = torch.Generator().manual_seed(100 + rank)
grads = [torch.randn(V, P, w, generator=g) for w in sharding.GRAD_WIDTHS]
buf = sharding.pack_gradients(grads)
assert buf.shape == (P, 43)
sharding.allreduce_gradients(buf)
exp = 0
for r in range(world):
    gr = torch.Generator().manual_seed(100 + r)
    exp = exp + sharding.pack_gradients([torch.randn(V, P, w, generator=gr) for w in sharding.GRAD_WIDTHS])
assert torch.allclose(buf, exp, atol=1e-5), (buf - exp).abs().max()
parts = sharding.unpack_gradients(buf)
assert [p.shape[1] for p in parts] == list(sharding.GRAD_WIDTHS)
views = sharding.shard_views(16, rank, world)
all_views = [None] * world
dist.all_gather_object(all_views, views)
assert sorted(v for vs in all_views for v in vs) == list(range(16))
dist.barrier(); dist.destroy_process_group()
print("worker ok", rank)
"""


def test_gradient_allreduce_world_size_2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29577")
        procs.append(subprocess.Popen([sys.executable, str(script), ROOT], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    for p in procs:
        out, _ = p.communicate(timeout=240)
        assert p.returncode == 0, out
        assert "worker ok" in out


def test_synthetic_inputs_are_deterministic_and_well_formed():
    from guava_renderer_b200 import synthetic
    a = synthetic.guava_cloud(1000, 42)
    b = synthetic.guava_cloud(1000, 42)
    for k in a:
        assert torch.equal(a[k], b[k])
    assert torch.allclose(a["rotation"].norm(dim=1), torch.ones(1000), atol=1e-6)
    assert a["opacity"].min() >= 0.002 and a["opacity"].max() <= 0.999
    cam = synthetic.camera(3, 8)
    # view matrix is a rigid transform (transposed convention): R R^T = I
    R = cam["viewmatrix"][:3, :3]
    assert torch.allclose(R @ R.T, torch.eye(3), atol=1e-5)
    # the cloud centre projects to the image centre with depth 22
    p = torch.tensor([0.0, 0.0, 0.0, 1.0]) @ cam["viewmatrix"]
    assert abs(float(p[2]) - 22.0) < 1e-3


def test_collective_footprint_cap_keeps_an_explicit_setting(monkeypatch):
    from guava_renderer_b200 import sharding
    monkeypatch.delenv("NCCL_MAX_CTAS", raising=False)
    sharding.limit_collective_footprint()
    assert os.environ["NCCL_MAX_CTAS"] == "4"
    monkeypatch.setenv("NCCL_MAX_CTAS", "16")
    sharding.limit_collective_footprint()
    assert os.environ["NCCL_MAX_CTAS"] == "16"


def test_compiled_torch_binding_is_built_and_exports_the_reference_entry_points():
    """lib/_guava_torch.so (csrc/torch_binding.cpp) is the default `_C`: the reference's three pybind functions
    (DGR32/ext.cpp:15-18) plus the batched pair.  Loading needs no GPU; calling does."""
    import guava_renderer_b200 as grb
    from guava_renderer_b200 import rasterizer
    assert os.path.isfile(rasterizer.TORCH_SO_PATH)
    if rasterizer.BINDING == "compiled":
        for name in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible",
                     "rasterize_gaussians_batch", "rasterize_gaussians_batch_backward",
                     "rasterize_gaussians_batch_backward_summed"):
            assert callable(getattr(grb._C, name)), name
    import diff_gaussian_rasterization_32 as dgr
    assert dgr._C is grb._C


def test_pose_avatar_is_well_formed_and_renders_like_the_benchmark_cloud():
    """synthetic.pose_avatar / pose_of (bench.py --workload C3pose): index ranges, unit quaternions, and -- deformed
    by the CPU restatement of ubody_gaussian.py:245-289 -- world-space sizes of the benchmark cloud."""
    from guava_renderer_b200 import synthetic
    from oracle import deform_oracle as do
    V, P = 2, 9000
    av = synthetic.pose_avatar(P, V, seed=4, rows=30, cols=40)
    Nv, F = av["vertices"].shape[0], av["faces"].shape[0]
    assert Nv == 1200 and F == 2 * 29 * 40 and av["uv_binding_face"].shape == (P - Nv,)
    assert int(av["faces"].max()) < Nv and int(av["faces"].min()) >= 0
    assert int(av["uv_binding_face"].max()) < F and int(av["uv_binding_face"].min()) >= 0
    assert torch.allclose(av["uv_face_bary"].sum(-1), torch.ones(P - Nv), atol=1e-6)
    assert torch.allclose(av["uv_rotation"].norm(dim=-1), torch.ones(V, P - Nv), atol=1e-5)
    pv, pt = synthetic.pose_of(av["vertices"], 3, V)
    assert pv.shape == (V, Nv, 3) and pt.shape == (V, Nv, 4, 4)
    R = pt[0, 0, :3, :3]
    assert torch.allclose(R @ R.T, torch.eye(3), atol=1e-6)
    out = do.deform(pv, pt, av["faces"], av["uv_binding_face"], av["uv_face_bary"], **{k: av[k] for k in (
        "smplx_rotation", "smplx_scaling", "smplx_opacity", "smplx_features_color", "uv_local_xyz", "uv_rotation",
        "uv_scaling", "uv_opacity", "uv_features_color")})
    assert out["xyz"].shape == (V, P, 3) and torch.isfinite(out["xyz"]).all()
    s0 = 0.0054 * (1e5 / P) ** 0.5
    assert 0.8 * s0 < float(out["scaling"][:, Nv:].median()) < 1.25 * s0          # face-bound Gaussians keep the target size
    assert torch.allclose(out["rotation"].norm(dim=-1), torch.ones(V, P), atol=1e-4)
