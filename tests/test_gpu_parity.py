"""GPU parity tests (run on the B200 box: `pytest -m gpu`).

Every test drives the CUDA path through the reference-shaped interface
(`_C.rasterize_gaussians`, `GaussianRasterizer_32`, the batched entry), i.e.
through the C ABI of libguava_rasterizer_b200.so, and checks it against

  (A) the unmodified reference CUDA build (oracle/_ref, built from
      /root/reference by oracle/build_ref.py) on identical inputs -- the gate named by
      BASELINE.json: sorted keys+values, tile ranges, radii, n_contrib bit-exact;
      features / inverse depth <= 1e-5 absolute; gradients <= 1e-4 relative;
  (B) the CPU oracle (oracle/splat_oracle.c): bit-exact for everything that does not pass
      through the hardware exp2 (radii, means2D, depths, conics, keys, lists, ranges),
      tolerance-based for the compositing outputs;
  (C) size-independent properties at the full benchmark sizes.
"""
import os

import numpy as np
import pytest
import torch

from tests import util

pytestmark = pytest.mark.gpu

import guava_renderer_b200 as grb  # noqa: E402
import diff_gaussian_rasterization_32 as ours  # noqa: E402

DEV = "cuda:0"
CAP = 1 << 22


@pytest.fixture(autouse=True)
def _options():
    grb.set_option("emit_sorted_keys", 1)
    grb.set_option("capacity_hint", CAP)
    grb.set_option("exact_binning", 0)
    yield
    grb.set_option("emit_sorted_keys", 0)
    grb.set_option("capacity_hint", 0)


@pytest.fixture(scope="module")
def ref():
    m = util.import_reference()
    if m is None:
        pytest.skip("oracle/_ref (reference CUDA build) not present")
    return m


def _ours_forward(inp, antialiasing=False, scale_modifier=1.0):
    R, color, radii, geom, binning, img, invd = util.c_forward(ours, inp, DEV, antialiasing, scale_modifier)
    torch.cuda.synchronize()
    dec = util.decode_ours(geom, binning, img, 1, inp["xyz"].shape[0], inp["W"], inp["H"], CAP)
    dec.update(R=R, color=color.cpu().numpy(), radii=radii.cpu().numpy(), invdepth=invd.cpu().numpy())
    return dec


CASES = [
    dict(P=10_000, H=512, W=512, seed=1000, scale_mult=1.0),            # BASELINE config 1
    dict(P=3_000, H=200, W=152, seed=21, scale_mult=1.5),               # ragged image, partial tiles
    dict(P=20_000, H=128, W=256, seed=33, scale_mult=3.0),              # long per-tile lists (> 2048)
    dict(P=1, H=64, W=64, seed=4, scale_mult=0.02),                     # single Gaussian
    dict(P=257, H=17, W=33, seed=5, scale_mult=0.2),                    # tiny odd image
    dict(P=40_000, H=32, W=32, seed=44, scale_mult=1.0),                # 4 tiles, > 16384 instances per tile
]


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.mark.parametrize("case", CASES, ids=lambda c: f"P{c['P']}_{c['H']}x{c['W']}")
@pytest.mark.parametrize("antialiasing", [False, True])
def test_forward_vs_cpu_oracle(case, antialiasing):
    inp = util.inputs_for(**case)
    inp["bg"] = torch.linspace(-1, 1, 32)
    f = util.oracle_forward(inp, antialiasing=antialiasing)
    o = _ours_forward(inp, antialiasing)
    vis = f["radii"] > 0
    assert o["R"] == f["num_rendered"] == o["num_rendered"]
    assert np.array_equal(o["radii"], f["radii"])
    assert np.array_equal(o["tiles_touched"], f["tiles_touched"])
    assert np.array_equal(_bits(o["depths"][vis]), _bits(f["depths"][vis]))
    assert np.array_equal(_bits(o["means2D"][vis]), _bits(f["means2D"][vis]))
    assert np.array_equal(_bits(o["conic_opacity"][vis]), _bits(f["conic_opacity"][vis]))
    assert np.array_equal(o["ranges"], f["ranges"])
    assert np.array_equal(o["point_list"], f["point_list"])
    assert np.array_equal(o["keys"], f["keys"])
    # compositing goes through MUFU.EX2 on the GPU and exp2() on the CPU
    assert np.mean(o["n_contrib"] != f["n_contrib"]) < 2e-3
    d = np.abs(o["color"] - f["color"])
    assert np.mean(d > 1e-5) < 2e-3 and np.median(d) < 1e-6, d.max()
    assert np.mean(np.abs(o["invdepth"] - f["invdepth"]) > 1e-5) < 2e-3


@pytest.fixture(params=[0, 1], ids=["tile_buckets", "onesweep"])
def sort_mode(request):
    """Both binning paths must give the reference's keys, values and ranges: 0 = per-tile buckets sorted
    in shared memory (binning.cu), 1 = scan + duplicateWithKeys + onesweep radix sort + identifyTileRanges
    (onesweep.cu; the reference's own chain, rasterizer_impl.cu:280-320)."""
    grb.set_option("sort_mode", request.param)
    yield request.param
    grb.set_option("sort_mode", -1)   # default: chosen per shape from the previous call's longest tile list


@pytest.mark.parametrize("case", CASES, ids=lambda c: f"P{c['P']}_{c['H']}x{c['W']}")
@pytest.mark.parametrize("antialiasing", [False, True])
def test_forward_vs_reference_build(ref, case, antialiasing, sort_mode):
    inp = util.inputs_for(**case)
    inp["bg"] = torch.linspace(-1, 1, 32)
    o = _ours_forward(inp, antialiasing)
    if sort_mode == 1:   # GeometryState::point_offsets = InclusiveSum(tiles_touched), rasterizer_impl.cu:280
        assert np.array_equal(o["point_offsets"], np.cumsum(o["tiles_touched"], dtype=np.uint64).astype(np.uint32))
    R, color, radii, geom, binning, img, invd = util.c_forward(ref, inp, DEV, antialiasing)
    torch.cuda.synchronize()
    r = util.decode_reference(geom, binning, img, inp["xyz"].shape[0], inp["H"], inp["W"], R)
    radii = radii.cpu().numpy()
    vis = radii > 0
    assert o["R"] == R
    assert np.array_equal(o["radii"], radii)                                   # bit-exact gates
    assert np.array_equal(o["keys"], r["keys"])
    assert np.array_equal(o["point_list"], r["point_list"])
    assert np.array_equal(o["ranges"], r["ranges"])
    assert np.array_equal(o["n_contrib"], r["n_contrib"])
    assert np.array_equal(o["tiles_touched"], r["tiles_touched"])              # extra diagnostics
    assert np.array_equal(_bits(o["depths"][vis]), _bits(r["depths"][vis]))
    assert np.array_equal(_bits(o["means2D"][vis]), _bits(r["means2D"][vis]))
    assert np.array_equal(_bits(o["conic_opacity"][vis]), _bits(r["conic_opacity"][vis]))
    assert np.array_equal(_bits(o["final_T"]), _bits(r["final_T"]))
    assert np.abs(o["color"] - color.cpu().numpy()).max() <= 1e-5              # tolerance gates
    assert np.abs(o["invdepth"] - invd.cpu().numpy()).max() <= 1e-5


def _run_autograd(mod, inp, dL_dcolor, dL_dinv, antialiasing=False):
    """The call pattern of models/UbodyAvatar/gaussian_render.py:30-63 for one view."""
    leaf = {k: inp[k].to(DEV).clone().requires_grad_(True) for k in ("xyz", "features", "opacity", "scaling", "rotation")}
    mean_2d = torch.zeros_like(leaf["xyz"], requires_grad=True)
    rast = mod.GaussianRasterizer_32(raster_settings=util.settings_for(mod, inp, DEV, antialiasing))
    color, radii, invd = rast(means3D=leaf["xyz"], means2D=mean_2d, shs=None, colors_precomp=leaf["features"],
                              opacities=leaf["opacity"], scales=leaf["scaling"], rotations=leaf["rotation"],
                              cov3D_precomp=None)
    loss = (color * dL_dcolor.to(DEV)).sum()
    if dL_dinv is not None:
        loss = loss + (invd * dL_dinv.to(DEV)).sum()
    loss.backward()
    torch.cuda.synchronize()
    g = {k: v.grad.cpu().numpy() for k, v in leaf.items()}
    g["means2D"] = mean_2d.grad.cpu().numpy()
    return color.detach().cpu().numpy(), g


GRAD_CASES = [CASES[0], CASES[1], CASES[2]]


@pytest.mark.parametrize("case", GRAD_CASES, ids=lambda c: f"P{c['P']}_{c['H']}x{c['W']}")
@pytest.mark.parametrize("with_invdepth", [False, True])
def test_backward_vs_reference_build(ref, case, with_invdepth):
    inp = util.inputs_for(**case)
    inp["bg"] = torch.linspace(-1, 1, 32)
    g = torch.Generator().manual_seed(9)
    dL_dcolor = torch.randn(32, case["H"], case["W"], generator=g)
    dL_dinv = torch.randn(1, case["H"], case["W"], generator=g) if with_invdepth else None
    _, go = _run_autograd(ours, inp, dL_dcolor, dL_dinv)
    _, gr = _run_autograd(ref, inp, dL_dcolor, dL_dinv)
    for k in gr:
        assert util.rel_err(go[k], gr[k]) <= 1e-4, (k, util.rel_err(go[k], gr[k]))
        util.grad_gate(k, go[k], gr[k])       # per element: |a-b| <= 1e-4 |b| + 1e-6 max|b|


@pytest.mark.parametrize("opts", [dict(taken_masks=1), dict(tma_staging=1), dict(taken_masks=1, tma_staging=1)],
                         ids=["taken_masks", "tma_staging", "taken_masks+tma_staging"])
@pytest.mark.parametrize("case", GRAD_CASES, ids=lambda c: f"P{c['P']}_{c['H']}x{c['W']}")
def test_compositor_options_give_the_same_results(case, opts, sort_mode):
    """Options of the compositors that must not change a result.  taken_masks: the forward marks, one byte per (list
    entry, 8x4 block), the entries each block got nothing from, and the backward skips them.  tma_staging: the
    survivors' feature rows are staged by cp.async.bulk.tensor gather4 instead of LDGSTS (same bytes).  Image
    bit-identical, gradients equal to reduction order, in both binning modes."""
    inp = util.inputs_for(**case)
    inp["bg"] = torch.linspace(-1, 1, 32)
    g = torch.Generator().manual_seed(9)
    dL_dcolor = torch.randn(32, case["H"], case["W"], generator=g)
    dL_dinv = torch.randn(1, case["H"], case["W"], generator=g)
    base = {k: grb.get_option(k) for k in opts}
    for k in opts:
        grb.set_option(k, 0)
    try:
        c0, g0 = _run_autograd(ours, inp, dL_dcolor, dL_dinv)
        o0 = _ours_forward(inp)
        for k, v in opts.items():
            grb.set_option(k, v)
        c1, g1 = _run_autograd(ours, inp, dL_dcolor, dL_dinv)
        o1 = _ours_forward(inp)
    finally:
        for k, v in base.items():
            grb.set_option(k, v)
    assert np.array_equal(c0, c1)
    for k in ("point_list", "keys", "ranges", "n_contrib", "radii", "color", "invdepth"):
        assert np.array_equal(o0[k], o1[k]), k
    assert np.array_equal(_bits(o0["final_T"]), _bits(o1["final_T"]))
    if opts.get("taken_masks"):
        assert o1["block_masks"] == 1 and "taken_marks" in o1 and "taken_marks" not in o0
    for k in g0:     # two runs differ by the order of their floating-point reductions
        assert util.rel_err(g1[k], g0[k]) <= 5e-5, (k, util.rel_err(g1[k], g0[k]))


@pytest.mark.parametrize("antialiasing", [False, True])
def test_backward_vs_cpu_oracle(antialiasing):
    import oracle
    case = CASES[1]
    inp = util.inputs_for(**case)
    inp["bg"] = torch.linspace(-1, 1, 32)
    g = torch.Generator().manual_seed(9)
    dL_dcolor = torch.randn(32, case["H"], case["W"], generator=g)
    dL_dinv = torch.randn(1, case["H"], case["W"], generator=g)
    _, go = _run_autograd(ours, inp, dL_dcolor, dL_dinv, antialiasing)
    f = util.oracle_forward(inp, antialiasing=antialiasing)
    gc = oracle.backward(f, dL_dcolor.numpy(), dL_dinv.numpy())
    names = dict(xyz="dL_dmeans3D", features="dL_dcolors", opacity="dL_dopacity", scaling="dL_dscales",
                 rotation="dL_drotations", means2D="dL_dmeans2D")
    for k, n in names.items():
        e = util.rel_err(go[k], gc[n].reshape(go[k].shape))
        assert e <= 1e-3, (k, e)


def test_batched_entry_equals_per_view_loop(sort_mode):
    """One launch chain over V views == V single-view calls (bitwise forward, tolerance backward)."""
    V, P, H, W = 3, 5000, 160, 176
    views = [util.inputs_for(P, H, W, seed=700 + v, view=v, num_views=V, scale_mult=1.3) for v in range(V)]
    g = torch.Generator().manual_seed(1)
    dL = torch.randn(V, 32, H, W, generator=g).to(DEV)

    stack = lambda k: torch.stack([x[k] for x in views], 0).to(DEV)
    leaf = {k: stack(k).requires_grad_(True) for k in ("xyz", "features", "opacity", "scaling", "rotation")}
    cam = {
        "world_view_transform": stack("viewmatrix"), "full_proj_transform": stack("projmatrix"),
        "tanfovx": torch.tensor([x["tanfovx"] for x in views]), "tanfovy": torch.tensor([x["tanfovy"] for x in views]),
        "image_height": torch.tensor([H] * V), "image_width": torch.tensor([W] * V),
        "camera_center": stack("campos"),
    }
    out = grb.render_views(dict(xyz=leaf["xyz"], opacity=leaf["opacity"], scaling=leaf["scaling"],
                                rotation=leaf["rotation"], features_color=leaf["features"]), cam, bg=0.25)
    (out["rendered_images"] * dL).sum().backward()
    torch.cuda.synchronize()

    for v in range(V):
        inp = dict(views[v]); inp["bg"] = torch.full((32,), 0.25)
        color, gv = _run_autograd(ours, inp, dL[v].cpu(), None)
        assert np.array_equal(out["rendered_images"][v].detach().cpu().numpy(), color)
        o = _ours_forward(inp)
        assert np.array_equal(out["radiis"][v].cpu().numpy(), o["radii"])
        for k in ("xyz", "features", "opacity", "scaling", "rotation"):
            assert util.rel_err(leaf[k].grad[v].cpu().numpy(), gv[k]) <= 1e-4, k
        assert util.rel_err(out["viewspace_points"].grad[v].cpu().numpy(), gv["means2D"]) <= 1e-4


def test_capacity_overflow_is_recovered(sort_mode):
    inp = util.inputs_for(**CASES[1])
    a = _ours_forward(inp)
    grb.set_option("capacity_hint", 512)          # far below num_rendered: forces the re-run path
    R, color, radii, geom, binning, img, invd = util.c_forward(ours, inp, DEV)
    torch.cuda.synchronize()
    assert R == a["R"] > 512
    assert np.array_equal(color.cpu().numpy(), a["color"])
    grb.set_option("capacity_hint", 0)
    grb.set_option("exact_binning", 1)            # the reference's wait-then-allocate order
    R2, color2, *_ = util.c_forward(ours, inp, DEV)
    assert R2 == R and np.array_equal(color2.cpu().numpy(), a["color"])
    grb.set_option("exact_binning", 0)


def test_mark_visible_and_empty_inputs():
    import oracle
    inp = util.inputs_for(4096, 64, 64, seed=8)
    inp["xyz"][::3, 2] -= 40.0                     # a third of the points behind the near plane
    rast = ours.GaussianRasterizer_32(util.settings_for(ours, inp, DEV))
    vis = rast.markVisible(inp["xyz"].to(DEV)).cpu().numpy()
    assert np.array_equal(vis, oracle.mark_visible(inp["xyz"].numpy(), inp["viewmatrix"].numpy(), inp["projmatrix"].numpy()))
    o = _ours_forward(inp)
    f = util.oracle_forward(inp)
    assert np.array_equal(o["radii"], f["radii"]) and (o["radii"][::3] == 0).all()
    # P == 0 (rasterize_points.cu:88): zeros, rendered = 0
    e = torch.Tensor([])
    z = torch.zeros(0, 3, device=DEV)
    R, color, radii, *_ = ours._C.rasterize_gaussians(
        inp["bg"].to(DEV), z, torch.zeros(0, 32, device=DEV), torch.zeros(0, 1, device=DEV), z, torch.zeros(0, 4, device=DEV),
        1.0, e, inp["viewmatrix"].to(DEV), inp["projmatrix"].to(DEV), inp["tanfovx"], inp["tanfovy"], 64, 64, e, 0,
        inp["campos"].to(DEV), False, False, False)
    assert R == 0 and color.shape == (32, 64, 64) and not color.any() and radii.numel() == 0


def test_error_behaviour_matches_reference():
    inp = util.inputs_for(16, 32, 32, seed=1)
    rast = ours.GaussianRasterizer_32(util.settings_for(ours, inp, DEV))
    d = lambda k: inp[k].to(DEV)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        rast(means3D=d("xyz"), means2D=d("xyz"), opacities=d("opacity"), scales=d("scaling"), rotations=d("rotation"))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        rast(means3D=d("xyz"), means2D=d("xyz"), opacities=d("opacity"), colors_precomp=d("features"))
    with pytest.raises(RuntimeError, match="means3D must have dimensions"):
        rast(means3D=d("xyz").reshape(-1), means2D=d("xyz"), opacities=d("opacity"), colors_precomp=d("features"),
             scales=d("scaling"), rotations=d("rotation"))
    with pytest.raises(RuntimeError, match="For non-RGB, provide precomputed Gaussian colors"):
        rast(means3D=d("xyz"), means2D=d("xyz"), opacities=d("opacity"), shs=torch.zeros(16, 1, 3, device=DEV),
             scales=d("scaling"), rotations=d("rotation"))


def test_cov3d_precomp_path():
    inp = util.inputs_for(2000, 96, 96, seed=12, scale_mult=0.6)
    f = util.oracle_forward(inp)
    e = torch.Tensor([])
    d = lambda k: inp[k].to(DEV)
    R, color, radii, *_ = ours._C.rasterize_gaussians(
        d("bg"), d("xyz"), d("features"), d("opacity"), e, e, 1.0, torch.from_numpy(f["cov3D"]).to(DEV),
        d("viewmatrix"), d("projmatrix"), inp["tanfovx"], inp["tanfovy"], 96, 96, e, 0, d("campos"), False, False, False)
    assert R == f["num_rendered"] and np.array_equal(radii.cpu().numpy(), f["radii"])


def test_full_size_properties():
    """BASELINE config 2 shape (100k Gaussians, 512x512, 8 views, one launch chain): properties that
    do not need an oracle run at this size."""
    V, P, H, W = 8, 100_000, 512, 512
    from guava_renderer_b200 import synthetic
    data = synthetic.make_views(2, P, H, W, list(range(V)), V)
    tanfov = torch.tensor([[data["tanfovx"], data["tanfovy"]]] * V, device=DEV)
    st = grb.BatchRasterizationSettings(H, W, tanfov, torch.zeros(V, 32, device=DEV), 1.0,
                                        data["viewmatrix"].to(DEV), data["projmatrix"].to(DEV))
    grb.set_option("capacity_hint", 0)
    m2 = torch.zeros(V, P, 3, device=DEV)
    color, radii, invd = grb.rasterize_gaussians_batch(
        data["xyz"].to(DEV), m2, data["features_color"].to(DEV), data["opacity"].to(DEV),
        data["scaling"].to(DEV), data["rotation"].to(DEV), st)
    torch.cuda.synchronize()
    assert torch.isfinite(color).all() and torch.isfinite(invd).all()
    assert (radii >= 0).all() and (radii > 0).float().mean() > 0.9
    # coverage of the synthetic avatar (~39-45 % of the frame)
    cov = (color.abs().sum(1) > 0).float().mean().item()
    assert 0.25 < cov < 0.7, cov
    # view 0 of the batch equals a single-view call on the same inputs
    inp = dict(xyz=data["xyz"][0], features=data["features_color"][0], opacity=data["opacity"][0],
               scaling=data["scaling"][0], rotation=data["rotation"][0], viewmatrix=data["viewmatrix"][0],
               projmatrix=data["projmatrix"][0], campos=data["campos"][0], tanfovx=data["tanfovx"],
               tanfovy=data["tanfovy"], H=H, W=W, bg=torch.zeros(32))
    grb.set_option("capacity_hint", CAP)
    o = _ours_forward(inp)
    assert np.array_equal(o["color"], color[0].cpu().numpy())
    keys = o["keys"]
    assert np.all(keys[1:] >= keys[:-1])                                  # sortedness
    lens = o["ranges"][:, 1].astype(np.int64) - o["ranges"][:, 0]
    assert lens.sum() == o["R"] == int(o["tiles_touched"].sum())         # a checksum of checksums
    same = keys[1:] == keys[:-1]
    assert np.all(o["point_list"][1:][same] > o["point_list"][:-1][same])  # ties keep index order
    # the global onesweep sort gives the same keys, values, ranges and image at full size
    grb.set_option("sort_mode", 0)
    o = _ours_forward(inp)
    grb.set_option("sort_mode", 1)
    try:
        o1 = _ours_forward(inp)
    finally:
        grb.set_option("sort_mode", -1)
    for k in ("keys", "point_list", "ranges", "n_contrib", "color"):
        assert np.array_equal(o1[k], o[k]), k


import glob as _glob
import os as _os

_GOLDENS = sorted(_glob.glob(_os.path.join(_os.path.dirname(__file__), "golden", "*.npz")))


@pytest.mark.parametrize("path", _GOLDENS, ids=[_os.path.basename(p) for p in _GOLDENS])
def test_cuda_path_against_committed_reference_golden(path, sort_mode):
    """The committed outputs of the unmodified reference build (tests/golden/make_golden.py): the same
    gates as test_forward_vs_reference_build / test_backward_vs_reference_build, without needing oracle/_ref."""
    from guava_renderer_b200 import synthetic
    z = np.load(path)
    P, H, W = int(z["P"]), int(z["H"]), int(z["W"])
    aa = bool(z["antialiasing"])
    inp = util.inputs_for(P, H, W, int(z["seed"]), scale_mult=float(z["scale_mult"]))
    inp["bg"] = torch.from_numpy(z["bg"])
    o = _ours_forward(inp, aa)
    vis = z["radii"] > 0
    assert o["R"] == int(z["num_rendered"])
    assert np.array_equal(o["radii"], z["radii"])
    assert np.array_equal(o["keys"], z["keys"])
    assert np.array_equal(o["point_list"], z["point_list"])
    assert np.array_equal(o["ranges"], z["ranges"])
    assert np.array_equal(o["n_contrib"], z["n_contrib"])
    assert np.array_equal(_bits(o["final_T"]), _bits(z["final_T"]))
    assert np.array_equal(_bits(o["means2D"][vis]), _bits(z["means2D"][vis]))
    assert np.array_equal(_bits(o["conic_opacity"][vis]), _bits(z["conic_opacity"][vis]))
    assert np.abs(o["color"] - z["color"]).max() <= 1e-5
    assert np.abs(o["invdepth"] - z["invdepth"]).max() <= 1e-5
    dc, _ = synthetic.upstream_grads(H, W, int(z["dL_dcolor_seed"]))
    _, g = _run_autograd(ours, inp, dc * float(z["dL_scale"]), None, aa)
    names = dict(xyz="dL_dmeans3D", features="dL_dcolors", opacity="dL_dopacity", scaling="dL_dscales",
                 rotation="dL_drotations", means2D="dL_dmeans2D")
    for k, n in names.items():
        assert util.rel_err(g[k], z[n].reshape(g[k].shape)) <= 1e-4, (k, util.rel_err(g[k], z[n].reshape(g[k].shape)))


def test_pack_view_gradients_and_graphed_step():
    """The fused [V,P,*] -> [P,43] packing kernel equals sum+cat, and a CUDA-graph replay of the whole
    step equals the eager batched call (forward bitwise; gradients to atomic-order tolerance)."""
    from guava_renderer_b200 import sharding
    V, P, H, W = 2, 6000, 96, 128
    views = [util.inputs_for(P, H, W, seed=900 + v, view=v, num_views=V, scale_mult=1.2) for v in range(V)]
    stack = lambda k: torch.stack([x[k] for x in views], 0).to(DEV)
    g = torch.Generator().manual_seed(2)
    dL = torch.randn(V, 32, H, W, generator=g).to(DEV)
    tanfov = torch.tensor([[x["tanfovx"], x["tanfovy"]] for x in views], device=DEV)
    st = grb.BatchRasterizationSettings(H, W, tanfov, torch.zeros(V, 32, device=DEV), 1.0,
                                        stack("viewmatrix"), stack("projmatrix"))
    grb.set_option("capacity_hint", 0)
    keys = ("xyz", "scaling", "rotation", "opacity", "features")
    leaf = {k: stack(k).requires_grad_(True) for k in keys}
    m2 = torch.zeros(V, P, 3, device=DEV, requires_grad=True)
    color, radii, invd = grb.rasterize_gaussians_batch(leaf["xyz"], m2, leaf["features"], leaf["opacity"],
                                                       leaf["scaling"], leaf["rotation"], st)
    color.backward(dL)
    grads = [leaf[k].grad for k in keys]
    packed = sharding.pack_gradients(grads)
    ref = torch.cat([x.sum(0) for x in grads], dim=1)
    assert packed.shape == (P, 43) and torch.allclose(packed, ref, rtol=1e-6, atol=1e-7 * float(ref.abs().max()))

    step = grb.GraphedRenderStep(stack("xyz"), stack("features"), stack("opacity"), stack("scaling"),
                                 stack("rotation"), st, dL_dcolor=dL)
    for _ in range(2):
        step.replay()
    step.check()
    assert step.launches_per_replay >= 8
    assert torch.equal(step.color, color.detach()) and torch.equal(step.radii, radii)
    assert util.rel_err(step.packed.cpu().numpy(), ref.cpu().numpy()) <= 1e-4
    assert abs(float(step.loss) - float((color.detach() * dL).sum())) <= 1e-3 * abs(float(step.loss)) + 1e-3


BIG = [
    dict(name="C4", P=300_000, H=1024, W=1024, seed=4000, scale_mult=1.0),    # BASELINE config 4 (one frame)
    dict(name="C5", P=1_000_000, H=1080, W=1920, seed=5000, scale_mult=4.0),  # BASELINE config 5
]


@pytest.mark.parametrize("case", BIG, ids=lambda c: c["name"])
def test_full_size_configs_vs_reference_build(ref, case, sort_mode):
    """BASELINE configs 4 and 5 at full size against the unmodified reference build: the bit-exact set
    (keys, lists, ranges, radii, n_contrib, final_T), the image tolerances and all gradients."""
    cap = 1 << 24
    grb.set_option("capacity_hint", cap)
    P, H, W = case["P"], case["H"], case["W"]
    inp = util.inputs_for(P, H, W, case["seed"], scale_mult=case["scale_mult"])
    R, color, radii, geom, binning, img, invd = util.c_forward(ours, inp, DEV)
    torch.cuda.synchronize()
    o = util.decode_ours(geom, binning, img, 1, P, W, H, cap)
    Rr, color_r, radii_r, geom_r, binning_r, img_r, invd_r = util.c_forward(ref, inp, DEV)
    torch.cuda.synchronize()
    r = util.decode_reference(geom_r, binning_r, img_r, P, H, W, Rr)
    assert R == Rr and R > 2 * P
    assert torch.equal(radii, radii_r)
    assert np.array_equal(o["keys"], r["keys"])
    assert np.array_equal(o["point_list"], r["point_list"])
    assert np.array_equal(o["ranges"], r["ranges"])
    assert np.array_equal(o["n_contrib"], r["n_contrib"])
    assert np.array_equal(_bits(o["final_T"]), _bits(r["final_T"]))
    assert float((color - color_r).abs().max()) <= 1e-5
    assert float((invd - invd_r).abs().max()) <= 1e-5
    del o, r, geom, binning, img, geom_r, binning_r, img_r
    from guava_renderer_b200 import synthetic
    dc, _ = synthetic.upstream_grads(H, W, 77)
    _, go = _run_autograd(ours, inp, dc * 1e4, None)
    _, gr = _run_autograd(ref, inp, dc * 1e4, None)
    for k in gr:
        assert util.rel_err(go[k], gr[k]) <= 1e-4, (k, util.rel_err(go[k], gr[k]))
        util.grad_gate(k, go[k], gr[k])


def test_refiner_handoff_is_the_same_image_in_channels_last_bf16():
    """SURVEY 8(f) row 3 (gaussian_render.py:70-85): the epilogue's channels-last bf16 copy equals
    permute + round-to-nearest-even cast of the fp32 image, bit for bit; raw/extra renders are views."""
    V, P, H, W = 2, 6000, 136, 200
    views = [util.inputs_for(P, H, W, seed=900 + v, view=v, num_views=V, scale_mult=1.2) for v in range(V)]
    stack = lambda k: torch.stack([x[k] for x in views], 0).to(DEV)
    cam = {"world_view_transform": stack("viewmatrix"), "full_proj_transform": stack("projmatrix"),
           "tanfovx": torch.tensor([x["tanfovx"] for x in views]), "tanfovy": torch.tensor([x["tanfovy"] for x in views]),
           "image_height": H, "image_width": W}
    assets = dict(xyz=stack("xyz"), opacity=stack("opacity"), scaling=stack("scaling"), rotation=stack("rotation"),
                  features_color=stack("features"))
    out = grb.render_views(assets, cam, bg=0.5, refiner_input=True)
    torch.cuda.synchronize()
    img = out["rendered_images"]
    ri = out["refiner_input"]
    assert ri.shape == img.shape and ri.dtype == torch.bfloat16
    assert ri.is_contiguous(memory_format=torch.channels_last)
    assert torch.equal(ri.view(torch.int16), img.to(torch.bfloat16).view(torch.int16))
    assert out["raw_renders"].data_ptr() == img.data_ptr() and out["raw_renders"].shape == (V, 3, H, W)
    assert out["extra_renders"].data_ptr() == img[:, 3:4].data_ptr()
    # the request is consumed by one call: the next call must not write to the old buffer
    ri.zero_()
    out2 = grb.render_views(assets, cam, bg=0.5)
    torch.cuda.synchronize()
    assert "refiner_input" not in out2 and int(ri.view(torch.int16).abs().max()) == 0


def test_equal_depth_ties_keep_index_order(sort_mode):
    """Gaussians with bit-identical depth in one tile must come out in ascending index order (the reference's
    stable sort over an index-ascending emission).  Every Gaussian is duplicated three times, so every tile
    list is full of ties, at list lengths that reach every size class of the per-tile sort."""
    base = util.inputs_for(2500, 96, 128, seed=61, scale_mult=2.0)
    inp = dict(base)
    for k in ("xyz", "features", "opacity", "scaling", "rotation"):
        inp[k] = torch.cat([base[k], base[k], base[k]], 0).contiguous()
    o = _ours_forward(inp)
    f = util.oracle_forward(inp)
    lens = f["ranges"][:, 1].astype(np.int64) - f["ranges"][:, 0]
    assert lens.max() > 256 and (lens > 256).sum() >= 4
    assert np.array_equal(o["ranges"], f["ranges"])
    assert np.array_equal(o["keys"], f["keys"])
    assert np.array_equal(o["point_list"], f["point_list"])
    k = o["keys"]
    same = k[1:] == k[:-1]
    assert same.mean() > 0.5 and np.all(o["point_list"][1:][same] > o["point_list"][:-1][same])


def test_invdepth_only_loss_and_repeated_backward(ref):
    """Upstream gradients autograd leaves undefined (an output the loss does not use) are not
    materialised; a loss on invdepth alone matches the reference build, and a second backward through
    the same batched graph (retain_graph) starts from a fresh gradient block."""
    case = CASES[1]
    inp = util.inputs_for(**case)
    g = torch.Generator().manual_seed(3)
    dL_dinv = torch.randn(1, case["H"], case["W"], generator=g).to(DEV)

    def single(mod):
        leaf = {k: inp[k].to(DEV).clone().requires_grad_(True) for k in ("xyz", "features", "opacity", "scaling", "rotation")}
        m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
        rast = mod.GaussianRasterizer_32(raster_settings=util.settings_for(mod, inp, DEV, False))
        color, radii, invd = rast(means3D=leaf["xyz"], means2D=m2, shs=None, colors_precomp=leaf["features"],
                                  opacities=leaf["opacity"], scales=leaf["scaling"], rotations=leaf["rotation"],
                                  cov3D_precomp=None)
        (invd * dL_dinv).sum().backward()
        torch.cuda.synchronize()
        return {k: v.grad.cpu().numpy() for k, v in leaf.items()}

    go, gr = single(ours), single(ref)
    for k in ("xyz", "opacity", "scaling", "rotation"):
        assert util.rel_err(go[k], gr[k]) <= 1e-4, (k, util.rel_err(go[k], gr[k]))
    assert not go["features"].any() and not gr["features"].any()

    # batched entry: the same loss, then the same backward twice
    leaf = {k: inp[k].to(DEV).clone()[None].requires_grad_(True) for k in ("xyz", "features", "opacity", "scaling", "rotation")}
    m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
    st = grb.BatchRasterizationSettings(case["H"], case["W"], torch.tensor([[inp["tanfovx"], inp["tanfovy"]]], device=DEV),
                                        inp["bg"].to(DEV)[None], 1.0, inp["viewmatrix"].to(DEV)[None],
                                        inp["projmatrix"].to(DEV)[None])
    color, radii, invd = grb.rasterize_gaussians_batch(leaf["xyz"], m2, leaf["features"], leaf["opacity"],
                                                       leaf["scaling"], leaf["rotation"], st)
    loss = (invd[0] * dL_dinv).sum()
    loss.backward(retain_graph=True)
    first = {k: v.grad.clone() for k, v in leaf.items()}
    for k in ("xyz", "opacity", "scaling", "rotation"):
        assert util.rel_err(first[k][0].cpu().numpy(), gr[k]) <= 1e-4, k
    for v in leaf.values():
        v.grad = None
    loss.backward()
    torch.cuda.synchronize()
    for k in ("xyz", "opacity", "scaling", "rotation"):
        assert util.rel_err(leaf[k].grad.cpu().numpy(), first[k].cpu().numpy()) <= 1e-4, k


@pytest.mark.parametrize("V,P", [(1, 1), (2, 4099), (3, 1001), (4, 257), (5, 33), (8, 130)])
def test_pack_view_gradients_all_batch_sizes(V, P):
    """Every view-count branch of the packing kernel (templated 1/2/4/8 and the runtime loop), row counts
    that do not fill the last warp's four rows: equal to the view-ordered sum, bit for bit."""
    from guava_renderer_b200 import sharding
    g = torch.Generator().manual_seed(V * 1000 + P)
    grads = [torch.randn(V, P, w, generator=g).to(DEV) for w in sharding.GRAD_WIDTHS]
    packed = sharding.pack_gradients(grads)
    cols = []
    for x in grads:
        acc = torch.zeros_like(x[0])
        for v in range(V):
            acc = acc + x[v]
        cols.append(acc)
    assert torch.equal(packed, torch.cat(cols, dim=1))


def test_graphed_step_reports_capacity_overflow_through_its_status_words():
    """A captured forward whose binning capacity is too small skips its instance kernels and says so in
    the step's pinned status words (written by the tile scan, no copy node); a normal capture does not."""
    V, P, H, W = 2, 4000, 96, 96
    views = [util.inputs_for(P, H, W, seed=950 + v, view=v, num_views=V, scale_mult=1.5) for v in range(V)]
    stack = lambda k: torch.stack([x[k] for x in views], 0).to(DEV)
    tanfov = torch.tensor([[x["tanfovx"], x["tanfovy"]] for x in views], device=DEV)
    st = grb.BatchRasterizationSettings(H, W, tanfov, torch.zeros(V, 32, device=DEV), 1.0,
                                        stack("viewmatrix"), stack("projmatrix"))
    args = (stack("xyz"), stack("features"), stack("opacity"), stack("scaling"), stack("rotation"), st)
    ok = grb.GraphedRenderStep(*args)
    ok.replay()
    ok.check()
    R = int(ok._status[0])
    assert R > 64 and int(ok._status[1]) > 0
    grb.set_option("capacity_hint", 64)
    try:
        small = grb.GraphedRenderStep(*args)
        small.replay()
        with pytest.raises(RuntimeError, match="capacity"):
            small.check()
        assert int(small._status[0]) == R
    finally:
        grb.set_option("capacity_hint", 0)
    ok.replay()
    ok.check()


@pytest.mark.parametrize("n", [1, 3, 4, 1027, 2 * 32 * 96 * 128, 5_000_003])
def test_image_dot_matches_float64_and_is_deterministic(n):
    from guava_renderer_b200.graphed import image_dot
    g = torch.Generator().manual_seed(n)
    a = torch.randn(n, generator=g).to(DEV)
    b = torch.randn(n, generator=g).to(DEV)
    buf = torch.zeros(64, device=DEV)
    first = float(image_dot(a, b, buf))
    again = float(image_dot(a, b, buf))          # the scratch words are left ready for reuse
    want = float((a.double() * b.double()).sum())
    assert first == again
    assert abs(first - want) <= 1e-5 * float((a.double() * b.double()).abs().sum()) + 1e-6


def test_list_longer_than_the_launched_sort_is_recovered():
    """While a shape's tile lists stay short the long-bucket sort is not launched; a call of the same shape
    whose lists outgrow the register-resident sort is detected by the tile scan and redone -- same image,
    bit for bit, as with exact binning (which always launches both sorts)."""
    P, H, W = 6000, 64, 64
    short = util.inputs_for(P, H, W, seed=77, scale_mult=0.3)
    long_ = util.inputs_for(P, H, W, seed=78, scale_mult=8.0)
    grb.set_option("sort_mode", 0)
    try:
        for _ in range(2):
            _ours_forward(short)                       # the shape is now known with short lists
        got = _ours_forward(long_)
        r = got["ranges"].reshape(-1, 2)
        assert int((r[:, 1] - r[:, 0]).max()) > 2048   # the case is what it claims to be
        grb.set_option("exact_binning", 1)
        want = _ours_forward(long_)
    finally:
        grb.set_option("exact_binning", 0)
        grb.set_option("sort_mode", -1)
    for k in ("color", "invdepth", "radii", "n_contrib", "point_list", "ranges"):
        assert np.array_equal(got[k], want[k]), k


# ------------------------------------------------------------------ rows a20, a8/a17 of SURVEY.md section 8
def test_debug_mode_gives_the_same_results(ref):
    """`debug=True` (CHECK_CUDA, auxiliary.h:178-185: synchronise and check after every stage) must not change a
    result: forward bitwise equal to the non-debug call, gradients equal to atomic-order tolerance, and equal to
    the reference build run with ITS debug flag."""
    case = CASES[1]
    inp = util.inputs_for(**case)
    inp["bg"] = torch.linspace(-1, 1, 32)
    g = torch.Generator().manual_seed(11)
    dL = torch.randn(32, case["H"], case["W"], generator=g)

    def run(mod, debug):
        leaf = {k: inp[k].to(DEV).clone().requires_grad_(True) for k in ("xyz", "features", "opacity", "scaling", "rotation")}
        m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
        rast = mod.GaussianRasterizer_32(raster_settings=util.settings_for(mod, inp, DEV, False, debug=debug))
        color, radii, invd = rast(means3D=leaf["xyz"], means2D=m2, shs=None, colors_precomp=leaf["features"],
                                  opacities=leaf["opacity"], scales=leaf["scaling"], rotations=leaf["rotation"],
                                  cov3D_precomp=None)
        (color * dL.to(DEV)).sum().backward()
        torch.cuda.synchronize()
        return color.detach().cpu().numpy(), radii.cpu().numpy(), {k: v.grad.cpu().numpy() for k, v in leaf.items()}

    c0, r0, g0 = run(ours, False)
    c1, r1, g1 = run(ours, True)
    cr, rr, gr = run(ref, True)
    assert np.array_equal(c0, c1) and np.array_equal(r0, r1) and np.array_equal(r1, rr)
    assert np.abs(c1 - cr).max() <= 1e-5
    for k in g0:
        assert util.rel_err(g1[k], g0[k]) <= 5e-5, k      # own run against own run: reduction-order jitter
        util.grad_gate(k, g1[k], gr[k])


_TRAP_SCRIPT = r"""
import os, sys, torch
sys.path.insert(0, {root!r})
from tests import util
import diff_gaussian_rasterization_32 as ours
inp = util.inputs_for(512, 64, 64, seed=8)
inp["xyz"][::3, 2] -= 40.0            # a third of the points behind the near plane: in_frustum() rejects them
rs = util.settings_for(ours, inp, "cuda:0", False, debug=True)._replace(prefiltered=True)
d = lambda k: inp[k].to("cuda:0")
try:
    ours.GaussianRasterizer_32(rs)(means3D=d("xyz"), means2D=d("xyz"), opacities=d("opacity"), colors_precomp=d("features"),
                                  scales=d("scaling"), rotations=d("rotation"))
    torch.cuda.synchronize()
except RuntimeError as e:
    print("RAISED:", e, flush=True)
    os._exit(0 if "preprocess" in str(e) else 3)      # the context is poisoned: skip the interpreter's CUDA teardown
os._exit(4)
"""


def test_prefiltered_trap_is_reported_by_debug_mode():
    """auxiliary.h:166-172: with `prefiltered` set, a point that fails the frustum test prints a message and
    executes __trap(); in debug mode the stage check turns that into an exception naming the stage
    (CHECK_CUDA).  The trap poisons the CUDA context, so this runs in its own process."""
    import subprocess
    import sys
    r = subprocess.run([sys.executable, "-c", _TRAP_SCRIPT.format(root=util.ROOT)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.returncode, r.stdout[-400:], r.stderr[-400:])
    assert "RAISED:" in r.stdout and "CUDA ERROR" in r.stdout


def test_cov3d_precomp_backward_vs_reference_build(ref):
    """The cov3D_precomp input path end to end (forward.cu:206-215, backward.cu:147-326): images and the
    gradients the reference returns for it -- dL_dcov3D above all (__init__.py:137) -- against the reference
    build.  The covariances are the reference's own (GeometryState::cov3D of a scales+rotations call)."""
    case = dict(P=3000, H=96, W=128, seed=12, scale_mult=0.8)
    inp = util.inputs_for(**case)
    inp["bg"] = torch.linspace(0, 1, 32)
    R, _, _, geom, binning, img, _ = util.c_forward(ref, inp, DEV)
    cov = torch.from_numpy(util.decode_reference(geom, binning, img, case["P"], case["H"], case["W"], R)["cov3D"]).to(DEV)
    g = torch.Generator().manual_seed(5)
    dL = torch.randn(32, case["H"], case["W"], generator=g).to(DEV)
    dLi = torch.randn(1, case["H"], case["W"], generator=g).to(DEV)

    def run(mod):
        leaf = {k: inp[k].to(DEV).clone().requires_grad_(True) for k in ("xyz", "features", "opacity")}
        c3 = cov.clone().requires_grad_(True)
        m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
        rast = mod.GaussianRasterizer_32(raster_settings=util.settings_for(mod, inp, DEV))
        color, radii, invd = rast(means3D=leaf["xyz"], means2D=m2, shs=None, colors_precomp=leaf["features"],
                                  opacities=leaf["opacity"], scales=None, rotations=None, cov3D_precomp=c3)
        ((color * dL).sum() + (invd * dLi).sum()).backward()
        torch.cuda.synchronize()
        out = {k: v.grad.cpu().numpy() for k, v in leaf.items()}
        out["cov3D"] = c3.grad.cpu().numpy()
        out["means2D"] = m2.grad.cpu().numpy()
        return color.detach().cpu().numpy(), radii.cpu().numpy(), out

    co, ro, go = run(ours)
    cr, rr, gr = run(ref)
    assert np.array_equal(ro, rr) and np.abs(co - cr).max() <= 1e-5
    assert np.abs(gr["cov3D"]).max() > 0
    for k in gr:
        util.grad_gate(k, go[k], gr[k])


def test_footprint_culling_is_output_neutral_on_the_device():
    """Brute force on the device (guava_selftest_cull): for random splats against random 8x4 blocks, whenever the
    compositors' footprint test rejects a splat, the reference's alpha test (forward.cu:350-362), evaluated with
    the compositors' own arithmetic, fails at all 32 pixel centres.  Splat shapes cover the benchmark's (sigma
    0.5..30 px, any orientation, opacity 0.002..0.999) plus degenerate conics; means lie inside, beside and far
    from the block."""
    import ctypes
    lib = grb._lib_mod.load()
    n = 2_000_000
    g = torch.Generator().manual_seed(123)
    sig = torch.exp(torch.rand(n, 2, generator=g) * 4.1 - 0.7)                 # 0.5 .. 30 px
    th = torch.rand(n, generator=g) * 3.14159265
    c, s_ = torch.cos(th), torch.sin(th)
    a = (c / sig[:, 0]) ** 2 + (s_ / sig[:, 1]) ** 2
    cc = (s_ / sig[:, 0]) ** 2 + (c / sig[:, 1]) ** 2
    b = c * s_ * (1 / sig[:, 0] ** 2 - 1 / sig[:, 1] ** 2)
    op = torch.sigmoid(1.0 + 2.0 * torch.randn(n, generator=g)).clamp(0.002, 0.999)
    org = torch.stack([torch.randint(0, 64, (n,), generator=g) * 8.0, torch.randint(0, 128, (n,), generator=g) * 4.0], 1)
    # distance of the mean from the block: within a few sigma of the boundary, where the decision is made
    reach = sig.max(1).values * torch.sqrt(2 * torch.log(255 * op).clamp_min(0.0) + 1e-3)
    off = (torch.rand(n, 2, generator=g) * 2 - 1) * (reach[:, None] * 1.4 + 6.0)
    mean = org + torch.tensor([3.5, 1.5]) + off
    splats = torch.stack([mean[:, 0], mean[:, 1], a, b, cc, op], 1).contiguous()
    splats[:1000, 2] = 0.0                      # degenerate / not positive definite: must be kept
    splats[1000:2000, 3] = 4.0 * torch.sqrt(splats[1000:2000, 2] * splats[1000:2000, 4])   # B^2 = 16 A C: indefinite
    splats[2000:3000, 4] = float("nan")
    sp, og = splats.to(DEV), org.contiguous().to(DEV)
    rej = torch.zeros(n, dtype=torch.uint8, device=DEV)
    pas = torch.zeros(n, dtype=torch.uint8, device=DEV)
    rc = lib.guava_selftest_cull(n, sp.data_ptr(), og.data_ptr(), rej.data_ptr(), pas.data_ptr(),
                                 ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, grb._lib_mod.last_error()
    torch.cuda.synchronize()
    rej, pas = rej.cpu().numpy(), pas.cpu().numpy()
    assert not np.any((rej != 0) & (pas != 0)), int(np.sum((rej != 0) & (pas != 0)))
    assert not rej[:3000].any()                                   # degenerate conics and NaNs are always kept
    # the sample exercises the test: a good share rejected, a good share kept, and the test is tight
    frac_rej = float((rej != 0).mean())
    kept_untouched = float(((rej == 0) & (pas == 0)).mean())
    assert 0.2 < frac_rej < 0.95, frac_rej
    assert kept_untouched < 0.6, kept_untouched


def test_graphed_training_step_that_overflows_leaves_zero_gradients_and_reports_it():
    """A replayed TRAINING step whose captured binning capacity is too small: the forward's instance kernels skip
    themselves, and so must the backward compositor and the backward preprocess -- they would otherwise walk lists
    that were never written.  No CUDA error, all gradients zero, and check() raises."""
    V, P, H, W = 2, 4000, 96, 96
    views = [util.inputs_for(P, H, W, seed=950 + v, view=v, num_views=V, scale_mult=1.5) for v in range(V)]
    stack = lambda k: torch.stack([x[k] for x in views], 0).to(DEV)
    tanfov = torch.tensor([[x["tanfovx"], x["tanfovy"]] for x in views], device=DEV)
    st = grb.BatchRasterizationSettings(H, W, tanfov, torch.zeros(V, 32, device=DEV), 1.0,
                                        stack("viewmatrix"), stack("projmatrix"))
    g = torch.Generator().manual_seed(4)
    dL = torch.randn(V, 32, H, W, generator=g).to(DEV)
    args = (stack("xyz"), stack("features"), stack("opacity"), stack("scaling"), stack("rotation"), st)
    ok = grb.GraphedRenderStep(*args, dL_dcolor=dL)
    ok.replay()
    ok.check()
    assert float(ok.packed.abs().sum()) > 0
    grb.set_option("capacity_hint", 64)
    try:
        small = grb.GraphedRenderStep(*args, dL_dcolor=dL)
        for _ in range(3):
            small.replay()
        torch.cuda.synchronize()                                  # no sticky CUDA error
        assert float(small.packed.abs().sum()) == 0.0
        for k, v in small.grads.items():
            assert float(v.abs().sum()) == 0.0, k
        with pytest.raises(RuntimeError, match="capacity"):
            small.check()
    finally:
        grb.set_option("capacity_hint", 0)
    ok.replay()
    ok.check()
    assert float(ok.packed.abs().sum()) > 0


_CTYPES_SCRIPT = r"""
import os, sys, torch, numpy as np
sys.path.insert(0, {root!r})
from tests import util
import guava_renderer_b200 as grb
import diff_gaussian_rasterization_32 as ours
assert grb.rasterizer.BINDING == os.environ["GUAVA_BINDING"]
inp = util.inputs_for(3000, 200, 152, seed=21, scale_mult=1.5)
inp["bg"] = torch.linspace(-1, 1, 32)
g = torch.Generator().manual_seed(9)
dL = torch.randn(32, 200, 152, generator=g).to("cuda:0")
leaf = {{k: inp[k].to("cuda:0").clone().requires_grad_(True) for k in ("xyz", "features", "opacity", "scaling", "rotation")}}
m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
rast = ours.GaussianRasterizer_32(raster_settings=util.settings_for(ours, inp, "cuda:0"))
color, radii, invd = rast(means3D=leaf["xyz"], means2D=m2, shs=None, colors_precomp=leaf["features"], opacities=leaf["opacity"],
                          scales=leaf["scaling"], rotations=leaf["rotation"], cov3D_precomp=None)
(color * dL).sum().backward()
vis = rast.markVisible(leaf["xyz"].detach())
V = 2
st = grb.BatchRasterizationSettings(200, 152, torch.tensor([[inp["tanfovx"], inp["tanfovy"]]] * V, device="cuda:0"),
                                    inp["bg"].to("cuda:0")[None].repeat(V, 1), 1.0, inp["viewmatrix"].to("cuda:0")[None].repeat(V, 1, 1),
                                    inp["projmatrix"].to("cuda:0")[None].repeat(V, 1, 1))
bl = {{k: inp[k].to("cuda:0").clone().requires_grad_(True) for k in ("xyz", "features", "opacity", "scaling", "rotation")}}
bc, br, bi = grb.rasterize_gaussians_batch(bl["xyz"], torch.zeros(V, 3000, 3, device="cuda:0"), bl["features"], bl["opacity"],
                                           bl["scaling"], bl["rotation"], st)
bc.backward(torch.stack([dL, dL]))
torch.cuda.synchronize()
np.savez(sys.argv[1], color=color.detach().cpu().numpy(), radii=radii.cpu().numpy(), invd=invd.detach().cpu().numpy(),
         vis=vis.cpu().numpy(), bcolor=bc.detach().cpu().numpy(), **{{"g_" + k: v.grad.cpu().numpy() for k, v in leaf.items()}},
         **{{"b_" + k: v.grad.cpu().numpy() for k, v in bl.items()}}, g_means2D=m2.grad.cpu().numpy())
"""


def test_compiled_and_ctypes_bindings_agree(tmp_path):
    """The default `_C` is the compiled torch binding (csrc/torch_binding.cpp, what replaces DGR32/ext.cpp +
    rasterize_points.cu); GUAVA_BINDING=ctypes selects the ctypes binding of the same C ABI.  The reference's call
    sequence (gaussian_render.py:37-67), markVisible and the batched entry (shared cloud, 2 views) through each, in
    separate processes: forward outputs bit for bit, gradients to atomic-order tolerance."""
    import subprocess
    import sys
    out = {}
    for b in ("compiled", "ctypes"):
        path = str(tmp_path / (b + ".npz"))
        env = dict(os.environ, GUAVA_BINDING=b)
        r = subprocess.run([sys.executable, "-c", _CTYPES_SCRIPT.format(root=util.ROOT), path], capture_output=True, text=True,
                           timeout=600, env=env)
        assert r.returncode == 0, (b, r.stderr[-800:])
        out[b] = np.load(path)
    a, c = out["compiled"], out["ctypes"]
    for k in ("color", "radii", "invd", "vis", "bcolor"):
        assert np.array_equal(a[k], c[k]), k
    assert np.array_equal(a["bcolor"][0], a["color"]) and np.array_equal(a["bcolor"][1], a["color"])
    for k in a.files:
        if k.startswith(("g_", "b_")):     # two runs differ by the order of their floating-point reductions
            assert util.rel_err(a[k], c[k]) <= 5e-5, k
    for k in ("xyz", "features", "opacity", "scaling", "rotation"):     # shared cloud: the two views' gradients are summed
        assert util.rel_err(a["b_" + k], 2.0 * a["g_" + k]) <= 5e-5, k


@pytest.mark.parametrize("fwd_opt,bwd_opt", [(0, 1), (1, 0)])
def test_taken_masks_option_may_change_between_forward_and_backward(fwd_opt, bwd_opt):
    """The forward notes in the call's header whether it marked; a backward that expects marks after a forward that
    wrote none walks every entry, a backward that ignores them runs its own footprint test: same gradients."""
    case = CASES[1]
    inp = util.inputs_for(**case)
    inp["bg"] = torch.linspace(-1, 1, 32)
    g = torch.Generator().manual_seed(9)
    dL = torch.randn(32, case["H"], case["W"], generator=g)
    _, want = _run_autograd(ours, inp, dL, None)
    leaf = {k: inp[k].to(DEV).clone().requires_grad_(True) for k in ("xyz", "features", "opacity", "scaling", "rotation")}
    m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
    base = grb.get_option("taken_masks")
    try:
        grb.set_option("taken_masks", fwd_opt)
        rast = ours.GaussianRasterizer_32(raster_settings=util.settings_for(ours, inp, DEV))
        color, radii, invd = rast(means3D=leaf["xyz"], means2D=m2, shs=None, colors_precomp=leaf["features"],
                                  opacities=leaf["opacity"], scales=leaf["scaling"], rotations=leaf["rotation"], cov3D_precomp=None)
        loss = (color * dL.to(DEV)).sum()
        torch.cuda.synchronize()
        grb.set_option("taken_masks", bwd_opt)
        loss.backward()
        torch.cuda.synchronize()
    finally:
        grb.set_option("taken_masks", base)
    for k, v in leaf.items():
        assert util.rel_err(v.grad.cpu().numpy(), want[k]) <= 5e-5, k


@pytest.mark.parametrize("how", ["rerun_after_overflow", "exact_binning"])
def test_backward_after_the_binning_buffer_was_resized(how):
    """The backward finds the forward's taken marks behind point_list from the capacity the call's header records:
    gradients after a forward whose binning buffer was re-sized (estimate too small -> tail re-enqueued with the exact
    size; or the reference's wait-then-allocate order) equal those of a forward with ample capacity."""
    case = CASES[1]
    inp = util.inputs_for(**case)
    inp["bg"] = torch.linspace(-1, 1, 32)
    g = torch.Generator().manual_seed(9)
    dL = torch.randn(32, case["H"], case["W"], generator=g)
    dLi = torch.randn(1, case["H"], case["W"], generator=g)
    c0, want = _run_autograd(ours, inp, dL, dLi)
    try:
        if how == "exact_binning":
            grb.set_option("exact_binning", 1)
        else:
            grb.set_option("capacity_hint", 512)
        c1, got = _run_autograd(ours, inp, dL, dLi)
    finally:
        grb.set_option("exact_binning", 0)
        grb.set_option("capacity_hint", CAP)
    assert np.array_equal(c0, c1)
    for k in want:
        assert util.rel_err(got[k], want[k]) <= 5e-5, (k, util.rel_err(got[k], want[k]))


def test_inference_calls_of_the_batched_entry_record_no_taken_marks():
    """guava_call_options::forward_only: a batched forward that no backward can follow (no input requires a gradient)
    skips the marks and their 8 bytes per instance; the image is the same bit for bit."""
    V, P, H, W = 2, 5000, 96, 112
    views = [util.inputs_for(P, H, W, seed=700 + v, view=v, num_views=V, scale_mult=1.3) for v in range(V)]
    stack = lambda k: torch.stack([x[k] for x in views], 0).to(DEV)
    tanfov = torch.tensor([[x["tanfovx"], x["tanfovy"]] for x in views], device=DEV)
    st = grb.BatchRasterizationSettings(H, W, tanfov, torch.zeros(V, 32, device=DEV), 1.0, stack("viewmatrix"), stack("projmatrix"))
    args = lambda rg: (stack("xyz").requires_grad_(rg), torch.zeros(V, P, 3, device=DEV), stack("features"), stack("opacity"),
                       stack("scaling"), stack("rotation"), st)
    out = {}      # (the fixture pins the binning capacity: the two buffers differ by the marks only)
    for rg in (False, True):
        color, radii, invd = grb.rasterize_gaussians_batch(*args(rg))
        torch.cuda.synchronize()
        geom, binning, img = grb.rasterizer._tls.last_buffers
        o = util.decode_ours(geom, binning, img, V, P, W, H)
        out[rg] = (color.detach().clone(), o["block_masks"], binning.numel())
    assert out[False][1] == 0 and out[True][1] == 1
    assert out[False][2] < out[True][2]
    assert torch.equal(out[False][0], out[True][0])


@pytest.mark.parametrize("binding", ["compiled", "ctypes"])
@pytest.mark.parametrize("shared", [False, True], ids=["per_view_clouds", "shared_cloud"])
@pytest.mark.parametrize("antialiasing", [False, True], ids=["plain", "antialiasing"])
def test_summed_backward_equals_the_sum_of_the_per_view_gradients(binding, shared, antialiasing, monkeypatch):
    """guava_rasterizer_backward_batch_summed (sum over the views inside the two backward kernels, straight into the
    packed [P,43] rows) against the per-view entry followed by pack_gradients, with an inverse-depth gradient, for
    per-view clouds and for one cloud shared by all views (view strides 0: the sum is that tensor's gradient), through
    both bindings; then the dense form (out_row_stride = 0) through the C ABI."""
    import ctypes
    from guava_renderer_b200 import rasterizer as rz
    if binding == "compiled" and rz.BINDING != "compiled":
        pytest.skip("this process runs with GUAVA_BINDING=ctypes")
    monkeypatch.setattr(rz, "BINDING", binding)      # the batched entry picks its binding per call
    V, P, H, W = 3, (5001 if shared else 5000), 80, 112      # 5001: the packed buffer's size is not a multiple of 16 bytes
    views = [util.inputs_for(P, H, W, seed=700 + (0 if shared else v), view=v, num_views=V, scale_mult=1.3) for v in range(V)]
    stack = lambda k: torch.stack([x[k] for x in views], 0).to(DEV)
    cloud = (lambda k: views[0][k].to(DEV)) if shared else stack
    g = torch.Generator().manual_seed(4)
    dL = torch.randn(V, 32, H, W, generator=g).to(DEV)
    dLi = torch.randn(V, 1, H, W, generator=g).to(DEV)
    tanfov = torch.tensor([[x["tanfovx"], x["tanfovy"]] for x in views], device=DEV)
    st = grb.BatchRasterizationSettings(H, W, tanfov, torch.linspace(-1, 1, 32, device=DEV).repeat(V, 1).contiguous(), 1.0,
                                        stack("viewmatrix"), stack("projmatrix"), antialiasing=antialiasing)
    grb.set_option("capacity_hint", 0)
    keys = ("xyz", "scaling", "rotation", "opacity", "features")

    def run(summed_out):
        leaf = {k: cloud(k).clone().requires_grad_(True) for k in keys}
        m2 = torch.zeros(V, P, 3, device=DEV, requires_grad=True)
        color, radii, invd = grb.rasterize_gaussians_batch(leaf["xyz"], m2, leaf["features"], leaf["opacity"],
                                                           leaf["scaling"], leaf["rotation"], st, summed_out=summed_out)
        ((color * dL).sum() + (invd * dLi).sum()).backward()
        torch.cuda.synchronize()
        return leaf, color.detach()

    leaf, color0 = run(None)
    grads = [leaf[k].grad for k in keys]
    want = torch.cat([x if x.dim() == 2 else x.sum(0) for x in grads], dim=1).cpu().numpy()
    packed = torch.full((P, 43), float("nan"), device=DEV)      # the forward zero-fills it (guava_call_options::zero_fill)
    leaf1, color1 = run(packed)
    assert torch.equal(color0, color1)
    assert all(leaf1[k].grad is None for k in keys)          # the sum is the only output of that backward
    got = packed.cpu().numpy()
    assert np.abs(want[:, 3:10]).max() > 0 and np.abs(want[:, 10]).max() > 0
    for name, sl in (("xyz", slice(0, 3)), ("scaling", slice(3, 6)), ("rotation", slice(6, 10)), ("opacity", slice(10, 11)),
                     ("features", slice(11, 43))):
        assert util.rel_err(got[:, sl], want[:, sl]) <= 5e-5, name
        util.grad_gate("summed " + name, got[:, sl], want[:, sl])
    if binding != "ctypes" or antialiasing:
        return
    # dense outputs: five [P,w] tensors, same call with out_row_stride = 0 on the buffers of one more forward
    lib = grb._lib_mod.load()
    leaf2 = {k: cloud(k).clone().requires_grad_(True) for k in keys}
    m2 = torch.zeros(V, P, 3, device=DEV, requires_grad=True)
    color, radii, invd = grb.rasterize_gaussians_batch(leaf2["xyz"], m2, leaf2["features"], leaf2["opacity"],
                                                       leaf2["scaling"], leaf2["rotation"], st)
    geom, binning, img = rz._tls.last_buffers
    outs = {k: torch.zeros(P, w, device=DEV) for k, w in zip(keys, (3, 3, 4, 1, 32))}
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    vs = lambda k, w: 0 if shared else P * w
    x = {k: leaf2[k].detach().contiguous() for k in keys}
    rc = lib.guava_rasterizer_backward_batch_summed(
        V, P, 0, p(st.bg), W, H, p(x["xyz"]), vs("xyz", 3), p(x["features"]), vs("features", 32), p(x["opacity"]),
        vs("opacity", 1), p(x["scaling"]), vs("scaling", 3), 1.0, p(x["rotation"]), vs("rotation", 4),
        p(st.viewmatrix), p(st.projmatrix), p(tanfov), p(radii), p(geom), p(binning), p(img), p(dL), p(dLi),
        p(outs["xyz"]), p(outs["scaling"]), p(outs["rotation"]), p(outs["opacity"]), p(outs["features"]), 0, 0, 0,
        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, grb._lib_mod.last_error()
    torch.cuda.synchronize()
    dense = torch.cat([outs[k] for k in keys], dim=1).cpu().numpy()
    assert util.rel_err(dense, want) <= 5e-5
    rc = lib.guava_rasterizer_backward_batch_summed(
        V, P, 0, p(st.bg), W, H, p(x["xyz"]), vs("xyz", 3), p(x["features"]), vs("features", 32), p(x["opacity"]),
        vs("opacity", 1), p(x["scaling"]), vs("scaling", 3), 1.0, p(x["rotation"]), vs("rotation", 4),
        p(st.viewmatrix), p(st.projmatrix), p(tanfov), p(radii), p(geom), p(binning), p(img), p(dL), p(dLi),
        p(outs["xyz"]), p(outs["scaling"]), p(outs["rotation"]), p(outs["opacity"]), p(outs["features"]), 17, 0, 0,
        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == grb._lib_mod.GUAVA_ERR_INVALID_ARGUMENT and "out_row_stride" in grb._lib_mod.last_error()


def _autograd_on(device, inp, dL, stream=None, delayed_fill=False):
    """One forward + backward through GaussianRasterizer_32 with every tensor on `device`.  delayed_fill: the inputs
    are written by a copy that sits on `stream` behind a ~20 ms spin kernel -- only work ordered on that stream sees them."""
    keys = ("xyz", "features", "opacity", "scaling", "rotation")
    with torch.cuda.device(device):
        ctx = torch.cuda.stream(stream) if stream is not None else torch.cuda.stream(torch.cuda.current_stream(device))
        src = {k: inp[k].to(device) for k in keys}
        torch.cuda.synchronize(device)
        with ctx:
            leaf = {k: torch.zeros_like(src[k]) for k in keys}
            if delayed_fill:
                torch.cuda._sleep(40_000_000)
            for k in keys:
                leaf[k].copy_(src[k], non_blocking=True)
                leaf[k].requires_grad_(True)
            m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
            rast = ours.GaussianRasterizer_32(raster_settings=util.settings_for(ours, inp, device))
            color, radii, invd = rast(means3D=leaf["xyz"], means2D=m2, shs=None, colors_precomp=leaf["features"],
                                      opacities=leaf["opacity"], scales=leaf["scaling"], rotations=leaf["rotation"],
                                      cov3D_precomp=None)
            (color * dL.to(device)).sum().backward()
        torch.cuda.synchronize(device)
    return color.detach().cpu(), radii.cpu(), {k: v.grad.cpu().numpy() for k, v in leaf.items()}


def test_calls_are_ordered_on_the_callers_current_stream():
    """The reference launches on the legacy default stream (no stream argument anywhere in rasterize_points.cu); the
    C ABI takes the stream and the bindings pass torch's current one.  Inputs that only exist on a side stream --
    written behind a spin kernel, no synchronisation before the call -- must be what the forward AND the backward see."""
    inp = util.inputs_for(P=4000, H=96, W=96, seed=70, scale_mult=1.0)
    dL = torch.randn(32, 96, 96, generator=torch.Generator().manual_seed(1))
    c0, r0, g0 = _autograd_on(DEV, inp, dL)
    side = torch.cuda.Stream(device=DEV)
    c1, r1, g1 = _autograd_on(DEV, inp, dL, stream=side, delayed_fill=True)
    assert int((r0 > 0).sum()) > 1000
    assert torch.equal(c0, c1) and torch.equal(r0, r1)
    for k in g0:
        assert util.rel_err(g1[k], g0[k]) <= 5e-5, k


def _two_gpus(fn):
    """Collected only where two GPUs are visible (`gpurun --gpus 2`); on a one-GPU box the test does not exist rather
    than showing up as skipped (profiles/r02_v19_two_gpu_tests.log holds its two-GPU run)."""
    return fn if torch.cuda.is_available() and torch.cuda.device_count() >= 2 else None


@_two_gpus
def test_calls_on_a_device_that_is_not_the_current_one():
    """Tensors on cuda:1 while cuda:0 is the current device (the reference has no device guard): same image, same
    radii, same gradients as on cuda:0 -- through the default binding and through the ctypes binding of the batched
    entry (its allocator callbacks, pinned status words and per-device state)."""
    inp = util.inputs_for(P=4000, H=96, W=96, seed=71, scale_mult=1.0)
    dL = torch.randn(32, 96, 96, generator=torch.Generator().manual_seed(1))
    torch.cuda.set_device(0)
    c0, r0, g0 = _autograd_on("cuda:0", inp, dL)
    # not inside `with torch.cuda.device(1)`: the call itself has to find the tensors' device
    keys = ("xyz", "features", "opacity", "scaling", "rotation")
    leaf = {k: inp[k].to("cuda:1").requires_grad_(True) for k in keys}
    m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
    rast = ours.GaussianRasterizer_32(raster_settings=util.settings_for(ours, inp, "cuda:1"))
    color, radii, invd = rast(means3D=leaf["xyz"], means2D=m2, shs=None, colors_precomp=leaf["features"],
                              opacities=leaf["opacity"], scales=leaf["scaling"], rotations=leaf["rotation"], cov3D_precomp=None)
    (color * dL.to("cuda:1")).sum().backward()
    torch.cuda.synchronize("cuda:1")
    assert torch.cuda.current_device() == 0
    assert torch.equal(color.detach().cpu(), c0) and torch.equal(radii.cpu(), r0)
    for k in g0:
        assert util.rel_err(leaf[k].grad.cpu().numpy(), g0[k]) <= 5e-5, k
    # the batched entry on cuda:1
    st = grb.BatchRasterizationSettings(96, 96, torch.tensor([[inp["tanfovx"], inp["tanfovy"]]], device="cuda:1"),
                                        inp["bg"][None].to("cuda:1"), 1.0, inp["viewmatrix"][None].to("cuda:1"),
                                        inp["projmatrix"][None].to("cuda:1"))
    x = {k: inp[k].to("cuda:1") for k in keys}
    cb, rb, _ = grb.rasterize_gaussians_batch(x["xyz"], torch.zeros(1, 4000, 3, device="cuda:1"), x["features"], x["opacity"],
                                              x["scaling"], x["rotation"], st)
    torch.cuda.synchronize("cuda:1")
    assert torch.equal(cb[0].cpu(), c0) and torch.equal(rb[0].cpu(), r0)
