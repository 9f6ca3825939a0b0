"""Seeded random shapes against the unmodified reference build: view counts 1..8 (the summed backward's lane path for
powers of two, its atomic path otherwise), ragged images, tile counts on both sides of the one-block scan's limit,
sparse and dense clouds, both binning modes, antialiasing on and off.  Per case: the bit-exact forward gates per view
(radii, point lists, keys, ranges, n_contrib, final_T, means2D, conics), 1e-5 images, the per-element gradient gate for
the batched per-view backward, and for the view-summed backward against the sum of the reference's per-view gradients.
"""
import os

import numpy as np
import pytest
import torch

from tests import util
from tests.test_gpu_bench_configs import KEYS, _compare_forward_per_view

pytestmark = pytest.mark.gpu

import guava_renderer_b200 as grb  # noqa: E402

DEV = "cuda:0"


@pytest.fixture(scope="module")
def ref():
    m = util.import_reference()
    if m is None:
        pytest.skip("oracle/_ref (reference CUDA build) not present")
    return m


@pytest.fixture(autouse=True)
def _options():
    grb.set_option("emit_sorted_keys", 1)
    grb.set_option("capacity_hint", 0)
    yield
    grb.set_option("emit_sorted_keys", 0)
    grb.set_option("sort_mode", -1)


def _gate(name, a, b, b_again=None):
    """util.grad_gate, whose two fractions (elements outside 1e-4 / 1e-3 relative) are fractions of LARGE tensors: in a
    tensor of a few thousand elements one element of accumulation-order noise on a cancelling sum is already more than
    1e-4 of them, so small tensors are allowed two such elements (the tensor-scale 1e-4 bound holds regardless; the
    worst case seen in 90 random shapes was two elements of 16 152 at a tensor-scale error of 7e-6).
    `b_again`: the reference's gradient from a second run on the same inputs -- the reference is not reproducible run
    to run (39 atomics per pair in arbitrary order), and where its own two runs disagree (a few huge splats, each sum
    over ~1e5 pixels) this implementation is allowed three times that."""
    n = max(int(np.asarray(a).size), 1)
    viol_max, out3_max = max(util.GRAD_VIOL_MAX, 2.5 / n), max(util.GRAD_OUT3_MAX, 2.5 / n)   # two elements in a small tensor
    if b_again is not None:
        own = util.grad_stats(b_again, b)
        viol_max, out3_max = max(viol_max, 3.0 * own["viol"]), max(out3_max, 3.0 * own["out3"])
    return util.grad_gate(name, a, b, viol_max=viol_max, out3_max=out3_max)


def _cases():
    rng = np.random.default_rng(int(os.environ.get("GUAVA_TEST_SEED", "20261105")))   # another seed: another ten shapes
    big = os.environ.get("GUAVA_TEST_BIG", "") == "1"     # larger clouds and images: long tile lists, every sort path
    hi, pmax = (700, 60_000) if big else (330, 9000)
    out = []
    for i in range(10):
        V = int(rng.choice([1, 2, 3, 4, 5, 8]))
        H = int(rng.integers(20, hi))
        W = int(rng.integers(20, hi))
        if i == 7:                       # more than 2048 tiles in total: the chained scan with its fill blocks
            V, H, W = 4, 400, 368
        if i == 2:
            V = 8                        # eight views in adjacent lanes of the summed backward
        if i == 5:
            V = 3                        # not a power of two: its atomic path
        P = int(rng.integers(1, pmax))
        out.append(dict(V=V, H=H, W=W, P=P, scale_mult=float(rng.choice([0.3, 1.0, 2.5, 5.0] if big else [0.3, 1.0, 2.5])), seed=int(rng.integers(1, 10_000)),
                        antialiasing=bool(rng.integers(0, 2)), sort_mode=int(rng.choice([0, 1, -1]))))
    return out


@pytest.mark.parametrize("case", _cases(), ids=lambda c: f"V{c['V']}_P{c['P']}_{c['H']}x{c['W']}_m{c['sort_mode']}")
def test_random_shape_against_reference_build(ref, case):
    V, H, W, P, aa = case["V"], case["H"], case["W"], case["P"], case["antialiasing"]
    grb.set_option("sort_mode", case["sort_mode"])
    views = [util.inputs_for(P, H, W, seed=case["seed"] + v, view=v, num_views=max(V, 2), scale_mult=case["scale_mult"])
             for v in range(V)]
    for x in views:
        x["bg"] = torch.linspace(-1, 1, 32)
    g = torch.Generator().manual_seed(case["seed"])
    dL = torch.randn(V, 32, H, W, generator=g).to(DEV)

    # reference: one call per view (gaussian_render.py:37-67), antialiasing as the case says
    fw, gref, gref2 = [], [], []
    for v, inp in enumerate(views):
        R, color, radii, geom, binning, img, invd = util.c_forward(ref, inp, DEV, aa)
        torch.cuda.synchronize()
        fw.append((R, color.cpu().numpy(), radii.cpu().numpy(), util.decode_reference(geom, binning, img, P, H, W, R)))
        leaf = {k: inp[k].to(DEV).clone().requires_grad_(True) for k in KEYS}
        m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
        rast = ref.GaussianRasterizer_32(raster_settings=util.settings_for(ref, inp, DEV, aa))
        c, _, _ = rast(means3D=leaf["xyz"], means2D=m2, shs=None, colors_precomp=leaf["features"], opacities=leaf["opacity"],
                       scales=leaf["scaling"], rotations=leaf["rotation"], cov3D_precomp=None)
        c.backward(dL[v], retain_graph=True)
        torch.cuda.synchronize()
        gref.append({k: t.grad.cpu().numpy().copy() for k, t in leaf.items()})
        for t in leaf.values():
            t.grad = None
        c.backward(dL[v])           # the reference once more: its own run-to-run deviation (see _gate)
        torch.cuda.synchronize()
        gref2.append({k: t.grad.cpu().numpy() for k, t in leaf.items()})

    stack = lambda k: torch.stack([x[k] for x in views], 0).to(DEV)
    tanfov = torch.tensor([[x["tanfovx"], x["tanfovy"]] for x in views], device=DEV)
    st = grb.BatchRasterizationSettings(H, W, tanfov, stack("bg"), 1.0, stack("viewmatrix"), stack("projmatrix"),
                                        antialiasing=aa)

    def run(summed_out):
        leaf = {k: stack(k).requires_grad_(True) for k in KEYS}
        m2 = torch.zeros(V, P, 3, device=DEV, requires_grad=True)
        color, radii, invd = grb.rasterize_gaussians_batch(leaf["xyz"], m2, leaf["features"], leaf["opacity"], leaf["scaling"],
                                                           leaf["rotation"], st, summed_out=summed_out)
        torch.cuda.synchronize()
        geom, binning, img = grb.rasterizer._tls.last_buffers
        o = util.decode_ours(geom, binning, img, V, P, W, H)
        o.update(R=o["num_rendered"], color=color.detach().cpu().numpy(), radii=radii.cpu().numpy())
        color.backward(dL)
        torch.cuda.synchronize()
        return o, leaf

    o, leaf = run(None)
    _compare_forward_per_view(o, fw, V, P, H, W)
    for v in range(V):
        for k in KEYS:
            _gate(f"{k}[{v}]", leaf[k].grad[v].cpu().numpy(), gref[v][k], gref2[v][k])

    packed = torch.full((P, 43), float("nan"), device=DEV)
    o2, leaf2 = run(packed)
    assert np.array_equal(o2["color"], o["color"]) and all(leaf2[k].grad is None for k in KEYS)
    order = ("xyz", "scaling", "rotation", "opacity", "features")
    want = np.concatenate([sum(gref[v][k] for v in range(V)) for k in order], 1)
    want2 = np.concatenate([sum(gref2[v][k] for v in range(V)) for k in order], 1)
    got = packed.cpu().numpy()
    assert np.isfinite(got).all()
    for name, sl in (("xyz", slice(0, 3)), ("scaling", slice(3, 6)), ("rotation", slice(6, 10)), ("opacity", slice(10, 11)),
                     ("features", slice(11, 43))):
        _gate("summed " + name, got[:, sl], want[:, sl], want2[:, sl])
