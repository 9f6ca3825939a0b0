"""GPU parity of the fused deformation (csrc/deform.cu, through the C ABI) against the CPU restatement of
ubody_gaussian.py:245-289 (oracle/deform_oracle.py) on the same seeded inputs: forward within 1e-5 of the
output scale, gradients within 1e-4 relative (fp32; the tolerances north_star states for the rasterizer)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

KEYS = ("xyz", "rotation", "scaling", "opacity", "features_color")
LEAVES = ("smplx_rotation", "smplx_scaling", "smplx_opacity", "smplx_features_color",
          "uv_local_xyz", "uv_rotation", "uv_scaling", "uv_opacity", "uv_features_color")


def _run(a, dev, weights):
    import guava_renderer_b200 as grb
    from oracle import deform_oracle as do
    mesh = ("vertices", "ver_transform_mat", "faces", "uv_binding_face", "uv_face_bary")
    cpu = {k: (v.clone().requires_grad_(True) if k in LEAVES else v) for k, v in a.items()}
    ref = do.deform(**cpu)
    sum((ref[k] * weights[k]).sum() for k in KEYS).backward()
    gpu = {k: (v.to(dev).requires_grad_(True) if k in LEAVES else v.to(dev)) for k, v in a.items()}
    n0 = grb.launch_count()
    out = grb.deform_gaussians(*[gpu[k] for k in mesh], **{k: gpu[k] for k in LEAVES})
    sum((out[k] * weights[k].to(dev)).sum() for k in KEYS).backward()
    torch.cuda.synchronize()
    assert grb.launch_count() - n0 >= 3
    return ref, cpu, out, gpu


@pytest.mark.parametrize("shape", [(2, 300, 500, 2000), (1, 10475, 20908, 60000), (3, 7, 5, 1), (1, 64, 10, 0)],
                         ids=["small", "smplx_mesh", "tiny", "no_uv"])
def test_deform_matches_reference_restatement(shape):
    from oracle import deform_oracle as do
    B, Nv, F, N = shape
    a = do.synthetic_avatar(B, Nv, F, N, seed=11)
    g = torch.Generator().manual_seed(5)
    P = Nv + N
    weights = {k: torch.randn(B, P, w, generator=g) for k, w in zip(KEYS, (3, 4, 3, 1, 32))}
    ref, cpu, out, gpu = _run(a, "cuda:0", weights)
    for k in KEYS:
        o, r = out[k].detach().cpu(), ref[k].detach()
        if k == "rotation":   # q and -q are the same rotation; a decision-branch tie may flip the sign
            flip = torch.sign((o * r).sum(-1, keepdim=True))
            assert (flip < 0).float().mean() < 1e-3
            o = o * flip
        assert float((o - r).abs().max()) <= 1e-5 * max(1.0, float(r.abs().max())), k
    assert out["smplx_xyz_deform"] is gpu["vertices"] or torch.equal(out["smplx_xyz_deform"], gpu["vertices"])
    for k in LEAVES:
        gg, gr = gpu[k].grad.cpu(), cpu[k].grad
        if gr.numel() == 0:
            continue
        bad = ((gg - gr).abs() > 1e-4 * float(gr.abs().max()) + 1e-7).float().mean()
        assert bad < 1e-3, (k, float(bad), float((gg - gr).abs().max()))


def test_deform_feeds_the_rasterizer_and_rejects_host_tensors():
    import guava_renderer_b200 as grb
    from oracle import deform_oracle as do
    a = do.synthetic_avatar(1, 200, 300, 3000, seed=2)
    mesh = ("vertices", "ver_transform_mat", "faces", "uv_binding_face", "uv_face_bary")
    with pytest.raises(grb.GuavaLibraryError):
        grb.deform_gaussians(*[a[k] for k in mesh], **{k: a[k] for k in LEAVES})
    dev = "cuda:0"
    gpu = {k: v.to(dev) for k, v in a.items()}
    gpu["uv_local_xyz"].requires_grad_(True)
    assets = grb.deform_gaussians(*[gpu[k] for k in mesh], **{k: gpu[k] for k in LEAVES})
    from tests import util
    cam = util.inputs_for(10, 64, 64, seed=1)
    cams = {"world_view_transform": cam["viewmatrix"][None].to(dev), "full_proj_transform": cam["projmatrix"][None].to(dev),
            "tanfovx": cam["tanfovx"], "tanfovy": cam["tanfovy"], "image_height": 64, "image_width": 64}
    out = grb.render_views(assets, cams)
    out["rendered_images"].sum().backward()
    assert torch.isfinite(out["rendered_images"]).all() and gpu["uv_local_xyz"].grad is not None


def test_deform_kernels_against_golden_from_the_reference_source():
    """The CUDA path against tests/golden/deform/deform_ref.npz -- outputs of the reference's own source for this row
    (see tests/golden/make_deform_golden.py; roma's helpers restated)."""
    import os
    import guava_renderer_b200 as grb
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "deform", "deform_ref.npz"))
    a = {k[3:]: torch.from_numpy(z[k]).to("cuda:0") for k in z.files if k.startswith("in_")}
    mesh = ("vertices", "ver_transform_mat", "faces", "uv_binding_face", "uv_face_bary")
    out = grb.deform_gaussians(*[a[k] for k in mesh], **{k: a[k] for k in LEAVES})
    torch.cuda.synchronize()
    for k in KEYS:
        o, r = out[k].detach().cpu(), torch.from_numpy(z["out_" + k])
        if k == "rotation":   # q and -q are the same rotation
            o = o * torch.sign((o * r).sum(-1, keepdim=True))
        assert float((o - r).abs().max()) <= 1e-5 * max(1.0, float(r.abs().max())), k


def test_graphed_step_with_a_deformation_producer_equals_the_eager_chain():
    """GraphedRenderStep(producer=...): deform_gaussians -> rasterizer forward -> loss -> rasterizer backward ->
    deformation backward captured as one graph and replayed == the same chain run eagerly (image bit for bit,
    gradients on the avatar attributes to atomic-order tolerance); a second replay after the posed mesh changed
    in place follows the new pose."""
    import guava_renderer_b200 as grb
    from guava_renderer_b200 import synthetic
    from tests import util
    dev = "cuda:0"
    V, P, H, W = 2, 14000, 96, 128
    av = synthetic.pose_avatar(P, V, seed=9, rows=40, cols=60)
    Nv = av["vertices"].shape[0]
    mesh = ("faces", "uv_binding_face", "uv_face_bary")
    g = {k: av[k].to(dev) for k in mesh + LEAVES}
    cams = [synthetic.camera(v, 16, 1.0 / 24 * W / H, 1.0 / 24) for v in range(V)]
    tanfov = torch.tensor([[c["tanfovx"], c["tanfovy"]] for c in cams], device=dev)
    st = grb.BatchRasterizationSettings(H, W, tanfov, torch.zeros(V, 32, device=dev), 1.0,
                                        torch.stack([c["viewmatrix"] for c in cams]).to(dev),
                                        torch.stack([c["projmatrix"] for c in cams]).to(dev))
    gen = torch.Generator().manual_seed(3)
    dL = torch.randn(V, 32, H, W, generator=gen).to(dev)
    pv, pt = synthetic.pose_of(av["vertices"], 0, V)
    verts, xf = pv.to(dev), pt.to(dev)

    class Producer:
        def __call__(self):
            self.leaves = {k: g[k].detach().requires_grad_(torch.is_grad_enabled()) for k in LEAVES}
            o = grb.deform_gaussians(verts, xf, g["faces"], g["uv_binding_face"], g["uv_face_bary"], **self.leaves)
            return dict(xyz=o["xyz"], features=o["features_color"], opacity=o["opacity"], scaling=o["scaling"],
                        rotation=o["rotation"])

    def eager():
        p = Producer()
        d = p()
        m2 = torch.zeros(V, P, 3, device=dev, requires_grad=True)
        color, radii, invd = grb.rasterize_gaussians_batch(d["xyz"], m2, d["features"], d["opacity"], d["scaling"], d["rotation"], st)
        color.backward(dL)
        torch.cuda.synchronize()
        return color.detach().clone(), {k: v.grad.clone() for k, v in p.leaves.items()}

    prod = Producer()
    step = grb.GraphedRenderStep(None, None, None, None, None, st, dL_dcolor=dL, producer=prod)
    for pose in (0, 5):
        pv, pt = synthetic.pose_of(av["vertices"], pose, V)
        verts.copy_(pv.to(dev)); xf.copy_(pt.to(dev))
        step.replay()
        step.check()
        want_color, want = eager()
        assert torch.equal(step.color.detach(), want_color)
        assert float(want_color.abs().sum()) > 0
        for k in LEAVES:
            got = prod.leaves[k].grad
            assert got is not None, k
            assert util.rel_err(got.cpu().numpy(), want[k].cpu().numpy()) <= 5e-5, k
        assert step.packed.shape == (P, 43) and float(step.packed.abs().sum()) > 0
