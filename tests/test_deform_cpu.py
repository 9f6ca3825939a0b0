"""CPU checks of the deformation oracle (oracle/deform_oracle.py): the restated roma helpers satisfy the
identities their published definitions imply, and the restatement of ubody_gaussian.py:245-289 is
self-consistent.  (roma itself is not installed: see the oracle header.)"""
import numpy as np
import torch

from oracle import deform_oracle as do


def test_rotmat_to_unitquat_roundtrip_all_branches():
    g = torch.Generator().manual_seed(0)
    q = torch.nn.functional.normalize(torch.randn(4000, 4, generator=g, dtype=torch.float64), dim=-1)
    # force every decision branch: large x, y, z or w component
    for k in range(4):
        q[k * 500:(k + 1) * 500, k] += 3.0
    q = torch.nn.functional.normalize(q, dim=-1)
    R = do.unitquat_to_rotmat(q)
    q2 = do.rotmat_to_unitquat(R)
    assert torch.allclose(q2.norm(dim=-1), torch.ones(4000, dtype=torch.float64))
    sign = torch.sign((q * q2).sum(-1, keepdim=True))
    assert torch.allclose(q2 * sign, q, atol=1e-12)
    assert torch.allclose(do.unitquat_to_rotmat(q2), R, atol=1e-12)


def test_restated_roma_helpers_against_scipy():
    """roma is not installed, but the algorithm its `rotmat_to_unitquat` uses is SciPy's
    (`scipy.spatial.transform.Rotation.from_matrix`: decision matrix of the diagonal and the trace, XYZW, no sign
    canonicalisation) and SciPy is: the restatement must reproduce it component for component, sign included, in every
    decision branch; the quaternion product against SciPy's composition of rotations (up to the sign of q)."""
    from scipy.spatial.transform import Rotation as Rot
    g = torch.Generator().manual_seed(4)
    q = torch.nn.functional.normalize(torch.randn(4000, 4, generator=g, dtype=torch.float64), dim=-1)
    for k in range(4):
        q[k * 500:(k + 1) * 500, k] += 3.0
    q = torch.nn.functional.normalize(q, dim=-1)
    R = do.unitquat_to_rotmat(q)
    mine = do.rotmat_to_unitquat(R).numpy()
    theirs = Rot.from_matrix(R.numpy()).as_quat()          # XYZW
    assert np.abs(mine - theirs).max() <= 1e-15
    assert np.abs(do.unitquat_to_rotmat(q).numpy() - Rot.from_quat(q.numpy()).as_matrix()).max() <= 1e-14
    p = torch.nn.functional.normalize(torch.randn(4000, 4, generator=g, dtype=torch.float64), dim=-1)
    prod = do.quat_product(p, q).numpy()
    comp = (Rot.from_quat(p.numpy()) * Rot.from_quat(q.numpy())).as_quat()
    sign = np.sign((prod * comp).sum(-1, keepdims=True))
    assert np.abs(prod * sign - comp).max() <= 1e-14


def test_quat_product_is_rotation_composition():
    g = torch.Generator().manual_seed(1)
    p = torch.nn.functional.normalize(torch.randn(100, 4, generator=g, dtype=torch.float64), dim=-1)
    q = torch.nn.functional.normalize(torch.randn(100, 4, generator=g, dtype=torch.float64), dim=-1)
    R = do.unitquat_to_rotmat(do.quat_product(p, q))
    assert torch.allclose(R, do.unitquat_to_rotmat(p) @ do.unitquat_to_rotmat(q), atol=1e-12)
    w = do.quat_xyzw_to_wxyz(p)
    assert torch.equal(do.quat_wxyz_to_xyzw(w), p) and torch.equal(w[:, 0], p[:, 3])


def test_face_frames_are_orthonormal_and_deform_is_rigid_per_face():
    a = do.synthetic_avatar(2, 300, 500, 2000, seed=3)
    R, s = do.compute_face_orientation(a["vertices"], a["faces"])
    eye = torch.eye(3).expand_as(R)
    assert torch.allclose(R.transpose(-1, -2) @ R, eye, atol=1e-4)
    assert (s > 0).all() and s.shape == (2, 500, 1)
    out = do.deform(**a)
    P = 300 + 2000
    assert out["xyz"].shape == (2, P, 3) and out["rotation"].shape == (2, P, 4)
    assert out["features_color"].shape == (2, P, 32) and out["opacity"].shape == (2, P, 1)
    # zero local offset puts a UV Gaussian at the barycentric point of its face
    a0 = dict(a, uv_local_xyz=torch.zeros_like(a["uv_local_xyz"]))
    c = do.deform(**a0)["xyz"][:, 300:]
    fv = a["vertices"][:, a["faces"].long()][:, a["uv_binding_face"].long()]
    assert torch.allclose(c, (a["uv_face_bary"][None, :, :, None] * fv).sum(2), atol=1e-6)
    # unit input quaternions stay unit
    assert torch.allclose(out["rotation"].norm(dim=-1), torch.ones(2, P), atol=1e-5)


def _deform_golden():
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "deform", "deform_ref.npz"))
    a = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("in_")}
    return z, a


def test_oracle_against_golden_from_the_reference_source():
    """tests/golden/deform/deform_ref.npz holds what the REFERENCE'S OWN SOURCE produces (utils/graphics_utils.py:52-80 and
    the body of Ubody_Gaussian.forward, ubody_gaussian.py:245-289, extracted with ast and executed by
    tests/golden/make_deform_golden.py; roma's four helpers bound to the restatement).  The oracle's line-by-line
    restatement must reproduce it."""
    z, a = _deform_golden()
    out = do.deform(**a)
    for k in ("xyz", "rotation", "scaling", "opacity", "features_color"):
        want = torch.from_numpy(z["out_" + k])
        assert out[k].shape == want.shape
        assert float((out[k] - want).abs().max()) <= 1e-6 * max(1.0, float(want.abs().max())), k
    R, s = do.compute_face_orientation(a["vertices"], a["faces"])
    assert float((R - torch.from_numpy(z["face_orientation"])).abs().max()) <= 1e-6
    assert float((s - torch.from_numpy(z["face_scale"])).abs().max()) <= 1e-6
