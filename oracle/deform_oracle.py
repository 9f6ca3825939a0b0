"""CPU restatement of GUAVA's deformation step -- TEST INFRASTRUCTURE ONLY (imported by tests/ and
tools/ timing scripts; the product path is csrc/deform.cu and never touches this file).

Follows models/UbodyAvatar/ubody_gaussian.py:245-289 line by line with plain torch ops, so that
`torch.autograd` provides the reference gradients.  `compute_face_orientation`, `safe_normalize`,
`length`, `dot` follow utils/graphics_utils.py:52-80.  The quaternion helpers restate roma 1.5.3
(requirements.txt:6; not vendored in /root/reference, not installed here => "parity unpinned" against roma itself;
pinned against SciPy instead, whose Rotation.from_matrix is the algorithm roma's rotmat_to_unitquat adapts -- component
for component, sign included, in every decision branch -- and against SciPy's composition of rotations for the
product, plus the algebraic checks, tests/test_deform_cpu.py):
  rotmat_to_unitquat  SciPy's Rotation.from_matrix decision (argmax over diag and trace), XYZW
  quat_product        Hamilton product, XYZW
  quat_xyzw_to_wxyz / quat_wxyz_to_xyzw
"""
import torch


def dot(x, y):
    return torch.sum(x * y, -1, keepdim=True)


def length(x, eps=1e-20):
    return torch.sqrt(torch.clamp(dot(x, x), min=eps))


def safe_normalize(x, eps=1e-20):
    return x / length(x, eps)


def compute_face_orientation(verts, faces):
    """utils/graphics_utils.py:61-80 with return_scale=True."""
    i0, i1, i2 = faces[..., 0].long(), faces[..., 1].long(), faces[..., 2].long()
    v0, v1, v2 = verts[..., i0, :], verts[..., i1, :], verts[..., i2, :]
    a0 = safe_normalize(v1 - v0)
    a1 = safe_normalize(torch.cross(a0, v2 - v0, dim=-1))
    a2 = -safe_normalize(torch.cross(a1, a0, dim=-1))
    orientation = torch.cat([a0[..., None], a1[..., None], a2[..., None]], dim=-1)
    s0 = length(v1 - v0)
    s1 = dot(a2, (v2 - v0)).abs()
    return orientation, (s0 + s1) / 2


def rotmat_to_unitquat(R):
    """roma.rotmat_to_unitquat (XYZW)."""
    shape = R.shape[:-2]
    m = R.reshape(-1, 3, 3)
    n = m.shape[0]
    dec = torch.empty((n, 4), dtype=m.dtype, device=m.device)
    dec[:, :3] = m.diagonal(dim1=1, dim2=2)
    dec[:, 3] = dec[:, :3].sum(dim=1)
    choices = dec.argmax(dim=1)
    quat = torch.empty((n, 4), dtype=m.dtype, device=m.device)
    ind = torch.nonzero(choices != 3, as_tuple=True)[0]
    i = choices[ind]
    j = (i + 1) % 3
    k = (j + 1) % 3
    quat[ind, i] = 1 - dec[ind, 3] + 2 * m[ind, i, i]
    quat[ind, j] = m[ind, j, i] + m[ind, i, j]
    quat[ind, k] = m[ind, k, i] + m[ind, i, k]
    quat[ind, 3] = m[ind, k, j] - m[ind, j, k]
    ind = torch.nonzero(choices == 3, as_tuple=True)[0]
    quat[ind, 0] = m[ind, 2, 1] - m[ind, 1, 2]
    quat[ind, 1] = m[ind, 0, 2] - m[ind, 2, 0]
    quat[ind, 2] = m[ind, 1, 0] - m[ind, 0, 1]
    quat[ind, 3] = 1 + dec[ind, 3]
    quat = quat / torch.norm(quat, dim=1)[:, None]
    return quat.reshape(*shape, 4)


def quat_product(p, q):
    """roma.quat_product (XYZW)."""
    vector = p[..., None, 3] * q[..., :3] + q[..., None, 3] * p[..., :3] + torch.cross(p[..., :3], q[..., :3], dim=-1)
    last = p[..., 3] * q[..., 3] - torch.sum(p[..., :3] * q[..., :3], dim=-1)
    return torch.cat((vector, last[..., None]), dim=-1)


def quat_xyzw_to_wxyz(q):
    return torch.cat((q[..., 3:4], q[..., 0:3]), dim=-1)


def quat_wxyz_to_xyzw(q):
    return torch.cat((q[..., 1:4], q[..., 0:1]), dim=-1)


def unitquat_to_rotmat(q):
    """XYZW -> rotation matrix (used only by the algebraic checks)."""
    x, y, z, w = q.unbind(-1)
    return torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
        2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
        2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1).reshape(*q.shape[:-1], 3, 3)


def deform(vertices, ver_transform_mat, faces, uv_binding_face, uv_face_bary, smplx_rotation, smplx_scaling,
           smplx_opacity, smplx_features_color, uv_local_xyz, uv_rotation, uv_scaling, uv_opacity, uv_features_color):
    """ubody_gaussian.py:245-289."""
    batch_size = vertices.shape[0]
    binding = uv_binding_face.long()
    d_deform_rot_xyzw = rotmat_to_unitquat(ver_transform_mat[:, :, :3, :3])                                   # :253
    smplx_rotation_deform = torch.nn.functional.normalize(
        quat_xyzw_to_wxyz(quat_product(d_deform_rot_xyzw, quat_wxyz_to_xyzw(smplx_rotation))), dim=-1)       # :254
    face_orien_mat, face_scaling = compute_face_orientation(vertices, faces)                                 # :257
    face_orien_quat = quat_xyzw_to_wxyz(rotmat_to_unitquat(face_orien_mat))                                   # :258
    face_vertices = vertices[:, faces.long()]                                                                 # :260
    face_vertices_nn = face_vertices[:, binding]                                                              # :261
    face_bary = uv_face_bary[None].expand(batch_size, -1, -1)                                                 # :262
    face_center_nn = torch.einsum('bnk,bnkj->bnj', face_bary, face_vertices_nn)                               # :263
    face_scaling_nn = face_scaling[:, binding]                                                                # :264
    xyz = torch.einsum('bnij,bnj->bni', face_orien_mat[:, binding], uv_local_xyz)                             # :266
    uv_xyz_deform = xyz * face_scaling_nn + face_center_nn                                                    # :267
    face_orien_quat = face_orien_quat[:, binding]                                                             # :269
    uv_rotation_deform = quat_xyzw_to_wxyz(quat_product(quat_wxyz_to_xyzw(face_orien_quat),
                                                        quat_wxyz_to_xyzw(uv_rotation)))                     # :270
    uv_scaling_deform = uv_scaling * face_scaling_nn                                                          # :271
    return {                                                                                                  # :274-288
        "xyz": torch.cat([vertices, uv_xyz_deform], dim=1),
        "rotation": torch.cat([smplx_rotation_deform, uv_rotation_deform], dim=1),
        "scaling": torch.cat([smplx_scaling, uv_scaling_deform], dim=1),
        "opacity": torch.cat([smplx_opacity, uv_opacity], dim=1),
        "features_color": torch.cat([smplx_features_color, uv_features_color], dim=1),
    }


def synthetic_avatar(B, Nv, F, N, seed):
    """Seeded stand-in for an SMPLX mesh + GUAVA attributes of the given sizes (GUAVA: Nv=10475,
    F=20908, N~190k): random non-degenerate triangles over a point cloud, random rigid vertex transforms."""
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    vertices = r(B, Nv, 3) * 0.3
    faces = torch.stack([torch.randperm(Nv, generator=g)[:3] for _ in range(min(F, 64))])
    if F > 64:   # cheap distinct-index triples for the rest
        a = torch.randint(0, Nv, (F - 64,), generator=g)
        faces = torch.cat([faces, torch.stack([a, (a + 1 + torch.randint(0, Nv - 2, (F - 64,), generator=g) % (Nv - 2)) % Nv,
                                               (a + Nv - 1) % Nv], 1)])
        faces[64:, 1] = torch.where(faces[64:, 1] == faces[64:, 2], (faces[64:, 1] + 1) % Nv, faces[64:, 1])
        faces[64:, 1] = torch.where(faces[64:, 1] == faces[64:, 0], (faces[64:, 1] + 2) % Nv, faces[64:, 1])
    q = torch.nn.functional.normalize(r(B, Nv, 4), dim=-1)
    T = torch.zeros(B, Nv, 4, 4)
    T[..., :3, :3] = unitquat_to_rotmat(q)
    T[..., :3, 3] = r(B, Nv, 3) * 0.1
    T[..., 3, 3] = 1
    bary = torch.rand(N, 3, generator=g)
    bary = bary / bary.sum(-1, keepdim=True)
    return dict(
        vertices=vertices, ver_transform_mat=T, faces=faces.int(), uv_binding_face=torch.randint(0, F, (N,), generator=g).int(),
        uv_face_bary=bary,
        smplx_rotation=torch.nn.functional.normalize(r(B, Nv, 4), dim=-1), smplx_scaling=torch.exp(r(B, Nv, 3) * 0.3) * 0.01,
        smplx_opacity=torch.sigmoid(r(B, Nv, 1)), smplx_features_color=r(B, Nv, 32),
        uv_local_xyz=r(B, N, 3) * 0.2, uv_rotation=torch.nn.functional.normalize(r(B, N, 4), dim=-1),
        uv_scaling=torch.exp(r(B, N, 3) * 0.3) * 0.01, uv_opacity=torch.sigmoid(r(B, N, 1)), uv_features_color=r(B, N, 32))
