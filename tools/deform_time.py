"""Fused deformation (csrc/deform.cu) vs the torch restatement of ubody_gaussian.py:245-289 run eagerly on the
same GPU, at GUAVA's sizes (SMPLX mesh Nv=10475, F=20908; ~190k UV Gaussians; batch 2).  Run on the GPU box."""
import json, sys, torch
sys.path.insert(0, '.')
import guava_renderer_b200 as grb
from oracle import deform_oracle as do
dev = 'cuda:0'
B, Nv, F, N = 2, 10475, 20908, 190000
a = {k: v.to(dev) for k, v in do.synthetic_avatar(B, Nv, F, N, seed=1).items()}
mesh = ("vertices", "ver_transform_mat", "faces", "uv_binding_face", "uv_face_bary")
leaves = [k for k in a if k not in mesh]
def timed(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n
def rot_dev(R):  # the oracle's rotmat_to_unitquat allocates on the CPU; device version of the same ops for timing
    m = R.reshape(-1, 3, 3); dec = torch.cat([m.diagonal(dim1=1, dim2=2), m.diagonal(dim1=1, dim2=2).sum(1, keepdim=True)], 1)
    c = dec.argmax(1); quat = torch.empty(m.shape[0], 4, device=R.device)
    ind = torch.nonzero(c != 3, as_tuple=True)[0]; i = c[ind]; j = (i + 1) % 3; k = (j + 1) % 3
    quat[ind, i] = 1 - dec[ind, 3] + 2 * m[ind, i, i]; quat[ind, j] = m[ind, j, i] + m[ind, i, j]
    quat[ind, k] = m[ind, k, i] + m[ind, i, k]; quat[ind, 3] = m[ind, k, j] - m[ind, j, k]
    ind = torch.nonzero(c == 3, as_tuple=True)[0]
    quat[ind, 0] = m[ind, 2, 1] - m[ind, 1, 2]; quat[ind, 1] = m[ind, 0, 2] - m[ind, 2, 0]
    quat[ind, 2] = m[ind, 1, 0] - m[ind, 0, 1]; quat[ind, 3] = 1 + dec[ind, 3]
    return (quat / quat.norm(dim=1)[:, None]).reshape(*R.shape[:-2], 4)
do.rotmat_to_unitquat = rot_dev
row = dict(B=B, Nv=Nv, F=F, N=N)
with torch.no_grad():
    row["fused_fwd_ms"] = round(timed(lambda: grb.deform_gaussians(*[a[k] for k in mesh], **{k: a[k] for k in leaves})), 4)
    row["torch_eager_fwd_ms"] = round(timed(lambda: do.deform(**a)), 4)
g = {k: (v.clone().requires_grad_(True) if k in leaves else v) for k, v in a.items()}
def fb(fn):
    out = fn()
    (out["xyz"].sum() + out["rotation"].sum() + out["scaling"].sum()).backward()
row["fused_fwd_bwd_ms"] = round(timed(lambda: fb(lambda: grb.deform_gaussians(*[g[k] for k in mesh], **{k: g[k] for k in leaves}))), 4)
row["torch_eager_fwd_bwd_ms"] = round(timed(lambda: fb(lambda: do.deform(**g))), 4)
P = Nv + N
row["algorithmic_MB_fwd"] = round(B * (P * (40 + 40 + 256) + F * 64 * 2 + Nv * 64) / 1e6, 1)
row["fused_fwd_GBps"] = round(row["algorithmic_MB_fwd"] / row["fused_fwd_ms"], 1)
print(json.dumps(row))
