"""Host-side cost of the drop-in per-view call (forward + backward) -- cProfile on the GPU box."""
import cProfile, pstats, sys, time, io
import torch
sys.path.insert(0, '.')
from tests import util
import diff_gaussian_rasterization_32 as ours
DEV = 'cuda:0'
P, H, W = 10000, 512, 512
inp = util.inputs_for(P, H, W, 1000, view=0, num_views=2)
d = {k: (t.to(DEV) if torch.is_tensor(t) else t) for k, t in inp.items()}
st = util.settings_for(ours, inp, DEV, False)
dL = torch.randn(32, H, W, device=DEV)


def step():
    leaf = {k: d[k].detach().requires_grad_(True) for k in ("xyz", "features", "opacity", "scaling", "rotation")}
    m2 = torch.zeros_like(leaf["xyz"], requires_grad=True)
    rast = ours.GaussianRasterizer_32(raster_settings=st)
    color, radii, invd = rast(means3D=leaf["xyz"], means2D=m2, shs=None, colors_precomp=leaf["features"],
                              opacities=leaf["opacity"], scales=leaf["scaling"], rotations=leaf["rotation"], cov3D_precomp=None)
    color.backward(dL)


for _ in range(20):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(200):
    step()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host enqueue per step {1e3*(t1-t0)/200:.4f} ms; incl. drain {1e3*(t2-t0)/200:.4f} ms")
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    step()
pr.disable()
torch.cuda.synchronize()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(28)
print(s.getvalue()[:6000])
