"""Binning time of both sort modes (tile buckets vs onesweep) at the BASELINE shapes; run on the GPU box."""
import json, sys, torch
sys.path.insert(0, '.')
from tests import util
import guava_renderer_b200 as grb
DEV = 'cuda:0'
for name, (P, H, W, V, sm) in dict(C3=(100000, 512, 512, 2, 1.0), C3b=(202000, 512, 512, 2, 1.0), C4=(300000, 1024, 1024, 2, 1.0), C5=(1000000, 1080, 1920, 1, 4.0)).items():
    views = [util.inputs_for(P, H, W, 4000 + v, view=v, num_views=4, scale_mult=sm) for v in range(V)]
    st = grb.BatchRasterizationSettings(H, W, torch.tensor([[x["tanfovx"], x["tanfovy"]] for x in views], device=DEV),
                                        torch.zeros(V, 32, device=DEV), 1.0, torch.stack([x["viewmatrix"] for x in views]).to(DEV),
                                        torch.stack([x["projmatrix"] for x in views]).to(DEV))
    stack = lambda k: torch.stack([x[k] for x in views]).to(DEV)
    args = (stack("xyz"), torch.zeros(V, P, 3, device=DEV), stack("features"), stack("opacity"), stack("scaling"), stack("rotation"), st)
    row = dict(config=name, P=P, H=H, W=W, views=V)
    for mode in (0, 1):
        grb.set_option("sort_mode", mode)
        with torch.no_grad():
            for _ in range(3): grb.rasterize_gaussians_batch(*args)
            torch.cuda.synchronize()
            grb.set_option("profile_stages", 1); grb._lib_mod.stage_times_reset()
            for _ in range(10): grb.rasterize_gaussians_batch(*args)
            torch.cuda.synchronize(); grb.set_option("profile_stages", 0)
        t = {k: v[0] / max(v[1], 1) for k, v in grb._lib_mod.stage_times().items()}
        row["buckets_ms" if mode == 0 else "onesweep_ms"] = round(t["scatter"] + t["tile_sort"], 4) if mode == 0 else round(t["tile_sort"], 4)
        row[f"forward_total_ms_mode{mode}"] = round(sum(t[k] for k in ("preprocess", "tile_scan", "scatter", "tile_sort", "render_forward")), 4)
    grb.set_option("sort_mode", -1)
    print(json.dumps(row))
