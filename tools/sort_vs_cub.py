#!/usr/bin/env python
"""Binning of this library against the reference's CUB calls on the same keys ("for sort and scan
the bar is CUB 2.8.2's sm_100 onesweep / decoupled-look-back scan on the same box").

For each shape: run our forward in both sort modes with per-stage CUDA-event timing, rebuild the reference's
UNSORTED (key, value) list in its emission order (rasterizer_impl.cu:70-111: Gaussian index ascending, rect walked
row-major) from our rects and depths, and time `cub::DeviceRadixSort::SortPairs(..., 0, 32 + bit)`
(rasterizer_impl.cu:303-311) on it with tools/ubench/sort_vs_cub.cu.  CUB's result must equal our sorted keys and
point_list bit for bit.  One JSON line per shape.

    python tools/sort_vs_cub.py --build-only      # nvcc cross-compile (no GPU needed)
    python tools/sort_vs_cub.py > gpurun_out/sort_vs_cub.jsonl
"""
import ctypes
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SRC = os.path.join(ROOT, "tools", "ubench", "sort_vs_cub.cu")
SO = os.path.join(ROOT, "tools", "ubench", "_build", "libsort_vs_cub.so")


def build():
    os.makedirs(os.path.dirname(SO), exist_ok=True)
    if not os.path.isfile(SO) or os.path.getmtime(SO) < os.path.getmtime(SRC):
        subprocess.check_call(["nvcc", "-shared", "-Xcompiler", "-fPIC", "-O3", "-std=c++17", "-gencode",
                               "arch=compute_100a,code=sm_100a", "-lineinfo", SRC, "-o", SO])
    return SO


def main():
    build()
    if "--build-only" in sys.argv:
        print(SO)
        return
    import numpy as np
    import torch
    import guava_renderer_b200 as grb
    from guava_renderer_b200 import synthetic
    from tests import util
    lib = ctypes.CDLL(SO)
    lib.cub_sort_pairs_ms.restype = ctypes.c_float
    lib.cub_sort_pairs_ms.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int] * 4 + [ctypes.c_void_p]
    lib.cub_inclusive_sum_ms.restype = ctypes.c_float
    lib.cub_inclusive_sum_ms.argtypes = [ctypes.c_void_p] * 2 + [ctypes.c_int] * 3 + [ctypes.c_void_p]
    dev = "cuda:0"
    shapes = [("C3 100k x2 views", 30, 100_000, 512, 512, 2, 16, 1.0), ("C3 202k x2 views", 30, 202_000, 512, 512, 2, 16, 1.0),
              ("C4 300k 1024^2", 40, 300_000, 1024, 1024, 1, 1, 1.0), ("C5 1M 1080p", 50, 1_000_000, 1080, 1920, 1, 1, 4.0)]
    grb.set_option("emit_sorted_keys", 1)
    for name, cfg, P, H, W, V, nv, sm in shapes:
        d = synthetic.make_views(cfg, P, H, W, list(range(V)), nv, scale_mult=sm)
        tanfov = torch.tensor([[d["tanfovx"], d["tanfovy"]]] * V, device=dev)
        st = grb.BatchRasterizationSettings(H, W, tanfov, torch.zeros(V, 32, device=dev), 1.0,
                                            d["viewmatrix"].to(dev), d["projmatrix"].to(dev))
        args = [d[k].to(dev) for k in ("xyz",)] + [torch.zeros(V, P, 3, device=dev)] + \
               [d[k].to(dev) for k in ("features_color", "opacity", "scaling", "rotation")] + [st]
        row = {"shape": name, "P": P, "H": H, "W": W, "views": V}
        dec = None
        for mode in (0, 1):
            grb.set_option("sort_mode", mode)
            with torch.no_grad():
                for _ in range(3):
                    grb.rasterize_gaussians_batch(*args)
                torch.cuda.synchronize()
                grb.set_option("profile_stages", 1)
                grb._lib_mod.stage_times_reset()
                for _ in range(10):
                    grb.rasterize_gaussians_batch(*args)
                torch.cuda.synchronize()
                grb.set_option("profile_stages", 0)
            t = {k: v[0] / max(v[1], 1) for k, v in grb._lib_mod.stage_times().items()}
            if mode == 0:
                row["ours_buckets_scatter_ms"] = round(t["scatter"], 4)
                row["ours_buckets_tile_sort_ms"] = round(t["tile_sort"], 4)
            else:
                row["ours_onesweep_chain_ms"] = round(t["tile_sort"], 4)   # emission (+scan, +histograms) + passes
            row["tile_scan_ms"] = round(t["tile_scan"], 4)
            geom, binning, img = grb.rasterizer._tls.last_buffers
            dec = util.decode_ours(geom, binning, img, V, P, W, H)
        R = dec["num_rendered"]
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        bit = int(V * tiles).bit_length()            # getHigherMsb, rasterizer_impl.cu:35-50 -- over all views of the batch
        # the reference's unsorted list in its emission order
        rects = torch.from_numpy(dec["rects"].astype(np.int64)).to(dev)
        depth_bits = torch.from_numpy(dec["depths"].view(np.uint32).astype(np.int64)).to(dev)
        w = rects[:, 2] - rects[:, 0]
        cnt = (w * (rects[:, 3] - rects[:, 1])).clamp_min(0)
        n = V * P
        owner = torch.repeat_interleave(torch.arange(n, device=dev), cnt)
        excl = torch.cumsum(cnt, 0) - cnt
        local = torch.arange(int(cnt.sum()), device=dev) - excl[owner]
        yy, xx = local // w[owner], local % w[owner]
        gx = (W + 15) // 16
        tile = (owner // P) * tiles + (rects[owner, 1] + yy) * gx + rects[owner, 0] + xx
        keys_in = ((tile << 32) | depth_bits[owner]).contiguous()             # int64 bit pattern == uint64 key
        vals_in = (owner % P).to(torch.int32).contiguous()
        assert keys_in.numel() == R
        keys_out, vals_out = torch.empty_like(keys_in), torch.empty_like(vals_in)
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        ms = lib.cub_sort_pairs_ms(keys_in.data_ptr(), keys_out.data_ptr(), vals_in.data_ptr(), vals_out.data_ptr(), R,
                                   32 + bit, 3, 20, stream)
        tt = torch.from_numpy(dec["tiles_touched"].astype(np.int32)).to(dev)
        ms_scan = lib.cub_inclusive_sum_ms(tt.data_ptr(), torch.empty_like(tt).data_ptr(), n, 3, 20, stream)
        torch.cuda.synchronize()
        ko = keys_out.cpu().numpy().view(np.uint64)
        ours_k = dec["keys"].copy()
        if dec["sort_mode"] == 0:       # bucket path: the key's high word is the tile within its view
            pass
        same_vals = bool(np.array_equal(vals_out.cpu().numpy().view(np.uint32), dec["point_list"]))
        same_depth = bool(np.array_equal(ko & np.uint64(0xFFFFFFFF), ours_k & np.uint64(0xFFFFFFFF)))
        same_tile = bool(np.array_equal(ko >> np.uint64(32), ours_k >> np.uint64(32)))
        rg = dec["ranges"].astype(np.int64)
        row.update(R=R, longest_tile_list=int((rg[:, 1] - rg[:, 0]).max()), sorted_bits=32 + bit, cub_sort_pairs_ms=round(ms, 4), cub_inclusive_sum_ms=round(ms_scan, 4),
                   cub_result_equals_ours={"values": same_vals, "depth_bits": same_depth, "tile": same_tile},
                   hbm_ms_of_one_pass_24B_per_pair=round(24 * R / 6548.8e9 * 1e3, 4))
        print(json.dumps(row), flush=True)
        del keys_in, vals_in, keys_out, vals_out, owner, local
    grb.set_option("sort_mode", -1)
    grb.set_option("emit_sorted_keys", 0)


if __name__ == "__main__":
    main()
