#!/usr/bin/env python
"""Work model of the compositors at different work-unit shapes (CPU, no GPU needed): for the benchmark's C3 input
(view 0), how many (pixel block, splat) pairs survive a footprint test when a warp owns an 8x4 block (one pixel per
lane, what ships) against an 8x8 block (two pixels per lane, the variant the round-1 verdict proposed)?  A splat's
footprint is the bounding box of its alpha >= 1/255 ellipse (the same over-estimate for both shapes).  Uses the C
oracle's preprocess for means2D / conic / opacity (tools/ may use oracle/: not a product path)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import importlib.util  # noqa: E402
spec = importlib.util.spec_from_file_location("_syn", os.path.join(ROOT, "guava_renderer_b200", "synthetic.py"))
syn = importlib.util.module_from_spec(spec); spec.loader.exec_module(syn)
import oracle  # noqa: E402


def main():
    P, H, W = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000, 512, 512
    d = syn.make_views(30, P, H, W, [0], 16)
    f = oracle.forward(d["xyz"][0].numpy(), d["features_color"][0].numpy(), d["opacity"][0].numpy(), d["scaling"][0].numpy(),
                       d["rotation"][0].numpy(), d["viewmatrix"][0].numpy(), d["projmatrix"][0].numpy(), d["tanfovx"],
                       d["tanfovy"], H, W, np.zeros(32, np.float32))
    vis = f["radii"] > 0
    m = f["means2D"][vis].astype(np.float64)
    co = f["conic_opacity"][vis].astype(np.float64)
    A, B, C, op = co[:, 0], co[:, 1], co[:, 2], co[:, 3]
    L = np.log(np.maximum(255.0 * op, 1.0))
    det = np.maximum(A * C - B * B, 1e-12)
    hx, hy = np.sqrt(2 * L * C / det), np.sqrt(2 * L * A / det)

    def blocks(bw, bh):
        x0 = np.clip(np.ceil((m[:, 0] - hx)), 0, W - 1); x1 = np.clip(np.floor(m[:, 0] + hx), 0, W - 1)
        y0 = np.clip(np.ceil((m[:, 1] - hy)), 0, H - 1); y1 = np.clip(np.floor(m[:, 1] + hy), 0, H - 1)
        ok = (x1 >= x0) & (y1 >= y0) & (L > 0)
        nx = (x1 // bw - x0 // bw + 1) * ok
        ny = (y1 // bh - y0 // bh + 1) * ok
        px = ((x1 - x0 + 1) * (y1 - y0 + 1) * ok).sum()
        return float((nx * ny).sum()), float(px)

    p84, px = blocks(8, 4)
    p88, _ = blocks(8, 8)
    p1616, _ = blocks(16, 16)
    out = {"P": P, "visible": int(vis.sum()), "pairs_8x4": p84, "pairs_8x8": p88, "pairs_16x16_tile": p1616,
           "bbox_pixels": px, "lane_use_8x4": px / (p84 * 32), "lane_use_8x8": px / (p88 * 64),
           "pairs_8x8_x2_over_pairs_8x4": 2 * p88 / p84}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
