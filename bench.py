#!/usr/bin/env python
"""Benchmark of the rasterizer hot path.  Contract: see the task statement; one JSON line on rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl own|reference] [--workload C3|C2|C4]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workloads (config.workload), all synthetic (guava_renderer_b200/synthetic.py):

  C3 (default, the configuration BASELINE.json's metric is quoted on) -- GUAVA training step,
      100k-Gaussian 32-channel avatar at 512x512, forward + backward, 16 views sharded over 8 GPUs
      => 2 views per GPU per step (weak scaling: N GPUs render 2N views per step), followed for
      N > 1 by one NCCL all-reduce(sum) of the packed per-Gaussian gradients [P,43] fp32.
  C2  -- inference shape: 8 views x 100k Gaussians, 512x512, forward only, one launch chain (N = 1;
      for N > 1 every rank renders its own 8 views: weak).
  C3pose -- the C3 step driven the way GUAVA drives it (N = 1 only; a second, clearly labelled end-to-end line, the
      headline stays C3): the avatar's per-Gaussian attributes are resident, what reaches the device every step is the
      POSE -- deformed mesh vertices [V,Nv,3] and per-vertex transforms [V,Nv,4,4], 1.6 MB -- and the step is
      deformation (models/UbodyAvatar/ubody_gaussian.py:245-289) -> rasterizer forward -> loss -> rasterizer
      backward -> deformation backward, gradients landing on the avatar attributes.
  C4  -- cross-reenactment video: 300k Gaussians, 1024x1024, 240 frames, forward only, frames split
      contiguously over the N ranks (strong scaling: the job is the 240 frames); the cloud is resident,
      each frame has its own positions (base + 0.02 sin(2 pi f/240 + phi_i)); no cross-GPU traffic.

A step = one pass of the hot path over one batch of views (C4: over the rank's share of the frames):
  value : inputs (Gaussians, cameras, upstream dL/dcolor) resident in HBM;
  e2e   : the same step through the public per-step call with HOST inputs: the step's inputs are copied
          from pinned host memory, and the step's result (loss + gradient checksum, or an image checksum
          for the forward-only workloads) is read back, inside the timed region.
`--impl reference` runs the UNMODIFIED reference CUDA build (oracle/_ref) through its own Python API under
the same sharding, inputs, loss recipe and collective; that arm imports nothing of the product.
`views_sets` distinct input sets are rotated so that consecutive steps do not re-use L2-resident inputs.
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "Mpix/s fwd+bwd, 32-ch 512x512 100k-Gaussian avatar"
UNIT = "Mpix/s"

WORKLOADS = {
    # name: P, H, W, views per GPU and step, backward, scaling
    "C3": dict(P=100_000, H=512, W=512, views=2, backward=True, scaling="weak", config_id=3),
    "C2": dict(P=100_000, H=512, W=512, views=8, backward=False, scaling="weak", config_id=2),
    "C4": dict(P=300_000, H=1024, W=1024, views=240, backward=False, scaling="strong", config_id=4),
    "C3pose": dict(P=100_000, H=512, W=512, views=2, backward=True, scaling="weak", config_id=3),
}
C4_CHUNK = 8      # frames per launch chain


def load_synthetic():
    """guava_renderer_b200/synthetic.py loaded BY PATH: pure torch, and the reference arm must not import
    the product package (its __init__ maps libguava_rasterizer_b200.so)."""
    name = "_bench_synthetic"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "guava_renderer_b200", "synthetic.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def shard_views(num_views, rank, world):
    """Contiguous blocks, as guava_renderer_b200.sharding.shard_views (restated: see load_synthetic)."""
    base, rem = divmod(num_views, world)
    start = rank * base + min(rank, rem)
    return list(range(start, start + base + (1 if rank < rem else 0)))


# ----------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed regions (NVML, ~1 ms period; the timed region of
    the default run is tens of milliseconds, far below nvidia-smi's start-up time)."""

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.samples, self.reasons = [], set()
        self.active = threading.Event()
        self.stop_flag = False
        self.h = None
        self.max_mhz = None
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.t = threading.Thread(target=self._loop, daemon=True)
            self.t.start()
        except Exception as e:   # noqa: BLE001
            self.err = repr(e)
            self.h = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while not self.stop_flag:
            if not self.active.wait(0.05):
                continue
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(get_reasons(self.h))
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:   # noqa: BLE001
                pass
            time.sleep(0.001)

    def begin(self):
        self.active.set()

    def end(self):
        self.active.clear()

    def result(self):
        self.stop_flag = True
        if self.h is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: " + getattr(self, "err", "?")],
                    "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def smi_index(device_index):
    cvd = os.environ.get("CUDA_VISIBLE_DEVICES", "")
    try:
        return int(cvd.split(",")[device_index]) if cvd else device_index
    except (ValueError, IndexError):
        return device_index


# ------------------------------------------------------------------- workload
PACK_ORDER = ("feat", "xyz", "sc", "rot", "op", "cams")
TENSOR_KEYS = ("xyz", "sc", "rot", "op", "feat")      # [V,P,3|3|4|1|32] = 43 floats per Gaussian
WIDTHS = dict(xyz=3, sc=3, rot=4, op=1, feat=32)


def carve(packed, V, P, fill=None):
    """Views [V,P,w] (and cams [V,32]) into one flat float buffer, in PACK_ORDER."""
    out, o = {}, 0
    for k in PACK_ORDER:
        shape = (V, 32) if k == "cams" else (V, P, WIDTHS[k])
        n = shape[0] * shape[1] * (shape[2] if len(shape) == 3 else 1)
        out[k] = packed[o:o + n].view(shape)
        if fill is not None:
            out[k].copy_(fill[k].reshape(shape))
        o += n
    return out


def build_inputs(config_id, P, H, W, views, num_views, n_sets, pin):
    """n_sets distinct input sets for the rank's views; host tensors (optionally one pinned block per set)."""
    synthetic = load_synthetic()
    sets = []
    for s in range(n_sets):
        d = synthetic.make_views(config_id * 10 + s, P, H, W, views, num_views)
        cams = torch.cat([d["viewmatrix"].reshape(len(views), 16), d["projmatrix"].reshape(len(views), 16)], 1).contiguous()
        e = dict(xyz=d["xyz"], sc=d["scaling"], rot=d["rotation"], op=d["opacity"], feat=d["features_color"], cams=cams)
        if pin:   # one pinned block per set (feat first: its rows must stay 16-byte aligned) -> one H2D copy per step
            packed = torch.empty(sum(v.numel() for v in e.values())).pin_memory()
            e = carve(packed, len(views), P, fill=e)
            e["packed"] = packed
        e.update(tanfovx=d["tanfovx"], tanfovy=d["tanfovy"], campos=d["campos"])
        sets.append(e)
    return sets


def build_c4(P, H, W, frames, pin):
    """C4: one resident cloud, per-frame positions base + 0.02 sin(2 pi f/240 + phi_i), a fixed
    camera (render_motion.py:177-214).  Returns the shared tensors and the rank's [F,P,3] positions (host)."""
    synthetic = load_synthetic()
    c = synthetic.guava_cloud(P, 4000, 1.0)
    g = torch.Generator().manual_seed(4001)
    phi = torch.rand(P, 3, generator=g) * (2.0 * np.pi)
    f = torch.tensor(frames, dtype=torch.float32).view(-1, 1, 1)
    pos = (c["xyz"][None] + 0.02 * torch.sin(2.0 * np.pi * f / 240.0 + phi[None])).contiguous()
    if pin:
        pos = pos.pin_memory()
    cam = synthetic.camera(0, 1, 1.0 / 24 * W / H, 1.0 / 24)
    return c, pos, cam


def upstream(V, H, W, device):
    synthetic = load_synthetic()
    gs = [synthetic.upstream_grads(H, W, 3000 + v)[0] for v in range(V)]
    return torch.stack(gs, 0).to(device)


def workload_name(args):
    wl = WORKLOADS[args.workload]
    if args.workload == "C3":
        return (f"C3 GUAVA training step: {args.P // 1000}k-Gaussian 32-ch avatar, {args.H}x{args.W}, fwd+bwd, "
                f"{args.views_per_gpu} views/GPU/step (16 views over 8 GPUs), NCCL all-reduce of [P,43] grads for N>1")
    if args.workload == "C3pose":
        return (f"C3pose GUAVA training step from the pose: {args.P // 1000}k-Gaussian avatar resident (10458 vertex + "
                f"{args.P - 10458} face-bound Gaussians), per step {args.views_per_gpu} posed meshes host->device, deformation + "
                f"rasterizer fwd+bwd + deformation bwd, {args.H}x{args.W}")
    if args.workload == "C2":
        return (f"C2 GUAVA inference: {args.P // 1000}k-Gaussian 32-ch avatar, {args.H}x{args.W}, forward only, "
                f"{args.views_per_gpu} views batched per GPU and step")
    return (f"C4 cross-reenactment video: {args.P // 1000}k Gaussians, {args.H}x{args.W}, forward only, "
            f"{wl['views']} frames split by frame over the GPUs, {C4_CHUNK} frames per launch chain")


def config_of(args, world):
    """The `config` object -- the SAME function for both arms, so the two lines carry identical dicts."""
    wl = WORKLOADS[args.workload]
    gv = wl["views"] if args.workload == "C4" else args.views_per_gpu * world
    return {"workload": workload_name(args), "P": args.P, "H": args.H, "W": args.W,
            "views_per_gpu": (wl["views"] // world if args.workload == "C4" else args.views_per_gpu),
            "global_views_per_step": gv, "parallelism": f"view-sharded x{world}",
            "l2": (f"inputs rotate over {args.view_sets} distinct view sets; per-step working set > 126 MB L2"
                   if args.workload != "C4" else "every frame has its own positions; per-step working set > 126 MB L2")}


def base_line(args, world, impl=None):
    wl = WORKLOADS[args.workload]
    line = {"metric": METRIC if args.workload == "C3" else METRIC + f" [workload {args.workload}: see config]",
            "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True,
            "scaling": wl["scaling"], "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": config_of(args, world)}
    if impl:
        line["impl"] = impl
    return line


# -------------------------------------------------------------------- own arm
def run_own(args, rank, world, device):
    import guava_renderer_b200 as grb
    from guava_renderer_b200 import sharding
    import torch.distributed as dist

    if args.workload == "C3pose":
        return run_own_pose(args, rank, world, device)
    if args.workload != "C3":
        return run_own_forward_only(args, rank, world, device)

    P, H, W, V = args.P, args.H, args.W, args.views_per_gpu
    clocks = ClockSampler(smi_index(device.index))
    views = sharding.shard_views(V * world, rank, world)
    sets_host = build_inputs(3, P, H, W, views, V * world, args.view_sets, pin=True)
    dL = upstream(V, H, W, device)
    tanfov = torch.tensor([[sets_host[0]["tanfovx"], sets_host[0]["tanfovy"]]] * V, device=device)
    bg = torch.zeros(V, 32, device=device)

    # device-resident copies for `value`
    sets_dev = []
    for s in sets_host:
        e = {k: s[k].to(device) for k in TENSOR_KEYS}
        cams = s["cams"].to(device)
        e.update(vm=cams[:, :16].reshape(V, 4, 4).contiguous(), pm=cams[:, 16:].reshape(V, 4, 4).contiguous())
        sets_dev.append(e)
    gbuf = torch.zeros(P, 43, device=device)

    def step_resident_eager(i):
        s = sets_dev[i % len(sets_dev)]
        leaves = [s[k].detach().requires_grad_(True) for k in ("xyz", "sc", "rot", "op", "feat")]
        st = grb.BatchRasterizationSettings(H, W, tanfov, bg, 1.0, s["vm"], s["pm"])
        m2 = torch.zeros(V, P, 3, device=device, requires_grad=True)
        if args.per_view_grads:
            color, radii, invd = grb.rasterize_gaussians_batch(leaves[0], m2, leaves[4], leaves[3], leaves[1], leaves[2], st)
            color.backward(dL)
            sharding.pack_gradients([leaves[0].grad, leaves[1].grad, leaves[2].grad, leaves[3].grad, leaves[4].grad], out=gbuf)
        else:       # the same work as the replayed step: the backward sums the views into the all-reduce buffer
            color, radii, invd = grb.rasterize_gaussians_batch(leaves[0], m2, leaves[4], leaves[3], leaves[1], leaves[2], st,
                                                               summed_out=gbuf)
            color.backward(dL)
        sharding.allreduce_gradients(gbuf)
        return color

    # The product's step call: the whole launch chain (forward, loss, backward, gradient packing) of one
    # input set captured once as a CUDA graph (guava_renderer_b200.GraphedRenderStep) and replayed.
    def settings_of(vm, pm):
        return grb.BatchRasterizationSettings(H, W, tanfov, bg, 1.0, vm, pm)

    graphed = not args.eager
    if graphed:
        # summed: the backward adds the views' parameter gradients straight into the step's [P,43] all-reduce buffer
        # (guava_rasterizer_backward_batch_summed) -- no [V,P,43] intermediates, no packing kernel
        g_res = [grb.GraphedRenderStep(s["xyz"], s["feat"], s["op"], s["sc"], s["rot"], settings_of(s["vm"], s["pm"]),
                                       dL_dcolor=dL, summed=not args.per_view_grads)
                 for s in sets_dev]      # each step owns its [P,43] buffer

    # N > 1: the all-reduce of step k is issued asynchronously (NCCL's stream) and joined one step later,
    # so it runs under the binning kernels of step k+1 ("overlap with the next
    # micro-batch's forward").  Everything is drained inside the timed region.
    pending = []

    def join(keep):
        while len(pending) > keep:
            h, fin = pending.pop(0)
            if fin is not None:
                fin(h)
            elif h is not None:
                h.wait()

    def step_resident(i):
        if not graphed:
            return step_resident_eager(i)
        g = g_res[i % len(g_res)]
        g.replay()
        pending.append((sharding.allreduce_gradients(g.packed, async_op=True), None))
        join(1)

    # host-input arm for `e2e`: H2D on a copy stream into a ring of NSTAGE staging buffers, issued two steps
    # ahead (the copy engine then runs back to back and a slow copy does not stall the next step)
    NSTAGE = 3
    copy_stream = torch.cuda.Stream(device=device)
    stage = []
    for _ in range(NSTAGE):
        packed = torch.empty(V * P * 43 + V * 32, device=device)
        stage.append(dict(carve(packed, V, P), packed=packed, ev=torch.cuda.Event()))
    result_host = torch.empty(2, dtype=torch.float32).pin_memory()
    h2d_bytes = V * P * 43 * 4 + V * 32 * 4
    d2h_bytes = 8

    copy_times = []

    def prefetch(i):
        b = stage[i % NSTAGE]
        s = sets_host[i % len(sets_host)]
        compute = torch.cuda.current_stream(device)
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_stream(compute)   # buffer i%NSTAGE is free once step i-NSTAGE (already enqueued) is done
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record(copy_stream)
            b["packed"].copy_(s["packed"], non_blocking=True)   # the step's [V,P,43] Gaussians + cameras
            t1.record(copy_stream)
            copy_times.append((t0, t1))
            b["ev"].record(copy_stream)

    if graphed:
        for i, b in enumerate(stage):      # valid contents while the steps are captured
            b["packed"].copy_(sets_host[i % len(sets_host)]["packed"])
        g_e2e = [grb.GraphedRenderStep(b["xyz"], b["feat"], b["op"], b["sc"], b["rot"],
                                       settings_of(b["cams"][:, :16].reshape(V, 4, 4), b["cams"][:, 16:].reshape(V, 4, 4)),
                                       dL_dcolor=dL, summed=not args.per_view_grads) for b in stage]

    # The step's result (loss + checksum of the reduced gradients -> pinned host memory) is produced on its
    # own stream, ordered after the step (N = 1) or after its all-reduce (N > 1), so that it runs under the
    # next step's binning kernels instead of between two steps; the step that reuses the gradient buffer
    # waits for it, and the timed region ends only after the last result has reached the host.
    result_stream = torch.cuda.Stream(device=device)
    if graphed:
        for g in g_e2e:
            g.done_ev, g.fin_ev = torch.cuda.Event(), torch.cuda.Event()

    trace = {}

    def trace_point(i):      # --trace-e2e: device timestamps of one step's phases (debug aid, not a bench number)
        return trace.setdefault(i, {k: torch.cuda.Event(enable_timing=True) for k in ("step0", "step1", "reduced", "result")})

    def step_e2e(i, last):
        b = stage[i % NSTAGE]
        compute = torch.cuda.current_stream(device)
        compute.wait_event(b["ev"])
        if graphed:
            g = g_e2e[i % NSTAGE]
            compute.wait_event(g.fin_ev)          # the result of step i-NSTAGE has been read out of g.packed
            tr = trace_point(i) if args.trace_e2e else None
            if tr:
                tr["step0"].record(compute)
            g.replay()
            g.done_ev.record(compute)
            if tr:
                tr["step1"].record(compute)

            def finish(h, g=g, tr=tr):
                with torch.cuda.stream(result_stream):
                    if h is not None:
                        h.wait()
                    else:
                        result_stream.wait_event(g.done_ev)
                    if tr:
                        tr["reduced"].record(result_stream)
                    res = torch.stack([g.loss, torch.linalg.vector_norm(g.packed.view(-1), 1)])
                    result_host.copy_(res, non_blocking=True)
                    g.fin_ev.record(result_stream)
                    if tr:
                        tr["result"].record(result_stream)
            pending.append((sharding.allreduce_gradients(g.packed, async_op=True), finish))
            join(1)
            return None
        leaves = [b[k].detach().requires_grad_(True) for k in TENSOR_KEYS]
        st = grb.BatchRasterizationSettings(H, W, tanfov, bg, 1.0, b["cams"][:, :16].reshape(V, 4, 4),
                                            b["cams"][:, 16:].reshape(V, 4, 4))
        m2 = torch.zeros(V, P, 3, device=device, requires_grad=True)
        color, radii, invd = grb.rasterize_gaussians_batch(leaves[0], m2, leaves[4], leaves[3], leaves[1], leaves[2], st)
        loss = grb.graphed.image_dot(color.detach(), dL)
        color.backward(dL)
        sharding.pack_gradients([l.grad for l in leaves], out=gbuf)
        sharding.allreduce_gradients(gbuf)
        res = torch.stack([loss, gbuf.abs().sum()])
        result_host.copy_(res, non_blocking=True)
        return res

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    # ---- resident timing
    for i in range(args.warmup):
        step_resident(i)
    join(0)
    barrier()
    n0 = grb.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clocks.begin()
    e0.record()
    for i in range(args.steps):
        step_resident(args.warmup + i)
    join(0)
    e1.record()
    barrier()
    clocks.end()
    launches = grb.launch_count() - n0
    if graphed:   # replays do not pass through the library's launch counter: count what each replay launches
        launches = sum(g_res[(args.warmup + i) % len(g_res)].launches_per_replay for i in range(args.steps))
    ms_res = e0.elapsed_time(e1)

    # ---- e2e timing (host inputs)
    prefetch(0)
    prefetch(1)
    for i in range(args.warmup):
        prefetch(i + 2)
        step_e2e(i, False)
    join(0)
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    del copy_times[:]
    t_wall0 = time.perf_counter()
    clocks.begin()
    e2.record()
    for i in range(args.steps):
        k = args.warmup + i
        prefetch(k + 2)                       # later steps' H2D overlaps this step's kernels
        step_e2e(k, i == args.steps - 1)
    join(0)
    torch.cuda.current_stream(device).wait_stream(result_stream)
    e3.record()
    t_enq = (time.perf_counter() - t_wall0) * 1e3   # host time to enqueue the timed steps
    barrier()
    clocks.end()
    _ = float(result_host[0])                 # the D2H result is consumed on the host
    t_wall = (time.perf_counter() - t_wall0) * 1e3
    ms_e2e = e2.elapsed_time(e3)
    clk = clocks.result()
    copy_stream.synchronize()
    ms_h2d_loaded = sum(a.elapsed_time(b) for a, b in copy_times) / max(len(copy_times), 1)
    if args.trace_e2e:
        torch.cuda.synchronize(device)
        rows = []
        first = args.warmup
        for j, (c0_, c1_) in enumerate(copy_times):
            rows.append({"copy_issue": j, "t0": round(e2.elapsed_time(c0_), 4), "t1": round(e2.elapsed_time(c1_), 4)})
        for i in sorted(trace):
            if i >= first:
                rows.append({"step": i - first, **{k: round(e2.elapsed_time(ev), 4) for k, ev in trace[i].items()}})
        with open(f"{args.trace_e2e}.rank{rank}.json", "w") as f:
            json.dump(rows, f, indent=0)

    # the step's host->device copy timed alone (what `e2e` would cost if the kernels were free)
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(copy_stream):
        c0.record(copy_stream)
        for i in range(10):
            stage[i % NSTAGE]["packed"].copy_(sets_host[i % len(sets_host)]["packed"], non_blocking=True)
        c1.record(copy_stream)
    copy_stream.synchronize()
    ms_h2d_alone = c0.elapsed_time(c1) / 10

    # ---- per-kernel times + algorithmic bytes for the roofline object (separate, untimed pass)
    grb.set_option("profile_stages", 1)
    grb._lib_mod.stage_times_reset()
    prof_steps = 5
    for i in range(prof_steps):
        step_resident_eager(i)
    torch.cuda.synchronize(device)
    if graphed:
        for g in g_res + g_e2e:
            g.check()      # no replay overflowed its captured binning capacity
    grb.set_option("profile_stages", 0)
    stages = grb._lib_mod.stage_times()
    roof = roofline_object(stages, sets_dev[0], tanfov, bg, P, H, W, V, device, backward=True)

    t = torch.tensor([ms_res, ms_e2e], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_res, ms_e2e = float(t[0]), float(t[1])
    pix = V * world * H * W * args.steps
    return dict(ms_res=ms_res, ms_e2e=ms_e2e, pix=pix, launches=int(launches), clocks=clk, roofline=roof,
                h2d=h2d_bytes, d2h=d2h_bytes, wall_ms_e2e=t_wall, enqueue_ms_e2e=t_enq, ms_h2d_alone=ms_h2d_alone,
                ms_h2d_loaded=ms_h2d_loaded,
                step_call="eager op-by-op" if args.eager else
                ("guava_renderer_b200.GraphedRenderStep (CUDA graph replay)" if args.per_view_grads else
                 "guava_renderer_b200.GraphedRenderStep(summed=True) (CUDA graph replay; views summed inside the backward)"))


ATTR_KEYS = ("smplx_rotation", "smplx_scaling", "smplx_opacity", "smplx_features_color",
             "uv_local_xyz", "uv_rotation", "uv_scaling", "uv_opacity", "uv_features_color")
POSE_SETS = 4


def build_pose_inputs(P, V, H, W, pin):
    """The resident avatar, POSE_SETS posed meshes (one packed block each: vertices [V,Nv,3] then transforms
    [V,Nv,16]) and the V cameras of the C3 orbit."""
    synthetic = load_synthetic()
    av = synthetic.pose_avatar(P, V, seed=3100)
    Nv = av["vertices"].shape[0]
    poses = []
    for i in range(POSE_SETS):
        pv, pt = synthetic.pose_of(av["vertices"], i, V)
        blk = torch.cat([pv.reshape(-1), pt.reshape(-1)])
        poses.append(blk.pin_memory() if pin else blk)
    cams = [synthetic.camera(v, 16, 1.0 / 24 * W / H, 1.0 / 24) for v in range(V)]
    return av, Nv, poses, cams


def run_own_pose(args, rank, world, device):
    """C3pose: pose in, gradients on the avatar attributes out.  The whole step -- deform_gaussians, rasterizer
    forward, loss, rasterizer backward, deformation backward, gradient packing -- is one captured CUDA graph per
    staging buffer (GraphedRenderStep with a producer)."""
    import guava_renderer_b200 as grb
    if world != 1:
        raise SystemExit("--workload C3pose is a single-GPU line")
    P, H, W, V = args.P, args.H, args.W, args.views_per_gpu
    clocks = ClockSampler(smi_index(device.index))
    av, Nv, poses_host, cams = build_pose_inputs(P, V, H, W, pin=True)
    N = P - Nv
    dL = upstream(V, H, W, device)
    tanfov = torch.tensor([[cams[0]["tanfovx"], cams[0]["tanfovy"]]] * V, device=device)
    bg = torch.zeros(V, 32, device=device)
    st = grb.BatchRasterizationSettings(H, W, tanfov, bg, 1.0, torch.stack([c["viewmatrix"] for c in cams]).to(device),
                                        torch.stack([c["projmatrix"] for c in cams]).to(device))
    mesh = {k: av[k].to(device) for k in ("faces", "uv_binding_face", "uv_face_bary")}
    attrs = {k: av[k].to(device) for k in ATTR_KEYS}

    class Producer:
        """deform_gaussians on the posed mesh in `buf`; fresh attribute leaves per call (see GraphedRenderStep)."""
        def __init__(self, buf):
            self.vertices = buf[:V * Nv * 3].view(V, Nv, 3)
            self.transforms = buf[V * Nv * 3:].view(V, Nv, 4, 4)
            self.leaves = {}

        def __call__(self):
            self.leaves = {k: t.detach().requires_grad_(torch.is_grad_enabled()) for k, t in attrs.items()}
            out = grb.deform_gaussians(self.vertices, self.transforms, mesh["faces"], mesh["uv_binding_face"],
                                       mesh["uv_face_bary"], **self.leaves)
            return dict(xyz=out["xyz"], features=out["features_color"], opacity=out["opacity"], scaling=out["scaling"],
                        rotation=out["rotation"])

    NSTAGE = 3
    stage = []
    for i in range(NSTAGE):
        buf = torch.empty(V * Nv * 19, device=device)
        buf.copy_(poses_host[i % POSE_SETS])
        prod = Producer(buf)
        stage.append(dict(buf=buf, prod=prod, ev=torch.cuda.Event(), done=torch.cuda.Event(), fin=torch.cuda.Event(),
                          step=grb.GraphedRenderStep(None, None, None, None, None, st, dL_dcolor=dL, producer=prod,
                                                     pack=False)))      # the step ends in the avatar attributes' gradients
    copy_stream = torch.cuda.Stream(device=device)
    result_stream = torch.cuda.Stream(device=device)     # the step's result is reduced and read back beside the next step
    result_host = torch.empty(3, dtype=torch.float32).pin_memory()
    h2d_bytes, d2h_bytes = V * Nv * 19 * 4, 12

    def prefetch(i):
        b = stage[i % NSTAGE]
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_stream(torch.cuda.current_stream(device))
            b["buf"].copy_(poses_host[i % POSE_SETS], non_blocking=True)
            b["ev"].record(copy_stream)

    def step(i, host_inputs):
        b = stage[i % NSTAGE]
        if host_inputs:
            torch.cuda.current_stream(device).wait_event(b["ev"])
        g = b["step"]
        compute = torch.cuda.current_stream(device)
        if host_inputs:
            compute.wait_event(b["fin"])      # the result of the step that last used these buffers has been read
        g.replay()
        if host_inputs:   # the step's result: loss, |d/d(avatar attributes)|_1, |d/d(uv_local_xyz)|_1 -> pinned host memory
            b["done"].record(compute)
            with torch.cuda.stream(result_stream):
                result_stream.wait_event(b["done"])
                lv = b["prod"].leaves      # the reference arm's recipe: loss, |d/d(all attributes)|_1, |d/d(uv_local_xyz)|_1
                res = torch.stack([g.loss, sum(torch.linalg.vector_norm(lv[k].grad.view(-1), 1) for k in ATTR_KEYS),
                                   torch.linalg.vector_norm(lv["uv_local_xyz"].grad.view(-1), 1)])
                result_host.copy_(res, non_blocking=True)
                b["fin"].record(result_stream)

    out = {}
    for name, host_inputs in (("res", False), ("e2e", True)):
        if host_inputs:
            prefetch(0); prefetch(1)
        for i in range(args.warmup):
            if host_inputs:
                prefetch(i + 2)
            step(i, host_inputs)
        torch.cuda.synchronize(device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        clocks.begin()
        e0.record()
        for i in range(args.steps):
            if host_inputs:
                prefetch(args.warmup + i + 2)
            step(args.warmup + i, host_inputs)
        torch.cuda.current_stream(device).wait_stream(result_stream)   # the last result has reached the host
        e1.record()
        torch.cuda.synchronize(device)
        clocks.end()
        out[name] = e0.elapsed_time(e1)
    vals = [float(x) for x in result_host]
    for b in stage:
        b["step"].check()
    assert all(np.isfinite(vals)) and vals[1] > 0 and vals[2] > 0, vals
    # roofline object: the rasterizer's kernels on one deformed input set (eager, per-stage events)
    with torch.no_grad():
        d = stage[0]["prod"]()
    prof = dict(xyz=d["xyz"], feat=d["features"], op=d["opacity"], sc=d["scaling"], rot=d["rotation"], vm=st.viewmatrix,
                pm=st.projmatrix)
    grb.set_option("profile_stages", 1)
    grb._lib_mod.stage_times_reset()
    for _ in range(3):
        leaves = [prof[k].detach().requires_grad_(True) for k in ("xyz", "feat", "op", "sc", "rot")]
        m2 = torch.zeros(V, P, 3, device=device, requires_grad=True)
        color, _, _ = grb.rasterize_gaussians_batch(leaves[0], m2, leaves[1], leaves[2], leaves[3], leaves[4], st)
        color.backward(dL)
    torch.cuda.synchronize(device)
    grb.set_option("profile_stages", 0)
    roof = roofline_object(grb._lib_mod.stage_times(), prof, tanfov, bg, P, H, W, V, device, backward=True)
    return dict(ms_res=out["res"], ms_e2e=out["e2e"], pix=V * H * W * args.steps,
                launches=int(stage[0]["step"].launches_per_replay * args.steps), clocks=clocks.result(), roofline=roof,
                h2d=h2d_bytes, d2h=d2h_bytes,
                step_call="guava_renderer_b200.GraphedRenderStep with producer=deform_gaussians (CUDA graph replay)")


def run_own_forward_only(args, rank, world, device):
    """C2 (8 views batched per step) and C4 (the rank's share of 240 frames per step, C4_CHUNK frames per launch
    chain, one resident cloud with per-frame positions)."""
    import guava_renderer_b200 as grb
    import torch.distributed as dist
    P, H, W = args.P, args.H, args.W
    clocks = ClockSampler(smi_index(device.index))
    result_host = torch.empty(1, dtype=torch.float32).pin_memory()
    copy_stream = torch.cuda.Stream(device=device)

    if args.workload == "C2":
        V = args.views_per_gpu
        views = shard_views(V * world, rank, world)
        sets_host = build_inputs(2, P, H, W, views, V * world, args.view_sets, pin=True)
        tanfov = torch.tensor([[sets_host[0]["tanfovx"], sets_host[0]["tanfovy"]]] * V, device=device)
        bg = torch.zeros(V, 32, device=device)
        NSTAGE = 3
        stage = []
        for i in range(NSTAGE):
            packed = torch.empty(V * P * 43 + V * 32, device=device)
            packed.copy_(sets_host[i % len(sets_host)]["packed"])
            stage.append(dict(carve(packed, V, P), packed=packed, ev=torch.cuda.Event()))
        steps_g = [grb.GraphedRenderStep(b["xyz"], b["feat"], b["op"], b["sc"], b["rot"],
                                         grb.BatchRasterizationSettings(H, W, tanfov, bg, 1.0,
                                                                        b["cams"][:, :16].reshape(V, 4, 4),
                                                                        b["cams"][:, 16:].reshape(V, 4, 4)))
                   for b in stage]
        h2d_bytes, pix_per_step = V * P * 43 * 4 + V * 32 * 4, V * H * W
        launches_per_step = steps_g[0].launches_per_replay

        def prefetch(i):
            b = stage[i % NSTAGE]
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_stream(torch.cuda.current_stream(device))
                b["packed"].copy_(sets_host[i % len(sets_host)]["packed"], non_blocking=True)
                b["ev"].record(copy_stream)

        def step(i, host_inputs):
            if host_inputs:
                torch.cuda.current_stream(device).wait_event(stage[i % NSTAGE]["ev"])
            g = steps_g[i % NSTAGE]
            g.replay()
            if host_inputs:   # the step's result: an image checksum -> pinned host memory
                result_host.copy_(g.color.view(-1)[::4099].sum().reshape(1), non_blocking=True)

        def check():
            for g in steps_g:
                g.check()
        prof_set = dict(xyz=stage[0]["xyz"], feat=stage[0]["feat"], op=stage[0]["op"], sc=stage[0]["sc"], rot=stage[0]["rot"],
                        vm=stage[0]["cams"][:, :16].reshape(V, 4, 4), pm=stage[0]["cams"][:, 16:].reshape(V, 4, 4))
        Vp = V
    else:   # C4
        frames = shard_views(WORKLOADS["C4"]["views"], rank, world)
        c, pos_host, cam = build_c4(P, H, W, frames, pin=True)
        F = len(frames)
        V = C4_CHUNK
        shared = {k: c[k].to(device) for k in ("scaling", "rotation", "opacity", "features_color")}
        pos_dev = pos_host.to(device)                                   # [F,P,3] resident for `value`
        pos_stage = [torch.empty(V, P, 3, device=device) for _ in range(3)]
        stage_ev = [torch.cuda.Event() for _ in range(3)]
        tanfov = torch.tensor([[cam["tanfovx"], cam["tanfovy"]]] * V, device=device)
        bg = torch.zeros(V, 32, device=device)
        vm = cam["viewmatrix"].to(device)[None].repeat(V, 1, 1).contiguous()
        pm = cam["projmatrix"].to(device)[None].repeat(V, 1, 1).contiguous()
        st = grb.BatchRasterizationSettings(H, W, tanfov, bg, 1.0, vm, pm)
        m2 = torch.zeros(V, P, 3, device=device)
        nchunks = (F + V - 1) // V
        h2d_bytes, pix_per_step = F * P * 12, F * H * W
        acc = torch.zeros(1, device=device)

        def render_chunk(xyz, n):
            stn = st if n == V else grb.BatchRasterizationSettings(H, W, tanfov[:n], bg[:n], 1.0, vm[:n], pm[:n])
            with torch.no_grad():
                color, radii, invd = grb.rasterize_gaussians_batch(xyz[:n], m2[:n], shared["features_color"], shared["opacity"],
                                                                   shared["scaling"], shared["rotation"], stn)
            return color

        n_before = grb.launch_count()
        render_chunk(pos_dev[:V], min(V, F))
        launches_per_step = (grb.launch_count() - n_before) * nchunks

        def prefetch(i):
            pass

        def step(i, host_inputs):
            acc.zero_()
            for ch in range(nchunks):
                n = min(V, F - ch * V)
                if host_inputs:   # the chunk's positions from pinned host memory, double-buffered on the copy stream
                    k = ch % 3
                    with torch.cuda.stream(copy_stream):
                        copy_stream.wait_stream(torch.cuda.current_stream(device))
                        pos_stage[k][:n].copy_(pos_host[ch * V:ch * V + n], non_blocking=True)
                        stage_ev[k].record(copy_stream)
                    torch.cuda.current_stream(device).wait_event(stage_ev[k])
                    color = render_chunk(pos_stage[k], n)
                    acc.add_(color.view(-1)[::4099].sum())
                else:
                    render_chunk(pos_dev[ch * V:ch * V + n], n)
            if host_inputs:
                result_host.copy_(acc, non_blocking=True)

        def check():
            pass
        prof_set = dict(xyz=pos_dev[:V], feat=shared["features_color"], op=shared["opacity"], sc=shared["scaling"],
                        rot=shared["rotation"], vm=vm, pm=pm)
        Vp = V

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    out = {}
    for name, host_inputs in (("res", False), ("e2e", True)):
        if host_inputs:
            prefetch(0); prefetch(1)
        for i in range(args.warmup):
            if host_inputs:
                prefetch(i + 2)
            step(i, host_inputs)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        clocks.begin()
        e0.record()
        for i in range(args.steps):
            if host_inputs:
                prefetch(args.warmup + i + 2)
            step(args.warmup + i, host_inputs)
        e1.record()
        barrier()
        clocks.end()
        out[name] = e0.elapsed_time(e1)
    _ = float(result_host[0])
    check()
    grb.set_option("profile_stages", 1)
    grb._lib_mod.stage_times_reset()
    with torch.no_grad():
        for _ in range(3):
            grb.rasterize_gaussians_batch(prof_set["xyz"], torch.zeros(Vp, P, 3, device=device), prof_set["feat"], prof_set["op"],
                                          prof_set["sc"], prof_set["rot"],
                                          grb.BatchRasterizationSettings(H, W, tanfov, bg, 1.0, prof_set["vm"], prof_set["pm"]))
    torch.cuda.synchronize(device)
    grb.set_option("profile_stages", 0)
    roof = roofline_object(grb._lib_mod.stage_times(), prof_set, tanfov, bg, P, H, W, Vp, device, backward=False)
    t = torch.tensor([out["res"], out["e2e"]], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if args.workload == "C4":
        pix = WORKLOADS["C4"]["views"] * H * W * args.steps        # the job: all 240 frames, whatever N is
    else:
        pix = pix_per_step * world * args.steps
    return dict(ms_res=float(t[0]), ms_e2e=float(t[1]), pix=pix, launches=int(launches_per_step * args.steps),
                clocks=clocks.result(), roofline=roof, h2d=h2d_bytes, d2h=4,
                step_call=("guava_renderer_b200.GraphedRenderStep (CUDA graph replay)" if args.workload == "C2"
                           else f"rasterize_gaussians_batch, {C4_CHUNK} frames per call, eager"))


def roofline_object(stages, s, tanfov, bg, P, H, W, V, device, backward):
    """HBM roofline of the dominant kernel: algorithmic bytes per launch (SURVEY.md 8(d),
    DESIGN.md section 5) / average launch duration measured above with CUDA events."""
    import guava_renderer_b200 as grb
    from guava_renderer_b200 import rasterizer as rz
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    per_launch = {k: (ms / n if n else 0.0) for k, (ms, n) in stages.items()}
    dom = max(per_launch, key=lambda k: stages[k][0])
    # R, Rc from the run's own buffers: a forward whose three buffers are handed back
    st = grb.BatchRasterizationSettings(H, W, tanfov, bg, 1.0, s["vm"], s["pm"])
    with torch.no_grad():
        grb.rasterize_gaussians_batch(s["xyz"], torch.zeros(V, P, 3, device=device), s["feat"], s["op"], s["sc"], s["rot"], st)
    torch.cuda.synchronize()
    img = rz._tls.last_buffers[2]
    L = grb.buffer_layout(V, P, W, H, 0)
    npix = V * H * W
    hdr = img[L.num_rendered:L.num_rendered + 32].view(torch.int32)
    Rtot = int(hdr[0].item())
    ncontrib = img[L.n_contrib:L.n_contrib + npix * 4].view(torch.int32).view(V, H // 16, 16, W // 16, 16)
    Rc = int(ncontrib.amax(dim=(2, 4)).sum().item())
    pairs = int(ncontrib.sum().item())
    Nt = V * (H // 16) * (W // 16)
    n = V * P
    alg = {
        "preprocess": 44 * n + 48 * n,                                   # inputs + (rec 32, depth 4, rect 8, touched 4)
        "tile_scan": 16 * Nt,
        "scatter": 12 * n + 8 * Rtot,
        "tile_sort": 8 * Nt + 8 * Rtot + 4 * Rtot,
        "render_forward": 164 * Rc + 8 * Nt + 140 * npix,                # id 4 + rec 32 + features 128 per instance
        "render_backward": 164 * Rc + 8 * Nt + 140 * npix + 164 * n,
        "preprocess_backward": 100 * n + 64 * n,
    }
    dur_s = per_launch[dom] * 1e-3
    achieved = alg[dom] / dur_s / 1e9 if dur_s > 0 else 0.0
    # DRAM bytes per launch of that kernel from the committed `ncu --set full` capture of this command
    traffic, issue, tpath = None, None, os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.isfile(tpath):
        t = json.load(open(tpath)).get(dom)
        if t and t.get("P") == P and t.get("views") == V and t.get("H") == H and t.get("W") == W:
            traffic = int(t["dram_bytes_read"] + t["dram_bytes_write"])
            if "issue_active_pct" in t:   # what actually bounds the kernel (same ncu capture): issue slots, not HBM
                issue = {"issue_slots_busy_pct": t["issue_active_pct"], "ipc": t["ipc"], "peak_ipc": 4.0,
                         "source": "profiles/" + t["source"]}
    return {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
            "frac": round(achieved / peak, 4), "traffic": traffic, "issue": issue, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": int(alg[dom]), "launch_ms": round(per_launch[dom], 4),
            "units": {"R": Rtot, "Rc": Rc, "pairs": pairs, "views": V, "P": P},
            "per_kernel_ms": {k: round(v, 4) for k, v in per_launch.items()},
            "per_kernel_frac": {k: (round(alg[k] / (per_launch[k] * 1e-3) / 1e9 / peak, 4) if per_launch[k] > 0 else None)
                                for k in per_launch}}


# ------------------------------------------------------------- cpu baseline
def cpu_baseline(args, max_views=2):
    """The oracle port (C + OpenMP, all host threads) on a bounded sample of the same workload.  The reference
    ships no CPU rasterizer (BASELINE.json); this is the CPU restatement of its equations, kind "port"."""
    import oracle
    synthetic = load_synthetic()
    P, H, W = args.P, args.H, args.W
    backward = WORKLOADS[args.workload]["backward"]
    oracle.build()
    cores = oracle.num_threads()
    d = synthetic.make_views(WORKLOADS[args.workload]["config_id"], P, H, W, list(range(max_views)), 16)
    t0 = time.perf_counter()
    for v in range(max_views):
        f = oracle.forward(d["xyz"][v].numpy(), d["features_color"][v].numpy(), d["opacity"][v].numpy(),
                           d["scaling"][v].numpy(), d["rotation"][v].numpy(), d["viewmatrix"][v].numpy(),
                           d["projmatrix"][v].numpy(), d["tanfovx"], d["tanfovy"], H, W, np.zeros(32, np.float32))
        if backward:
            dc, di = synthetic.upstream_grads(H, W, 3000 + v)
            oracle.backward(f, dc.numpy(), di.numpy())
    dt = time.perf_counter() - t0
    return {"value": round(max_views * H * W / dt / 1e6, 4), "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{max_views} views of the bench workload (P={P}, {H}x{W}, {'fwd+bwd' if backward else 'fwd'}), "
                      "oracle/splat_oracle.c with OpenMP"}


# --------------------------------------------------------- reference arm
def import_reference():
    """The unmodified reference build (oracle/_ref) by path, or None when it has not been built."""
    d = os.path.join(ROOT, "oracle", "_ref", "diff_gaussian_rasterization_32")
    if not os.path.isfile(os.path.join(d, "_C.so")):
        return None
    spec = importlib.util.spec_from_file_location("ref_diff_gaussian_rasterization_32", os.path.join(d, "__init__.py"),
                                                  submodule_search_locations=[d])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_diff_gaussian_rasterization_32"] = mod
    spec.loader.exec_module(mod)
    return mod


def run_reference(args, rank, world, local_rank):
    """Reference arm.  The reference ships no CPU rasterizer (BASELINE.json), so "the reference's own
    implementation of the path" is its CUDA extension: when oracle/_ref (the unmodified build) loads and a GPU
    is visible, EVERY rank times it through its own Python API with the per-view loop of
    models/UbodyAvatar/gaussian_render.py:37-67 under the same view sharding, the same loss recipe
    (backward(dL) + a dot product for the value) and -- N > 1 -- one all-reduce of the [P,43] gradients per
    step, as its trainer's DDP step does (main/trainer.py:40-43,95).  `value` = inputs resident, `e2e` = host
    inputs copied in and the result read back every step (the own arm's meanings).  Without a GPU or the
    build, rank 0 times the CPU oracle port.  Nothing of the product is imported here."""
    ref = import_reference() if torch.cuda.is_available() else None
    base = base_line(args, world, impl="reference")
    P, H, W = args.P, args.H, args.W
    wl = WORKLOADS[args.workload]
    if ref is None:
        if rank != 0:
            return None
        cpu = cpu_baseline(args, max_views=2)
        v = cpu["value"]
        base.update(value=v, ms_per_step=round(config_of(args, world)["global_views_per_step"] * H * W / (v * 1e6) * 1e3, 3),
                    cpu_baseline=cpu, e2e={"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                    reference_kind="oracle port on host cores (reference CUDA build not loadable here)")
        return base

    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    clocks = ClockSampler(smi_index(device.index))
    backward = wl["backward"]
    bgv = torch.zeros(32, device=device)
    result_host = torch.empty(2, dtype=torch.float32).pin_memory()

    if args.workload == "C4":
        frames = shard_views(wl["views"], rank, world)
        c, pos_host, cam = build_c4(P, H, W, frames, pin=True)
        F = len(frames)
        shared = {k: c[k].to(device) for k in ("scaling", "rotation", "opacity", "features_color")}
        pos_dev = pos_host.to(device)
        pos_stage = [torch.empty(C4_CHUNK, P, 3, device=device) for _ in range(2)]
        rs = ref.GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=bgv, scale_modifier=1.0,
            viewmatrix=cam["viewmatrix"].to(device), projmatrix=cam["projmatrix"].to(device), sh_degree=0,
            campos=cam["campos"].to(device), prefiltered=False, debug=False, antialiasing=False)
        rast = ref.GaussianRasterizer_32(raster_settings=rs)
        m2 = torch.zeros(P, 3, device=device)
        h2d, d2h = F * P * 12, 4
        pix_per_step = None
        acc = torch.zeros(1, device=device)

        def step(i, host_inputs):
            acc.zero_()
            with torch.no_grad():
                for f0 in range(0, F, C4_CHUNK):          # main/test.py:70-74: one frame after the other
                    n = min(C4_CHUNK, F - f0)
                    if host_inputs:
                        buf = pos_stage[(f0 // C4_CHUNK) % 2]
                        buf[:n].copy_(pos_host[f0:f0 + n], non_blocking=True)
                    else:
                        buf = pos_dev[f0:f0 + n]
                    for j in range(n):
                        img, radii, depth = rast(means3D=buf[j], means2D=m2, shs=None, colors_precomp=shared["features_color"],
                                                 opacities=shared["opacity"], scales=shared["scaling"],
                                                 rotations=shared["rotation"], cov3D_precomp=None)
                        if host_inputs:
                            acc.add_(img.view(-1)[::4099].sum())
            if host_inputs:
                result_host[:1].copy_(acc, non_blocking=True)
    elif args.workload == "C3pose":
        # the reference's own path for this step: Ubody_Gaussian.forward (ubody_gaussian.py:245-289) as eager torch ops
        # on the GPU -- its line-by-line restatement oracle/deform_oracle.py, the module itself needs pytorch3d / roma --
        # then the per-view loop over the unmodified rasterizer build, backward through both
        if world != 1:
            raise SystemExit("--workload C3pose is a single-GPU line")
        from oracle import deform_oracle
        V = args.views_per_gpu
        av, Nv, poses_host, cams = build_pose_inputs(P, V, H, W, pin=True)
        dL = upstream(V, H, W, device)
        mesh = {k: av[k].to(device) for k in ("faces", "uv_binding_face", "uv_face_bary")}
        attrs = {k: av[k].to(device) for k in ATTR_KEYS}
        pose_dev = [poses_host[i % POSE_SETS].to(device) for i in range(2)]
        cam_dev = [{k: (c[k].to(device) if torch.is_tensor(c[k]) else c[k]) for k in c} for c in cams]
        result_host = torch.empty(3, dtype=torch.float32).pin_memory()
        h2d, d2h = V * Nv * 19 * 4, 12

        def step(i, host_inputs):
            buf = pose_dev[i % 2]
            if host_inputs:
                buf.copy_(poses_host[i % POSE_SETS], non_blocking=True)
            leaves = {k: t.detach().requires_grad_(True) for k, t in attrs.items()}
            d = deform_oracle.deform(buf[:V * Nv * 3].view(V, Nv, 3), buf[V * Nv * 3:].view(V, Nv, 4, 4), mesh["faces"],
                                     mesh["uv_binding_face"], mesh["uv_face_bary"], **leaves)
            mean_2d = torch.zeros(V, P, 3, device=device, requires_grad=True)
            imgs = []
            for bi in range(V):          # gaussian_render.py:37-67
                c = cam_dev[bi]
                rs = ref.GaussianRasterizationSettings(
                    image_height=H, image_width=W, tanfovx=c["tanfovx"], tanfovy=c["tanfovy"], bg=bgv, scale_modifier=1.0,
                    viewmatrix=c["viewmatrix"], projmatrix=c["projmatrix"], sh_degree=0, campos=c["campos"],
                    prefiltered=False, debug=False, antialiasing=False)
                img, radii, depth = ref.GaussianRasterizer_32(raster_settings=rs)(
                    means3D=d["xyz"][bi], means2D=mean_2d[bi], shs=None, colors_precomp=d["features_color"][bi],
                    opacities=d["opacity"][bi], scales=d["scaling"][bi], rotations=d["rotation"][bi], cov3D_precomp=None)
                imgs.append(img)
            color = torch.stack(imgs, 0)
            loss = torch.dot(color.detach().view(-1), dL.view(-1))
            color.backward(dL)
            if host_inputs:
                res = torch.stack([loss, sum(torch.linalg.vector_norm(leaves[k].grad.view(-1), 1) for k in ATTR_KEYS),
                                   torch.linalg.vector_norm(leaves["uv_local_xyz"].grad.view(-1), 1)])
                result_host.copy_(res, non_blocking=True)
    else:
        V = args.views_per_gpu
        views = shard_views(V * world, rank, world)
        sets_host = build_inputs(wl["config_id"], P, H, W, views, V * world, args.view_sets, pin=True)
        dL = upstream(V, H, W, device) if backward else None
        stage = [dict({k: torch.empty(V, P, WIDTHS[k], device=device) for k in TENSOR_KEYS},
                      cams=torch.empty(V, 32, device=device)) for _ in range(2)]
        campos_dev = [s["campos"].to(device) for s in sets_host]          # moved once, outside the step
        gbuf = torch.zeros(P, 43, device=device)
        h2d, d2h = V * P * 43 * 4 + V * 32 * 4, (8 if backward else 4)

        def step(i, host_inputs):
            b = stage[i % 2]
            si = i % len(sets_host)
            s = sets_host[si]
            if host_inputs:
                for k in TENSOR_KEYS + ("cams",):
                    b[k].copy_(s[k], non_blocking=True)
            leaves = [b[k].detach().requires_grad_(backward) for k in TENSOR_KEYS]
            mean_2d = torch.zeros(V, P, 3, device=device, requires_grad=backward)
            imgs = []
            with torch.set_grad_enabled(backward):
                for bi in range(V):          # gaussian_render.py:37-67
                    rs = ref.GaussianRasterizationSettings(
                        image_height=H, image_width=W, tanfovx=s["tanfovx"], tanfovy=s["tanfovy"], bg=bgv, scale_modifier=1.0,
                        viewmatrix=b["cams"][bi, :16].reshape(4, 4), projmatrix=b["cams"][bi, 16:].reshape(4, 4), sh_degree=0,
                        campos=campos_dev[si][bi], prefiltered=False, debug=False, antialiasing=False)
                    img, radii, depth = ref.GaussianRasterizer_32(raster_settings=rs)(
                        means3D=leaves[0][bi], means2D=mean_2d[bi], shs=None, colors_precomp=leaves[4][bi],
                        opacities=leaves[3][bi], scales=leaves[1][bi], rotations=leaves[2][bi], cov3D_precomp=None)
                    imgs.append(img)
                color = torch.stack(imgs, 0)   # gaussian_render.py:70 -- part of the reference's own loop
            if not backward:
                if host_inputs:
                    result_host[:1].copy_(color.view(-1)[::4099].sum().reshape(1), non_blocking=True)
                return
            # the own arm's loss recipe: loss value = <color, dL> as one dot product, dL handed straight to backward
            loss = torch.dot(color.detach().view(-1), dL.view(-1))
            color.backward(dL)
            torch.cat([leaves[0].grad.sum(0), leaves[1].grad.sum(0), leaves[2].grad.sum(0), leaves[3].grad.sum(0),
                       leaves[4].grad.sum(0)], dim=1, out=gbuf)
            if world > 1:
                dist.all_reduce(gbuf)          # what DDP does inside fabric.backward (main/trainer.py:95)
            if host_inputs:
                result_host.copy_(torch.stack([loss, torch.linalg.vector_norm(gbuf.view(-1), 1)]), non_blocking=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    out = {}
    for name, host_inputs in (("resident", False), ("e2e", True)):
        if args.workload not in ("C4", "C3pose"):
            for i in range(2):   # both staging buffers hold data for the resident pass
                for k in TENSOR_KEYS + ("cams",):
                    stage[i][k].copy_(sets_host[i % len(sets_host)][k])
        for i in range(args.warmup):
            step(i, host_inputs)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        clocks.begin()
        e0.record()
        for i in range(args.steps):
            step(args.warmup + i, host_inputs)
        e1.record()
        barrier()
        clocks.end()
        out[name] = e0.elapsed_time(e1)
    _ = float(result_host[0])
    t = torch.tensor([out["resident"], out["e2e"]], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return None
    ms_res, ms_e2e = float(t[0]), float(t[1])
    gv = config_of(args, world)["global_views_per_step"]
    pix = gv * H * W * args.steps
    v_res, v_e2e = pix / (ms_res * 1e-3) / 1e6, pix / (ms_e2e * 1e-3) / 1e6
    base.update(value=round(v_res, 2), ms_per_step=round(ms_res / args.steps, 4),
                e2e={"value": round(v_e2e, 2), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                     "ms_per_step": round(ms_e2e / args.steps, 4)},
                clocks=clocks.result(),
                notes={"reference_kind": "unmodified reference CUDA build (oracle/_ref) through its own Python API, per-view "
                                         "loop + torch.stack of gaussian_render.py:37-70, one process per GPU under the own "
                                         "arm's view sharding" + (", dist.all_reduce of the [P,43] gradients" if backward else ""),
                       "loss_recipe": "torch.dot(color, dL) for the value + color.backward(dL)" if backward else "forward only",
                       "timed": "value: inputs resident; e2e: pinned-host inputs copied in and the result read back every step"})
    if world == 1:
        base["cpu_baseline"] = cpu_baseline(args, max_views=2)
    return base


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--workload", default="C3", choices=sorted(WORKLOADS))
    ap.add_argument("--P", type=int, default=None)
    ap.add_argument("--H", type=int, default=None)
    ap.add_argument("--W", type=int, default=None)
    ap.add_argument("--views-per-gpu", type=int, default=None)
    ap.add_argument("--view-sets", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--trace-e2e", default="", help="write device timestamps of the e2e steps' phases to PATH.rank<r>.json")
    ap.add_argument("--no-numa-bind", action="store_true", help="do not pin each rank to its GPU's NUMA node")
    ap.add_argument("--eager", action="store_true", help="enqueue every step op by op instead of replaying CUDA graphs")
    ap.add_argument("--per-view-grads", action="store_true",
                    help="C3: materialise the [V,P,43] per-view gradients and pack them with a kernel, instead of summing "
                         "over the views inside the backward (GraphedRenderStep(summed=True), the default)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    wl = WORKLOADS[args.workload]
    args.P = args.P or wl["P"]
    args.H = args.H or wl["H"]
    args.W = args.W or wl["W"]
    args.views_per_gpu = args.views_per_gpu or (wl["views"] if args.workload != "C4" else C4_CHUNK)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        line = run_reference(args, rank, world, local_rank)
        if line is not None:
            print(json.dumps(line), flush=True)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py (own arm) needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    from guava_renderer_b200 import sharding
    numa_cpus = sharding.bind_host_to_gpu(local_rank) if world > 1 and not args.no_numa_bind else None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        sharding.limit_collective_footprint()
        dist.init_process_group("nccl", device_id=device)
    r = run_own(args, rank, world, device)
    if rank == 0:
        value = r["pix"] / (r["ms_res"] * 1e-3) / 1e6
        e2e = r["pix"] / (r["ms_e2e"] * 1e-3) / 1e6
        line = base_line(args, world)
        line.update({
            "value": round(value, 2), "ms_per_step": round(r["ms_res"] / args.steps, 4),
            "e2e": {"value": round(e2e, 2), "unit": UNIT, "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                    "ms_per_step": round(r["ms_e2e"] / args.steps, 4)},
            "gpu_launches": r["launches"],
            "clocks": r["clocks"],
            "roofline": r["roofline"],
            "notes": {"step_call": r["step_call"],
                      "host_numa_bind": (f"rank 0 pinned to {len(numa_cpus)} CPUs local to its GPU" if numa_cpus else "none")},
        })
        if "enqueue_ms_e2e" in r:
            line["e2e"].update(host_enqueue_ms_per_step=round(r["enqueue_ms_e2e"] / args.steps, 4),
                               h2d_ms_alone=round(r["ms_h2d_alone"], 4), h2d_ms_under_kernels=round(r["ms_h2d_loaded"], 4))
        if not args.no_cpu_baseline and world == 1:   # rank 0 at N = 1 only
            line["cpu_baseline"] = cpu_baseline(args, max_views=2)
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
