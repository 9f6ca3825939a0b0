"""CUDA-graph capture of a fixed-shape render step (forward [+ loss + backward + gradient packing]).

At GUAVA sizes one batched forward+backward is < 1 ms of GPU work, i.e. the same order as the host
time PyTorch needs to enqueue it (~60 tensor ops); the reference additionally blocks on a device->host
copy per view (rasterizer_impl.cu:284).  `GraphedRenderStep` records the whole launch chain once -- the
C-ABI forward in its no-host-wait mode (guava_call_options::no_host_wait), the loss, the C-ABI backward and the
packing of the per-Gaussian gradients -- and replays it with one call, so the step costs what its kernels
cost.  Inputs are STATIC device tensors (the graph reads their storage on every replay; refresh them
with `copy_`), exactly as for torch.cuda.graphs whole-network capture.

The binning buffer of a captured forward has a fixed capacity (learnt from eager warm-up runs on the
same inputs, +25 %).  Every replay's tile scan stores (num_rendered, longest list | overflow << 31)
straight into two pinned host words the step owns (no copy node: as the graph's last node a 4-byte
device->host copy cost 22 us per replay); `check()` raises if the latest replay needed more capacity,
after which the step must be re-captured.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch

from . import _lib
from . import rasterizer as _rz
from .rasterizer import BatchRasterizationSettings, rasterize_gaussians_batch
from .sharding import pack_gradients


def image_dot(a: torch.Tensor, b: torch.Tensor, out64: Optional[torch.Tensor] = None) -> torch.Tensor:
    """<a, b> as a 0-dim tensor, computed by the library's 16-CTA reduction (`guava_image_dot`) on the
    current stream.  `out64`: 64 zero-initialised floats to reuse across calls (CUDA-graph friendly)."""
    import ctypes
    if not (a.is_cuda and b.is_cuda) or a.dtype != torch.float32 or b.dtype != torch.float32 or a.numel() != b.numel():
        raise ValueError("image_dot needs two float32 CUDA tensors of equal size")
    a, b = a.contiguous(), b.contiguous()
    if out64 is None:
        out64 = torch.zeros(64, device=a.device)
    stream = ctypes.c_void_p(torch.cuda.current_stream(a.device).cuda_stream)
    rc = _lib.load().guava_image_dot(a.data_ptr(), b.data_ptr(), a.numel(), out64.data_ptr(), stream)
    if rc != _lib.GUAVA_OK:
        raise RuntimeError(_lib.last_error())
    return out64[0]


class GraphedRenderStep:
    """color, radii, invdepth = rasterize_gaussians_batch(...) [; loss; backward; pack] as one CUDA graph.

    means3D [V,P,3] | colors_precomp [V,P,32] | opacities [V,P,1] | scales [V,P,3] | rotations [V,P,4]
    (or [P,...] shared by all views), `settings` as for rasterize_gaussians_batch -- all static.
    Training mode: give `dL_dcolor` [V,32,H,W] (upstream gradient; loss = <color, dL_dcolor>) or
    `loss_fn(color, invdepth) -> scalar`.  After `replay()`:
        .color .radii .invdepth          forward outputs
        .loss                            scalar tensor (training mode)
        .grads                           dict xyz / scaling / rotation / opacity / features (per view)
        .packed                          [P,43] sum over views, the all-reduce payload (sharding.py)
    `summed=True`: the backward sums over the views inside its kernels, straight into `.packed`
    (guava_rasterizer_backward_batch_summed: no [V,P,43] intermediates, no packing kernel); `.grads` and
    `.viewspace_grad` are then not produced.  Not with a `producer` (its backward needs the per-view gradients).
    `pack=False`: `.packed` is not produced at all (e.g. with a `producer`, whose own leaves' gradients are the result).
    `producer` (instead of the five tensors, pass None for them): a callable that computes the rasterizer's inputs
    inside the captured step from its own static tensors -- e.g. `deform_gaussians` on a posed mesh (the step before
    the rasterizer in GUAVA, ubody_gaussian.py:245-289) -- and returns them as a dict xyz / features / opacity /
    scaling / rotation.  In training mode the backward then runs on through the producer's autograd graph into ITS
    leaves, which it must create afresh on every call (`t.detach().requires_grad_(True)`: a captured backward assigns
    `.grad`, it must not accumulate into a tensor left over from the warm-up runs).
    """

    def __init__(self, means3D, colors_precomp, opacities, scales, rotations, settings: BatchRasterizationSettings,
                 dL_dcolor: Optional[torch.Tensor] = None,
                 loss_fn: Optional[Callable[[torch.Tensor, torch.Tensor], torch.Tensor]] = None,
                 packed_out: Optional[torch.Tensor] = None, warmup: int = 2,
                 producer: Optional[Callable[[], Dict[str, torch.Tensor]]] = None, summed: bool = False,
                 pack: bool = True):
        if summed and producer is not None:
            raise ValueError("summed=True gives no per-view gradients for a producer's backward")
        self.summed = bool(summed)
        self.pack = bool(pack)      # False: `.packed` is not produced (a producer's step ends in ITS leaves' gradients)
        self.producer = producer
        if producer is not None:
            means3D = settings.viewmatrix       # device only
        if not means3D.is_cuda:
            raise RuntimeError("GraphedRenderStep needs CUDA tensors (there is no CPU path)")
        self.inputs = None if producer is not None else dict(xyz=means3D, features=colors_precomp, opacity=opacities,
                                                             scaling=scales, rotation=rotations)
        self.settings = settings
        self.dL_dcolor = dL_dcolor
        self.loss_fn = loss_fn
        self.training = dL_dcolor is not None or loss_fn is not None
        self.device = means3D.device
        self.packed = packed_out
        self.color = self.radii = self.invdepth = self.loss = None
        self.grads: Dict[str, torch.Tensor] = {}
        # (num_rendered, longest list | overflow << 31) of the latest replay: the tile scan stores them
        # straight into these pinned words (guava_call_options::status_words) -- no copy node in the graph
        self._status = torch.zeros(2, dtype=torch.int32).pin_memory()
        # means2D is a gradient sink whose values no kernel reads (gaussian_render.py:30-33): never filled
        self._side = torch.cuda.Stream(device=self.device)
        self._loss_buf = torch.zeros(64, device=self.device)      # guava_image_dot: [0] value, rest scratch
        self.graph = torch.cuda.CUDAGraph()

        # eager warm-up on a side stream: loads kernels, sets function attributes, learns the capacity
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._step()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)

        n0 = _lib.launch_count()
        with torch.cuda.graph(self.graph):
            self._step(no_host_wait=True)
        # kernels of libguava_rasterizer_b200.so that one replay launches
        self.launches_per_replay = _lib.launch_count() - n0

    # one step, eager or under capture
    def _step(self, no_host_wait: bool = False):
        V = int(self.settings.viewmatrix.shape[0])
        if self.producer is not None:
            with torch.set_grad_enabled(self.training):
                leaf = self.producer()
            if self.training:
                for v in leaf.values():
                    if v.requires_grad and not v.is_leaf:
                        v.retain_grad()      # the rasterizer-input gradients feed .grads / .packed
        elif self.training:
            leaf = {k: v.detach().requires_grad_(True) for k, v in self.inputs.items()}
        else:
            leaf = self.inputs
        P = int(leaf["xyz"].shape[-2])
        m2 = torch.empty((V, P, 3), device=self.device, requires_grad=self.training)
        summed_out = None
        if self.training and self.summed:
            if self.packed is None:
                self.packed = torch.empty(P, 43, dtype=torch.float32, device=self.device)
            summed_out = self.packed     # zero-filled by the forward (under its tile scan); the backward's kernels add into it
        with torch.set_grad_enabled(self.training):
            color, radii, invd = rasterize_gaussians_batch(leaf["xyz"], m2, leaf["features"], leaf["opacity"],
                                                           leaf["scaling"], leaf["rotation"], self.settings,
                                                           status_words=self._status, no_host_wait=no_host_wait,
                                                           summed_out=summed_out)
        self.color, self.radii, self.invdepth = color, radii, invd
        if not self.training:
            return
        if self.loss_fn is not None:
            loss = self.loss_fn(color, invd)
            loss.backward()
            self.loss = loss.detach()
        else:
            # loss = <color, dL_dcolor>: one dot-product for the value, and the upstream gradient handed
            # straight to the rasterizer's backward (no product tensor is materialised).  The value is
            # not an input of the backward, so its HBM-bound reduction runs on a side stream under the
            # issue-bound backward compositor -- on 16 CTAs (guava_image_dot): a full-grid library
            # reduction launched there held the SMs while the compositor's first wave arrived (-15 us);
            # the step's stream rejoins the side stream before the step ends.
            cur = torch.cuda.current_stream(self.device)
            self._side.wait_stream(cur)
            with torch.cuda.stream(self._side):
                self.loss = image_dot(color.detach(), self.dL_dcolor, self._loss_buf)
            color.backward(self.dL_dcolor)
        if self.summed:
            self.grads, self.viewspace_grad = {}, None
        else:
            self.grads = {k: v.grad for k, v in leaf.items()}
            self.viewspace_grad = m2.grad
            if self.pack:
                order = ("xyz", "scaling", "rotation", "opacity", "features")
                self.packed = pack_gradients([self.grads[k] for k in order], out=self.packed)
        if self.loss_fn is None:
            cur.wait_stream(self._side)

    def replay(self):
        self.graph.replay()
        return self

    def check(self):
        """Synchronise and raise if the latest replay overflowed the captured binning capacity."""
        torch.cuda.current_stream(self.device).synchronize()
        if int(self._status[1]) < 0:      # bit 31: the replay's num_rendered exceeded the captured capacity
            raise RuntimeError("GraphedRenderStep: a replay needed more binning capacity (or a longer tile list) than "
                               "was captured; re-create the step (or raise option 'capacity_hint')")
