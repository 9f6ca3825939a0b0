"""ctypes binding of libguava_rasterizer_b200.so (C ABI: include/guava_rasterizer.h).

There is no CPU or PyTorch fallback: if the shared object is missing or a
required symbol is absent, loading raises and every product entry point fails.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import (CFUNCTYPE, POINTER, Structure, c_char_p, c_float, c_int, c_longlong,
                    c_size_t, c_ulonglong, c_void_p)

_HERE = os.path.dirname(os.path.abspath(__file__))
# GUAVA_LIB_DIR: a directory holding another build of the two libraries (A/B runs of kernel variants, tools/gpu_ab_lib.sh)
SO_PATH = os.path.join(os.environ.get("GUAVA_LIB_DIR") or os.path.join(_HERE, "lib"), "libguava_rasterizer_b200.so")

ALLOC_FN = CFUNCTYPE(c_void_p, c_void_p, c_size_t)

EXPORTS = (
    "guava_rasterizer_forward",
    "guava_rasterizer_backward",
    "guava_rasterizer_mark_visible",
    "guava_rasterizer_forward_batch",
    "guava_rasterizer_backward_batch",
    "guava_rasterizer_backward_batch_summed",
    "guava_pack_view_gradients",
    "guava_image_dot",
    "guava_selftest_cull",
    "guava_deform_gaussians_forward",
    "guava_deform_gaussians_backward",
    "guava_deform_split_gradients",
    "guava_deform_gaussians_forward_static",
    "guava_buffer_layout",
    "guava_buffer_layout_of_call",
    "guava_set_option",
    "guava_get_option",
    "guava_launch_count",
    "guava_stage_time",
    "guava_stage_time_reset",
    "guava_last_error",
    "guava_version",
)

GUAVA_OK = 0
GUAVA_ERR_INVALID_ARGUMENT = 1
GUAVA_ERR_CUDA = 2
GUAVA_ERR_ALLOC = 3
GUAVA_ERR_UNSUPPORTED = 4


class Layout(Structure):
    _fields_ = [(n, c_size_t) for n in (
        "geom_bytes", "splat_rec", "depths", "rects", "tiles_touched", "grad_geo",
        "image_bytes", "final_T", "n_contrib", "ranges", "tile_cursor", "tile_order", "num_rendered",
        "binning_bytes", "point_list", "keys_unsorted", "keys_sorted",
        "point_offsets", "os_vals_alt", "os_scratch", "taken_marks")]


class CallOptions(Structure):
    """guava_call_options (include/guava_rasterizer.h): per-call options of the batched entry."""
    _fields_ = [("struct_size", c_size_t), ("refiner_output_nhwc_bf16", c_void_p), ("status_words", c_void_p),
                ("no_host_wait", c_int), ("forward_only", c_int), ("zero_fill", c_void_p), ("zero_fill_bytes", c_size_t)]

    def __init__(self, refiner_output=0, status_words=0, no_host_wait=False, forward_only=False, zero_fill=0,
                 zero_fill_bytes=0):
        super().__init__(ctypes.sizeof(CallOptions), refiner_output or None, status_words or None, int(bool(no_host_wait)),
                         int(bool(forward_only)), zero_fill or None, int(zero_fill_bytes))


class GuavaLibraryError(RuntimeError):
    pass


_lib = None


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(SO_PATH):
        raise GuavaLibraryError(
            f"{SO_PATH} is missing: the CUDA library has not been built "
            "(run `python guava_renderer_b200/build.py`). There is no CPU fallback.")
    lib = ctypes.CDLL(SO_PATH)
    for name in EXPORTS:
        if not hasattr(lib, name):
            raise GuavaLibraryError(f"{SO_PATH} does not export {name}")
    P = c_void_p
    lib.guava_rasterizer_forward.restype = c_int
    lib.guava_rasterizer_forward.argtypes = [
        ALLOC_FN, P, ALLOC_FN, P, ALLOC_FN, P,
        c_int, c_int, c_int, P, c_int, c_int,
        P, P, P, P, P, c_float, P, P, P, P, P,
        c_float, c_float, c_int, P, P, c_int, P, c_int, P, POINTER(c_int)]
    lib.guava_rasterizer_backward.restype = c_int
    lib.guava_rasterizer_backward.argtypes = [
        c_int, c_int, c_int, c_int, P, c_int, c_int,
        P, P, P, P, P, c_float, P, P, P, P, P, c_float, c_float,
        P, P, P, P, P, P,
        P, P, P, P, P, P, P, P, P, P,
        c_int, c_int, P]
    lib.guava_rasterizer_mark_visible.restype = c_int
    lib.guava_rasterizer_mark_visible.argtypes = [c_int, P, P, P, P, P]
    lib.guava_rasterizer_forward_batch.restype = c_int
    lib.guava_rasterizer_forward_batch.argtypes = [
        ALLOC_FN, P, ALLOC_FN, P, ALLOC_FN, P,
        c_int, c_int, P, c_int, c_int,
        P, c_size_t, P, c_size_t, P, c_size_t, P, c_size_t, c_float, P, c_size_t,
        P, P, P, P, P, c_int, P, c_int, P, POINTER(c_int), POINTER(CallOptions)]
    lib.guava_rasterizer_backward_batch.restype = c_int
    lib.guava_rasterizer_backward_batch.argtypes = [
        c_int, c_int, c_int, P, c_int, c_int,
        P, c_size_t, P, c_size_t, P, c_size_t, P, c_size_t, c_float, P, c_size_t,
        P, P, P, P, P, P, P, P, P,
        P, P, P, P, P, P, P, P, P,
        c_int, c_int, P]
    lib.guava_rasterizer_backward_batch_summed.restype = c_int
    lib.guava_rasterizer_backward_batch_summed.argtypes = [
        c_int, c_int, c_int, P, c_int, c_int,
        P, c_size_t, P, c_size_t, P, c_size_t, P, c_size_t, c_float, P, c_size_t,
        P, P, P, P, P, P, P,
        P, P, P, P, P, P, P, c_size_t,
        c_int, c_int, P]
    lib.guava_pack_view_gradients.restype = c_int
    lib.guava_pack_view_gradients.argtypes = [c_int, c_int, P, P, P, P, P, P, P]
    lib.guava_image_dot.restype = c_int
    lib.guava_image_dot.argtypes = [P, P, c_size_t, P, P]
    lib.guava_deform_gaussians_forward.restype = c_int
    lib.guava_deform_gaussians_forward.argtypes = [c_int] * 4 + [P] * 21
    lib.guava_deform_gaussians_forward_static.restype = c_int
    lib.guava_deform_gaussians_forward_static.argtypes = [c_int] * 4 + [P] * 20
    lib.guava_deform_gaussians_backward.restype = c_int
    lib.guava_deform_gaussians_backward.argtypes = [c_int] * 4 + [P] * 12
    lib.guava_deform_split_gradients.restype = c_int
    lib.guava_deform_split_gradients.argtypes = [c_int] * 3 + [P] * 9
    lib.guava_buffer_layout.restype = c_int
    lib.guava_buffer_layout.argtypes = [c_int, c_int, c_int, c_int, c_size_t, POINTER(Layout)]
    lib.guava_selftest_cull.restype = c_int
    lib.guava_selftest_cull.argtypes = [c_int, P, P, P, P, P]
    lib.guava_buffer_layout_of_call.restype = c_int
    lib.guava_buffer_layout_of_call.argtypes = [P, c_int, c_int, c_int, c_int, P, POINTER(Layout), POINTER(c_int),
                                                POINTER(c_int), POINTER(c_size_t)]
    lib.guava_set_option.restype = c_int
    lib.guava_set_option.argtypes = [c_char_p, c_longlong]
    lib.guava_get_option.restype = c_longlong
    lib.guava_get_option.argtypes = [c_char_p]
    lib.guava_launch_count.restype = c_ulonglong
    lib.guava_stage_time.restype = c_int
    lib.guava_stage_time.argtypes = [c_char_p, POINTER(ctypes.c_double), POINTER(c_ulonglong)]
    lib.guava_stage_time_reset.restype = None
    lib.guava_last_error.restype = c_char_p
    lib.guava_version.restype = c_char_p
    _lib = lib
    # development switches: GUAVA_OPT_<name>=<int> presets an option at load time
    for k, v in os.environ.items():
        if k.startswith("GUAVA_OPT_"):
            lib.guava_set_option(k[len("GUAVA_OPT_"):].lower().encode(), int(v))
    return lib


def last_error() -> str:
    return load().guava_last_error().decode("utf-8", "replace")


def set_option(name: str, value: int) -> None:
    rc = load().guava_set_option(name.encode(), int(value))
    if rc != GUAVA_OK:
        raise ValueError(f"unknown option {name!r}")


def get_option(name: str) -> int:
    return int(load().guava_get_option(name.encode()))


def launch_count() -> int:
    return int(load().guava_launch_count())


STAGES = ("preprocess", "tile_scan", "scatter", "tile_sort", "render_forward", "render_backward",
          "preprocess_backward")


def stage_times() -> dict:
    """{stage: (total_ms, calls)} accumulated while option "profile_stages" is on."""
    lib = load()
    out = {}
    for s in STAGES:
        ms, n = ctypes.c_double(0), c_ulonglong(0)
        lib.guava_stage_time(s.encode(), ctypes.byref(ms), ctypes.byref(n))
        out[s] = (ms.value, int(n.value))
    return out


def stage_times_reset() -> None:
    load().guava_stage_time_reset()


def version() -> str:
    return load().guava_version().decode()


def buffer_layout_of_call(image_buffer_ptr: int, V: int, P: int, W: int, H: int, stream: int = 0):
    """(Layout, sort_mode, block_masks, capacity) of a finished forward call, read from its image buffer's header."""
    L, mode, masks, cap = Layout(), c_int(0), c_int(0), c_size_t(0)
    rc = load().guava_buffer_layout_of_call(image_buffer_ptr, V, P, W, H, stream, ctypes.byref(L), ctypes.byref(mode),
                                            ctypes.byref(masks), ctypes.byref(cap))
    if rc != GUAVA_OK:
        raise ValueError(last_error())
    return L, int(mode.value), int(masks.value), int(cap.value)


def buffer_layout(V: int, P: int, W: int, H: int, capacity: int = 0) -> Layout:
    L = Layout()
    rc = load().guava_buffer_layout(V, P, W, H, capacity, ctypes.byref(L))
    if rc != GUAVA_OK:
        raise ValueError(last_error())
    return L
