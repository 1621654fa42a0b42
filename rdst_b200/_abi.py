"""ctypes binding of the C ABI in include/b200sort.h (libb200sort.so, built in tree).

This is the Python stand-in for the `unsafe extern "C"` block a Rust binding would carry
(see INTEGRATION.md).  There is no fallback: if the CUDA library is missing or a call
fails, an exception is raised — mirroring the reference, whose API is infallible and
panics on a broken invariant (reference src/radix_sort_builder.rs:22).
"""
from __future__ import annotations

import ctypes as C
import os

B200_MAX_LEVELS = 64

KEY_UINT, KEY_SINT, KEY_FLOAT, KEY_BYTES = 0, 1, 2, 3

ERR_NAMES = {
    0: "B200_OK",
    1: "B200_ERR_LEVELS_ZERO",
    2: "B200_ERR_BAD_DESC",
    3: "B200_ERR_CUDA",
    4: "B200_ERR_NO_DEVICE",
    5: "B200_ERR_ALLOC",
    6: "B200_ERR_BAD_ARG",
    7: "B200_ERR_LEVEL_RANGE",
    8: "B200_ERR_NCCL",
    9: "B200_ERR_TOO_LARGE",
}


class Level(C.Structure):
    _fields_ = [("byte_offset", C.c_uint16), ("xor_mask", C.c_uint8), ("_pad", C.c_uint8)]


class KeyDesc(C.Structure):
    _fields_ = [
        ("elem_size", C.c_uint32),
        ("elem_align", C.c_uint32),
        ("levels", C.c_uint32),
        ("kind", C.c_uint32),
        ("key_offset", C.c_uint32),
        ("key_width", C.c_uint32),
        ("level", Level * B200_MAX_LEVELS),
    ]


class SortOpts(C.Structure):
    _fields_ = [
        ("stable", C.c_uint32),
        ("variant", C.c_int32),
        ("device", C.c_int32),
        ("stream", C.c_void_p),
        ("route", C.c_uint32),
        ("_reserved", C.c_uint32),
    ]


HINT_AUTO, HINT_LSD, HINT_LOW_SCRATCH = 0, 1, 2  # b200_route_hint
STATIC_MAX_RECORDS, STATIC_MAX_BYTES = 2 << 20, 16 << 20  # b200_sort_device_static (include/b200sort.h)
B200_MAX_GPUS = 16


class Stats(C.Structure):
    _fields_ = [
        ("levels_run", C.c_uint32),
        ("levels_skipped", C.c_uint32),
        ("variant", C.c_uint32),
        ("copy_back", C.c_uint32),
        ("algorithmic_bytes", C.c_uint64),
        ("ms_hist", C.c_float),
        ("ms_passes", C.c_float),
        ("ms_h2d", C.c_float),
        ("ms_d2h", C.c_float),
        ("ms_total", C.c_float),
        ("tiles_per_pass", C.c_uint64),
        ("already_sorted", C.c_uint32),
        ("kernel_launches", C.c_uint32),
        ("hybrid", C.c_uint32),
        ("slow_segments", C.c_uint32),
        ("local_levels", C.c_uint32),
        ("ms_joint", C.c_float),
        ("ms_local", C.c_float),
        ("unstable", C.c_uint32),
        ("heavy_segments", C.c_uint32),
        ("heavy_rounds", C.c_uint32),
    ]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class MultiStats(C.Structure):
    """b200_multi_stats (include/b200sort.h)."""

    _fields_ = [
        ("ngpus", C.c_uint32),
        ("level", C.c_uint32),
        ("splitter_mode", C.c_uint32),
        ("kernel_launches", C.c_uint32),
        ("imbalance", C.c_double),
        ("max_sent_bytes", C.c_uint64),
        ("n_out", C.c_uint64 * B200_MAX_GPUS),
        ("ms_plan", C.c_float),
        ("ms_exchange", C.c_float),
        ("ms_local_sort", C.c_float),
        ("ms_h2d", C.c_float),
        ("ms_d2h", C.c_float),
        ("ms_total", C.c_float),
        ("local0", Stats),
    ]

    def as_dict(self):
        d = {n: getattr(self, n) for n, _ in self._fields_ if n not in ("n_out", "local0")}
        d["n_out"] = [int(x) for x in self.n_out[: self.ngpus]]
        d["local0"] = self.local0.as_dict()
        return d


class B200SortError(RuntimeError):
    def __init__(self, code, detail=""):
        self.code = code
        name = ERR_NAMES.get(code, str(code))
        super().__init__(f"libb200sort: {name}" + (f" ({detail})" if detail else ""))


class LevelsZeroError(B200SortError, AssertionError):
    """RadixKey::LEVELS == 0 — the reference panics with
    "RadixKey must have at least 1 level" (reference src/radix_sort_builder.rs:22)."""


LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libb200sort.so")

_lib = None

# name -> (restype, argtypes): every symbol include/b200sort.h declares
class Route(C.Structure):
    """b200_route (include/b200sort.h): the route the planner starts on."""

    _fields_ = [
        ("route", C.c_uint32),
        ("active_levels", C.c_uint32),
        ("levels_run", C.c_uint32),
        ("msd_level", C.c_uint32),
        ("max_bucket", C.c_uint64),
    ]


ROUTE_NAMES = {0: "lsd", 1: "hybrid", 2: "msd1", 3: "small", 4: "trivial"}


class WindowPlan(C.Structure):
    """b200_window_plan (include/b200sort.h): the digits the hybrid route would sort on."""

    _fields_ = [
        ("eligible", C.c_uint32),
        ("aligned", C.c_uint32),
        ("vary", C.c_uint64),
        ("digit_lo", C.c_uint32 * 2),
        ("digit_width", C.c_uint32 * 2),
        ("nlocal", C.c_uint32),
        ("local_lo", C.c_uint32 * 8),
        ("local_width", C.c_uint32 * 8),
    ]


SYMBOLS = {
    "b200_sort_host": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(KeyDesc), C.POINTER(SortOpts), C.POINTER(Stats)]),
    "b200_sort_device": (
        C.c_int,
        [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(KeyDesc), C.POINTER(SortOpts), C.POINTER(Stats)],
    ),
    "b200_sort_device_hinted": (
        C.c_int,
        [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(KeyDesc), C.c_uint64, C.POINTER(SortOpts), C.POINTER(Stats)],
    ),
    "b200_sort_device_presplit": (
        C.c_int,
        [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(KeyDesc), C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(SortOpts),
         C.POINTER(Stats), C.POINTER(C.c_void_p)],
    ),
    "b200_sort_segments_device": (
        C.c_int,
        [C.c_void_p, C.c_size_t, C.POINTER(C.c_uint64), C.c_size_t, C.POINTER(KeyDesc), C.POINTER(SortOpts), C.POINTER(Stats)],
    ),
    "b200_sort_device_static": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(KeyDesc), C.POINTER(SortOpts)]),
    "b200_histogram_device": (
        C.c_int,
        [C.c_void_p, C.c_size_t, C.POINTER(KeyDesc), C.POINTER(SortOpts), C.POINTER(C.c_uint64)],
    ),
    "b200_pass_device": (
        C.c_int,
        [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(KeyDesc), C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(SortOpts)],
    ),
    "b200_partition_device": (
        C.c_int,
        [
            C.c_void_p,
            C.c_void_p,
            C.c_size_t,
            C.POINTER(KeyDesc),
            C.c_uint32,
            C.POINTER(C.c_uint64),
            C.POINTER(C.c_uint32),
            C.c_uint32,
            C.POINTER(C.c_uint64),
            C.POINTER(SortOpts),
        ],
    ),
    "b200_partition_peers_device": (
        C.c_int,
        [
            C.c_void_p,
            C.POINTER(C.c_void_p),
            C.c_size_t,
            C.POINTER(KeyDesc),
            C.c_uint32,
            C.POINTER(C.c_uint64),
            C.POINTER(C.c_uint32),
            C.c_uint32,
            C.POINTER(C.c_uint64),
            C.POINTER(SortOpts),
        ],
    ),
    "b200_key_prefix_supported": (C.c_int, [C.POINTER(KeyDesc)]),
    "b200_key_prefix_sample_device": (
        C.c_int,
        [C.c_void_p, C.c_size_t, C.POINTER(KeyDesc), C.c_size_t, C.c_size_t, C.POINTER(C.c_uint64), C.POINTER(SortOpts)],
    ),
    "b200_splitter_count_device": (
        C.c_int,
        [C.c_void_p, C.c_size_t, C.POINTER(KeyDesc), C.POINTER(C.c_uint64), C.c_uint32, C.POINTER(C.c_uint64),
         C.POINTER(C.c_uint64), C.POINTER(SortOpts)],
    ),
    "b200_partition_splitters_device": (
        C.c_int,
        [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.c_size_t, C.POINTER(KeyDesc), C.POINTER(C.c_uint64),
         C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(SortOpts)],
    ),
    "b200_ipc_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t, C.c_void_p]),
    "b200_ipc_open": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "b200_ipc_close": (C.c_int, [C.c_void_p]),
    "b200_ipc_free": (C.c_int, [C.c_void_p]),
    "b200_sort_multi_device": (
        C.c_int,
        [C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.c_int, C.POINTER(C.c_int), C.POINTER(KeyDesc), C.POINTER(SortOpts),
         C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.POINTER(MultiStats)],
    ),
    "b200_sort_multi_host": (
        C.c_int,
        [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_int), C.POINTER(KeyDesc), C.POINTER(SortOpts), C.POINTER(MultiStats)],
    ),
    "b200_multi_release": (C.c_int, []),
    "b200_multi_plan_digits": (
        C.c_int,
        [C.POINTER(C.c_uint64), C.c_uint32, C.c_uint32, C.POINTER(C.c_int32), C.POINTER(C.c_uint32), C.POINTER(C.c_int64),
         C.POINTER(C.c_double)],
    ),
    "b200_multi_choose_sample_splitters": (
        C.c_int,
        [C.POINTER(C.c_uint64), C.POINTER(C.c_double), C.c_size_t, C.c_uint32, C.POINTER(C.c_uint64)],
    ),
    "b200_multi_plan_splitters": (
        C.c_int,
        [C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_uint64), C.c_uint32, C.c_uint32,
         C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_uint32)],
    ),
    "b200_copy_h2d": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]),
    "b200_copy_d2h": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]),
    "b200_get_level": (C.c_int, [C.c_void_p, C.POINTER(KeyDesc), C.c_uint32]),
    "b200_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "b200_host_free": (C.c_int, [C.c_void_p]),
    "b200_release_scratch": (C.c_int, [C.c_int]),
    "b200_hybrid_plan": (C.c_int, [C.POINTER(KeyDesc), C.POINTER(C.c_uint64), C.c_size_t, C.POINTER(WindowPlan)]),
    "b200_route_plan": (C.c_int, [C.POINTER(KeyDesc), C.POINTER(C.c_uint64), C.c_size_t, C.POINTER(Route)]),
    "b200_variant_count": (C.c_int, []),
    "b200_variant_name": (C.c_char_p, [C.c_int]),
    "b200_strerror": (C.c_char_p, [C.c_int]),
    "b200_last_cuda_error": (C.c_char_p, []),
    "b200_version": (C.c_char_p, []),
}


def lib():
    """Load libb200sort.so (once).  Raises if it has not been built — there is no CPU path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C rdst_b200/csrc`. rdst_b200 has no CPU fallback."
            )
        l = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)  # AttributeError if the library does not export it
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(code):
    if code == 0:
        return
    detail = ""
    if code in (3, 9):
        detail = (lib().b200_last_cuda_error() or b"").decode()
    if code == 1:
        raise LevelsZeroError(code, "RadixKey must have at least 1 level")
    raise B200SortError(code, detail)


CUDA_STREAM_LEGACY = 0x1  # cudaStreamLegacy


def torch_stream_handle(torch):
    """cudaStream_t of torch's current stream.  torch reports its default stream as 0, which
    in b200_sort_opts means "the library's own stream"; the legacy default stream has the
    explicit handle cudaStreamLegacy, so the sort is ordered after work torch already queued."""
    h = torch.cuda.current_stream().cuda_stream
    return h if h else CUDA_STREAM_LEGACY
