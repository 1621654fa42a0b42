"""RadixSort / RadixSortBuilder — the reference's entry points over the C ABI.

Reference surface mirrored here (same names, argument meaning, error behaviour):
  RadixSort::radix_sort_unstable / radix_sort_builder   reference src/radix_sort.rs:4-45
  RadixSortBuilder::{with_parallel, with_low_mem_tuner, with_single_threaded_tuner,
                     with_tuner, sort}                  reference src/radix_sort_builder.rs:8-158
The sort is IN PLACE on the caller's buffer and returns None, like `&mut self` in the
reference.  Host buffers (numpy) go through b200_sort_host, device buffers (torch CUDA
tensors) through b200_sort_device on torch's current stream.  No CPU path exists.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi
from .radix_key import RadixKey, key_for_array
from .tuner import (ALGORITHM_ROUTE, HINT_LOW_SCRATCH, STABLE_ALGORITHMS, LowMemoryTuner, SingleThreadedTuner,
                    StandardTuner, Tuner, TuningParams)

_STANDARD = StandardTuner()
_LOW_MEM = LowMemoryTuner()
_SINGLE = SingleThreadedTuner()


def _stream_handle(torch):
    return _abi.torch_stream_handle(torch)


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _buffer_info(data, key):
    """-> (ptr, n_records, is_device, device_index, keepalive)"""
    if _is_torch(data):
        import torch

        if not data.is_contiguous():
            raise ValueError("tensor must be contiguous (the sort is in place)")
        nbytes = data.numel() * data.element_size()
        if nbytes % key.elem_size:
            raise ValueError("tensor byte size is not a multiple of the record size")
        n = nbytes // key.elem_size
        if data.is_cuda:
            return data.data_ptr(), n, True, data.device.index, data
        return data.data_ptr(), n, False, -1, data
    arr = data
    if not isinstance(arr, np.ndarray):
        raise TypeError("radix sort needs a numpy array or a torch tensor (in-place sort)")
    if not arr.flags.c_contiguous or not arr.flags.writeable:
        raise ValueError("array must be C-contiguous and writeable (the sort is in place)")
    if arr.nbytes % key.elem_size:
        raise ValueError("array byte size is not a multiple of the record size")
    return arr.ctypes.data, arr.nbytes // key.elem_size, False, -1, arr


class RadixSortBuilder:
    """reference src/radix_sort_builder.rs:8-12: (data, multi_threaded, tuner)."""

    def __init__(self, data, key: RadixKey | None = None):
        self._data = data
        self._key = key_for_array(data, key)
        # assert_ne!(T::LEVELS, 0, "RadixKey must have at least 1 level")  (:22) — raised at
        # construction, even for empty input, exactly like the reference
        if self._key.levels == 0:
            raise _abi.LevelsZeroError(1, "RadixKey must have at least 1 level")
        self._multi_threaded = True
        self._tuner: Tuner = _STANDARD
        self._variant = -1
        self._stable = False
        self.stats = None

    # ---- reference builder methods ----
    def with_parallel(self, parallel: bool):
        """:53-60.  Kept for source compatibility: the device always runs every CTA in parallel."""
        self._multi_threaded = bool(parallel)
        return self

    def with_low_mem_tuner(self):  # :86-90
        """LowMemoryTuner answers Ska / Regions for anything large (reference low_memory_tuner.rs:13-43): the
        engine's low-scratch route — LSD passes with the small-tile variant, scratch pool released after the
        call.  Peak scratch is still one copy of the array (there is no in-place pass on the device)."""
        self._tuner = _LOW_MEM
        return self

    def with_single_threaded_tuner(self):  # :106-110
        self._tuner = _SINGLE
        return self

    def with_tuner(self, tuner: Tuner):  # :128-132
        self._tuner = tuner
        return self

    # ---- engine-specific additions ----
    def with_variant(self, variant: int):
        """Force a tile / occupancy variant of the pass kernel (b200_variant_count())."""
        self._variant = int(variant)
        return self

    def with_stable(self, stable: bool = True):
        """Ask for rdst's stable order (what Lsb / LrLsb / MtLsb produce).  Records with a payload are sorted
        stably in every mode; records that are nothing but their key take the unstable kernels, whose
        result is bit-identical (equal keys are identical records)."""
        self._stable = bool(stable)
        return self

    def sort(self):  # :149-157
        ptr, n, on_device, dev, keep = _buffer_info(self._data, self._key)
        if n <= 1:  # "By definition, this is already sorted" (:151)
            return None
        lib = _abi.lib()
        desc = self._key.to_desc()
        opts = _abi.SortOpts()
        opts.stable = 1 if self._stable else 0
        opts.variant = self._variant
        opts.device = dev
        opts.stream = None
        stats = _abi.Stats()
        custom = self._tuner not in (_STANDARD, _LOW_MEM, _SINGLE)
        if custom:
            return self._sort_with_custom_tuner(lib, desc, opts, stats, n, on_device)
        opts.route = 0
        if self._tuner is _LOW_MEM and self._variant < 0:
            opts.route = HINT_LOW_SCRATCH
        if on_device:
            import torch

            with torch.cuda.device(dev):
                opts.stream = _stream_handle(torch)
                _abi.check(lib.b200_sort_device(ptr, None, n, C.byref(desc), C.byref(opts), C.byref(stats)))
        else:
            _abi.check(lib.b200_sort_host(ptr, n, C.byref(desc), C.byref(opts), C.byref(stats)))
        self.stats = stats.as_dict()
        return None

    def _sort_with_custom_tuner(self, lib, desc, opts, stats, n, on_device):
        """A user Tuner sees TuningParams for the top level and that level's counts
        (reference src/sorter.rs:67-76) — here taken from the K1 histogram."""
        import torch

        if on_device:
            t = self._data
        else:
            host = torch.from_numpy(self._data.view(np.uint8).reshape(-1)) if isinstance(self._data, np.ndarray) \
                else self._data.view(torch.uint8).reshape(-1)
            t = host.cuda()
        dev = t.device.index
        opts.device = dev
        with torch.cuda.device(dev):
            opts.stream = _stream_handle(torch)
            L = self._key.levels
            hist = (C.c_uint64 * (256 * L))()
            _abi.check(lib.b200_histogram_device(t.data_ptr(), n, C.byref(desc), C.byref(opts), hist))
            top = [hist[(L - 1) * 256 + b] for b in range(256)]
            p = TuningParams(threads=torch.cuda.get_device_properties(dev).multi_processor_count, level=L - 1,
                             total_levels=L, input_len=n, parent_len=None)
            algo = self._tuner.pick_algorithm(p, top)
            opts.route = ALGORITHM_ROUTE.get(algo, 0) if self._variant < 0 else 0
            if algo in STABLE_ALGORITHMS:
                opts.stable = 1
            _abi.check(lib.b200_sort_device(t.data_ptr(), None, n, C.byref(desc), C.byref(opts), C.byref(stats)))
            if not on_device:
                host.copy_(t)
        self.stats = stats.as_dict()
        self.stats["algorithm"] = algo.name
        self.stats["route_hint"] = int(opts.route)
        return None


def sort_multi_host(data, key: RadixKey | None = None, gpus=None):
    """Sort a host array over several GPUs of this box in ONE process (b200_sort_multi_host): the array is cut
    into len(gpus) shards, every GPU pulls its shard over its own PCIe link, the shards are range-partitioned
    over NVLink (fused partition + exchange) and sorted locally, and the sorted ranges are written back in
    place.  The multi-GPU form of `radix_sort_unstable` on a host `&mut [T]`; returns the b200_multi_stats dict."""
    key = key_for_array(data, key)
    if key.levels == 0:
        raise _abi.LevelsZeroError(1, "RadixKey must have at least 1 level")
    ptr, n, on_device, _, _keep = _buffer_info(data, key)
    if on_device:
        raise ValueError("sort_multi_host takes a host array")
    lib = _abi.lib()
    if gpus is None:
        import torch

        gpus = list(range(torch.cuda.device_count()))
    devs = (C.c_int * len(gpus))(*gpus)
    desc = key.to_desc()
    st = _abi.MultiStats()
    _abi.check(lib.b200_sort_multi_host(ptr, n, len(gpus), devs, C.byref(desc), None, C.byref(st)))
    return st.as_dict()


def sort_segments(data, seg_off, key: RadixKey | None = None):
    """Many independent sorts in one call (b200_sort_segments_device): every segment [seg_off[i], seg_off[i + 1]) of a
    torch CUDA tensor is sorted on its own, in place — what rdst's recursion does with the buckets of an MSB split
    (reference src/sorter.rs:121-157).  `seg_off`: ascending record offsets (host), len = segments + 1.
    Returns the b200_stats dict."""
    key = key_for_array(data, key)
    if key.levels == 0:
        raise _abi.LevelsZeroError(1, "RadixKey must have at least 1 level")
    ptr, n, on_device, dev, _keep = _buffer_info(data, key)
    if not on_device:
        raise ValueError("sort_segments takes a device tensor")
    off = np.ascontiguousarray(seg_off, dtype=np.uint64)
    if off.ndim != 1 or len(off) < 1:
        raise ValueError("seg_off: a 1-d array of segments + 1 offsets")
    import torch

    lib = _abi.lib()
    desc = key.to_desc()
    opts = _abi.SortOpts()
    opts.variant, opts.device, opts.route = -1, dev, 0
    stats = _abi.Stats()
    with torch.cuda.device(dev):
        opts.stream = _stream_handle(torch)
        _abi.check(lib.b200_sort_segments_device(ptr, n, off.ctypes.data_as(C.POINTER(C.c_uint64)), len(off) - 1,
                                                 C.byref(desc), C.byref(opts), C.byref(stats)))
    return stats.as_dict()


def radix_sort_builder(data, key: RadixKey | None = None) -> RadixSortBuilder:
    """RadixSort::radix_sort_builder (reference src/radix_sort.rs:17-18)."""
    return RadixSortBuilder(data, key)


def radix_sort_unstable(data, key: RadixKey | None = None) -> None:
    """RadixSort::radix_sort_unstable = radix_sort_builder().sort() (reference src/radix_sort.rs:25-27)."""
    RadixSortBuilder(data, key).sort()


def radix_sort_stable(data, key: RadixKey | None = None) -> None:
    """rdst's stable order (a tuner returning Lsb / LrLsb / MtLsb at the top level)."""
    RadixSortBuilder(data, key).with_stable(True).sort()
