"""rdst_b200 — a B200-native (sm_100a) radix sort engine behind nessex/rdst's API surface.

Only the sort path lives here: the CUDA kernels + C ABI (csrc/, built to libb200sort.so)
and the host-side mirror of the reference interface (RadixKey, RadixSort entry points,
RadixSortBuilder, tuner).  There is no CPU fallback: importing works without a GPU, but
every sort call goes to the CUDA library and raises if it is missing or fails.
"""
from . import tuner
from ._abi import B200SortError, LevelsZeroError
from .radix_key import RadixKey
from .radix_sort import RadixSortBuilder, radix_sort_builder, radix_sort_stable, radix_sort_unstable, sort_multi_host, sort_segments

__all__ = [
    "RadixKey", "RadixSortBuilder", "radix_sort_builder", "radix_sort_unstable", "radix_sort_stable", "sort_multi_host", "sort_segments",
    "tuner", "B200SortError", "LevelsZeroError",
]
