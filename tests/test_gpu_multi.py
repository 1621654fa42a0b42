"""b200_sort_multi_host / b200_sort_multi_device — the sharded sort behind the C ABI, ONE process driving
several GPUs (peer access, fused partition + exchange, local sorts) — against the oracle's stable order on
the same inputs as the one-process-per-GPU check (tests/dist_check.py).  With a single device the entry points
still run (one shard, one part): that is what a 1-GPU test box exercises; with more they use every device and
every power of two below."""
import ctypes as C

import numpy as np
import pytest

import oracle_ffi as o
from dist_check import cases
from rdst_b200 import RadixKey, _abi, sort_multi_host

pytestmark = pytest.mark.gpu


def gpu_counts():
    import torch

    n = torch.cuda.device_count()
    return sorted({1, n} | {p for p in (2, 4, 8) if p <= n})


def test_multi_host_matches_the_oracle_on_every_case():
    for P in gpu_counts():
        for name, key, full in cases():
            exp = o.stable_sorted(full, key).tobytes()
            a = full.copy()
            st = sort_multi_host(a, key, gpus=list(range(P)))
            assert a.tobytes() == exp, (P, name, st)
            assert st["ngpus"] == P and sum(st["n_out"]) == len(full) // key.elem_size or st["level"] == 0xFFFFFFFF
            if name in ("zipf", "two_values") and P >= 2:
                assert st["splitter_mode"] == 1, (name, st)  # one digit outweighs a rank's share: sample-sort splitters
            if name == "all_equal":
                assert st["level"] == 0xFFFFFFFF  # nothing to partition on: the shards are the result


def test_multi_device_entry_returns_sorted_key_ranges():
    import torch

    lib = _abi.lib()
    rng = np.random.default_rng(99)
    for P in gpu_counts():
        for key, full in ((RadixKey.for_dtype("u8"), rng.integers(0, 2**64, 2_000_003, dtype="u8").view(np.uint8)),
                          (RadixKey.for_dtype("i4"), rng.integers(-2**31, 2**31 - 1, 1_000_001, dtype="i4").view(np.uint8)),
                          (RadixKey.for_dtype("u8"), np.zeros(8 * 5, dtype=np.uint8))):  # fewer records than GPUs is fine
            es = key.elem_size
            n = len(full) // es
            bounds = [n * g // P for g in range(P + 1)]
            shards = [torch.from_numpy(full[bounds[g] * es: bounds[g + 1] * es].copy()).to(f"cuda:{g}") for g in range(P)]
            ptrs = (C.c_void_p * P)(*[C.c_void_p(s.data_ptr() if s.numel() else 0) for s in shards])
            n_in = (C.c_size_t * P)(*[bounds[g + 1] - bounds[g] for g in range(P)])
            outp = (C.c_void_p * P)()
            n_out = (C.c_size_t * P)()
            devs = (C.c_int * P)(*range(P))
            st = _abi.MultiStats()
            d = key.to_desc()
            _abi.check(lib.b200_sort_multi_device(ptrs, n_in, P, devs, C.byref(d), None, outp, n_out, C.byref(st)))
            got = b""
            for g in range(P):
                if n_out[g]:
                    h = np.empty(n_out[g] * es, dtype=np.uint8)
                    _abi.check(lib.b200_copy_d2h(h.ctypes.data, outp[g], h.nbytes, g))
                    got += h.tobytes()
            assert got == o.stable_sorted(full, key).tobytes(), (P, key)
    assert lib.b200_multi_release() == 0
