"""b200_sort_device_presplit: the local sort of the sharded sort when the exchange has already split the records on
one level (rdst_b200/dist.py `presplit`): K0, K1 and the first MSD pass are skipped.  Here the array is partitioned with
the library's own pass (b200_pass_device), the way a rank's receive buffer arrives, and the result must be the oracle's."""
import ctypes as C

import numpy as np
import pytest

import oracle_ffi as o
from rdst_b200 import RadixKey, _abi

pytestmark = pytest.mark.gpu


def presplit_sort(a, key, level, keep=None, wrong_counts=False):
    """partition `a` on `level` (stable pass), then the presplit sort; returns (bytes, stats, result_is_tmp)"""
    import torch

    l = _abi.lib()
    d = key.to_desc()
    es = key.elem_size
    n = a.nbytes // es
    hist = np.zeros((key.levels, 256), dtype=np.uint64)
    src = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()
    dst = torch.empty_like(src)
    op = _abi.SortOpts()
    op.variant, op.device, op.route = -1, -1, 0
    op.stream = _abi.torch_stream_handle(torch)
    _abi.check(l.b200_histogram_device(src.data_ptr(), n, C.byref(d), C.byref(op), hist.ctypes.data_as(C.POINTER(C.c_uint64))))
    row = np.ascontiguousarray(hist[level])
    _abi.check(l.b200_pass_device(src.data_ptr(), dst.data_ptr(), n, C.byref(d), level, row.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(op)))
    if wrong_counts:
        row = row.copy()
        i, j = np.argsort(row)[-2:]
        row[i] += 1
        row[j] -= 1
    st = _abi.Stats()
    res = C.c_void_p()
    rc = l.b200_sort_device_presplit(dst.data_ptr(), src.data_ptr(), n, C.byref(d), level, row.ctypes.data_as(C.POINTER(C.c_uint64)),
                                     C.byref(op), C.byref(st), C.byref(res))
    torch.cuda.synchronize()
    if rc:
        return rc, None, None
    out = src if res.value == src.data_ptr() else dst
    assert res.value in (src.data_ptr(), dst.data_ptr())
    return out.cpu().numpy().tobytes(), st.as_dict(), res.value == src.data_ptr()


@pytest.mark.parametrize("dt", ["u8", "i8", "f8", "u4", "f4"])
def test_presplit_matches_the_oracle(dt):
    rng = np.random.default_rng(808 + np.dtype(dt).num)
    key = RadixKey.for_dtype(dt)
    n = 3_000_017
    a = rng.integers(0, 256, n * key.elem_size, dtype=np.uint8).view(dt).copy()
    got, st, in_tmp = presplit_sort(a, key, key.levels - 1)
    assert got == o.sorted_copy(a, key).tobytes(), dt
    assert st["hybrid"] == 1 and st["unstable"] & 1 and in_tmp, st     # pass 2 went dev_ptr -> dev_tmp, K3-U sorted there


def test_presplit_one_ranks_key_range_and_narrow_keys():
    rng = np.random.default_rng(5)
    key = RadixKey.for_dtype("u8")
    n = 2_500_000
    # a rank's range of an 8-way sharded sort: 32 of the 256 top digits occur
    a = rng.integers(3 << 61, 4 << 61, n, dtype="u8")
    got, st, _ = presplit_sort(a, key, 7)
    assert got == np.sort(a).tobytes() and st["hybrid"] == 1
    # levels 1..5 constant: only the two MSD levels and the lowest byte vary — every segment finds its own top bit
    b = (rng.integers(0, 1 << 16, n, dtype="u8") << np.uint64(48)) | rng.integers(0, 256, n, dtype="u8")
    got, st, _ = presplit_sort(b, key, 7)
    assert got == np.sort(b).tobytes() and st["hybrid"] == 1 and st["slow_segments"] == 0, st
    # level 6 constant too: the 256 buckets are the segments, far too large for an SM -> the general sort finishes
    c = (rng.integers(0, 256, n, dtype="u8") << np.uint64(56)) | rng.integers(0, 1 << 40, n, dtype="u8")
    got, st, _ = presplit_sort(c, key, 7)
    assert got == np.sort(c).tobytes()
    # heavily duplicated keys
    e = rng.integers(0, 2**64, n, dtype="u8")
    e[rng.random(n) < 0.3] = np.uint64(0xDEADBEEF12345678)
    got, st, _ = presplit_sort(e, key, 7)
    assert got == np.sort(e).tobytes()


def test_presplit_falls_back_for_other_shapes_and_rejects_wrong_counts():
    rng = np.random.default_rng(6)
    rec = np.dtype([("key", "<u4"), ("idx", "<u4")])
    key = RadixKey.field(rec, "key", 4)
    a = np.zeros(400_000, dtype=rec)
    a["key"] = rng.integers(0, 1 << 30, len(a))
    a["idx"] = np.arange(len(a))
    got, st, _ = presplit_sort(a, key, 3)                  # records with a payload: the general (stable) sort
    assert got == o.stable_sorted(a, key).tobytes()
    k8 = RadixKey.for_dtype("u8")
    b = rng.integers(0, 1 << 16, 1_000_000, dtype="u8")
    got, st, _ = presplit_sort(b, k8, 1)                   # level < 2: no room for a second digit and a segment sort
    assert got == np.sort(b).tobytes()
    b = rng.integers(0, 2**64, 1_000_000, dtype="u8")
    rc, _, _ = presplit_sort(b, k8, 7, wrong_counts=True)  # counts that do not describe the array
    assert rc == 3
