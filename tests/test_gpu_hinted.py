"""b200_sort_device_hinted: the device sort for a caller that knows which key-image bits can vary (the sharded sort's
receivers, rdst_b200/dist.py) — no K0, no K1.  Same result as b200_sort_device for exact hints and for supersets; inputs
the MSD route does not take fall through to the general path."""
import ctypes as C

import numpy as np
import pytest

import oracle_ffi as o
from rdst_b200 import RadixKey, _abi

pytestmark = pytest.mark.gpu


def sort_hinted(a, key, vary):
    import torch

    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()
    d = key.to_desc()
    op = _abi.SortOpts()
    op.variant, op.device, op.route = -1, -1, 0
    op.stream = _abi.torch_stream_handle(torch)
    st = _abi.Stats()
    n = a.nbytes // key.elem_size
    _abi.check(_abi.lib().b200_sort_device_hinted(t.data_ptr(), None, n, C.byref(d), C.c_uint64(vary), C.byref(op), C.byref(st)))
    torch.cuda.synchronize()
    return t.cpu().numpy().tobytes(), st.as_dict()


def image_vary(a, key):
    km = o.key_matrix(a, key).astype(np.uint64)
    img = np.zeros(km.shape[0], dtype=np.uint64)
    for l in range(key.levels):
        img |= km[:, l] << np.uint64(8 * l)
    return int(np.bitwise_or.reduce(img) & ~np.bitwise_and.reduce(img))


@pytest.mark.parametrize("dt", ["u8", "i8", "f8", "u4", "f4"])
def test_hinted_route_matches_the_oracle(dt, monkeypatch):
    monkeypatch.setenv("B200SORT_HYBRID", "1")     # the MSD route whatever the size
    rng = np.random.default_rng(70 + np.dtype(dt).num)
    key = RadixKey.for_dtype(dt)
    n = 1_500_003
    a = rng.integers(0, 256, n * key.elem_size, dtype=np.uint8).view(dt).copy()
    want = o.sorted_copy(a, key).tobytes()
    full = (1 << (8 * key.levels)) - 1
    for vary in (image_vary(a, key), full):
        got, st = sort_hinted(a, key, vary)
        assert got == want, (dt, hex(vary))
        assert st["hybrid"] == 1 and st["unstable"] & 1, st
        assert st["ms_hist"] == 0.0                   # no counting pass ran


def test_hinted_on_a_ranks_key_range_and_narrow_keys(monkeypatch):
    monkeypatch.setenv("B200SORT_HYBRID", "1")
    rng = np.random.default_rng(12)
    key = RadixKey.for_dtype("u8")
    n = 2_000_000
    a = rng.integers(5 << 61, 6 << 61, n, dtype="u8")                    # one rank's range of an 8-way split
    got, st = sort_hinted(a, key, (1 << 61) - 1)
    assert got == np.sort(a).tobytes() and st["hybrid"] == 1
    b = (rng.integers(0, 1 << 20, n, dtype="u8") << np.uint64(30)) | np.uint64(0x155)   # bits 30..49 vary
    for vary in (((1 << 20) - 1) << 30, (1 << 64) - 1, (1 << 50) - 1):
        got, st = sort_hinted(b, key, vary)
        assert got == np.sort(b).tobytes(), hex(vary)
    c = rng.integers(0, 2**64, n, dtype="u8")
    c[rng.random(n) < 0.4] = np.uint64(77)                               # a heavy key: oversized segment, pivot splits
    got, st = sort_hinted(c, key, (1 << 64) - 1)
    assert got == np.sort(c).tobytes()


def test_hinted_falls_through_for_inputs_the_route_does_not_take():
    rng = np.random.default_rng(13)
    k8 = RadixKey.for_dtype("u8")
    for n in (2, 1000, 300_000, 5_000_000):                              # small, one MSD pass, LSD passes
        a = rng.integers(0, 2**64, n, dtype="u8")
        got, st = sort_hinted(a, k8, (1 << 64) - 1)
        assert got == np.sort(a).tobytes(), n
    rec = np.dtype([("key", "<u4"), ("idx", "<u4")])
    key = RadixKey.field(rec, "key", 4)
    r = np.zeros(700_000, dtype=rec)
    r["key"] = rng.integers(0, 1 << 28, len(r))
    r["idx"] = np.arange(len(r))
    got, st = sort_hinted(r, key, (1 << 32) - 1)
    assert got == o.stable_sorted(r, key).tobytes()
    k16 = RadixKey.u128()
    w = rng.integers(0, 256, 16 * 100_000, dtype=np.uint8)
    got, st = sort_hinted(w, k16, 0xFFFF)                                # LEVELS > 8: the hint is ignored
    assert got == o.stable_sorted(w, k16).tobytes()
