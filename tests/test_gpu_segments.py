"""b200_sort_segments_device: many independent sorts in one call — the tail of rdst's MSB-first recursion, where every
bucket of a split is sorted on its own (reference src/sorter.rs:121-157, src/sorts/lsb_sort.rs:38-128,
src/sorts/comparative_sort.rs:29-119).  Every segment must come out exactly as the oracle sorts it alone, records
outside the segments must not move, and stable order must hold for records with a payload."""
import ctypes as C

import numpy as np
import pytest

import oracle_ffi as o
from rdst_b200 import RadixKey, _abi, sort_segments

CAP = 16_384
REC8 = np.dtype([("idx", "<u4"), ("key", "<i4")])
REC12 = np.dtype([("key", "<u4"), ("lo", "<u4"), ("hi", "<u4")])


def rand_bits(rng, dt, n):
    return rng.integers(0, 256, n * np.dtype(dt).itemsize, dtype=np.uint8).view(dt).copy()


def offsets_from_sizes(sizes, lead=0):
    off = np.zeros(len(sizes) + 1, dtype=np.uint64)
    off[0] = lead
    off[1:] = lead + np.cumsum(np.asarray(sizes, dtype=np.uint64))
    return off


def expected(a, off, key):
    out = a.copy()
    for i in range(len(off) - 1):
        s, e = int(off[i]), int(off[i + 1])
        if e - s > 1:
            out[s:e] = o.stable_sorted(np.ascontiguousarray(a[s:e]), key)
    return out


def run(a, off, key):
    import torch

    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()
    st = sort_segments(t, off, key)
    torch.cuda.synchronize()
    return t.cpu().numpy().view(a.dtype).reshape(a.shape), st


def test_segments_entry_point_validates_offsets_without_a_device():
    l = _abi.lib()
    d = RadixKey.for_dtype("u8").to_desc()
    fake = C.c_void_p(256)

    def call(off, n, ptr=fake):
        arr = np.asarray(off, dtype=np.uint64)
        return l.b200_sort_segments_device(ptr, n, arr.ctypes.data_as(C.POINTER(C.c_uint64)), len(arr) - 1, C.byref(d), None, None)

    assert call([0], 100) == 0                       # no segments
    assert call([0, 1, 2, 3], 100) == 0              # segments of one record: nothing to do, no device needed
    assert call([0, 10, 5], 100) == 6                # offsets must ascend
    assert call([0, 10, 101], 100) == 6              # ... and stay inside the array
    assert call([0, 10], 100, None) == 6             # null data
    assert l.b200_sort_segments_device(fake, 100, None, 3, C.byref(d), None, None) == 6
    z = RadixKey.for_dtype("u8", levels=0).to_desc()
    arr = np.array([0, 10], dtype=np.uint64)
    assert l.b200_sort_segments_device(fake, 100, arr.ctypes.data_as(C.POINTER(C.c_uint64)), 1, C.byref(z), None, None) == 1


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["u8", "u4", "i8", "f4", "f8", "i4"])
def test_ragged_segments_match_the_oracle_segment_by_segment(dt):
    rng = np.random.default_rng(500 + np.dtype(dt).num)
    key = RadixKey.for_dtype(dt)
    sizes = [0, 1, 2, 3, 0, 17, 31, 32, 33, 127, 128, 129, 1000, 4095, 4096, 4097, 8191, 8192, 8193, CAP - 1, CAP, CAP + 1,
             40_001, 5, 0, 0, 2, 100_003, 7] + list(rng.integers(0, 3000, 300))
    off = offsets_from_sizes(sizes, lead=3)           # records before the first and after the last segment stay put
    n = int(off[-1]) + 11
    a = rand_bits(rng, dt, n)
    got, st = run(a, off, key)
    assert got.tobytes() == expected(a, off, key).tobytes(), dt
    assert st["heavy_segments"] == 3 and st["tiles_per_pass"] == sum(1 for s in sizes if 2 <= s <= CAP)
    assert st["unstable"] == 1


@pytest.mark.gpu
@pytest.mark.parametrize("dt,bits", [("u8", 16), ("u8", 40), ("u4", 12), ("i8", 8)])
def test_buckets_of_an_msd_split_share_their_leading_bits(dt, bits):
    """The use the reference makes of it: split on the leading bits, then sort every bucket.  All keys of a bucket agree
    on those bits, so the on-chip sort must cut its digit below them (per-segment top bit) — no segment may fall back."""
    rng = np.random.default_rng(77)
    n, nb = 3_000_000, 4096
    width = 8 * np.dtype(dt).itemsize
    ut = np.dtype("u" + dt[1])
    lead = np.sort(rng.integers(0, nb, n))                                  # 4096 buckets on the 12 leading bits
    low = rng.integers(0, 1 << bits, n, dtype=np.uint64)                    # `bits` varying bits at the bottom, constant between
    img = ((lead.astype(np.uint64) << np.uint64(width - 12)) | low).astype(ut)
    if dt[0] == "i":
        img = img ^ ut.type(1 << (width - 1))                               # image -> two's complement
    a = np.ascontiguousarray(img.view(dt))
    off = np.concatenate([[0], np.cumsum(np.bincount(lead, minlength=nb))]).astype(np.uint64)
    got, st = run(a, off, RadixKey.for_dtype(dt))
    assert got.tobytes() == np.sort(a).tobytes()      # buckets in order + every bucket sorted = the array sorted
    assert st["slow_segments"] == 0 and st["heavy_segments"] == 0 and st["tiles_per_pass"] == nb, st


@pytest.mark.gpu
def test_duplicates_and_equal_segments():
    rng = np.random.default_rng(3)
    key = RadixKey.for_dtype("u8")
    sizes = [5000, 12_000, 300, 16_000, 9000, 2]
    off = offsets_from_sizes(sizes)
    a = rand_bits(rng, "u8", int(off[-1]))
    a[0:5000] = 42                                                  # one key
    a[5000:17_000] = rng.integers(0, 3, 12_000).astype("u8")        # three keys
    a[17_300:33_300] = rng.integers(0, 2**64, 16_000, dtype="u8") & np.uint64(0xFF)          # 256 keys in the low byte
    a[33_300:42_300] = (rng.integers(0, 2**64, 9000, dtype="u8") & np.uint64(0xFFFF)) << np.uint64(13)
    got, st = run(a, off, key)
    assert got.tobytes() == expected(a, off, key).tobytes(), st


@pytest.mark.gpu
def test_records_with_a_payload_are_sorted_stably_per_segment():
    rng = np.random.default_rng(9)
    key = RadixKey.field(REC8, "key", 4)
    sizes = list(rng.integers(0, 6000, 200)) + [CAP, 20_000]
    off = offsets_from_sizes(sizes, lead=1)
    n = int(off[-1]) + 2
    a = np.zeros(n, dtype=REC8)
    a["key"] = rng.integers(-50, 50, n)
    a["idx"] = np.arange(n)
    got, st = run(a, off, key)
    assert got.tobytes() == expected(a, off, key).tobytes()
    assert st["unstable"] == 0 and st["heavy_segments"] == 1


@pytest.mark.gpu
def test_shapes_the_on_chip_sort_does_not_take_still_sort_every_segment():
    rng = np.random.default_rng(10)
    sizes = [3000, 0, 1, 70_000, 513, 2]
    off = offsets_from_sizes(sizes, lead=5)
    n = int(off[-1]) + 1
    a12 = np.zeros(n, dtype=REC12)
    a12["key"] = rng.integers(0, 1000, n)
    a12["lo"] = np.arange(n)
    k12 = RadixKey.field(REC12, "key", 4)
    got, st = run(a12, off, k12)
    assert got.tobytes() == expected(a12, off, k12).tobytes()
    assert st["tiles_per_pass"] == 0 and st["heavy_segments"] == 4
    a2 = rand_bits(rng, "u2", n)
    got, _ = run(a2, off, RadixKey.for_dtype("u2"))
    assert got.tobytes() == expected(a2, off, RadixKey.for_dtype("u2")).tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["u8", "u4", "f4", "i8", "rec8"])
def test_tiny_segments_one_warp_each(case, monkeypatch):
    """Segments of up to 128 (key-only) / 256 (payload) records take K3-T, a warp per segment ranking by comparison —
    rdst's comparative_sort for small chunks (reference src/sorts/comparative_sort.rs:29-119).  Every size from 0 to 260,
    few distinct keys (ties must keep their order), then the same input with the kernel switched off."""
    rng = np.random.default_rng(41)
    sizes = list(range(0, 261)) + list(rng.integers(0, 200, 2000))
    off = offsets_from_sizes(sizes, lead=2)
    n = int(off[-1]) + 3
    if case == "rec8":
        key = RadixKey.field(REC8, "key", 4)
        a = np.zeros(n, dtype=REC8)
        a["key"] = rng.integers(-4, 4, n)
        a["idx"] = np.arange(n)
    else:
        key = RadixKey.for_dtype(case)
        a = rand_bits(rng, case, n)
        a[: n // 2] = a[rng.integers(0, 16, n // 2)]      # the first half: 16 distinct keys
    want = expected(a, off, key)
    got, st = run(a, off, key)
    assert got.tobytes() == want.tobytes(), case
    monkeypatch.setenv("B200SORT_TINY", "0")
    got2, st2 = run(a, off, key)
    assert got2.tobytes() == want.tobytes(), case
    assert st["kernel_launches"] == st2["kernel_launches"] + 1   # the extra launch is K3-T
