"""GPU parity: the CUDA path, called through the C ABI (via rdst_b200's ctypes mirror of
the reference API), against the oracle on the same seeded inputs.  Bit-exact everywhere:
this is integer / byte work.  Key-only sorts must equal the oracle byte for byte; records
with payload must equal rdst's STABLE order (oracle: the Lsb algorithm); at full
BASELINE sizes, size-independent properties are checked on the device."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle_ffi as o
from rdst_b200 import RadixKey, _abi, radix_sort_builder, radix_sort_stable, radix_sort_unstable, tuner

pytestmark = pytest.mark.gpu

G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_vectors.json")))
NPT = {"i32": "i4", "u64": "u8"}


def torch_mod():
    import torch

    assert torch.cuda.is_available()
    return torch


def gpu_sort_both_ways(a, key, stable=False, variant=-1):
    """Sort a copy through the host entry point and another through the device entry
    point; both must agree; returns the result."""
    torch = torch_mod()
    h = a.copy()
    b = radix_sort_builder(h, key).with_variant(variant)
    (b.with_stable(True) if stable else b).sort()
    raw = np.ascontiguousarray(a).view(np.uint8).reshape(-1)
    if raw.size:
        t = torch.from_numpy(raw.copy()).cuda()
        b2 = radix_sort_builder(t, key).with_variant(variant)
        (b2.with_stable(True) if stable else b2).sort()
        torch.cuda.synchronize()
        assert t.cpu().numpy().tobytes() == h.tobytes(), "host and device entry points disagree"
    return h


def rand_bits(rng, dt, n):
    return rng.integers(0, 256, n * np.dtype(dt).itemsize, dtype=np.uint8).view(dt).copy()


# ---------------------------------------------------------------- golden vectors
def test_reference_golden_vectors_on_gpu():
    for vec in G["sort_vectors"]:
        a = np.array(vec["input"], dtype=NPT[vec["type"]])
        radix_sort_unstable(a)
        assert a.tolist() == vec["expected"], vec["cite"]
    for vec in G["byte_array_vectors"]:
        a = np.array(vec["input"], dtype=np.uint8)
        radix_sort_unstable(a, RadixKey.byte_array(vec["n"]))
        assert a.tolist() == vec["expected"]
    for vec in G["custom_key_vectors"]:
        a = np.array(vec["input"], dtype=np.uint8)
        radix_sort_unstable(a, RadixKey.custom(vec["elem_size"], vec["level_offsets"]))
        assert a.tolist() == vec["expected"], vec["cite"]


def test_builder_variants_of_the_reference_api():
    """doctests of reference src/lib.rs:110-172 and src/radix_sort_builder.rs:43-147."""
    a = np.array([10, 15, 0, 22, 9], dtype="u8")
    radix_sort_builder(a).with_low_mem_tuner().sort()
    assert a.tolist() == [0, 9, 10, 15, 22]
    a = np.array([10, 15, 0, 22, 9], dtype="u8")
    radix_sort_builder(a).with_single_threaded_tuner().with_parallel(False).sort()
    assert a.tolist() == [0, 9, 10, 15, 22]

    class MyTuner(tuner.Tuner):
        def __init__(self):
            self.calls = []

        def pick_algorithm(self, p, counts):
            self.calls.append((p, list(counts)))
            return tuner.Algorithm.Ska if p.input_len >= 500_000 else tuner.Algorithm.Lsb

    t = MyTuner()
    a = np.array([10, 25, 9, 22, 6], dtype="u8")
    radix_sort_builder(a).with_tuner(t).sort()
    assert a.tolist() == [6, 9, 10, 22, 25]
    p, counts = t.calls[0]
    assert (p.level, p.total_levels, p.input_len, p.parent_len) == (7, 8, 5, None)
    assert counts[0] == 5 and sum(counts) == 5


# ---------------------------------------------------------------- key-only primitives
SIZES = [2, 10, 100, 128, 129, 400, 5_000, 10_000, 50_000, 100_000, 260_000, 300_000, 1_000_000]


@pytest.mark.parametrize("dt", ["u1", "u2", "u4", "u8", "i1", "i2", "i4", "i8", "f4", "f8"])
def test_primitive_size_ladder(dt):
    """Size ladder + bimodal shift of reference src/test_utils.rs:51-95 (all 10 key types;
    random BIT patterns, so floats include NaN / Inf / subnormal / -0.0 like block_rand)."""
    rng = np.random.default_rng(100 + np.dtype(dt).num)
    key = RadixKey.for_dtype(dt)
    bits = np.dtype(dt).itemsize * 8
    for n in SIZES:
        for shift in (0, bits // 2):
            a = rand_bits(rng, dt, n)
            if shift:
                u = a.view("u%d" % a.dtype.itemsize)
                u[: n // 2] >>= np.array(shift, dtype=u.dtype)
                u[n // 2:] <<= np.array(shift, dtype=u.dtype)
            got = gpu_sort_both_ways(a, key)
            assert got.tobytes() == o.sorted_copy(a, key).tobytes(), (dt, n, shift)


def test_u32_mask_patterns():
    """validate_u32_patterns, reference src/test_utils.rs:148-262."""
    rng = np.random.default_rng(5)
    key = RadixKey.for_dtype("u4")
    masks = [0xFFFF_FFFF, 0xFFFF_FF00, 0xFFFF_00FF, 0xFF00_FFFF, 0x00FF_FFFF, 0x0000_FFFF, 0xFFFF_0000,
             0x8000_0000, 0x1, 0xFFFF_FFFE, 0x7FFF_FFFF, 0xAAAA_AAAA, 0x5555_5555]
    bases = [np.full(128, 0xFFFFFFFF, dtype="u4"), rng.integers(0, 2**32, 128, dtype="u4"),
             rng.integers(0, 2**32, 128_000, dtype="u4"), rng.integers(0, 2**32, 4, dtype="u4")]
    for base in bases:
        for m in masks:
            a = base & np.uint32(m)
            assert gpu_sort_both_ways(a, key).tobytes() == o.sorted_copy(a, key).tobytes(), hex(m)
    a = np.zeros(128_000, dtype="u4")
    a[:5] = 0x7700_0000
    assert gpu_sort_both_ways(a, key).tobytes() == o.sorted_copy(a, key).tobytes()


@pytest.mark.parametrize("dist", ["all_equal", "sorted", "reverse", "mostly_sorted", "two_values", "zipf"])
def test_distributions_u64(dist):
    rng = np.random.default_rng(21)
    n = 777_777
    if dist == "all_equal":
        a = np.full(n, 0xDEADBEEFCAFEF00D, dtype="u8")
    elif dist == "sorted":
        a = np.sort(rng.integers(0, 2**63, n, dtype="u8"))
    elif dist == "reverse":
        a = np.sort(rng.integers(0, 2**63, n, dtype="u8"))[::-1].copy()
    elif dist == "mostly_sorted":
        a = np.sort(rng.integers(0, 2**63, n, dtype="u8"))
        a[-n // 10:] = rng.integers(0, 2**63, len(a[-n // 10:]), dtype="u8")
    elif dist == "two_values":
        a = rng.choice(np.array([5, 2**40 + 1], dtype="u8"), n)
    else:
        a = rng.zipf(1.1, n).astype("u8")
    key = RadixKey.for_dtype("u8")
    assert gpu_sort_both_ways(a, key).tobytes() == o.sorted_copy(a, key).tobytes()


def test_float_total_order_with_specials():
    """IEEE total order incl. NaN payloads, +-0, +-inf, subnormals (reference
    src/radix_key_impl.rs:162-185; src/radix_sort.rs:97-144)."""
    rng = np.random.default_rng(31)
    for dt, ut in (("f4", "u4"), ("f8", "u8")):
        key = RadixKey.for_dtype(dt)
        sp = [b["bits"] for b in G["f32_levels" if dt == "f4" else "f64_levels"]]
        a = np.concatenate([np.array(sp, dtype=ut), rand_bits(rng, ut, 200_000)]).view(dt)
        rng.shuffle(a.view(ut))
        got = gpu_sort_both_ways(a, key)
        assert got.tobytes() == o.sorted_copy(a, key).tobytes()
        for _ in range(200):  # the reference's 10-element random-bit cases
            b = rand_bits(rng, dt, 10)
            h = b.copy()
            radix_sort_unstable(h)
            assert h.tobytes() == o.sorted_copy(b, key).tobytes()


# ---------------------------------------------------------------- records with payload
REC12 = np.dtype([("key", "<u4"), ("idx", "<u8")])                      # BASELINE config 4
REC16 = np.dtype([("key", "<u8"), ("idx", "<u8")])
LARGE = np.dtype([("a", "<u8"), ("b", "<u8"), ("sort_key", "<f4"), ("c", "<u4"), ("d", "<u8"), ("e", "<u8")])  # 40 B
REC8 = np.dtype([("idx", "<u4"), ("key", "<i4")])
REC6 = np.dtype([("key", "<u2"), ("idx", "<u4")])                       # odd stride -> byte path


@pytest.mark.parametrize("rec,field,levels", [(REC12, "key", 3), (REC12, "key", 4), (REC16, "key", 8),
                                              (REC16, "key", 2), (LARGE, "sort_key", 4), (REC8, "key", 4),
                                              (REC6, "key", 2), (REC6, "key", 1)])
def test_struct_records_stable(rec, field, levels):
    rng = np.random.default_rng(41)
    key = RadixKey.field(rec, field, levels)
    for n in (129, 4_097, 100_000, 1_000_003):
        a = np.zeros(n, dtype=rec)
        raw = a.view(np.uint8).reshape(n, rec.itemsize)
        raw[:] = rng.integers(0, 256, raw.shape, dtype=np.uint8)  # garbage everywhere, incl. ignored key bytes
        kd = rec.fields[field][0]
        if kd.kind == "f":
            a[field] = rng.integers(-50, 50, n).astype(kd)  # duplicates + negatives
        else:
            a[field] = rng.integers(0, 1 << 11, n).astype(kd) | (rng.integers(0, 256, n).astype(kd) << (8 * levels) if levels < kd.itemsize else 0)
        got = gpu_sort_both_ways(a, key, stable=True)
        exp = o.stable_sorted(a, key)
        assert got.tobytes() == exp.tobytes(), (rec, field, levels, n)
        assert exp.tobytes() == o.numpy_stable_sorted(a, key).tobytes()


def test_unstable_mode_is_sorted_permutation():
    """Default (unstable) mode on payload records: sorted by key AND a permutation of the
    input.  (The engine happens to be stable; only this is promised.)"""
    rng = np.random.default_rng(43)
    n = 500_000
    a = np.zeros(n, dtype=REC12)
    a["key"] = rng.integers(0, 1 << 24, n)
    a["idx"] = np.arange(n)
    key = RadixKey.field(REC12, "key", 3)
    got = gpu_sort_both_ways(a, key)
    assert o.is_sorted(got, key)
    assert np.array_equal(np.sort(got["idx"]), a["idx"])
    assert np.array_equal(a["key"][got["idx"]], got["key"])


def test_partial_levels_on_primitives_ignore_high_bytes():
    rng = np.random.default_rng(47)
    for dt, levels in (("u4", 2), ("u8", 5), ("i8", 3), ("u8", 1)):
        key = RadixKey.for_dtype(dt, levels=levels)
        a = rand_bits(rng, dt, 300_000)
        got = gpu_sort_both_ways(a, key, stable=True)
        assert got.tobytes() == o.stable_sorted(a, key).tobytes(), (dt, levels)


@pytest.mark.parametrize("n_bytes", [3, 5, 7, 16, 20])
def test_byte_arrays_lexicographic(n_bytes):
    rng = np.random.default_rng(53)
    key = RadixKey.byte_array(n_bytes)
    a = rng.integers(0, 4, (50_001, n_bytes), dtype=np.uint8)
    got = gpu_sort_both_ways(a, key)
    exp = o.numpy_stable_sorted(a, key)
    assert got.tobytes() == exp.tobytes()
    if n_bytes in (3, 5, 7, 16, 20):
        assert exp.tobytes() == o.sorted_copy(a, key, tuner="single_algo", algo="Lsb").tobytes()


def test_u128_i128():
    rng = np.random.default_rng(59)
    a = rng.integers(0, 256, (200_003, 16), dtype=np.uint8)
    a[:, 9:] &= 1  # duplicates in the high bytes
    for key in (RadixKey.u128(), RadixKey.i128()):
        got = gpu_sort_both_ways(a, key)
        assert got.tobytes() == o.numpy_stable_sorted(a, key).tobytes()
        assert got.tobytes() == o.stable_sorted(a, key).tobytes()


def test_custom_multi_field_key():
    """reference src/lib.rs:78-104 MyStruct {key_1,key_2,key_3}: level 0 -> key_1 ..."""
    rng = np.random.default_rng(61)
    a = rng.integers(0, 7, (123_457, 4), dtype=np.uint8)
    key = RadixKey.custom(4, [0, 1, 2])
    got = gpu_sort_both_ways(a, key, stable=True)
    assert got.tobytes() == o.stable_sorted(a, key).tobytes()
    key = RadixKey.custom(4, [(3, 0x80), (0, 0xFF)])
    got = gpu_sort_both_ways(a, key, stable=True)
    assert got.tobytes() == o.stable_sorted(a, key).tobytes()


# ---------------------------------------------------------------- kernels one by one
def _dev(a):
    torch = torch_mod()
    return torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()


@pytest.mark.parametrize("dt,n", [("u4", 1), ("u4", 1_000_003), ("u8", 999_999), ("f4", 300_001), ("f8", 70_000),
                                  ("i4", 65_536), ("u2", 100_001), ("u1", 4_099)])
def test_k1_histogram_matches_get_counts(dt, n):
    """K1 vs get_counts_with_ends (reference src/sort_utils.rs:108-180) for every level."""
    rng = np.random.default_rng(67)
    a = rand_bits(rng, dt, n)
    key = RadixKey.for_dtype(dt)
    t = _dev(a)
    hist = (C.c_uint64 * (256 * key.levels))()
    d = key.to_desc()
    _abi.check(_abi.lib().b200_histogram_device(t.data_ptr(), n, C.byref(d), None, hist))
    for l in range(key.levels):
        counts, _, _, _ = o.get_counts(a, key, l)
        assert [hist[l * 256 + b] for b in range(256)] == counts.tolist(), (dt, l)


def test_k1_histogram_struct_and_skew():
    rng = np.random.default_rng(71)
    n = 400_000
    a = np.zeros(n, dtype=LARGE)
    a.view(np.uint8)[:] = rng.integers(0, 256, a.nbytes, dtype=np.uint8)
    a["sort_key"] = np.where(rng.random(n) < 0.9, 1.5, rng.standard_normal(n)).astype("f4")
    key = RadixKey.field(LARGE, "sort_key")
    t = _dev(a)
    hist = (C.c_uint64 * (256 * 4))()
    d = key.to_desc()
    _abi.check(_abi.lib().b200_histogram_device(t.data_ptr(), n, C.byref(d), None, hist))
    for l in range(4):
        assert [hist[l * 256 + b] for b in range(256)] == o.get_counts(a, key, l)[0].tolist()


@pytest.mark.parametrize("dt,level", [("u4", 0), ("u4", 3), ("u8", 5), ("f4", 3), ("i8", 7), ("u2", 1)])
def test_k2_single_pass_matches_out_of_place_sort(dt, level):
    """K2 vs out_of_place_sort (reference src/sorts/out_of_place_sort.rs:51-108)."""
    torch = torch_mod()
    rng = np.random.default_rng(73)
    n = 654_321
    a = rand_bits(rng, dt, n)
    key = RadixKey.for_dtype(dt)
    counts = o.get_counts(a, key, level)[0]
    src, dst = _dev(a), torch.empty(a.nbytes, dtype=torch.uint8, device="cuda")
    d = key.to_desc()
    _abi.check(_abi.lib().b200_pass_device(src.data_ptr(), dst.data_ptr(), n, C.byref(d), level,
                                           counts.ctypes.data_as(C.POINTER(C.c_uint64)), None))
    exp = o.out_of_place_pass(a, key, level)
    assert dst.cpu().numpy().tobytes() == exp.tobytes()
    assert exp.tobytes() == o.out_of_place_pass(a, key, level, lr=True).tobytes()  # LR variant is the same order


def test_partition_pass():
    torch = torch_mod()
    rng = np.random.default_rng(79)
    n = 500_009
    a = rng.integers(0, 2**64, n, dtype="u8")
    key = RadixKey.for_dtype("u8")
    level = 7
    counts = o.get_counts(a, key, level)[0]
    split = np.array([64, 128, 200], dtype=np.uint32)
    parts = 4
    src, dst = _dev(a), torch.empty(a.nbytes, dtype=torch.uint8, device="cuda")
    pc = (C.c_uint64 * parts)()
    d = key.to_desc()
    _abi.check(_abi.lib().b200_partition_device(src.data_ptr(), dst.data_ptr(), n, C.byref(d), level,
                                                counts.ctypes.data_as(C.POINTER(C.c_uint64)),
                                                split.ctypes.data_as(C.POINTER(C.c_uint32)), parts, pc, None))
    top = (a >> np.uint64(56)).astype(np.int64)
    part = np.searchsorted(split, top, side="right")
    exp = a[np.argsort(part, kind="stable")]
    assert dst.cpu().numpy().tobytes() == exp.tobytes()
    assert list(pc) == np.bincount(part, minlength=parts).tolist()


def _np_prefix(a, key):
    """numpy model of the 64-bit key prefix: top min(LEVELS, 8) levels, level LEVELS-1 most significant."""
    km = o.key_matrix(a.view(np.uint8).reshape(-1), key).astype(np.uint64)
    L = key.levels
    nl = min(L, 8)
    p = np.zeros(km.shape[0], dtype=np.uint64)
    for i in range(nl):
        p |= km[:, L - 1 - i] << np.uint64(8 * (nl - 1 - i))
    return p


def _splitter_cases():
    rng = np.random.default_rng(211)
    n = 300_007
    yield "u64", RadixKey.for_dtype("u8"), rng.integers(0, 2**64, n, dtype="u8")
    yield "u64_dups", RadixKey.for_dtype("u8"), rng.integers(0, 40, n, dtype="u8") * np.uint64(0x0101010101010101)
    yield "i64", RadixKey.for_dtype("i8"), rng.integers(-2**62, 2**62, n, dtype="i8")
    yield "u32", RadixKey.for_dtype("u4"), rng.integers(0, 2**32, n, dtype="u4")
    yield "f32", RadixKey.for_dtype("f4"), rand_bits(rng, "f4", n)
    yield "f64", RadixKey.for_dtype("f8"), rng.standard_normal(n)
    yield "u32_partial3", RadixKey.for_dtype("u4", levels=3), rng.integers(0, 2**32, n, dtype="u4")
    a = np.zeros(n, dtype=REC12)
    a["key"] = rng.integers(0, 1 << 20, n)
    a["idx"] = np.arange(n)
    yield "rec12", RadixKey.field(REC12, "key", 3), a
    b = np.zeros((n, 2), dtype="u8")
    b[:, 0] = rng.integers(0, 2**64, n, dtype="u8")
    b[:, 1] = rng.integers(0, 1000, n, dtype="u8")
    yield "u128", RadixKey.u128(), b.reshape(-1)


@pytest.mark.parametrize("name,key,a", list(_splitter_cases()), ids=[c[0] for c in _splitter_cases()])
def test_splitter_entry_points_match_numpy_model(name, key, a):
    """b200_key_prefix_sample_device / b200_splitter_count_device / b200_partition_splitters_device
    (the skewed-key range partition of the sharded sort) against a numpy model of the same rule."""
    from rdst_b200.dist import part_of_prefix

    torch = torch_mod()
    l = _abi.lib()
    d = key.to_desc()
    assert l.b200_key_prefix_supported(C.byref(d)) == 1
    n = a.nbytes // key.elem_size
    pref = _np_prefix(a, key)
    src = _dev(a)
    stride, count = 37, n // 37
    out = np.zeros(count, dtype=np.uint64)
    _abi.check(l.b200_key_prefix_sample_device(src.data_ptr(), n, C.byref(d), stride, count,
                                               out.ctypes.data_as(C.POINTER(C.c_uint64)), None))
    assert np.array_equal(out, pref[::stride][:count])
    srt = np.sort(pref)
    splitters = np.array([srt[n // 5], srt[n // 5], srt[n // 2], srt[(4 * n) // 5]], dtype=np.uint64)
    less, equal = np.zeros(4, dtype=np.uint64), np.zeros(4, dtype=np.uint64)
    _abi.check(l.b200_splitter_count_device(src.data_ptr(), n, C.byref(d), splitters.ctypes.data_as(C.POINTER(C.c_uint64)),
                                            4, less.ctypes.data_as(C.POINTER(C.c_uint64)),
                                            equal.ctypes.data_as(C.POINTER(C.c_uint64)), None))
    assert less.tolist() == [int((pref < s).sum()) for s in splitters]
    assert equal.tolist() == [int((pref == s).sum()) for s in splitters]
    for incl in (0b0000, 0b0101, 0b1001, 0b1111):
        part = part_of_prefix(pref, splitters, incl)
        pc = np.bincount(part, minlength=5).astype(np.uint64)
        dst = torch.zeros(a.nbytes, dtype=torch.uint8, device="cuda")
        _abi.check(l.b200_partition_splitters_device(src.data_ptr(), dst.data_ptr(), None, n, C.byref(d),
                                                     splitters.ctypes.data_as(C.POINTER(C.c_uint64)), 4, incl,
                                                     pc.ctypes.data_as(C.POINTER(C.c_uint64)), None))
        rows = a.view(np.uint8).reshape(n, key.elem_size)
        exp = rows[np.argsort(part, kind="stable")]
        assert dst.cpu().numpy().tobytes() == exp.tobytes(), (name, incl)
    # equal splitters whose tie flags would put the parts out of order are rejected
    pc = np.zeros(5, dtype=np.uint64)
    pc[0] = n
    rc = l.b200_partition_splitters_device(src.data_ptr(), dst.data_ptr(), None, n, C.byref(d),
                                           splitters.ctypes.data_as(C.POINTER(C.c_uint64)), 4, 0b0010,
                                           pc.ctypes.data_as(C.POINTER(C.c_uint64)), None)
    assert rc != 0


def test_splitter_mode_unsupported_keys_are_reported():
    key = RadixKey.byte_array(6)
    d = key.to_desc()
    assert _abi.lib().b200_key_prefix_supported(C.byref(d)) == 0


def test_every_variant_and_tile_boundaries():
    rng = np.random.default_rng(83)
    l = _abi.lib()
    for v in range(l.b200_variant_count()):
        for dt in ("u4", "u8"):
            key = RadixKey.for_dtype(dt)
            for n in (4095, 4096, 4097, 8191, 8192, 8193, 12287, 12288, 12289, 16383, 16384, 16385, 24576, 24577,
                      200_000):
                a = rand_bits(rng, dt, n)
                got = gpu_sort_both_ways(a, key, variant=v)
                assert got.tobytes() == np.sort(a).tobytes(), (v, dt, n)
        a = np.zeros(100_001, dtype=REC12)
        a["key"] = rng.integers(0, 1 << 16, len(a))
        a["idx"] = np.arange(len(a))
        key = RadixKey.field(REC12, "key", 3)
        assert gpu_sort_both_ways(a, key, stable=True, variant=v).tobytes() == o.stable_sorted(a, key).tobytes()


def test_stats_report_level_skipping():
    """Single-bucket levels are skipped (reference lsb_sort.rs:20-23, mt_lsb_sort.rs:170-174)."""
    rng = np.random.default_rng(89)
    a = rng.integers(0, 1 << 40, 300_000, dtype="u8")
    b = radix_sort_builder(a)
    b.sort()
    assert b.stats["levels_run"] == 5 and b.stats["levels_skipped"] == 3
    assert b.stats["algorithmic_bytes"] >= a.nbytes * (1 + 2 * 5)
    assert np.array_equal(a, np.sort(a))
    c = np.full(20_000, 7, dtype="u4")  # (up to 16 384 records the one-launch K3 path runs: no histogram)
    b = radix_sort_builder(c)
    b.sort()
    assert b.stats["levels_run"] == 0 and b.stats["levels_skipped"] == 4


# ---------------------------------------------------------------- BASELINE sizes (properties)
def _device_props_check(torch, t_sorted, t_in):
    assert bool((t_sorted[1:] >= t_sorted[:-1]).all())
    assert int(t_sorted.sum()) == int(t_in.sum())  # wrapping int64 sum: multiset checksum
    assert int((t_sorted * 0x9E3779B1 ^ (t_sorted >> 7)).sum()) == int((t_in * 0x9E3779B1 ^ (t_in >> 7)).sum())


def test_config1_u32_100m_properties():
    """BASELINE configs[0]: 100 M uniform u32 — sortedness, multiset checksums, idempotence,
    and exact equality with the oracle on a 2 M-record slice of the same generator."""
    torch = torch_mod()
    g = torch.Generator(device="cuda").manual_seed(0x5EED0001)
    n = 100_000_000
    ti = torch.randint(-2**31, 2**31, (n,), dtype=torch.int32, device="cuda", generator=g)
    ref = ti.to(torch.int64) & 0xFFFFFFFF
    keys = ti.clone()
    u32 = RadixKey.for_dtype("u4")
    radix_sort_unstable(keys, u32)  # the int32 storage holds u32 bit patterns
    torch.cuda.synchronize()
    s = keys.to(torch.int64) & 0xFFFFFFFF
    _device_props_check(torch, s, ref)
    del s, ref
    again = keys.clone()
    radix_sort_unstable(again, u32)
    assert torch.equal(again, keys)
    small = ti[:2_000_000].cpu().numpy().view("u4")
    h = small.copy()
    radix_sort_unstable(h)
    assert h.tobytes() == o.sorted_copy(small, RadixKey.for_dtype("u4")).tobytes()


def test_config2_u64_1b_properties():
    """BASELINE configs[1]: 1 B uniform u64 (8 GB + 8 GB scratch)."""
    torch = torch_mod()
    g = torch.Generator(device="cuda").manual_seed(0x5EED0002)
    n = 1_000_000_000
    t = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)
    chk = int(t.sum()), int((t ^ (t >> 9)).sum())
    radix_sort_unstable(t, RadixKey.for_dtype("u8"))  # the int64 storage holds u64 bit patterns
    torch.cuda.synchronize()
    # unsigned order: compare with the sign bit flipped
    u = t ^ (-2**63)
    ok = True
    step = 1 << 27
    for i in range(0, n - 1, step):
        j = min(n - 1, i + step)
        ok = ok and bool((u[i + 1:j + 1] >= u[i:j]).all())
    assert ok
    assert chk == (int(t.sum()), int((t ^ (t >> 9)).sum()))
    del u
    _abi.lib().b200_release_scratch(-1)


def test_config4_struct_100m_stable_property():
    """BASELINE configs[3] shape at 100 M: {u32 key (LEVELS=3), 8-byte payload = index}.
    Stability <=> within equal keys the payload indices ascend."""
    torch = torch_mod()
    g = torch.Generator(device="cuda").manual_seed(0x5EED0004)
    n = 100_000_000
    rec = torch.empty((n, 3), dtype=torch.int32, device="cuda")
    rec[:, 0] = torch.randint(0, 1 << 24, (n,), dtype=torch.int32, device="cuda", generator=g)
    rec[:, 1] = torch.arange(n, dtype=torch.int32, device="cuda")
    rec[:, 2] = 0
    radix_sort_stable(rec, RadixKey("uint", 12, 3, 0, 4))
    torch.cuda.synchronize()
    k, idx = rec[:, 0], rec[:, 1]
    assert bool((k[1:] >= k[:-1]).all())
    same = k[1:] == k[:-1]
    assert bool((idx[1:][same] > idx[:-1][same]).all())
    assert int(idx.to(torch.int64).sum()) == n * (n - 1) // 2
    _abi.lib().b200_release_scratch(-1)


def _float_image(torch, t):
    """Order-preserving integer image of IEEE bit patterns held in an int tensor (total order, reference
    src/radix_key_impl.rs:162-185): negative -> all bits flipped, else the sign bit flipped; compared as signed
    after flipping the sign bit back, i.e. plain `t` for non-negative and `~t ^ MIN` for negative patterns."""
    mn = -2**63 if t.dtype == torch.int64 else -2**31
    return torch.where(t < 0, ~t ^ mn, t)


def test_config3_f32_500m_properties():
    """BASELINE configs[2], first half: 500 M f32 random bit patterns (about half negative), NaN- and Inf-free.
    Sortedness under the float total order, multiset checksums, and exact equality with the oracle on a
    2 M-record slice of the same generator."""
    torch = torch_mod()
    g = torch.Generator(device="cuda").manual_seed(0x5EED0003)
    n = 500_000_000
    t = torch.randint(-2**31, 2**31, (n,), dtype=torch.int32, device="cuda", generator=g)
    t = torch.where(((t >> 23) & 0xFF) == 0xFF, t & -2139095041, t)  # clear the exponent of NaN / Inf patterns
    assert 0.45 < float((t < 0).float().mean()) < 0.55
    small = t[:2_000_000].cpu().numpy().view("f4")
    chk = int(t.to(torch.int64).sum()), int((t.to(torch.int64) * 0x9E3779B1 ^ (t.to(torch.int64) >> 7)).sum())
    key = RadixKey.for_dtype("f4")
    b = radix_sort_builder(t, key)
    b.sort()
    torch.cuda.synchronize()
    img = _float_image(torch, t)
    assert bool((img[1:] >= img[:-1]).all())
    del img
    assert chk == (int(t.to(torch.int64).sum()), int((t.to(torch.int64) * 0x9E3779B1 ^ (t.to(torch.int64) >> 7)).sum()))
    h = small.copy()
    radix_sort_unstable(h)
    assert h.tobytes() == o.sorted_copy(small, key).tobytes()
    _abi.lib().b200_release_scratch(-1)


def test_config3_i64_500m_properties():
    """BASELINE configs[2], second half: 500 M i64 including negatives."""
    torch = torch_mod()
    g = torch.Generator(device="cuda").manual_seed(0x5EED0013)
    n = 500_000_000
    t = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)
    small = t[:2_000_000].cpu().numpy().view("i8")
    chk = int(t.sum()), int((t ^ (t >> 9)).sum())
    key = RadixKey.for_dtype("i8")
    b = radix_sort_builder(t, key)
    b.sort()
    torch.cuda.synchronize()
    assert b.stats["hybrid"] == 1 and b.stats["unstable"] & 1, b.stats  # key-only route at this size
    assert bool((t[1:] >= t[:-1]).all())
    assert chk == (int(t.sum()), int((t ^ (t >> 9)).sum()))
    h = small.copy()
    radix_sort_unstable(h)
    assert h.tobytes() == o.sorted_copy(small, key).tobytes()
    _abi.lib().b200_release_scratch(-1)


def test_config4_struct_1b_stable_property():
    """BASELINE configs[3] at its stated size: 1 B records {u32 key (LEVELS = 3), 8-byte payload = input index},
    12 GB + 12 GB of scratch.  Stable <=> within equal keys the payload indices ascend; every index survives."""
    torch = torch_mod()
    g = torch.Generator(device="cuda").manual_seed(0x5EED0014)
    n = 1_000_000_000
    rec = torch.empty((n, 3), dtype=torch.int32, device="cuda")
    step = 1 << 28
    for i in range(0, n, step):  # in slices: no 4 GB temporaries next to the 12 GB array
        j = min(n, i + step)
        rec[i:j, 0] = torch.randint(0, 1 << 24, (j - i,), dtype=torch.int32, device="cuda", generator=g)
        rec[i:j, 1] = torch.arange(i, j, dtype=torch.int32, device="cuda")
    rec[:, 2] = 0
    b = radix_sort_builder(rec, RadixKey("uint", 12, 3, 0, 4)).with_stable(True)
    b.sort()
    torch.cuda.synchronize()
    assert b.stats["levels_run"] == 3 and b.stats["unstable"] == 0, b.stats
    ok, idx_sum = True, 0
    for i in range(0, n - 1, step):
        j = min(n - 1, i + step)
        k0, k1 = rec[i:j, 0], rec[i + 1:j + 1, 0]
        i0, i1 = rec[i:j, 1], rec[i + 1:j + 1, 1]
        ok = ok and bool((k1 >= k0).all()) and bool(((k1 > k0) | (i1 > i0)).all())
    for i in range(0, n, step):
        idx_sum += int(rec[i:min(n, i + step), 1].to(torch.int64).sum())
    assert ok
    assert idx_sum == n * (n - 1) // 2
    assert int(rec[:, 2].abs().max()) == 0
    del rec
    _abi.lib().b200_release_scratch(-1)


def test_tuner_answers_select_observable_routes():
    """The policy hook (reference src/tuner.rs:33-35, asked at src/sorter.rs:67-76): a user Tuner's Algorithm
    changes what the engine runs, and b200_stats shows it."""
    torch = torch_mod()

    class One(tuner.Tuner):
        def __init__(self, algo):
            self.algo, self.seen = algo, None

        def pick_algorithm(self, p, counts):
            self.seen = (p.level, p.total_levels, p.input_len, p.parent_len, len(counts), sum(counts))
            return self.algo

    rng = np.random.default_rng(5)
    n = 40_000_000  # above the key-only hybrid gate for u64
    a = rng.integers(0, 2**64, n, dtype="u8")
    exp = np.sort(a)
    got = {}
    for algo in (tuner.Algorithm.Scanning, tuner.Algorithm.Lsb, tuner.Algorithm.Ska):
        t = torch.from_numpy(a.copy()).cuda()
        tn = One(algo)
        b = radix_sort_builder(t.view(torch.uint8), RadixKey.for_dtype("u8")).with_tuner(tn)
        b.sort()
        torch.cuda.synchronize()
        assert np.array_equal(t.cpu().numpy(), exp), algo
        assert tn.seen == (7, 8, n, None, 256, n)
        got[algo] = b.stats
    assert got[tuner.Algorithm.Scanning]["hybrid"] == 1 and got[tuner.Algorithm.Scanning]["route_hint"] == 0
    assert got[tuner.Algorithm.Lsb]["hybrid"] == 0 and got[tuner.Algorithm.Lsb]["levels_run"] == 8
    assert got[tuner.Algorithm.Ska]["hybrid"] == 0 and got[tuner.Algorithm.Ska]["route_hint"] == 2
    assert got[tuner.Algorithm.Ska]["variant"] != got[tuner.Algorithm.Lsb]["variant"]  # the small-tile variant
    # with_low_mem_tuner(): same route through the builder, and the scratch pool is gone afterwards
    free0 = torch.cuda.mem_get_info()[0]
    t = torch.from_numpy(a.copy()).cuda()
    free1 = torch.cuda.mem_get_info()[0]
    b = radix_sort_builder(t.view(torch.uint8), RadixKey.for_dtype("u8")).with_low_mem_tuner()
    b.sort()
    torch.cuda.synchronize()
    assert np.array_equal(t.cpu().numpy(), exp)
    assert b.stats["hybrid"] == 0
    assert torch.cuda.mem_get_info()[0] >= free1 - (64 << 20), "the low-scratch route keeps no scratch pool"
    del free0


def test_host_entry_pinned_and_pageable_overlapped_paths():
    """b200_sort_host on inputs large enough for its overlapped timeline (K0 / K1 behind the H2D chunks; D2H behind
    the first slice of the segment sort): pinned and pageable memory, random and presorted, key-only and payload."""
    torch = torch_mod()
    rng = np.random.default_rng(6)
    n = 40_000_000
    a = rng.integers(0, 2**64, n, dtype="u8")
    exp = np.sort(a)
    for pinned in (False, True):
        h = torch.from_numpy(a.copy())
        if pinned:
            h = h.pin_memory()
        arr = h.numpy()
        b = radix_sort_builder(arr, RadixKey.for_dtype("u8"))
        b.sort()
        assert np.array_equal(arr, exp), pinned
        assert b.stats["hybrid"] == 1 and b.stats["ms_h2d"] > 0 and b.stats["ms_d2h"] >= 0
        b2 = radix_sort_builder(arr, RadixKey.for_dtype("u8"))  # presorted now: K0 behind the chunks says so
        b2.sort()
        assert b2.stats["already_sorted"] == 1 and np.array_equal(arr, exp)
    # one swapped pair exactly on a chunk boundary of the chunked K0 must be seen
    srt = exp.copy()
    cut = (128 << 20) // 8  # first chunk boundary in records
    srt[cut - 1], srt[cut] = srt[cut], srt[cut - 1]
    b3 = radix_sort_builder(srt, RadixKey.for_dtype("u8"))
    b3.sort()
    assert b3.stats["already_sorted"] == 0 and np.array_equal(srt, exp)
    # 12-byte records (chunk boundaries that are not powers of two), stable
    rec = np.dtype([("key", "<u4"), ("idx", "<u8")])
    r = np.zeros(12_000_000, dtype=rec)
    r["key"] = rng.integers(0, 1 << 20, len(r))
    r["idx"] = np.arange(len(r))
    key = RadixKey.field(rec, "key", 3)
    hr = r.copy()
    radix_sort_builder(hr, key).with_stable(True).sort()
    assert hr.tobytes() == o.stable_sorted(r, key).tobytes()


# ---------------------------------------------------------------------------------------------
# Hybrid path (two MSD passes + K3 shared-memory segment sort): must give the oracle's answer
# bit for bit on every input the LSD path handles.  B200SORT_HYBRID=1 forces it at small sizes.
# ---------------------------------------------------------------------------------------------
def _hybrid_cases():
    rng = np.random.default_rng(977)
    n = 1_500_007
    yield "u64_uniform", "u8", rng.integers(0, 2**64, n, dtype="u8"), True
    yield "i64_uniform", "i8", rng.integers(-2**63, 2**63 - 1, n, dtype="i8"), True
    yield "f64_normal", "f8", rng.standard_normal(n), None  # few sign/exponent values: segments may be too large
    yield "f64_bits", "f8", rand_bits(rng, "f8", n), True
    yield "u32_uniform", "u4", rng.integers(0, 2**32, n, dtype="u4"), True
    yield "f32_bits", "f4", rand_bits(rng, "f4", n), True
    yield "i32_uniform", "i4", rng.integers(-2**31, 2**31 - 1, n, dtype="i4"), True
    # few distinct values: long runs of equal fast-level prefixes -> slow path inside K3
    yield "u64_dups", "u8", rng.integers(0, 3000, n, dtype="u8") * np.uint64(0x0001000100010001), True
    # segments of ~6 k records (16 x 16 values in the two MSD digits): the big-segment regime of the 1 B-key run
    top = (rng.integers(0, 16, n, dtype="u8") << np.uint64(60)) | (rng.integers(0, 16, n, dtype="u8") << np.uint64(52))
    yield "u64_big_segments", "u8", top | rng.integers(0, 2**44, n, dtype="u8"), True
    yield "u64_big_segments_dups", "u8", top | (rng.integers(0, 50, n, dtype="u8") << np.uint64(20)), True
    # 4 x 4 values in the two MSD digits: segments of ~94 k records do not fit an SM -> falls back to LSD from the scratch buffer
    top4 = (rng.integers(0, 4, n, dtype="u8") << np.uint64(62)) | (rng.integers(0, 4, n, dtype="u8") << np.uint64(54))
    yield "u64_segments_too_large", "u8", top4 | rng.integers(0, 2**46, n, dtype="u8"), False
    # keys that do not fill their type: the MSD digits are bit windows below the top varying bit
    yield "u64_61_bits", "u8", rng.integers(0, 2**61, n, dtype="u8"), True
    yield "u64_45_bits", "u8", rng.integers(0, 2**45, n, dtype="u8"), True
    yield "u64_36_bits_window_across_words", "u8", rng.integers(0, 2**36, n, dtype="u8"), True
    yield "u64_rank_range", "u8", rng.integers(3 << 58, 5 << 58, n, dtype="u8"), True
    yield "i64_narrow", "i8", rng.integers(-2**40, 2**40, n, dtype="i8"), None
    yield "i64_positive_50_bits", "i8", rng.integers(0, 2**50, n, dtype="i8"), True
    yield "u32_27_bits", "u4", rng.integers(0, 2**27, n, dtype="u4"), True
    yield "f32_one_to_two", "f4", (rng.random(n, dtype="f4") + np.float32(1.0)), True
    yield "f64_unit_interval", "f8", rng.random(n), None
    yield "f64_negative_bits", "f8", -np.abs(rand_bits(rng, "f8", n)), None
    # trivial levels in the middle and on top (level skip inside the hybrid plan)
    yield "u64_masked", "u8", rng.integers(0, 2**64, n, dtype="u8") & np.uint64(0x00FFFF00FFFF00FF), True
    z = np.minimum(rng.zipf(1.1, n), 2**32).astype("u8")
    yield "u64_zipf_hashed", "u8", (z * np.uint64(0x9E3779B97F4A7C15)) ^ (z >> np.uint64(7)), None
    yield "u64_sorted", "u8", np.sort(rng.integers(0, 2**64, n, dtype="u8")), True
    yield "u64_reverse", "u8", np.sort(rng.integers(0, 2**64, n, dtype="u8"))[::-1].copy(), True
    yield "u64_tiny", "u8", rng.integers(0, 2**64, 1000, dtype="u8"), True
    # key-only route: two counters of K3-U's wide digit hold ~2.9 k UNEQUAL keys each -> the segment goes to the
    # slow list and the stable segment sort finishes it
    yield "u64_slow_list", "u8", top | (rng.integers(0, 2, n, dtype="u8") << np.uint64(40)) | rng.integers(0, 256, n, dtype="u8"), True
    # heavy duplicates inside big segments: long runs of EQUAL keys need no fix-up (warp check)
    yield "u64_big_segments_few_values", "u8", top | (rng.integers(0, 3, n, dtype="u8") * np.uint64(0x0000051234567891)), True
    # oversized segments (key-only route: split around pivot keys, heavy.cuh; stable route: falls back to LSD passes)
    heavy = rng.integers(0, 2**64, n, dtype="u8")
    heavy[rng.random(n) < 0.3] = np.uint64(0x123456789ABCDEF0)          # one key holds 30 % of the records
    yield "u64_one_heavy_key", "u8", heavy, None
    heavy2 = heavy.copy()
    heavy2[rng.random(n) < 0.1] = np.uint64(0x123456789ABCDEF0 + 5)      # a second heavy key in the same segment: two rounds
    heavy2[rng.random(n) < 0.05] = np.uint64(0xFEDC000000000001)
    yield "u64_heavy_keys_share_a_segment", "u8", heavy2, None
    cluster = rng.integers(0, 2**64, n, dtype="u8")
    m = rng.random(n) < 0.08                                             # 120 k DISTINCT keys in one segment: quicksort-like rounds
    cluster[m] = np.uint64(0xABCD000000000000) | (cluster[m] & np.uint64(0xFFFFFFFFFFFF))
    yield "u64_oversized_cluster", "u8", cluster, None
    yield "i64_one_heavy_key", "i8", heavy.view("i8"), None
    hf = rand_bits(rng, "f8", n)
    hf[rng.random(n) < 0.25] = -1.5
    yield "f64_one_heavy_key", "f8", hf, None
    h4 = rng.integers(0, 2**32, n, dtype="u4")
    h4[rng.random(n) < 0.4] = np.uint32(0xDEADBEEF)
    yield "u32_one_heavy_key", "u4", h4, None
    yield "u32_big_segments", "u4", ((rng.integers(0, 16, n, dtype="u4") << np.uint32(28)) | (rng.integers(0, 16, n, dtype="u4") << np.uint32(20))
                                     | rng.integers(0, 2**16, n, dtype="u4")), True


@pytest.mark.parametrize("unstable", [1, 0], ids=["key_only_route", "stable_route"])
@pytest.mark.parametrize("name,dt,a,expect_hybrid", list(_hybrid_cases()), ids=[c[0] for c in _hybrid_cases()])
def test_hybrid_path_matches_oracle(monkeypatch, name, dt, a, expect_hybrid, unstable):
    """Every case is a key-only array, so it runs twice: through the unstable flavour of the route (K2-U atomic
    ranking + K3-U wide-digit segment sort, the default for key-only records) and through the stable one
    (B200SORT_UNSTABLE=0: what records with a payload take).  Both must equal the oracle bit for bit."""
    torch = torch_mod()
    monkeypatch.setenv("B200SORT_HYBRID", "1")
    monkeypatch.setenv("B200SORT_UNSTABLE", str(unstable))
    # the route under test is the hybrid one: no early-out for the presorted case, no single-launch
    # path for the tiny one (tests/test_gpu_sorted_check.py and test_gpu_small_sort.py cover those)
    monkeypatch.setenv("B200SORT_SORTED_CHECK", "0")
    monkeypatch.setenv("B200SORT_SMALL", "0")
    key = RadixKey.for_dtype(dt)
    exp = o.sorted_copy(a, key)
    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()
    b = radix_sort_builder(t, key)
    b.sort()
    torch.cuda.synchronize()
    assert t.cpu().numpy().tobytes() == exp.tobytes(), name
    if expect_hybrid is not None:
        assert bool(b.stats["hybrid"]) == expect_hybrid, (name, b.stats)
    if b.stats["hybrid"]:
        assert bool(b.stats["unstable"] & 1) == bool(unstable), b.stats
    if name == ("u64_slow_list" if unstable else "u64_dups"):
        assert b.stats["slow_segments"] > 0, b.stats
    if unstable and name in ("u64_one_heavy_key", "u64_heavy_keys_share_a_segment", "u64_oversized_cluster", "i64_one_heavy_key",
                             "f64_one_heavy_key", "u32_one_heavy_key", "u64_zipf_hashed"):
        assert b.stats["hybrid"] == 1 and b.stats["heavy_segments"] >= 1 and b.stats["heavy_rounds"] >= 1, b.stats
        if name == "u64_heavy_keys_share_a_segment":
            assert b.stats["heavy_rounds"] >= 2, b.stats
        # cut the rounds short: the route gives up cleanly and the LSD passes finish the job
        monkeypatch.setenv("B200SORT_HEAVY_ROUNDS", "1" if name != "u64_one_heavy_key" else "0")
        t3 = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()
        b3 = radix_sort_builder(t3, key)
        b3.sort()
        torch.cuda.synchronize()
        assert t3.cpu().numpy().tobytes() == exp.tobytes(), name + " rounds cut short"
        monkeypatch.delenv("B200SORT_HEAVY_ROUNDS")
        monkeypatch.setenv("B200SORT_HEAVY", "0")  # the switch: oversized segments send the array to the LSD passes at once
        t4 = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()
        b4 = radix_sort_builder(t4, key)
        b4.sort()
        torch.cuda.synchronize()
        assert t4.cpu().numpy().tobytes() == exp.tobytes() and b4.stats["heavy_segments"] == 0, name + " heavy off"
        monkeypatch.delenv("B200SORT_HEAVY")
    if unstable and name in ("u64_dups", "u64_big_segments_few_values"):
        assert b.stats["slow_segments"] == 0, b.stats  # long runs of equal keys are not a reason to fall back
    # the host entry point takes the same route
    h = a.copy()
    radix_sort_unstable(h)
    assert h.tobytes() == exp.tobytes(), name
    if unstable:
        # the other size of K3-U (512 threads, one CTA per SM) and a narrow digit (long comparison runs)
        monkeypatch.setenv("B200SORT_K3U_SIZE", "2")
        monkeypatch.setenv("B200SORT_K3U_BMAX", "9")
    else:
        # both fast-level settings of K3 agree (nfast = 1 leaves more to the insertion fix-up)
        monkeypatch.setenv("B200SORT_HYBRID_NFAST", "1")
    t2 = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()
    radix_sort_builder(t2, key).sort()
    torch.cuda.synchronize()
    assert t2.cpu().numpy().tobytes() == exp.tobytes(), name + " second setting"


def test_hybrid_is_stable_for_key_plus_payload_of_one_word():
    """8-byte records {u32 key, u32 payload}: K3's fix-up and slow path must keep input order among equal keys."""
    import os

    torch = torch_mod()
    rec = np.dtype([("key", "<u4"), ("idx", "<u4")])
    rng = np.random.default_rng(31)
    a = np.zeros(1_200_000, dtype=rec)
    a["key"] = rng.integers(0, 1 << 31, len(a)) & 0xFFFF0FFF
    a["key"][::300] = 12345  # 4000 equal keys: one long run (slow path), still one segment
    a["idx"] = np.arange(len(a))
    key = RadixKey.field(rec, "key", 4)
    os.environ["B200SORT_HYBRID"] = "1"
    try:
        t = torch.from_numpy(a.view(np.uint8).copy()).cuda()
        b = radix_sort_builder(t, key).with_stable(True)
        b.sort()
        torch.cuda.synchronize()
    finally:
        del os.environ["B200SORT_HYBRID"]
    assert b.stats["hybrid"] == 1
    assert t.cpu().numpy().tobytes() == o.stable_sorted(a, key).tobytes()
