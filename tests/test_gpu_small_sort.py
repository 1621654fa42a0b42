"""The small-input path: arrays of up to 16 384 four- or eight-byte records are sorted by ONE
launch of K3 (the shared-memory segment sort) with the whole array as its only segment — the
GPU counterpart of rdst's small-chunk routes (reference src/sorter.rs:33-45,
src/sorts/comparative_sort.rs:29-119).  Bit-exact against the oracle, stable for key+payload."""
import numpy as np
import pytest

import oracle_ffi as o
from rdst_b200 import RadixKey, radix_sort_builder

pytestmark = pytest.mark.gpu

CAP = 16_384
REC8 = np.dtype([("idx", "<u4"), ("key", "<i4")])
REC8F = np.dtype([("key", "<f4"), ("idx", "<u4")])


def rand_bits(rng, dt, n):
    return rng.integers(0, 256, n * np.dtype(dt).itemsize, dtype=np.uint8).view(dt).copy()


def sort_both(a, key, stable=False):
    import torch

    h = a.copy()
    bh = radix_sort_builder(h, key).with_stable(stable)
    bh.sort()
    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()
    bd = radix_sort_builder(t, key).with_stable(stable)
    bd.sort()
    torch.cuda.synchronize()
    assert t.cpu().numpy().tobytes() == h.tobytes(), "host and device entry points disagree"
    return h, bh.stats, bd.stats


@pytest.mark.parametrize("dt", ["u4", "u8", "i4", "i8", "f4", "f8"])
def test_small_primitives_one_launch(dt):
    rng = np.random.default_rng(900 + np.dtype(dt).num)
    key = RadixKey.for_dtype(dt)
    for n in (2, 3, 15, 16, 17, 31, 32, 33, 511, 512, 513, 1_000, 4_097, 9_999, CAP - 1, CAP, CAP + 1):
        a = rand_bits(rng, dt, n)
        got, sh, sd = sort_both(a, key)
        assert got.tobytes() == o.sorted_copy(a, key).tobytes(), (dt, n)
        one = n <= CAP
        assert (sh["kernel_launches"] == 1) == one and (sd["kernel_launches"] == 1) == one, (dt, n, sh)
        if one:
            assert sh["levels_run"] == key.levels and sh["copy_back"] == 0


@pytest.mark.parametrize("dt", ["u4", "i8", "f4"])
def test_small_duplicates_narrow_ranges_and_order_patterns(dt):
    rng = np.random.default_rng(33)
    key = RadixKey.for_dtype(dt)
    n = 12_345
    cases = {
        "all_equal": np.full(n, 3, dtype=dt),
        "two_values": rng.choice(np.array([-5, 7]).astype(dt) if dt != "u4" else np.array([5, 1 << 30], dtype=dt), n),
        "narrow": rng.integers(0, 200, n).astype(dt),
        "sorted": np.sort(rng.integers(0, 1 << 20, n)).astype(dt),
        "reverse": np.sort(rng.integers(0, 1 << 20, n))[::-1].astype(dt),
    }
    for name, a in cases.items():
        got, sh, _ = sort_both(np.ascontiguousarray(a), key)
        assert sh["kernel_launches"] == 1
        assert got.tobytes() == o.sorted_copy(a, key).tobytes(), (dt, name)


@pytest.mark.parametrize("rec,field,levels", [(REC8, "key", 4), (REC8, "key", 2), (REC8F, "key", 4)])
def test_small_key_plus_payload_is_stable(rec, field, levels):
    rng = np.random.default_rng(61)
    key = RadixKey.field(rec, field, levels)
    for n in (129, 5_000, CAP):
        a = np.zeros(n, dtype=rec)
        a[field] = rng.integers(-40, 40, n).astype(rec.fields[field][0])  # heavy duplicates, negatives
        a["idx"] = np.arange(n)
        got, sh, _ = sort_both(a, key, stable=True)
        assert sh["kernel_launches"] == 1
        exp = o.stable_sorted(a, key)
        assert got.tobytes() == exp.tobytes() == o.numpy_stable_sorted(a, key).tobytes(), (rec, levels, n)


def test_small_partial_levels_ignore_high_bytes():
    rng = np.random.default_rng(8)
    for dt, levels in (("u4", 1), ("u4", 3), ("u8", 5), ("i8", 7)):
        key = RadixKey.for_dtype(dt, levels=levels)
        a = rand_bits(rng, dt, 10_001)
        got, sh, _ = sort_both(a, key, stable=True)
        assert sh["kernel_launches"] == 1 and sh["levels_run"] == levels
        assert got.tobytes() == o.stable_sorted(a, key).tobytes(), (dt, levels)


def test_other_shapes_keep_the_pass_path():
    """2-byte keys, 12-byte records and explicit variants are not K3 shapes: same answers, LSD passes."""
    rng = np.random.default_rng(9)
    a = rand_bits(rng, "u2", 5_000)
    got, sh, _ = sort_both(a, RadixKey.for_dtype("u2"))
    assert sh["kernel_launches"] > 1 and got.tobytes() == np.sort(a).tobytes()
    b = rand_bits(rng, "u8", 5_000)
    h = b.copy()
    bb = radix_sort_builder(h).with_variant(0)
    bb.sort()
    assert bb.stats["kernel_launches"] > 1 and np.array_equal(h, np.sort(b))


# ---------------------------------------------------------------------------------------------
# Mid-size inputs: one MSD pass on the top active level + K3 on its 256 buckets (stats.hybrid == 2)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dt", ["u4", "u8", "i4", "i8", "f4", "f8"])
def test_mid_size_one_msd_pass_then_k3(dt):
    rng = np.random.default_rng(1200 + np.dtype(dt).num)
    key = RadixKey.for_dtype(dt)
    for n in (CAP + 1, 50_000, 300_001, 1_000_003, 3_900_000):
        a = rand_bits(rng, dt, n)
        got, sh, sd = sort_both(a, key)
        assert got.tobytes() == o.sorted_copy(a, key).tobytes(), (dt, n)
        # five or more active levels always take it, four only from 2^20 records (launch-latency break-even)
        expect = 2 if (key.levels >= 5 or n >= (1 << 20)) else 0
        assert sh["hybrid"] == expect and sd["hybrid"] == expect, (dt, n, sh)
        assert sh["levels_run"] == key.levels
        if expect:
            assert sh["local_levels"] == key.levels - 1


def test_mid_size_route_falls_back_when_a_bucket_does_not_fit(monkeypatch):
    rng = np.random.default_rng(71)
    key = RadixKey.for_dtype("u8")
    # 40-bit keys: five active levels, 256 values on the top one
    a = rng.integers(0, 1 << 40, 200_000, dtype="u8")
    got, sh, _ = sort_both(a, key)
    assert sh["hybrid"] == 2 and sh["levels_run"] == 5 and sh["levels_skipped"] == 3 and sh["local_levels"] == 4
    assert got.tobytes() == np.sort(a).tobytes()
    # 36-bit keys: the top active level has 16 values -> 25 000 records per bucket: LSD passes
    b = rng.integers(0, 1 << 36, 400_000, dtype="u8")
    got, sh, _ = sort_both(b, key)
    assert sh["hybrid"] == 0 and sh["levels_run"] == 5
    assert got.tobytes() == np.sort(b).tobytes()
    # three active levels: LSD passes are cheaper
    c = rng.integers(0, 1 << 24, 100_000, dtype="u8")
    got, sh, _ = sort_both(c, key)
    assert sh["hybrid"] == 0 and got.tobytes() == np.sort(c).tobytes()
    # switched off: same bytes from the LSD passes
    monkeypatch.setenv("B200SORT_MSD1", "0")
    got, sh, _ = sort_both(a, key)
    assert sh["hybrid"] == 0 and got.tobytes() == np.sort(a).tobytes()


@pytest.mark.parametrize("nfast", ["0", "1", "2", None])
def test_mid_size_duplicates_skew_and_patterns(monkeypatch, nfast):
    """Every fast-level setting of K3 (insertion fix-up of runs, slow-path fallback on long runs)
    gives the oracle's bytes on duplicate-heavy, masked, reversed and nearly sorted inputs."""
    if nfast is not None:
        monkeypatch.setenv("B200SORT_HYBRID_NFAST", nfast)
    rng = np.random.default_rng(5)
    n = 700_001
    srt = np.sort(rng.integers(0, 2**64, n, dtype="u8"))
    nearly = srt.copy()
    nearly[-n // 10:] = rng.integers(0, 2**64, len(nearly[-n // 10:]), dtype="u8")
    cases = {
        "dups": (rng.integers(0, 3000, n, dtype="u8") * np.uint64(0x0001000100010001), "u8"),
        "masked": (rng.integers(0, 2**64, n, dtype="u8") & np.uint64(0xFF00FFFF00FF00FF), "u8"),
        "reverse": (srt[::-1].copy(), "u8"),
        "nearly_sorted": (nearly, "u8"),
        "i64_narrow": (rng.integers(-2**33, 2**33, n, dtype="i8"), "i8"),
        "f32_unit": (rng.random(n, dtype="f4") - np.float32(0.5), "f4"),
        "f64_normal": (rng.standard_normal(n), "f8"),
        "u32_dups": (rng.integers(0, 5000, n, dtype="u4") * np.uint32(0x00010001), "u4"),
    }
    for name, (a, dt) in cases.items():
        key = RadixKey.for_dtype(dt)
        got, sh, _ = sort_both(a, key)
        assert got.tobytes() == o.sorted_copy(a, key).tobytes(), (name, nfast, sh)


@pytest.mark.parametrize("rec,field,levels", [(REC8, "key", 4), (REC8, "key", 3), (REC8F, "key", 4)])
def test_mid_size_key_plus_payload_is_stable(rec, field, levels):
    rng = np.random.default_rng(62)
    key = RadixKey.field(rec, field, levels)
    for n in (40_000, 900_001):
        a = np.zeros(n, dtype=rec)
        kd = rec.fields[field][0]
        if kd.kind == "f":
            a[field] = (rng.integers(-3000, 3000, n) * 0.37).astype(kd)
        else:
            a[field] = rng.integers(-(1 << 22), 1 << 22, n).astype(kd)  # duplicates, negatives, 3+ active levels
        a["idx"] = np.arange(n)
        got, sh, _ = sort_both(a, key, stable=True)
        exp = o.stable_sorted(a, key)
        assert got.tobytes() == exp.tobytes() == o.numpy_stable_sorted(a, key).tobytes(), (rec, levels, n, sh)
