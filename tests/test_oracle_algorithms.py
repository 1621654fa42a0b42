"""The reference's per-algorithm suites replayed on the oracle: every
algorithm through the router (SingleAlgoTuner, reference src/test_utils.rs:40-49,264-278),
the size ladder (:63-95, trimmed for CPU time), bimodal shift inputs (:51-61) and the u32
bit-mask patterns (:148-262), checked against an independent numpy model."""
import numpy as np
import pytest

import oracle_ffi as o
from rdst_b200 import RadixKey

SIZES = [0, 1, 10, 100, 128, 129, 400, 5_000, 50_000, 260_000, 300_000, 1_000_000]
MASKS = [0xFFFF_FFFF, 0xFFFF_FF00, 0xFFFF_00FF, 0xFF00_FFFF, 0x00FF_FFFF, 0x0000_FFFF, 0xFFFF_0000,
         0x8000_0000, 0x1, 0xFFFF_FFFE, 0x7FFF_FFFF, 0xAAAA_AAAA, 0x5555_5555]


def gen_inputs(rng, dt, n, shift):
    """gen_inputs(n, shift): random bits, first half >>= shift, second half <<= shift."""
    a = rng.integers(0, 256, n * np.dtype(dt).itemsize, dtype=np.uint8).view(dt).copy()
    if shift:
        u = a.view("u%d" % a.dtype.itemsize)
        u[: n // 2] >>= np.array(shift, dtype=u.dtype)
        u[n // 2:] <<= np.array(shift, dtype=u.dtype)
    return a


@pytest.mark.parametrize("algo", list(o.ALGO))
@pytest.mark.parametrize("dt", ["u1", "u2", "u4", "u8"])
def test_sort_comparison_suite(algo, dt):
    rng = np.random.default_rng(hash((algo, dt)) & 0xFFFF)
    k = RadixKey.for_dtype(dt)
    bits = np.dtype(dt).itemsize * 8
    for n in SIZES:
        if algo == "Comparative" and n > 50_000:
            continue
        for shift in (0, bits // 2):
            a = gen_inputs(rng, dt, n, shift)
            got = o.sorted_copy(a, k, tuner="single_algo", algo=algo)
            assert got.tobytes() == np.sort(a, kind="stable").tobytes(), (algo, dt, n, shift)


@pytest.mark.parametrize("tuner", ["standard", "low_memory", "single_threaded"])
@pytest.mark.parametrize("dt", ["u4", "i4", "i8", "f4", "f8", "i2", "i1"])
def test_full_path_tuners(tuner, dt):
    rng = np.random.default_rng(3)
    k = RadixKey.for_dtype(dt)
    for n in (100, 5_000, 300_000, 6_000_000 if dt in ("u4", "f4") else 600_000):
        for shift in (0, np.dtype(dt).itemsize * 4):
            a = gen_inputs(rng, dt, n, shift)
            got = o.sorted_copy(a, k, tuner=tuner, multi_threaded=(tuner != "single_threaded"))
            assert got.tobytes() == o.numpy_stable_sorted(a, k).tobytes(), (tuner, dt, n)
            assert o.is_sorted(got, k)


@pytest.mark.parametrize("algo", ["Lsb", "LrLsb", "MtLsb", "Ska", "Recombinating", "Scanning", "Regions", "MtOop"])
def test_u32_patterns(algo):
    rng = np.random.default_rng(5)
    k = RadixKey.for_dtype("u4")
    bases = [np.full(128, 0xFFFFFFFF, dtype="u4"), rng.integers(0, 2**32, 128, dtype="u4"),
             rng.integers(0, 2**32, 128_000, dtype="u4"), rng.integers(0, 2**32, 4, dtype="u4")]
    for base in bases:
        for m in MASKS:
            a = base & np.uint32(m)
            got = o.sorted_copy(a, k, tuner="single_algo", algo=algo)
            assert got.tobytes() == np.sort(a).tobytes(), (algo, hex(m), len(a))
    # heavy skew: everything zero except the first five = 0x77 in the top byte
    a = np.zeros(128_000, dtype="u4")
    a[:5] = 0x7700_0000
    got = o.sorted_copy(a, k, tuner="single_algo", algo=algo)
    assert got.tobytes() == np.sort(a).tobytes()


def test_issue5_single_tile_400():  # reference src/sorts/mt_lsb_sort.rs:323-328
    rng = np.random.default_rng(9)
    a = rng.integers(0, 2**32, 400, dtype="u4")
    got = o.sorted_copy(a, RadixKey.for_dtype("u4"), tuner="single_algo", algo="MtLsb")
    assert got.tobytes() == np.sort(a).tobytes()


REC12 = np.dtype([("key", "<u4"), ("idx", "<u8")], align=False)


@pytest.mark.parametrize("algo", ["Lsb", "LrLsb", "MtLsb"])
def test_stable_algorithms_keep_payload_order(algo):
    """Stable order (reference lsb_sort.rs:5-7, mt_lsb_sort.rs:12-15, out_of_place_sort.rs:22-33)."""
    rng = np.random.default_rng(13)
    for n in (129, 5_000, 400_000):
        a = np.zeros(n, dtype=REC12)
        a["key"] = rng.integers(0, 1 << 12, n)  # many duplicates; LEVELS = 3 partial key
        a["idx"] = np.arange(n)
        k = RadixKey.field(REC12, "key", levels=3)
        got = o.sorted_copy(a, k, tuner="single_algo", algo=algo)
        exp = o.numpy_stable_sorted(a, k)
        assert got.tobytes() == exp.tobytes(), (algo, n)


def test_unstable_algorithms_sorted_and_permutation():
    rng = np.random.default_rng(17)
    n = 300_000
    a = np.zeros(n, dtype=REC12)
    a["key"] = rng.integers(0, 1 << 20, n)
    a["idx"] = np.arange(n)
    k = RadixKey.field(REC12, "key", levels=3)
    for algo in ("Ska", "Recombinating", "Scanning", "Regions", "MtOop"):
        got = o.sorted_copy(a, k, tuner="single_algo", algo=algo)
        assert o.is_sorted(got, k)
        assert np.array_equal(np.sort(got["idx"]), a["idx"])  # multiset permutation
        assert np.array_equal(a["key"][got["idx"]], got["key"])


def test_counts_and_passes():
    rng = np.random.default_rng(19)
    a = rng.integers(0, 2**32, 100_000, dtype="u4")
    k = RadixKey.for_dtype("u4")
    for l in range(4):
        counts, srt, first, last = o.get_counts(a, k, l)
        d = (a >> np.uint32(8 * l)) & np.uint32(0xFF)
        assert counts.tolist() == np.bincount(d, minlength=256).tolist()
        assert srt is False and first == d[0] and last == d[-1]
        for lr in (False, True):
            out = o.out_of_place_pass(a, k, l, lr)
            assert out.tobytes() == a[np.argsort(d, kind="stable")].tobytes()
    s = np.sort(a)
    assert o.get_counts(s, k, 3)[1] is True


def test_levels_zero_rejected():
    a = np.arange(4, dtype="u4")
    with pytest.raises(RuntimeError):
        o.sort(a, RadixKey.for_dtype("u4", levels=0))
