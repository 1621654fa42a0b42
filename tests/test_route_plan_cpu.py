"""CPU-only checks of the host planner's route choice (b200_route_plan, include/b200sort.h): the GPU
meaning of rdst's Tuner::pick_algorithm (reference src/tuner.rs:33-35; the reference's own rules
are size cut-offs and count thresholds, src/sorter.rs:33-45, src/tuners/standard_tuner.rs:13-63).
Histograms come from numpy on real arrays, or are synthetic for sizes no CPU test can hold.
No kernels run; tests/test_gpu_small_sort.py checks on a B200 that b200_stats reports the same routes."""
import ctypes as C

import numpy as np
import pytest

import oracle_ffi as o
from rdst_b200 import RadixKey, _abi

REC12 = np.dtype([("key", "<u4"), ("lo", "<u4"), ("hi", "<u4")])
REC8 = np.dtype([("idx", "<u4"), ("key", "<i4")])
CAP = 16_384


def level_histograms(arr, key):
    km = o.key_matrix(arr, key)
    return np.stack([np.bincount(km[:, l], minlength=256) for l in range(key.levels)]).astype(np.uint64)


def route(key, n, hist=None):
    d = key.to_desc()
    r = _abi.Route()
    hp = None
    if hist is not None:
        hist = np.ascontiguousarray(hist, dtype=np.uint64)
        hp = hist.ctypes.data_as(C.POINTER(C.c_uint64))
    _abi.check(_abi.lib().b200_route_plan(C.byref(d), hp, n, C.byref(r)))
    return _abi.ROUTE_NAMES[r.route], r


def route_of(arr, key):
    n = arr.nbytes // key.elem_size
    return route(key, n, level_histograms(arr, key))


def uniform_hist(levels, n, active):
    """Synthetic per-level histograms: `active` levels uniform over 256 buckets, the others one bucket."""
    h = np.zeros((levels, 256), dtype=np.uint64)
    for l in range(levels):
        if l in active:
            h[l, :] = n // 256
            h[l, 0] += n - (n // 256) * 256
        else:
            h[l, 0] = n
    return h


def test_trivial_and_small_inputs_need_no_histogram():
    assert route(RadixKey.for_dtype("u8"), 0)[0] == "trivial"
    assert route(RadixKey.for_dtype("u8"), 1)[0] == "trivial"  # reference src/radix_sort_builder.rs:151
    for dt in ("u4", "i4", "f4", "u8", "i8", "f8"):
        key = RadixKey.for_dtype(dt)
        for n in (2, 129, CAP):
            name, r = route(key, n)
            assert name == "small" and r.levels_run == key.levels, (dt, n)
    name, r = route(RadixKey.field(REC8, "key", 3), 5_000)
    assert name == "small" and r.levels_run == 3
    # shapes K3 does not take: 1- and 2-byte keys, 12-byte records, [u8; N], 16-byte words
    rng = np.random.default_rng(1)
    for key, a in ((RadixKey.for_dtype("u2"), rng.integers(0, 1 << 16, 5_000, dtype="u2")),
                   (RadixKey.for_dtype("u1"), rng.integers(0, 256, 5_000, dtype="u1")),
                   (RadixKey.field(REC12, "key", 4), rng.integers(0, 256, 12 * 5_000, dtype="u1")),
                   (RadixKey.byte_array(5), rng.integers(0, 256, 5 * 5_000, dtype="u1")),
                   (RadixKey.u128(), rng.integers(0, 256, 16 * 5_000, dtype="u1"))):
        name, r = route_of(a, key)
        assert name == "lsd", key
    with pytest.raises(Exception):
        route(RadixKey.for_dtype("u8"), CAP + 1)  # past the small path the histograms are required


def test_mid_size_route_rule():
    rng = np.random.default_rng(2)
    k8, k4 = RadixKey.for_dtype("u8"), RadixKey.for_dtype("u4")
    name, r = route_of(rng.integers(0, 2**64, 300_000, dtype="u8"), k8)
    assert name == "msd1" and r.msd_level == 7 and r.levels_run == 8 and 0 < r.max_bucket <= CAP
    # 40-bit keys: the MSD level is the top ACTIVE level, trivial levels are skipped
    name, r = route_of(rng.integers(0, 1 << 40, 200_000, dtype="u8"), k8)
    assert name == "msd1" and r.msd_level == 4 and r.active_levels == 5 and r.levels_run == 5
    # 36-bit keys: 16 values on the top active level -> 25 000 per bucket does not fit an SM
    name, r = route_of(rng.integers(0, 1 << 36, 400_000, dtype="u8"), k8)
    assert name == "lsd" and r.levels_run == 5
    # four active levels pay only from 2^20 records on; three never
    assert route_of(rng.integers(0, 2**32, 500_000, dtype="u4"), k4)[0] == "lsd"
    assert route_of(rng.integers(0, 2**32, (1 << 20) + 5, dtype="u4"), k4)[0] == "msd1"
    assert route_of(rng.integers(0, 1 << 24, 2_000_000, dtype="u4"), k4)[0] == "lsd"
    # signed and float keys: same rule on the transformed digits
    assert route_of(rng.integers(-2**63, 2**63 - 1, 100_000, dtype="i8"), RadixKey.for_dtype("i8"))[0] == "msd1"
    assert route_of(rng.integers(0, 256, 8 * 100_000, dtype="u1").view("f8"), RadixKey.for_dtype("f8"))[0] == "msd1"
    # a skewed top level (normal floats: few exponents) overflows a bucket -> passes
    assert route_of(rng.standard_normal(2_000_000), RadixKey.for_dtype("f8"))[0] == "lsd"
    # past ~4 M uniform keys a bucket no longer fits
    assert route(k8, 6_000_000, uniform_hist(8, 6_000_000, range(8)))[0] == "lsd"
    assert route(k8, 4_000_000, uniform_hist(8, 4_000_000, range(8)))[0] == "msd1"


def test_hybrid_route_rule_on_synthetic_histograms():
    k8, k4 = RadixKey.for_dtype("u8"), RadixKey.for_dtype("u4")
    n = 1_000_000_000
    name, r = route(k8, n, uniform_hist(8, n, range(8)))  # BASELINE configs[1]
    assert name == "hybrid" and r.levels_run == 8
    assert route(k8, n, uniform_hist(8, n, range(5)))[0] == "hybrid"          # 40-bit keys: five active levels
    # records that are nothing but their key take the unstable flavour of the route (K2-U + K3-U), which pays
    # from four active levels and far smaller inputs on
    assert route(k8, n, uniform_hist(8, n, range(4)))[0] == "hybrid"
    assert route(k8, n, uniform_hist(8, n, range(3)))[0] == "lsd"             # three: passes are cheaper
    assert route(k4, n, uniform_hist(4, n, range(4)))[0] == "hybrid"          # u32, 1 B keys
    assert route(k4, 100_000_000, uniform_hist(4, 100_000_000, range(4)))[0] == "lsd"   # u32 below 128 Mi records
    assert route(k4, 500_000_000, uniform_hist(4, 500_000_000, range(4)))[0] == "hybrid"
    assert route(k8, 200_000_000, uniform_hist(8, 200_000_000, range(8)))[0] == "hybrid"
    assert route(k8, 50_000_000, uniform_hist(8, 50_000_000, range(8)))[0] == "hybrid"
    assert route(k8, 25_000_000, uniform_hist(8, 25_000_000, range(8)))[0] == "lsd"     # u64 below 32 Mi records
    assert route(k8, 5_000_000_000, uniform_hist(8, 5_000_000_000, range(8)))[0] == "lsd"  # 32-bit offsets in K3
    # records with a payload stay on the stable flavour: five active levels, 256 Mi records
    rec8 = RadixKey.field(REC8, "key", 4)
    assert route(rec8, n, uniform_hist(4, n, range(4)))[0] == "lsd"
    rec16 = RadixKey("uint", 8, 6, 0, 8)  # u64 with LEVELS = 6: the two top bytes are payload
    assert route(rec16, n, uniform_hist(6, n, range(6)))[0] == "hybrid"
    assert route(rec16, 200_000_000, uniform_hist(6, 200_000_000, range(6)))[0] == "lsd"
    # two values on the top level: the expected segment (n / 2 / 256) does not fit an SM
    h = uniform_hist(8, n, range(8))
    h[7, :] = 0
    h[7, 0] = h[7, 1] = n // 2
    name, r = route(k8, n, h)
    assert name in ("hybrid", "lsd")  # bit windows re-centre the digits on the varying bits
    # skewed keys (rdst_b200/csrc/heavy.cuh): a tenth of the records share one key -> on both MSD digits one bucket
    # holds a tenth plus its uniform share, the expected segment is far past an SM, yet the route is kept for
    # key-only records (the oversized segment is split around a pivot) ...
    h = np.zeros((8, 256), dtype=np.uint64)
    h[:, :] = (n - n // 10) // 256
    h[:, 17] += n // 10
    h[:, 0] += np.uint64(n) - h[0].sum()
    assert int(h[0].sum()) == n
    assert route(k8, n, h)[0] == "hybrid"
    # ... but not for records with a payload (equal keys are then different records, nothing may be filled in) ...
    assert route(rec16, n, h[:6])[0] == "lsd"
    # ... and not for clustered keys: a bucket with half of the records on the second digit is not duplicates
    h2 = uniform_hist(8, n, range(8))
    h2[6, :] = (n // 2) // 256
    h2[6, 5] += np.uint64(n) - h2[6].sum()
    assert int(h2[6].sum()) == n and int(h2[6].max()) > n // 4
    assert route(k8, n, h2)[0] == "lsd"
    all_one = uniform_hist(8, n, [])
    name, r = route(k8, n, all_one)
    assert name == "lsd" and r.levels_run == 0 and r.active_levels == 0  # every level trivial: nothing to run


def test_environment_knobs_switch_routes_off(monkeypatch):
    rng = np.random.default_rng(3)
    k8 = RadixKey.for_dtype("u8")
    a = rng.integers(0, 2**64, 100_000, dtype="u8")
    assert route_of(a, k8)[0] == "msd1"
    monkeypatch.setenv("B200SORT_MSD1", "0")
    assert route_of(a, k8)[0] == "lsd"
    monkeypatch.delenv("B200SORT_MSD1")
    assert route(k8, 1_000)[0] == "small"
    monkeypatch.setenv("B200SORT_SMALL", "0")
    h = level_histograms(a[:1_000], k8)
    assert route(k8, 1_000, h)[0] in ("lsd", "msd1")
    monkeypatch.delenv("B200SORT_SMALL")
    monkeypatch.setenv("B200SORT_HYBRID", "0")  # "no K3 anywhere": passes only
    assert route(k8, 1_000, h)[0] == "lsd"
    assert route_of(a, k8)[0] == "lsd"
    n = 1_000_000_000
    assert route(k8, n, uniform_hist(8, n, range(8)))[0] == "lsd"
