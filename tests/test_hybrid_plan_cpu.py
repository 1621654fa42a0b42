"""CPU-only checks of the hybrid route's digit planner (b200_hybrid_plan, host logic of
rdst_b200/csrc/b200sort.cu::plan_windows): the bit windows it picks from the per-level histograms
must (1) cover every varying bit of the key image exactly once, most significant digit first, and
(2) order the records exactly as the key does — sorting by (digit 0, digit 1, local windows) with a
stable numpy sort reproduces the oracle's answer.  No kernels run."""
import ctypes as C

import numpy as np
import pytest

import oracle_ffi as o
from rdst_b200 import RadixKey, _abi


def key_image(arr, key):
    """64-bit order-preserving image of the key: level l at bits 8l..8l+7 (what RadixKey::get_level
    yields, reference src/radix_key.rs:1-5, concatenated)."""
    km = o.key_matrix(arr, key).astype(np.uint64)
    img = np.zeros(km.shape[0], dtype=np.uint64)
    for l in range(key.levels):
        img |= km[:, l] << np.uint64(8 * l)
    return img


def level_histograms(arr, key):
    km = o.key_matrix(arr, key)
    return np.stack([np.bincount(km[:, l], minlength=256) for l in range(key.levels)]).astype(np.uint64)


def plan(arr, key):
    d = key.to_desc()
    hist = np.ascontiguousarray(level_histograms(arr, key))
    wp = _abi.WindowPlan()
    n = arr.nbytes // key.elem_size
    _abi.check(_abi.lib().b200_hybrid_plan(C.byref(d), hist.ctypes.data_as(C.POINTER(C.c_uint64)), n, C.byref(wp)))
    return wp


def window(img, lo, width):
    return (img >> np.uint64(lo)) & np.uint64((1 << width) - 1)


def cases():
    rng = np.random.default_rng(4242)
    n = 20_000
    yield "u64_full", "u8", rng.integers(0, 2**64, n, dtype="u8"), (56, 48, True)
    yield "u64_61_bits", "u8", rng.integers(0, 2**61, n, dtype="u8"), (53, 45, False)
    yield "u64_45_bits", "u8", rng.integers(0, 2**45, n, dtype="u8"), (37, 29, False)
    yield "u64_40_bits", "u8", rng.integers(0, 2**40, n, dtype="u8"), (32, 24, True)
    yield "u64_36_bits_across_words", "u8", rng.integers(0, 2**36, n, dtype="u8"), (28, 20, False)
    yield "u64_rank_range", "u8", rng.integers(3 << 58, 5 << 58, n, dtype="u8"), (53, 45, False)
    # constant bits between the varying ones: the second digit starts at the next varying bit
    gap = (rng.integers(0, 2**8, n, dtype="u8") << np.uint64(52)) | rng.integers(0, 2**30, n, dtype="u8")
    yield "u64_gap", "u8", gap, (52, 22, False)
    yield "i64_full", "i8", rng.integers(-2**63, 2**63 - 1, n, dtype="i8"), (56, 48, True)
    yield "i64_positive_50_bits", "i8", rng.integers(0, 2**50, n, dtype="i8"), (42, 34, False)
    yield "i64_narrow", "i8", rng.integers(-2**40, 2**40, n, dtype="i8"), (56, 48, True)
    yield "u32_full", "u4", rng.integers(0, 2**32, n, dtype="u4"), (24, 16, True)
    yield "u32_27_bits", "u4", rng.integers(0, 2**27, n, dtype="u4"), (19, 11, False)
    yield "f32_one_to_two", "f4", (rng.random(n, dtype="f4") + np.float32(1.0)), (15, 7, False)
    yield "f64_unit_interval", "f8", rng.random(n), None
    yield "f64_normal", "f8", rng.standard_normal(n), None
    yield "f32_bits", "f4", rng.integers(0, 2**32, n, dtype="u4").view("f4"), (24, 16, True)


CASES = list(cases())


@pytest.mark.parametrize("name,dt,a,expect", CASES, ids=[c[0] for c in CASES])
def test_windows_cover_the_varying_bits_and_order_like_the_key(name, dt, a, expect):
    key = RadixKey.for_dtype(dt)
    wp = plan(a, key)
    assert wp.eligible == 1, name
    img = key_image(a, key)
    # the planner's idea of the varying bits is the truth
    vary = int(np.bitwise_or.reduce(img) & ~np.bitwise_and.reduce(img))
    assert wp.vary == vary
    d1 = (wp.digit_lo[0], wp.digit_width[0])
    d2 = (wp.digit_lo[1], wp.digit_width[1])
    local = [(wp.local_lo[i], wp.local_width[i]) for i in range(wp.nlocal)]
    if expect is not None:
        assert (d1[0], d2[0], bool(wp.aligned)) == expect, (name, d1, d2, wp.aligned)
    # digit 0 ends on the top varying bit; widths fit a 256-bin radix; windows descend without overlap
    assert d1[0] + d1[1] - 1 == vary.bit_length() - 1 and d1[1] == 8
    wins = [d1, d2] + local[::-1]  # most significant first
    for (lo, w), (lo2, w2) in zip(wins, wins[1:]):
        assert 1 <= w2 <= 8 and lo2 + w2 <= lo
    covered = 0
    for lo, w in wins:
        covered |= ((1 << w) - 1) << lo
    assert vary & ~covered == 0, "a varying bit is in no window"
    # sorting by the windows (stable, least significant first) == sorting by the key
    order = np.arange(len(a))
    for lo, w in wins[::-1]:
        order = order[np.argsort(window(img[order], lo, w), kind="stable")]
    got = a.view(np.uint8).reshape(len(a), -1)[order].reshape(-1)
    assert got.tobytes() == o.stable_sorted(a, key).view(np.uint8).tobytes()


def test_plan_says_no_when_the_route_cannot_apply():
    rng = np.random.default_rng(7)
    # a single varying byte: nothing below the first digit
    a = rng.integers(0, 256, 1000, dtype="u8")
    assert plan(a, RadixKey.for_dtype("u8")).eligible == 0
    # all keys equal
    assert plan(np.full(1000, 77, dtype="u8"), RadixKey.for_dtype("u8")).eligible == 0
    # 16-byte keys and 12-byte records are not covered by the route
    assert plan(np.zeros(2000, dtype="u8"), RadixKey.u128()).eligible == 0
    rec = np.dtype([("key", "<u4"), ("idx", "<u8")])
    r = np.zeros(1000, dtype=rec)
    r["key"] = rng.integers(0, 1 << 30, 1000)
    assert plan(r, RadixKey.field(rec, "key", 4)).eligible == 0


def test_key_plus_payload_record_plans_on_the_key_word_only():
    rec = np.dtype([("key", "<u4"), ("idx", "<u4")])
    rng = np.random.default_rng(9)
    r = np.zeros(5000, dtype=rec)
    r["key"] = rng.integers(0, 1 << 29, len(r))
    r["idx"] = np.arange(len(r))
    wp = plan(r, RadixKey.field(rec, "key", 4))
    assert wp.eligible == 1
    assert wp.digit_lo[0] == 21 and wp.digit_lo[1] == 13
    assert wp.vary < (1 << 32)
