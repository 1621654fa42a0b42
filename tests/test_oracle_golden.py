"""Pins the oracle (oracle/, the C restatement of the reference) against every known
answer the reference's own tests / doctests / examples hold for the sort path
(tests/golden/reference_vectors.json, transcribed + source-checked by
tests/golden/make_reference_vectors.py) and against an independent numpy model."""
import json
import os
import struct

import numpy as np
import pytest

import oracle_ffi as o
from rdst_b200 import RadixKey

G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_vectors.json")))
NPT = {"i32": "i4", "u64": "u8", "u8": "u1", "u32": "u4", "i64": "i8"}
ALL_TUNERS = [("standard", "Lsb"), ("low_memory", "Lsb"), ("single_threaded", "Lsb")] + \
             [("single_algo", a) for a in o.ALGO]


def test_fixture_was_checked_against_the_reference_source():
    assert G["source_checked"] is True


@pytest.mark.parametrize("vec", G["sort_vectors"], ids=lambda v: v["cite"])
@pytest.mark.parametrize("tuner,algo", ALL_TUNERS)
def test_sort_vectors(vec, tuner, algo):
    a = np.array(vec["input"], dtype=NPT[vec["type"]])
    o.sort(a, RadixKey.for_dtype(a.dtype), tuner=tuner, algo=algo)
    assert a.tolist() == vec["expected"]


@pytest.mark.parametrize("tuner,algo", ALL_TUNERS)
def test_byte_array_vectors(tuner, algo):
    for vec in G["byte_array_vectors"]:
        a = np.array(vec["input"], dtype=np.uint8)
        o.sort(a, RadixKey.byte_array(vec["n"]), tuner=tuner, algo=algo)
        assert a.tolist() == vec["expected"]


@pytest.mark.parametrize("vec", G["custom_key_vectors"], ids=lambda v: v["cite"])
def test_custom_key_vectors(vec):
    a = np.array(vec["input"], dtype=np.uint8)
    key = RadixKey.custom(vec["elem_size"], vec["level_offsets"])
    o.sort(a, key)  # 3 elements -> comparative sort on the declared levels (sorter.rs:35-38)
    assert a.tolist() == vec["expected"]
    b = np.array(vec["input"], dtype=np.uint8)
    o.sort(b, key, tuner="single_algo", algo="Lsb")
    assert b.tolist() == vec["expected"]


@pytest.mark.parametrize("vec", G["tile_counts_sorted"], ids=lambda v: f'{v["cite"]}-{v["data"]}-{v["tile_size"]}')
def test_tile_counts_sorted(vec):
    a = np.array(vec["data"], dtype=np.uint8)
    got, counts = o.tile_counts_sorted(a, RadixKey.for_dtype("u1"), vec["tile_size"], vec["level"])
    assert got == vec["sorted"]
    assert counts.tolist() == np.bincount(a, minlength=256).tolist()


def test_f32_level_known_answers():
    k = RadixKey.for_dtype("f4")
    for e in G["f32_levels"]:
        raw = struct.pack("<I", e["bits"])
        assert [o.get_level_builtin("f32", raw, l) for l in range(4)] == e["levels"]
        assert [o.get_level(raw, k, l) for l in range(4)] == e["levels"]  # descriptor form agrees
    for bits, hx in G["f32_mapped_hex"]["pairs"]:
        raw = struct.pack("<I", bits)
        assert "".join(f"{o.get_level_builtin('f32', raw, l):02x}" for l in (3, 2, 1, 0)) == hx


def test_f64_level_known_answers():
    k = RadixKey.for_dtype("f8")
    for e in G["f64_levels"]:
        raw = struct.pack("<Q", e["bits"])
        assert [o.get_level_builtin("f64", raw, l) for l in range(8)] == e["levels"]
        assert [o.get_level(raw, k, l) for l in range(8)] == e["levels"]


def test_int_level_known_answers():
    iv = G["int_levels"]
    for e in iv["u32"]:
        raw = struct.pack("<I", e["value"])
        assert [o.get_level_builtin("u32", raw, l) for l in range(4)] == e["levels"]
        assert [o.get_level(raw, RadixKey.for_dtype("u4"), l) for l in range(4)] == e["levels"]
    for e in iv["i32"]:
        raw = struct.pack("<i", e["value"])
        assert [o.get_level_builtin("i32", raw, l) for l in range(4)] == e["levels"]
        assert [o.get_level(raw, RadixKey.for_dtype("i4"), l) for l in range(4)] == e["levels"]
    for e in iv["i64"]:
        raw = struct.pack("<q", e["value"])
        assert [o.get_level_builtin("i64", raw, l) for l in range(8)] == e["levels"]
        assert [o.get_level(raw, RadixKey.for_dtype("i8"), l) for l in range(8)] == e["levels"]


def test_wide_and_small_builtin_levels():
    rng = np.random.default_rng(7)
    for _ in range(200):
        raw = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
        v = int.from_bytes(raw, "little")
        for l in range(16):
            assert o.get_level_builtin("u128", raw, l) == (v >> (8 * l)) & 0xFF
            assert o.get_level_builtin("i128", raw, l) == ((v ^ (1 << 127)) >> (8 * l)) & 0xFF
            assert o.get_level(raw, RadixKey.u128(), l) == (v >> (8 * l)) & 0xFF
            assert o.get_level(raw, RadixKey.i128(), l) == ((v ^ (1 << 127)) >> (8 * l)) & 0xFF
        for t, w, sx in (("u8", 1, 0), ("i8", 1, 0x80), ("u16", 2, 0), ("i16", 2, 0x8000)):
            x = int.from_bytes(raw[:w], "little")
            for l in range(w):
                assert o.get_level_builtin(t, raw[:w], l) == ((x ^ sx) >> (8 * l)) & 0xFF
    assert o.get_level_builtin("u32", b"\0\0\0\0", 4) == -1  # level >= LEVELS (radix_key.rs:18)


def test_f32_random_bits_match_total_cmp():
    """reference src/radix_sort.rs:206-218: 1000 x 10 random-bit f32/f64 vs total_cmp order."""
    rng = np.random.default_rng(11)
    for dt, it in (("f4", "i4"), ("f8", "i8")):
        k = RadixKey.for_dtype(dt)
        for _ in range(300):
            bits = rng.integers(0, 256, 10 * np.dtype(dt).itemsize, dtype=np.uint8).view(dt)
            got = o.sorted_copy(bits, k)
            s = bits.view(it).astype(object)
            w = np.dtype(dt).itemsize * 8
            keyed = sorted(range(10), key=lambda i: int(s[i]) ^ (((int(s[i]) >> (w - 1)) & ((1 << w) - 1)) >> 1))
            exp = bits[keyed]
            assert got.tobytes() == exp.tobytes()
