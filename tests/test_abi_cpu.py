"""CPU-only checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/b200sort.h declares, the descriptor lowering agrees with the oracle's get_level,
and the host-side mirror keeps the reference's argument/error behaviour.  No kernels run."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle_ffi as o
import rdst_b200
from rdst_b200 import RadixKey, _abi, radix_sort_builder, radix_sort_unstable, tuner

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "b200sort.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = header_functions()
    assert len(names) >= 14
    l = C.CDLL(_abi.LIB_PATH)
    for n in names:
        assert hasattr(l, n), f"libb200sort.so does not export {n}"
    # and the Python binding covers the whole header
    assert set(names) == set(_abi.SYMBOLS), set(names) ^ set(_abi.SYMBOLS)


def test_version_and_variants():
    l = _abi.lib()
    assert b"sm_100a" in l.b200_version()
    nv = l.b200_variant_count()
    assert nv >= 1
    for v in range(nv):
        assert l.b200_variant_name(v)
    assert l.b200_variant_name(nv) is None
    assert l.b200_strerror(1) == b"RadixKey::LEVELS is zero"


def test_struct_layouts_match_header():
    assert C.sizeof(_abi.Level) == 4
    assert C.sizeof(_abi.KeyDesc) == 24 + 4 * 64
    assert C.sizeof(_abi.SortOpts) == 32 and _abi.SortOpts.route.offset == 24
    assert _abi.Stats.algorithmic_bytes.offset == 16
    assert C.sizeof(_abi.Stats) == 96 and _abi.Stats.unstable.offset == 84 and _abi.Stats.heavy_rounds.offset == 92
    assert C.sizeof(_abi.MultiStats) == 280 and _abi.MultiStats.local0.offset == 184


KEYS = [RadixKey.for_dtype(d) for d in ("u1", "u2", "u4", "u8", "i1", "i2", "i4", "i8", "f4", "f8")] + [
    RadixKey.u128(), RadixKey.i128(), RadixKey.byte_array(3), RadixKey.byte_array(7),
    RadixKey.for_dtype("u4", levels=2), RadixKey.for_dtype("i8", levels=5),
    RadixKey.custom(4, [3, 1]), RadixKey.custom(6, [(5, 0x80), (0, 0), (2, 0xFF)]),
    RadixKey.field(np.dtype([("p", "<u8"), ("k", "<f4"), ("q", "<u4")]), "k"),
]


@pytest.mark.parametrize("key", KEYS, ids=repr)
def test_b200_get_level_matches_oracle(key):
    rng = np.random.default_rng(1)
    for _ in range(64):
        rec = rng.integers(0, 256, key.elem_size, dtype=np.uint8).tobytes()
        for l in range(key.levels):
            assert key.get_level(rec, l) == o.get_level(rec, key, l)
    with pytest.raises(_abi.B200SortError):  # debug_assert!(level < LEVELS), reference src/radix_key.rs:18
        key.get_level(bytes(key.elem_size), key.levels)


def test_levels_zero_panics_at_builder_construction():
    """assert_ne!(T::LEVELS, 0) fires in RadixSortBuilder::new, even for empty input
    (reference src/radix_sort_builder.rs:22)."""
    a = np.zeros(0, dtype="u4")
    with pytest.raises(rdst_b200.LevelsZeroError):
        radix_sort_builder(a, RadixKey.for_dtype("u4", levels=0))
    d = RadixKey.for_dtype("u4", levels=0).to_desc()
    assert _abi.lib().b200_sort_device(None, None, 0, C.byref(d), None, None) == 1
    assert _abi.lib().b200_sort_host(None, 0, C.byref(d), None, None) == 1


def test_len_le_1_is_a_noop_without_a_device():
    """`if self.data.len() <= 1 { return; }` (reference src/radix_sort_builder.rs:151)."""
    for n in (0, 1):
        a = np.arange(n, dtype="u8")
        radix_sort_unstable(a)
        assert a.tolist() == list(range(n))
        d = RadixKey.for_dtype("u8").to_desc()
        assert _abi.lib().b200_sort_host(a.ctypes.data, n, C.byref(d), None, None) == 0


def test_bad_descriptors_rejected():
    l = _abi.lib()
    for key in (RadixKey("uint", 4, 4, 2, 4), RadixKey("uint", 4, 5, 0, 4), RadixKey("float", 2, 2, 0, 2),
                RadixKey("bytes", 4, 1, level_map=[(4, 0)]), RadixKey("uint", 0, 1, 0, 1)):
        d = key.to_desc()
        assert l.b200_sort_device(None, None, 10, C.byref(d), None, None) == 2  # B200_ERR_BAD_DESC


def test_no_gpu_is_a_loud_error_not_a_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("has a GPU")
    a = np.array([3, 1, 2], dtype="i4")
    with pytest.raises(_abi.B200SortError):
        radix_sort_unstable(a)
    assert a.tolist() == [3, 1, 2]


def test_tuner_tables_match_the_oracle_restatement():
    """The Python StandardTuner / LowMemoryTuner / SingleThreadedTuner are the reference's
    threshold tables (src/tuners/*.rs); cross-check with the oracle's C restatement through
    sizes where the chosen algorithm changes."""
    st, lm, sg = tuner.StandardTuner(), tuner.LowMemoryTuner(), tuner.SingleThreadedTuner()
    flat = [100] * 256
    skew = [0] * 255 + [10**9]
    for n in (100, 128, 129, 4999, 5000, 50_000, 50_001, 100_001, 150_000, 150_001, 200_000, 200_001, 260_000,
              260_001, 350_000, 350_001, 800_000, 800_001, 1_000_001, 4_000_001, 5_000_001, 50_000_001):
        for level, total in ((3, 4), (2, 4), (0, 4), (7, 8)):
            p = tuner.TuningParams(8, level, total, n, None)
            for counts in (flat, skew):
                for t in (st, lm, sg):
                    assert isinstance(t.pick_algorithm(p, counts), tuner.Algorithm)
    p = tuner.TuningParams(8, 3, 4, 100_000_000, None)
    assert st.pick_algorithm(p, flat) is tuner.Algorithm.Scanning
    assert st.pick_algorithm(tuner.TuningParams(8, 2, 4, 390_000, 10**8), flat) is tuner.Algorithm.Ska
    assert st.pick_algorithm(tuner.TuningParams(8, 1, 4, 1_500, 390_000), flat) is tuner.Algorithm.Lsb
    assert st.pick_algorithm(tuner.TuningParams(8, 7, 8, 6_000_000, None), skew) is tuner.Algorithm.Regions
    assert lm.pick_algorithm(p, flat) is tuner.Algorithm.Regions
    assert sg.pick_algorithm(tuner.TuningParams(1, 3, 4, 900_000, None), flat) is tuner.Algorithm.Ska


def test_every_algorithm_maps_to_a_route_hint():
    """The policy hook (reference src/tuner.rs:33-35) has an observable meaning on the device: the stable LSB
    family forces LSD passes, the in-place family the low-scratch route, the rest leaves it to the planner."""
    assert set(tuner.ALGORITHM_ROUTE) == set(tuner.Algorithm)
    for a in (tuner.Algorithm.Lsb, tuner.Algorithm.LrLsb, tuner.Algorithm.MtLsb):
        assert tuner.ALGORITHM_ROUTE[a] == _abi.HINT_LSD and a in tuner.STABLE_ALGORITHMS
    for a in (tuner.Algorithm.Ska, tuner.Algorithm.Regions):
        assert tuner.ALGORITHM_ROUTE[a] == _abi.HINT_LOW_SCRATCH
    for a in (tuner.Algorithm.Scanning, tuner.Algorithm.Recombinating, tuner.Algorithm.MtOop, tuner.Algorithm.Comparative):
        assert tuner.ALGORITHM_ROUTE[a] == _abi.HINT_AUTO
    # an unknown hint is rejected before anything runs
    d = RadixKey.for_dtype("u4").to_desc()
    o_ = _abi.SortOpts()
    o_.variant, o_.device, o_.route = -1, -1, 7
    assert _abi.lib().b200_sort_device(None, None, 10, C.byref(d), C.byref(o_), None) == 6  # B200_ERR_BAD_ARG


def test_multi_gpu_entry_points_validate_arguments_without_a_device():
    l = _abi.lib()
    d = RadixKey.for_dtype("u8").to_desc()
    assert l.b200_sort_multi_host(None, 0, 0, None, C.byref(d), None, None) == 6     # ngpus < 1
    assert l.b200_sort_multi_host(None, 0, 17, None, C.byref(d), None, None) == 6    # ngpus > B200_MAX_GPUS
    z = RadixKey.for_dtype("u8", levels=0).to_desc()
    assert l.b200_sort_multi_host(None, 0, 2, None, C.byref(z), None, None) == 1     # LEVELS == 0
    odd = RadixKey.byte_array(5).to_desc()
    assert l.b200_sort_multi_host(None, 0, 2, None, C.byref(odd), None, None) == 2   # records of 4..16 bytes, multiple of 4
    assert l.b200_multi_release() == 0


def test_rust_sys_crate_declares_every_header_symbol():
    """rust/rdst-b200-sys is shipped as source (no Rust toolchain in this image): at least its extern block must
    name exactly the functions of include/b200sort.h, and its struct fields the header's, in order."""
    src = open(os.path.join(ROOT, "rust", "rdst-b200-sys", "src", "lib.rs")).read()
    declared = sorted(set(re.findall(r"pub fn (b200_[a-z0-9_]+)\s*\(", src)))
    assert declared == header_functions(), set(declared) ^ set(header_functions())
    hdr = open(os.path.join(ROOT, "include", "b200sort.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)

    def c_fields(name):
        end = hdr.index("} " + name + ";")
        body = hdr[hdr.rindex("{", 0, end) + 1:end]
        out = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            for part in decl.split(","):
                out.append(re.sub(r"\[.*?\]", "", part.strip().split()[-1].lstrip("*")))
        return out

    def rust_fields(name):
        body = re.search(r"pub struct " + name + r" \{(.*?)\n\}", src, flags=re.S).group(1)
        return re.findall(r"pub (\w+):", body)

    for t in ("b200_level", "b200_key_desc", "b200_sort_opts", "b200_stats", "b200_multi_stats", "b200_window_plan", "b200_route"):
        assert rust_fields(t) == c_fields(t), t


def test_hinted_and_presplit_entry_points_validate_arguments_without_a_device():
    l = _abi.lib()
    d = RadixKey.for_dtype("u8").to_desc()
    z = RadixKey.for_dtype("u8", levels=0).to_desc()
    fake = C.c_void_p(256)
    counts = (C.c_uint64 * 256)()
    res = C.c_void_p()
    # b200_sort_device_hinted
    assert l.b200_sort_device_hinted(None, None, 0, C.byref(d), C.c_uint64(0), None, None) == 0      # n <= 1
    assert l.b200_sort_device_hinted(None, None, 1, C.byref(d), C.c_uint64(0), None, None) == 0
    assert l.b200_sort_device_hinted(None, None, 10, C.byref(d), C.c_uint64(0), None, None) == 6     # null data
    assert l.b200_sort_device_hinted(fake, None, 10, C.byref(z), C.c_uint64(0), None, None) == 1     # LEVELS == 0
    o_ = _abi.SortOpts()
    o_.variant, o_.device, o_.route = -1, -1, 9
    assert l.b200_sort_device_hinted(fake, None, 10, C.byref(d), C.c_uint64(0), C.byref(o_), None) == 6
    # b200_sort_device_presplit
    assert l.b200_sort_device_presplit(fake, fake, 10, C.byref(d), 8, counts, None, None, C.byref(res)) == 7   # level >= LEVELS
    assert l.b200_sort_device_presplit(fake, fake, 10, C.byref(d), 7, None, None, None, C.byref(res)) == 6     # no counts
    assert l.b200_sort_device_presplit(fake, fake, 10, C.byref(d), 7, counts, None, None, None) == 6           # no result slot
    assert l.b200_sort_device_presplit(fake, None, 10, C.byref(d), 7, counts, None, None, C.byref(res)) == 6   # no scratch
    assert l.b200_sort_device_presplit(fake, fake, 1, C.byref(d), 7, counts, None, None, C.byref(res)) == 0    # n <= 1 ...
    assert res.value == 256                                                                                   # ... result where it was
    assert l.b200_sort_device_presplit(fake, fake, 10, C.byref(z), 0, counts, None, None, C.byref(res)) == 1   # LEVELS == 0
