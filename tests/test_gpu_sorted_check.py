"""K0, the already-sorted early-out (rdst's already_sorted shortcut: reference
src/sort_utils.rs:125-136, src/sorter.rs:59-65).  An input that is non-decreasing by its key
comes back untouched with stats.already_sorted == 1; a single out-of-order pair anywhere —
first pair, last pair, chunk boundaries of the check kernel — must take the full sort and
match the oracle bit for bit."""
import numpy as np
import pytest

import oracle_ffi as o
from rdst_b200 import RadixKey, radix_sort_builder

pytestmark = pytest.mark.gpu

REC12 = np.dtype([("key", "<u4"), ("lo", "<u4"), ("hi", "<u4")])
REC6 = np.dtype([("key", "<u2"), ("idx", "<u4")])  # odd stride: byte path of the check kernel
LARGE = np.dtype([("a", "<u8"), ("b", "<u8"), ("sort_key", "<f4"), ("c", "<u4"), ("d", "<u8"), ("e", "<u8")])
SYNC_MIN = 4 << 20
SMALL_CAP = 16_384  # 4- and 8-byte records up to this many take the one-launch K3 path instead (no K0)


def small_path(a, key, n):
    return n <= SMALL_CAP and key.elem_size in (4, 8) and key.levels <= 8


def rand_bits(rng, dt, n):
    return rng.integers(0, 256, n * np.dtype(dt).itemsize, dtype=np.uint8).view(dt).copy()


def sort_both(a, key):
    """Host and device entry points on copies of `a`; returns (host result, host stats, device stats)."""
    import torch

    h = a.copy()
    bh = radix_sort_builder(h, key).with_stable(True)
    bh.sort()
    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()
    bd = radix_sort_builder(t, key).with_stable(True)
    bd.sort()
    torch.cuda.synchronize()
    assert t.cpu().numpy().tobytes() == h.tobytes(), "host and device entry points disagree"
    return h, bh.stats, bd.stats


@pytest.mark.parametrize("dt", ["u1", "u2", "u4", "u8", "i1", "i2", "i4", "i8", "f4", "f8"])
def test_presorted_primitives_return_early(dt):
    """Random BIT patterns (floats: NaNs, infinities, -0.0, subnormals) put in rdst's order by the
    oracle; the GPU must recognise the order for every key transform."""
    rng = np.random.default_rng(500 + np.dtype(dt).num)
    key = RadixKey.for_dtype(dt)
    for n in (2, 3, 1_025, 70_001, 70_002, 70_003, 131_072):
        a = o.sorted_copy(rand_bits(rng, dt, n), key)
        got, sh, sd = sort_both(a, key)
        assert got.tobytes() == a.tobytes()
        if small_path(a, key, n):
            assert sh["kernel_launches"] == 1 and sd["kernel_launches"] == 1 and sh["levels_run"] == key.levels
            continue
        assert sh["already_sorted"] == 1 and sd["already_sorted"] == 1, (dt, n)
        assert sh["levels_run"] == 0 and sh["levels_skipped"] == key.levels
        # the last pair (past the last full 16-byte vector for the odd lengths) out of order
        b = a.copy()
        b[n - 1], b[n - 2] = a[n - 2], a[n - 1]
        if b.tobytes() != a.tobytes():
            got, sh, sd = sort_both(b, key)
            assert got.tobytes() == a.tobytes(), (dt, n)
            assert sh["already_sorted"] == 0 and sd["already_sorted"] == 0, (dt, n)


@pytest.mark.parametrize("kernel", ["vector", "scalar"])
@pytest.mark.parametrize("dt", ["u4", "i8", "f4", "f8"])
def test_one_descending_pair_anywhere_takes_the_full_sort(monkeypatch, dt, kernel):
    """Strictly increasing keys with exactly one adjacent swap at the positions where the check
    kernels' vector / lane / thread / chunk / CTA boundaries fall (B200SORT_SORTED_CHECK=2 keeps
    key-only arrays on the scalar kernel that records with payload use)."""
    if kernel == "scalar":
        if dt in ("f4", "f8"):
            pytest.skip("the scalar kernel's float transform is covered by the record tests; keeps the suite short")
        monkeypatch.setenv("B200SORT_SORTED_CHECK", "2")
    n = 5_000_000
    key = RadixKey.for_dtype(dt)
    base = (np.arange(n, dtype="i8") - n // 2) * 3
    base = base.astype(dt) if np.dtype(dt).kind != "u" else (base + 2 * n).astype(dt)
    assert o.is_sorted(base, key)
    # 2^j - 1 and 2^j: the last / first key of every vector, warp, thread-step, chunk and grid stride of
    # both check kernels (16-byte vectors of 2 or 4 keys, 32 lanes, 256 threads, 4 steps, <= 1184 CTAs)
    # plus the vector kernel's chunk (256 threads x 4 vectors) and grid stride (8 CTAs x 148 SMs of them)
    chunk = 256 * 4
    marks = sorted({m for j in range(22) for m in ((1 << j) - 1, 1 << j)} |
                   {148 * 8 * 1_024 - 1, 148 * 8 * 1_024, n // 2, n - 3, n - 2} |
                   {chunk * k * g + d for k in (2, 4) for g in (1, 2, 1184) for d in (-1, 0)})
    for pos in marks:
        if pos + 1 >= n:
            continue
        a = base.copy()
        a[pos], a[pos + 1] = base[pos + 1], base[pos]
        got, sh, sd = sort_both(a, key)
        assert sh["already_sorted"] == 0 and sd["already_sorted"] == 0, (dt, pos)
        assert got.tobytes() == base.tobytes(), (dt, pos)


def test_presorted_records_and_byte_path_keys():
    rng = np.random.default_rng(77)
    cases = [(REC12, RadixKey.field(REC12, "key", 3)), (REC12, RadixKey.field(REC12, "key", 4)),
             (REC6, RadixKey.field(REC6, "key", 2)), (LARGE, RadixKey.field(LARGE, "sort_key", 4))]
    for rec, key in cases:
        for n in (50_001,):
            a = np.zeros(n, dtype=rec)
            raw = a.view(np.uint8).reshape(n, rec.itemsize)
            raw[:] = rng.integers(0, 256, raw.shape, dtype=np.uint8)
            s = o.stable_sorted(a, key)
            got, sh, sd = sort_both(s, key)
            assert got.tobytes() == s.tobytes()
            assert sh["already_sorted"] == 1 and sd["already_sorted"] == 1, (rec, n)
            # reversed: not sorted (unless every key is equal), full sort, stable order of the reversed input
            r = s[::-1].copy()
            got, sh, _ = sort_both(r, key)
            assert sh["already_sorted"] == 0
            assert got.tobytes() == o.stable_sorted(r, key).tobytes()
    for nb in (3, 16):  # [u8; N]: lexicographic, byte path
        key = RadixKey.byte_array(nb)
        a = rng.integers(0, 4, (20_000, nb), dtype=np.uint8)
        s = o.sorted_copy(a.reshape(-1), key)
        got, sh, sd = sort_both(s, key)
        assert got.tobytes() == s.tobytes() and sh["already_sorted"] == 1 and sd["already_sorted"] == 1
    for key in (RadixKey.u128(), RadixKey.i128()):  # 16-byte words: byte path
        a = rng.integers(0, 256, 16 * 30_000, dtype=np.uint8)
        s = o.sorted_copy(a, key)
        got, sh, sd = sort_both(s, key)
        assert got.tobytes() == s.tobytes() and sh["already_sorted"] == 1 and sd["already_sorted"] == 1


def test_partial_levels_ignore_the_high_bytes_in_the_order_check():
    """LEVELS = 2 on u32 (reference src/lib.rs:58-76): only the low two bytes are the key, so
    descending high bytes do not make the input unsorted — and the stable result is the input."""
    n = 100_000
    key = RadixKey.for_dtype("u4", levels=2)
    low = np.sort(np.random.default_rng(5).integers(0, 1 << 16, n, dtype="u4"))
    a = low | (np.arange(n, 0, -1, dtype="u4") << 16)
    got, sh, sd = sort_both(a, key)
    assert sh["already_sorted"] == 1 and sd["already_sorted"] == 1
    assert got.tobytes() == a.tobytes() == o.stable_sorted(a, key).tobytes()
    b = a.copy()
    b[n - 1] = b[0]  # the smallest key last
    got, sh, _ = sort_both(b, key)
    assert sh["already_sorted"] == 0 or low[0] == low[-2]
    assert got.tobytes() == o.stable_sorted(b, key).tobytes()


def test_all_equal_and_negative_zero():
    a = np.full(20_000, 7, dtype="u4")
    _, sh, sd = sort_both(a, RadixKey.for_dtype("u4"))
    assert sh["already_sorted"] == 1 and sd["already_sorted"] == 1
    pad = np.arange(1, 20_001, dtype="f4")  # past the small-input path
    f = np.concatenate([np.array([-0.0, 0.0, 0.0], dtype="f4"), pad])
    got, sh, _ = sort_both(f, RadixKey.for_dtype("f4"))
    assert sh["already_sorted"] == 1 and got.tobytes() == f.tobytes()
    f = np.concatenate([np.array([0.0, -0.0], dtype="f4"), pad])  # -0.0 sorts before +0.0 (radix_key_impl.rs:162-173)
    got, sh, _ = sort_both(f, RadixKey.for_dtype("f4"))
    assert sh["already_sorted"] == 0
    assert got.tobytes() == np.concatenate([np.array([-0.0, 0.0], dtype="f4"), pad]).tobytes()


def test_large_presorted_input_skips_k1_on_the_device():
    """K1 is gated on K0's flag on the device: a presorted input costs one read of the array (K1's
    CTAs leave at once), whatever its size; late, early and middle violations take the full sort."""
    import torch

    n = SYNC_MIN + 12_345
    t = torch.arange(n, dtype=torch.int64, device="cuda") * 5 - 7 * n
    ref = t.clone()
    raw = t.view(torch.uint8)
    b = radix_sort_builder(raw, RadixKey.for_dtype("i8"))
    b.sort()
    torch.cuda.synchronize()
    assert b.stats["already_sorted"] == 1 and b.stats["levels_run"] == 0
    assert torch.equal(t, ref)
    for pos in (0, n // 3, n - 2):  # one late, one early and one middle violation
        t = ref.clone()
        t[pos], t[pos + 1] = ref[pos + 1], ref[pos]
        b = radix_sort_builder(t.view(torch.uint8), RadixKey.for_dtype("i8"))
        b.sort()
        torch.cuda.synchronize()
        assert b.stats["already_sorted"] == 0 and b.stats["levels_run"] >= 1
        assert torch.equal(t, ref), pos
    # host entry point: nothing is copied back when the input was in order
    h = ref[:SYNC_MIN + 1].cpu().numpy()
    keep = h.copy()
    bh = radix_sort_builder(h, RadixKey.for_dtype("i8"))
    bh.sort()
    assert bh.stats["already_sorted"] == 1 and np.array_equal(h, keep)
