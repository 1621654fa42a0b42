"""b200_sort_device_static (include/b200sort.h): the device sort whose scratch is a file-scope
`static __device__` array, for callers that cannot call cudaMalloc.  No counterpart in the reference.
Checked through the C ABI against b200_sort_device on the same inputs
(bit-exact, like every other parity test here), on calls in a row that share the array, and against the oracle.

Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

import oracle_ffi as o
from rdst_b200 import RadixKey, _abi

REC12 = np.dtype([("key", "<u4"), ("lo", "<u4"), ("hi", "<u4")])
MAX_RECORDS, MAX_BYTES = _abi.STATIC_MAX_RECORDS, _abi.STATIC_MAX_BYTES


def rand_bits(rng, dt, n):
    return rng.integers(0, 256, n * np.dtype(dt).itemsize, dtype=np.uint8).view(dt).copy()


def opts_for(torch, stream=None):
    op = _abi.SortOpts()
    op.stable, op.variant, op.device, op.route = 0, -1, -1, _abi.HINT_AUTO
    op.stream = stream if stream is not None else _abi.torch_stream_handle(torch)
    return op


def sort_static(torch, a, key, op=None):
    """a: numpy records -> bytes of the result of the no-allocation entry point"""
    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()
    d = key.to_desc()
    n = a.nbytes // key.elem_size
    _abi.check(_abi.lib().b200_sort_device_static(t.data_ptr(), n, C.byref(d), C.byref(op) if op is not None else None))
    torch.cuda.synchronize()  # the call only enqueues
    return t.cpu().numpy().tobytes()


def sort_existing(torch, a, key):
    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()
    d = key.to_desc()
    op = opts_for(torch)
    n = a.nbytes // key.elem_size
    _abi.check(_abi.lib().b200_sort_device(t.data_ptr(), None, n, C.byref(d), C.byref(op), None))
    torch.cuda.synchronize()
    return t.cpu().numpy().tobytes()


def test_static_entry_point_validates_before_touching_a_device():
    l = _abi.lib()
    d8 = RadixKey.for_dtype("u8").to_desc()
    assert l.b200_sort_device_static(None, 0, C.byref(d8), None) == 0       # n <= 1: nothing to do
    assert l.b200_sort_device_static(None, 1, C.byref(d8), None) == 0
    assert l.b200_sort_device_static(None, 10, C.byref(d8), None) == 6      # null pointer
    z = RadixKey.for_dtype("u8", levels=0).to_desc()
    assert l.b200_sort_device_static(None, 10, C.byref(z), None) == 1       # LEVELS == 0
    # beyond the static scratch: rejected with a message that names the limits, whatever the pointer
    fake = C.c_void_p(256)
    assert l.b200_sort_device_static(fake, MAX_BYTES // 8 + 1, C.byref(d8), None) == 9
    msg = l.b200_last_cuda_error().decode()
    assert "exceed the static scratch" in msg and str(MAX_BYTES) in msg and str(MAX_RECORDS) in msg
    d1 = RadixKey.for_dtype("u1").to_desc()
    assert l.b200_sort_device_static(fake, MAX_RECORDS + 1, C.byref(d1), None) == 9   # record count limit
    assert l.b200_strerror(9) == b"more records than the static scratch holds"
    op = _abi.SortOpts()
    op.variant, op.device, op.route = -1, -1, _abi.HINT_LSD
    assert l.b200_sort_device_static(fake, 10, C.byref(d8), C.byref(op)) == 6          # no route hints here


@pytest.mark.gpu
def test_static_matches_the_allocating_entry_point_on_calls_in_a_row():
    import torch

    rng = np.random.default_rng(2026)
    k8, k4 = RadixKey.for_dtype("u8"), RadixKey.for_dtype("u4")
    # (records, key): consecutive calls share the static array — large then small, passes then the one-launch
    # route and back, odd level counts (copy back), the byte path, a payload record, narrow keys (levels that a
    # histogram read-back would have skipped), the largest sizes allowed
    cases = [
        (rand_bits(rng, "u8", 1_000_000), k8),
        (rand_bits(rng, "u8", 1_000_000), k8),                      # the same shape again: stale look-back state
        (rand_bits(rng, "u8", 300_001), k8),
        (rand_bits(rng, "u8", 16_384), k8),                         # one launch of K3
        (rand_bits(rng, "u4", 2_000), k4),
        (rand_bits(rng, "u4", 2_000), k4),
        (rand_bits(rng, "u4", MAX_RECORDS), k4),                    # record limit
        (rand_bits(rng, "u8", MAX_BYTES // 8), k8),                 # byte limit
        (rand_bits(rng, "i8", 77_777), RadixKey.for_dtype("i8")),
        (rand_bits(rng, "f4", 500_000), RadixKey.for_dtype("f4")),
        (rng.standard_normal(250_000), RadixKey.for_dtype("f8")),
        (rng.integers(0, 1 << 20, 400_000, dtype="u8"), k8),        # five trivial levels, all of them run
        (rng.integers(0, 7, 100_000, dtype="u4"), k4),              # heavy duplicates
        (rand_bits(rng, "u1", 3 * 100_001).view(np.uint8), RadixKey.byte_array(3)),   # byte path, LEVELS = 3 (odd)
        (rand_bits(rng, "u2", 200_000), RadixKey.for_dtype("u2")),
        (rand_bits(rng, "u1", 16 * 50_000).view(np.uint8), RadixKey.u128()),
    ]
    a12 = np.zeros(300_000, dtype=REC12)
    a12["key"] = rng.integers(0, 1000, len(a12))
    a12["lo"] = np.arange(len(a12))
    cases.append((a12, RadixKey.field(REC12, "key", 4)))            # stable: payload order is observable
    for i, (a, key) in enumerate(cases):
        got = sort_static(torch, a, key, opts_for(torch))
        assert got == sort_existing(torch, a, key), (i, key)
        if a.nbytes <= 4 << 20:
            assert got == o.stable_sorted(a, key).tobytes(), (i, key)


@pytest.mark.gpu
def test_static_twice_in_a_row_same_input_and_null_opts():
    import torch

    rng = np.random.default_rng(7)
    key = RadixKey.for_dtype("u8")
    a = rand_bits(rng, "u8", 654_321)
    want = sort_existing(torch, a, key)
    first = sort_static(torch, a, key, None)      # opts = NULL: legacy default stream, current device
    second = sort_static(torch, a, key, None)
    assert first == want and second == want
    # an explicit variant forces the passes even for an array that fits one SM
    op = opts_for(torch)
    op.variant = 12
    b = rand_bits(rng, "u8", 5_000)
    assert sort_static(torch, b, key, op) == sort_existing(torch, b, key)
    # a side stream of the caller's
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        assert sort_static(torch, a, key, opts_for(torch, s.cuda_stream)) == want


@pytest.mark.gpu
def test_static_allocates_no_device_memory():
    import torch

    rng = np.random.default_rng(8)
    key = RadixKey.for_dtype("u8")
    a = rand_bits(rng, "u8", 500_000)
    t = torch.from_numpy(a.view(np.uint8).copy()).cuda()
    keep = t.clone()
    d = key.to_desc()
    op = opts_for(torch)
    l = _abi.lib()
    _abi.check(l.b200_sort_device_static(t.data_ptr(), len(a), C.byref(d), C.byref(op)))   # loads the kernels
    torch.cuda.synchronize()
    t.copy_(keep)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    _abi.check(l.b200_sort_device_static(t.data_ptr(), len(a), C.byref(d), C.byref(op)))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0
    assert t.cpu().numpy().tobytes() == np.sort(a).tobytes()
    # too large: refused, the data is untouched
    big = torch.zeros(MAX_BYTES + 8, dtype=torch.uint8, device="cuda")
    assert l.b200_sort_device_static(big.data_ptr(), MAX_BYTES // 8 + 1, C.byref(d), C.byref(op)) == 9
