"""Sharded sort across the GPUs of one box: one process per GPU (torch.distributed, NCCL
over NVLink / NVSwitch), MSD range partition on the most significant non-trivial level,
one all-to-all of records, then an independent local LSD sort per GPU.

The reference has no distributed path (one `&mut [T]` in host RAM); this is the widening
BASELINE.json configs[4] names.  Input: rank g holds a contiguous shard of the logical
array.  Output: rank g holds the g-th key range, sorted; concatenating the outputs in rank
order gives the sorted array.  Stable: records travel in (source rank, source position)
order and every local step is stable.

The data path runs through an `engine` with three operations — all-level histogram (K1),
stable range partition (K2 with a 256 -> P remap) and local sort — whose product
implementation is the CUDA C ABI.  (CPU tests of the host logic inject their own engine;
nothing here falls back to a CPU path.)
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi
from .radix_key import RadixKey


# --------------------------------------------------------------------------- host logic
def pick_partition_level(global_hist: np.ndarray) -> int:
    """Most significant level whose global histogram is not a single bucket (rdst's
    trivial-level skip, reference src/sorts/ska_sort.rs:52-54, applied to the splitters).
    -1 when every level is trivial (all keys equal)."""
    n = int(global_hist[0].sum())
    for level in range(global_hist.shape[0] - 1, -1, -1):
        if int(global_hist[level].max()) != n:
            return level
    return -1


def choose_splitters(row: np.ndarray, parts: int) -> np.ndarray:
    """P-1 digit boundaries on one 256-bin row: part g = digits in [split[g-1], split[g]).
    Boundary g is placed where the running count is closest to g * N / P."""
    total = int(row.sum())
    cum = np.concatenate([[0], np.cumsum(row.astype(np.int64))])  # cum[d] = records with digit < d, d in 0..256
    split = np.empty(parts - 1, dtype=np.uint32)
    lo = 0
    for g in range(1, parts):
        target = total * g / parts
        d = lo + int(np.argmin(np.abs(cum[lo:] - target)))  # first minimiser at or after the previous boundary
        split[g - 1] = d
        lo = d
    return split


def part_of_digit(split: np.ndarray, parts: int) -> np.ndarray:
    """256-entry digit -> part table (what the partition kernel's remap holds)."""
    return np.searchsorted(split, np.arange(256), side="right").astype(np.int64).clip(0, parts - 1)


def send_matrix(rows: np.ndarray, split: np.ndarray, parts: int) -> np.ndarray:
    """rows[r] = rank r's 256-bin row on the partition level -> M[r, g] = records rank r
    sends to rank g.  Every rank computes the whole matrix from the all-gathered rows, so
    no count exchange is needed before the all-to-all."""
    pod = part_of_digit(split, parts)
    m = np.zeros((rows.shape[0], parts), dtype=np.int64)
    for g in range(parts):
        m[:, g] = rows[:, pod == g].sum(axis=1)
    return m


# --------------------------------------------------------------------------- CUDA engine
class CudaEngine:
    """The product engine: K1 / K2-partition / local sort through the C ABI on the
    current CUDA device and torch's current stream."""

    def __init__(self, key: RadixKey):
        import torch

        self.torch = torch
        self.key = key
        self.desc = key.to_desc()
        self.lib = _abi.lib()
        self.launches = 0

    def _opts(self):
        o = _abi.SortOpts()
        o.stable = 1
        o.variant = -1
        o.device = self.torch.cuda.current_device()
        o.stream = _abi.torch_stream_handle(self.torch)
        return o

    def empty(self, n_records):
        return self.torch.empty(max(1, n_records) * self.key.elem_size, dtype=self.torch.uint8, device="cuda")

    def histogram(self, buf, n):
        L = self.key.levels
        hist = np.zeros((L, 256), dtype=np.uint64)
        o = self._opts()
        _abi.check(self.lib.b200_histogram_device(buf.data_ptr(), n, C.byref(self.desc), C.byref(o),
                                                  hist.ctypes.data_as(C.POINTER(C.c_uint64))))
        self.launches += 1
        return hist

    def partition(self, src, dst, n, level, row, split, parts):
        o = self._opts()
        pc = (C.c_uint64 * parts)()
        row = np.ascontiguousarray(row, dtype=np.uint64)
        split = np.ascontiguousarray(split, dtype=np.uint32)
        _abi.check(self.lib.b200_partition_device(src.data_ptr(), dst.data_ptr(), n, C.byref(self.desc), level,
                                                  row.ctypes.data_as(C.POINTER(C.c_uint64)),
                                                  split.ctypes.data_as(C.POINTER(C.c_uint32)), parts, pc, C.byref(o)))
        self.launches += 1
        return np.array(list(pc), dtype=np.int64)

    def sort(self, buf, n, tmp=None):
        o = self._opts()
        st = _abi.Stats()
        _abi.check(self.lib.b200_sort_device(buf.data_ptr(), tmp.data_ptr() if tmp is not None else None, n,
                                             C.byref(self.desc), C.byref(o), C.byref(st)))
        self.launches += st.kernel_launches
        return st.as_dict()

    def to_comm(self, arr_np):
        return self.torch.from_numpy(arr_np).cuda()

    def exchange(self, send, recv, send_counts, recv_counts, group):
        import torch.distributed as dist

        es = self.key.elem_size
        n_send, n_recv = int(send_counts.sum()), int(recv_counts.sum())
        dist.all_to_all_single(recv[: n_recv * es].view(-1, es), send[: n_send * es].view(-1, es),
                               output_split_sizes=[int(x) for x in recv_counts],
                               input_split_sizes=[int(x) for x in send_counts], group=group)


# --------------------------------------------------------------------------- the sort
def sharded_sort(local, n_local: int, key: RadixKey, engine=None, group=None, recv_slack: float = 1.3,
                 scratch=None, timings=None):
    """Sort the logical array whose rank-r shard is `local` (a flat uint8 device buffer of
    n_local records).  Returns (buffer, n_out): this rank's sorted key range.
    `scratch` may carry pre-allocated {"part", "recv", "tmp"} buffers (bench reuses them)."""
    import torch.distributed as dist

    eng = engine or CudaEngine(key)
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world == 1:
        st = eng.sort(local, n_local, scratch.get("tmp") if scratch else None)
        if timings is not None:
            timings["local"] = st
        return local, n_local

    # 1. K1 on the local shard, all levels in one read
    hist = eng.histogram(local, n_local)
    # 2. global histogram + every rank's rows: one small all-gather ([P, L, 256] i64)
    mine = eng.to_comm(hist.astype(np.int64))
    gathered = [mine.new_empty(mine.shape) for _ in range(world)]
    dist.all_gather(gathered, mine, group=group)
    all_hist = np.stack([g.cpu().numpy() for g in gathered])  # [P, L, 256]
    global_hist = all_hist.sum(axis=0)
    level = pick_partition_level(global_hist)
    if level < 0:
        return local, n_local  # every key equal: any arrangement is sorted
    # 3. splitters on digit boundaries of that level
    split = choose_splitters(global_hist[level], world)
    m = send_matrix(all_hist[:, level, :], split, world)
    send_counts, recv_counts = m[rank], m[:, rank]
    n_recv = int(recv_counts.sum())
    # 4. stable local partition into P contiguous send regions, then the all-to-all
    part = scratch["part"] if scratch else eng.empty(n_local)
    cap = int(max(n_recv, 1))
    recv = scratch["recv"] if scratch else eng.empty(cap)
    if recv.numel() < n_recv * key.elem_size:
        raise _abi.B200SortError(5, f"receive buffer too small for {n_recv} records (skewed splitters)")
    pc = eng.partition(local, part, n_local, level, hist[level], split, world)
    assert np.array_equal(pc, send_counts), "partition counts disagree with the gathered histograms"
    eng.exchange(part, recv, send_counts, recv_counts, group)
    # 5. independent local LSD sort of the received range (levels above `level` are
    #    single-bucket per rank only by luck; K1 decides what to skip)
    st = eng.sort(recv, n_recv, scratch.get("tmp") if scratch else None)
    if timings is not None:
        timings.update({"level": level, "split": split.tolist(), "n_recv": n_recv, "local": st,
                        "sent_bytes": int((send_counts.sum() - send_counts[rank]) * key.elem_size)})
    return recv, n_recv
