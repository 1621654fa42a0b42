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

    def partition_peers(self, src, dst_ptrs, n, level, row, split, parts):
        o = self._opts()
        pc = (C.c_uint64 * parts)()
        row = np.ascontiguousarray(row, dtype=np.uint64)
        split = np.ascontiguousarray(split, dtype=np.uint32)
        ptrs = (C.c_void_p * parts)(*[C.c_void_p(int(x)) for x in dst_ptrs])
        _abi.check(self.lib.b200_partition_peers_device(src.data_ptr(), ptrs, n, C.byref(self.desc), level,
                                                        row.ctypes.data_as(C.POINTER(C.c_uint64)),
                                                        split.ctypes.data_as(C.POINTER(C.c_uint32)), parts, pc,
                                                        C.byref(o)))
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


class PeerBuffers:
    """Receive buffers shared between the ranks of one box through CUDA IPC, so that the
    partition kernel of every rank can store its runs straight into the owners' buffers over
    NVLink (fused partition + all-to-all).  One cudaMalloc per rank, handles all-gathered once."""

    def __init__(self, capacity_bytes: int, group=None):
        import torch
        import torch.distributed as dist

        self.torch, self.group = torch, group
        self.lib = _abi.lib()
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.capacity = int(capacity_bytes)
        ptr = C.c_void_p()
        handle = (C.c_uint8 * 64)()
        _abi.check(self.lib.b200_ipc_alloc(C.byref(ptr), self.capacity, handle))
        self.local_ptr = ptr.value
        mine = torch.tensor(list(handle), dtype=torch.uint8, device="cuda")
        allh = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(allh, mine, group=group)
        self.ptrs = []
        for r in range(self.world):
            if r == self.rank:
                self.ptrs.append(self.local_ptr)
                continue
            hb = (C.c_uint8 * 64)(*allh[r].cpu().tolist())
            pp = C.c_void_p()
            _abi.check(self.lib.b200_ipc_open(hb, C.byref(pp)))
            self.ptrs.append(pp.value)
        self._flag = torch.zeros(1, dtype=torch.int32, device="cuda")
        # a flat uint8 view of the local buffer for the local sort / the caller
        self.local = _device_view(torch, self.local_ptr, self.capacity)

    def fence(self):
        """Stream-ordered barrier across ranks: a tiny all-reduce completes on a rank's stream
        only when every rank's stream has reached it."""
        import torch.distributed as dist

        dist.all_reduce(self._flag, group=self.group)

    def close(self):
        for r, p in enumerate(self.ptrs):
            if r != self.rank and p:
                self.lib.b200_ipc_close(p)
        self.lib.b200_ipc_free(self.local_ptr)
        self.ptrs = []


def _device_view(torch, ptr, nbytes):
    """torch uint8 tensor over raw device memory (no ownership)."""

    class _Holder:
        pass

    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}
    return torch.as_tensor(h, device="cuda")


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

    ev = None
    if timings is not None and hasattr(eng, "torch"):
        ev = [eng.torch.cuda.Event(enable_timing=True) for _ in range(5)]
        ev[0].record()
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
    fused = bool(scratch and scratch.get("peers") is not None)
    part = recv = None
    if not fused:
        part = scratch["part"] if scratch else eng.empty(n_local)
        recv = scratch["recv"] if scratch else eng.empty(int(max(n_recv, 1)))
        if recv.numel() < n_recv * key.elem_size:
            raise _abi.B200SortError(5, f"receive buffer too small for {n_recv} records (skewed splitters)")
    peers = scratch.get("peers") if scratch else None
    if ev:
        ev[1].record()
    if peers is not None:
        # fused: this rank's records for rank g start, in g's buffer, after those of the lower ranks
        if peers.capacity < n_recv * key.elem_size:
            raise _abi.B200SortError(5, f"peer receive buffer too small for {n_recv} records")
        starts = np.concatenate([np.zeros((1, world), dtype=np.int64), np.cumsum(m, axis=0)])[rank]  # [g]
        dst = [peers.ptrs[g] + int(starts[g]) * key.elem_size for g in range(world)]
        peers.fence()  # every rank is done with the previous contents of its receive buffer
        pc = eng.partition_peers(local, dst, n_local, level, hist[level], split, world)
        assert np.array_equal(pc, send_counts), "partition counts disagree with the gathered histograms"
        if ev:
            ev[2].record()
        peers.fence()  # every rank's partition kernel has finished: the received range is complete
        recv = peers.local
        if ev:
            ev[3].record()
    else:
        pc = eng.partition(local, part, n_local, level, hist[level], split, world)
        assert np.array_equal(pc, send_counts), "partition counts disagree with the gathered histograms"
        if ev:
            ev[2].record()
        eng.exchange(part, recv, send_counts, recv_counts, group)
        if ev:
            ev[3].record()
    # 5. independent local LSD sort of the received range (levels above `level` are
    #    single-bucket per rank only by luck; K1 decides what to skip)
    st = eng.sort(recv, n_recv, scratch.get("tmp") if scratch else None)
    if timings is not None:
        timings.update({"level": level, "split": split.tolist(), "n_recv": n_recv, "local": st,
                        "sent_bytes": int((send_counts.sum() - send_counts[rank]) * key.elem_size)})
        if ev:
            ev[4].record()
            ev[4].synchronize()
            timings["phase_ms"] = {"hist_allgather_plan": ev[0].elapsed_time(ev[1]), "partition": ev[1].elapsed_time(ev[2]),
                                   "all_to_all": ev[2].elapsed_time(ev[3]), "local_sort": ev[3].elapsed_time(ev[4])}
    return recv, n_recv
