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
import os

import numpy as np

from . import _abi
from .radix_key import RadixKey


# --------------------------------------------------------------------------- host logic
# The exchange planner lives in the C library (csrc/multi.cu: b200_multi_plan_*), shared with the single-process
# entry points b200_sort_multi_*; plan_digits / plan_sample_splitters / plan_splitters below call it.  The numpy
# functions that follow are the MODEL of the same rules: the CPU test double of tests/test_dist_gloo.py runs on
# them, and tests/test_multi_plan_cpu.py checks that the C planner agrees with them case by case.
def plan_digits(all_hist: np.ndarray):
    """[P, L, 256] K1 histograms -> (level, split[P-1], send matrix [P, P], imbalance) from the C planner."""
    lib = _abi.lib()
    h = np.ascontiguousarray(all_hist, dtype=np.uint64)
    P, L = h.shape[0], h.shape[1]
    level = C.c_int32()
    split = np.zeros(max(1, P - 1), dtype=np.uint32)
    m = np.zeros((P, P), dtype=np.int64)
    imb = C.c_double(1.0)
    _abi.check(lib.b200_multi_plan_digits(h.ctypes.data_as(C.POINTER(C.c_uint64)), P, L, C.byref(level),
                                          split.ctypes.data_as(C.POINTER(C.c_uint32)),
                                          m.ctypes.data_as(C.POINTER(C.c_int64)), C.byref(imb)))
    return int(level.value), split[: P - 1], m, float(imb.value)


def plan_sample_splitters(samples: np.ndarray, weights: np.ndarray, parts: int) -> np.ndarray:
    lib = _abi.lib()
    v = np.ascontiguousarray(samples, dtype=np.uint64)
    w = np.ascontiguousarray(weights, dtype=np.float64)
    out = np.zeros(parts - 1, dtype=np.uint64)
    _abi.check(lib.b200_multi_choose_sample_splitters(v.ctypes.data_as(C.POINTER(C.c_uint64)),
                                                      w.ctypes.data_as(C.POINTER(C.c_double)), len(v), parts,
                                                      out.ctypes.data_as(C.POINTER(C.c_uint64))))
    return out


def plan_splitters(less: np.ndarray, equal: np.ndarray, n_per_rank: np.ndarray, splitters: np.ndarray, parts: int,
                   exact_ties: bool = True):
    """-> (tie thresholds t[P-1], send matrix [R, P], incl masks [R]) from the C planner."""
    lib = _abi.lib()
    le = np.ascontiguousarray(less, dtype=np.int64)
    eq = np.ascontiguousarray(equal, dtype=np.int64)
    npr = np.ascontiguousarray(n_per_rank, dtype=np.int64)
    sp = np.ascontiguousarray(splitters, dtype=np.uint64)
    R = le.shape[0]
    t = np.zeros(parts - 1, dtype=np.int64)
    m = np.zeros((R, parts), dtype=np.int64)
    masks = np.zeros(R, dtype=np.uint32)
    i64 = C.POINTER(C.c_int64)
    _abi.check(lib.b200_multi_plan_splitters(le.ctypes.data_as(i64), eq.ctypes.data_as(i64), npr.ctypes.data_as(i64),
                                             sp.ctypes.data_as(C.POINTER(C.c_uint64)), R, parts, 1 if exact_ties else 0,
                                             t.ctypes.data_as(i64), m.ctypes.data_as(i64),
                                             masks.ctypes.data_as(C.POINTER(C.c_uint32))))
    return t, m, masks


_BIT_OF_DIGIT = ((np.arange(256)[:, None] >> np.arange(8)[None, :]) & 1).astype(np.int64)  # [256, 8]


def global_vary_bits(global_hist: np.ndarray) -> int:
    """Bits of the key image (level l = bits 8l..8l+7, the first 8 levels) in which some two records of the whole job
    differ, from the summed per-level histograms [L, 256]."""
    g = np.asarray(global_hist, dtype=np.int64)[:8]
    n = int(g[0].sum())
    ones = g @ _BIT_OF_DIGIT  # [L, 8]: records whose digit on level l has bit b set
    varies = (ones > 0) & (ones < n)
    vary = 0
    for l, b in zip(*np.nonzero(varies)):
        vary |= 1 << (8 * int(l) + int(b))
    return vary


def range_vary_bits(global_vary: int, lo_key: int, hi_key: int) -> int:
    """Bits that can vary among keys of [lo_key, hi_key]: everything above the highest bit in which the two bounds
    differ is the same for the whole range.  A superset of the truth, which is all b200_sort_device_hinted needs."""
    x = int(lo_key) ^ int(hi_key)
    return global_vary & ((1 << x.bit_length()) - 1)


def presplit_layout(rows: np.ndarray, split: np.ndarray, parts: int, rank: int):
    """Where rank `rank` stores its records of every digit when the exchange delivers each digit of the partition level
    as its own run (`presplit`): rows = every rank's 256-bin row on that level [P, 256].  Returns (owner[256] = the rank
    that owns digit d, offset[256] = first record of THIS rank's run of digit d inside the owner's receive buffer,
    counts[256] = the bucket sizes of the array this rank receives).  A receive buffer is laid out bucket by bucket in
    digit order, inside a bucket sender by sender in rank order; the runs of all senders tile it without gaps."""
    tot = np.asarray(rows, dtype=np.int64)
    gtot = tot.sum(axis=0)
    owner = part_of_digit(split, parts)
    bucket_start = np.zeros(256, dtype=np.int64)
    for g in range(parts):
        idx = np.nonzero(owner == g)[0]
        if len(idx):
            bucket_start[idx] = np.concatenate([[0], np.cumsum(gtot[idx])[:-1]])
    offset = bucket_start + tot[:rank].sum(axis=0)
    counts = np.where(owner == rank, gtot, 0).astype(np.uint64)
    return owner, offset, counts


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


# ---- full-resolution splitters for skewed keys (sample-sort style) ----
def digit_split_imbalance(m: np.ndarray) -> float:
    """max records any rank would receive / the even share, for a send matrix M[r, g]."""
    total = int(m.sum())
    return float(m.sum(axis=0).max()) * m.shape[1] / total if total else 1.0


def choose_sample_splitters(samples: np.ndarray, weights: np.ndarray, parts: int) -> np.ndarray:
    """P-1 splitter prefixes = weighted quantiles g/P of the gathered samples (weights =
    records each sample stands for).  Duplicates are kept: a value heavier than one part
    yields several equal splitters, and the tie thresholds split it by source rank."""
    order = np.argsort(samples, kind="stable")
    v, w = samples[order], weights[order].astype(np.float64)
    cum = np.cumsum(w)
    total = cum[-1]
    idx = [min(len(v) - 1, int(np.searchsorted(cum, total * g / parts, side="left"))) for g in range(1, parts)]
    return v[idx].astype(np.uint64)


def splitter_left_counts(less: np.ndarray, equal: np.ndarray, t: np.ndarray) -> np.ndarray:
    """left[r, b] = rank r's records left of boundary b."""
    R = less.shape[0]
    ranks = np.arange(R)[:, None]
    return less + np.where(ranks < t[None, :], equal, 0)


def plan_splitter_exchange(less: np.ndarray, equal: np.ndarray, n_per_rank: np.ndarray, splitters: np.ndarray, parts: int,
                           exact_ties: bool = True):
    """-> (t [P-1], M [R, P]): tie thresholds and the send matrix M[r, g] for the splitter partition.
    Boundary b puts a record LEFT when prefix < s_b, or prefix == s_b and source rank < t_b; t_b (0..R)
    is the threshold whose left count is closest to the target (b + 1) * N / P, non-decreasing over
    equal splitters.  exact_ties=False (keys longer than the 8-level prefix): equal prefixes are not
    equal keys, so a tie group is never split between ranks (t_b is 0 or R)."""
    R, S = less.shape
    total = int(n_per_rank.sum())
    t = np.zeros(S, dtype=np.int64)
    for b in range(S):
        target = total * (b + 1) / parts
        base = int(less[:, b].sum())
        cand = base + np.concatenate([[0], np.cumsum(equal[:, b])])  # left count for t = 0..R
        if not exact_ties:
            cand = cand[[0, R]]
        tb = int(np.argmin(np.abs(cand - target)))
        if not exact_ties:
            tb *= R
        if b and splitters[b] == splitters[b - 1]:
            tb = max(tb, int(t[b - 1]))
        t[b] = tb
    left = splitter_left_counts(less, equal, t)
    edges = np.concatenate([np.zeros((R, 1), dtype=np.int64), left.astype(np.int64), n_per_rank.astype(np.int64)[:, None]], axis=1)
    m = np.diff(edges, axis=1)
    assert (m >= 0).all(), "splitter boundaries out of order"
    return t, m


def incl_mask_for_rank(t: np.ndarray, rank: int) -> int:
    """bit b set = on this rank a record EQUAL to splitter b goes right of boundary b."""
    mask = 0
    for b, tb in enumerate(t):
        if rank >= int(tb):
            mask |= 1 << b
    return mask


def part_of_prefix(prefix: np.ndarray, splitters: np.ndarray, incl_mask: int) -> np.ndarray:
    """numpy model of the partition rule (used by the CPU test double)."""
    part = np.zeros(len(prefix), dtype=np.int64)
    for b, sb in enumerate(splitters):
        right = prefix > sb
        if (incl_mask >> b) & 1:
            right |= prefix == sb
        part += right
    return part


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
        self.c_planner = True

    def _opts(self):
        o = _abi.SortOpts()
        o.stable = 1
        o.variant = -1
        o.device = self.torch.cuda.current_device()
        o.stream = _abi.torch_stream_handle(self.torch)
        o.route = 0
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

    # ---- splitter mode (skewed keys) ----
    def prefix_supported(self):
        return bool(self.lib.b200_key_prefix_supported(C.byref(self.desc)))

    def prefix_sample(self, buf, n, stride, count):
        out = np.zeros(count, dtype=np.uint64)
        if count:
            o = self._opts()
            _abi.check(self.lib.b200_key_prefix_sample_device(buf.data_ptr(), n, C.byref(self.desc), stride, count,
                                                              out.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(o)))
            self.launches += 1
        return out

    def splitter_count(self, buf, n, splitters):
        sp = np.ascontiguousarray(splitters, dtype=np.uint64)
        less, equal = np.zeros(len(sp), dtype=np.uint64), np.zeros(len(sp), dtype=np.uint64)
        o = self._opts()
        _abi.check(self.lib.b200_splitter_count_device(buf.data_ptr(), n, C.byref(self.desc),
                                                       sp.ctypes.data_as(C.POINTER(C.c_uint64)), len(sp),
                                                       less.ctypes.data_as(C.POINTER(C.c_uint64)),
                                                       equal.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(o)))
        self.launches += 1
        return less.astype(np.int64), equal.astype(np.int64)

    def partition_splitters(self, src, dst, dst_ptrs, n, splitters, incl_mask, part_counts):
        sp = np.ascontiguousarray(splitters, dtype=np.uint64)
        pc = np.ascontiguousarray(part_counts, dtype=np.uint64)
        ptrs = None
        if dst_ptrs is not None:
            ptrs = (C.c_void_p * len(dst_ptrs))(*[C.c_void_p(int(x)) for x in dst_ptrs])
        o = self._opts()
        _abi.check(self.lib.b200_partition_splitters_device(src.data_ptr(), dst.data_ptr() if dst is not None else None,
                                                            ptrs, n, C.byref(self.desc),
                                                            sp.ctypes.data_as(C.POINTER(C.c_uint64)), len(sp),
                                                            int(incl_mask), pc.ctypes.data_as(C.POINTER(C.c_uint64)),
                                                            C.byref(o)))
        self.launches += 1

    def sort(self, buf, n, tmp=None):
        o = self._opts()
        st = _abi.Stats()
        _abi.check(self.lib.b200_sort_device(buf.data_ptr(), tmp.data_ptr() if tmp is not None else None, n,
                                             C.byref(self.desc), C.byref(o), C.byref(st)))
        self.launches += st.kernel_launches
        return st.as_dict()

    def presplit_ok(self):
        """Records that are nothing but their key, 4 or 8 bytes: the shape the exchange may deliver split on all 256
        digits of the partition level, so that the local sort starts at its second MSD pass."""
        k = self.key
        return (k.kind in ("uint", "int", "float") and k.elem_size in (4, 8) and k.key_offset == 0
                and k.key_width == k.elem_size and k.levels == k.elem_size)

    def sort_presplit(self, buf, tmp, n, level, counts):
        """Local sort of an array that arrives partitioned on `level` (b200_sort_device_presplit).  Returns
        (the buffer holding the result: buf or tmp, stats)."""
        o = self._opts()
        st = _abi.Stats()
        cnt = np.ascontiguousarray(counts, dtype=np.uint64)
        res = C.c_void_p()
        _abi.check(self.lib.b200_sort_device_presplit(buf.data_ptr(), tmp.data_ptr(), n, C.byref(self.desc), int(level),
                                                      cnt.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(o), C.byref(st),
                                                      C.byref(res)))
        self.launches += st.kernel_launches
        return (tmp if res.value == tmp.data_ptr() else buf), st.as_dict()

    def sort_hinted(self, buf, n, tmp, vary_bits):
        """b200_sort_device_hinted: the local sort without its counting pass (the caller knows which bits can vary)."""
        o = self._opts()
        st = _abi.Stats()
        _abi.check(self.lib.b200_sort_device_hinted(buf.data_ptr(), tmp.data_ptr() if tmp is not None else None, n,
                                                    C.byref(self.desc), C.c_uint64(int(vary_bits)), C.byref(o), C.byref(st)))
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


def _pipelined_exchange(eng, key, local, scratch, peers, bounds, hist_c, level, split, world, rank, starts, send_counts):
    """Partition + exchange as a two-stage pipeline: the shard is partitioned chunk by chunk into
    local send regions (K2 on the SMs, HBM speed) while the copy engines push the previous chunk's
    regions into the owners' receive buffers over NVLink.  The fused kernel stores straight into the
    peers, but an SM whose store queue is full of NVLink writes also stalls its ranking warps
    (measured: 13.5 ms at N=8 = partition + drain, not max(partition, drain)); the copy engines need
    no SM at all.  Destination layout is the fused path's: by source rank, then in input order."""
    torch = eng.torch
    es = key.elem_size
    part = scratch["part"]
    streams = scratch.setdefault("_ce_streams", [torch.cuda.Stream() for _ in range(2)])
    main = torch.cuda.current_stream()
    sent = np.zeros(world, dtype=np.int64)  # records of this rank already placed per destination
    views = scratch.setdefault("_peer_views", {})
    for g in range(world):
        if g not in views:
            views[g] = peers.local if g == rank else _device_view(torch, peers.ptrs[g], peers.capacity)
    part_of = part_of_digit(split, world)
    for c in range(len(bounds) - 1):
        lo, hi = bounds[c], bounds[c + 1]
        # the chunk's records per destination, from its own histogram; this rank's own records go
        # straight into its receive buffer, the others into send regions packed at the chunk's place
        pc_plan = np.bincount(part_of, weights=hist_c[c][level].astype(np.float64), minlength=world).astype(np.int64)
        off = lo + np.concatenate([[0], np.cumsum(pc_plan)[:-1]])
        dst = [part.data_ptr() + int(off[g]) * es for g in range(world)]
        dst[rank] = peers.local_ptr + (int(starts[rank]) + int(sent[rank])) * es
        # returns when the chunk's send regions are complete (the entry point synchronises the stream)
        pc = eng.partition_peers(local[lo * es: hi * es], dst, hi - lo, level, hist_c[c][level], split, world)
        assert np.array_equal(pc, pc_plan), "partition counts disagree with the chunk histogram"
        sent[rank] += int(pc[rank])
        for k in range(world - 1):
            g = (rank + 1 + k) % world  # at any moment every destination hears from one source
            cnt = int(pc[g])
            if cnt == 0:
                continue
            d0 = (int(starts[g]) + int(sent[g])) * es
            with torch.cuda.stream(streams[k % len(streams)]):
                views[g][d0: d0 + cnt * es].copy_(part[int(off[g]) * es: (int(off[g]) + cnt) * es], non_blocking=True)
            sent[g] += cnt
    assert np.array_equal(sent, send_counts), "partition counts disagree with the gathered histograms"
    for st in streams:
        main.wait_stream(st)


# --------------------------------------------------------------------------- the sort
MAX_SPLITTERS = 15        # include/b200sort.h: nsplit <= 15
SAMPLES_PER_RANK = 8192   # key prefixes each rank contributes to the splitter choice


def sharded_sort(local, n_local: int, key: RadixKey, engine=None, group=None, recv_slack: float = 1.3,
                 scratch=None, timings=None, balance_tol: float = 1.25):
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
    # 1. K1 on the local shard, all levels in one read (chunk by chunk for the pipelined exchange,
    #    which needs every chunk's own histogram of the partition level)
    ce_chunks = int(scratch.get("ce_chunks", 0)) if scratch else 0
    pipelined = (ce_chunks > 1 and scratch.get("peers") is not None and scratch.get("part") is not None
                 and hasattr(eng, "torch") and n_local >= ce_chunks * 4096 and key.elem_size % 4 == 0
                 and key.elem_size <= 16)
    bounds = hist_c = None
    if pipelined:
        # interior chunk bounds on multiples of 16 records: every supported record size then starts a chunk on a
        # 16-byte boundary, which the word kernels and the fast histogram need
        bounds = [min(n_local, ((n_local * c) // ce_chunks) // 16 * 16) for c in range(ce_chunks)] + [n_local]
        es = key.elem_size
        hist_c = [eng.histogram(local[bounds[c] * es: bounds[c + 1] * es], bounds[c + 1] - bounds[c]) for c in range(ce_chunks)]
        hist = np.sum(hist_c, axis=0).astype(np.uint64)
    else:
        hist = eng.histogram(local, n_local)
    # 2. global histogram + every rank's rows: one small all-gather ([P, L, 256] i64)
    mine = eng.to_comm(hist.astype(np.int64))
    gathered = [mine.new_empty(mine.shape) for _ in range(world)]
    dist.all_gather(gathered, mine, group=group)
    all_hist = np.stack([g.cpu().numpy() for g in gathered])  # [P, L, 256]
    c_planner = getattr(eng, "c_planner", False)  # product engine: the planner of the C library
    if c_planner:
        level, split, m, _imb = plan_digits(all_hist)
    else:
        level = pick_partition_level(all_hist.sum(axis=0))
    if level < 0:
        return local, n_local  # every key equal: any arrangement is sorted
    # 3. splitters on digit boundaries of that level
    if not c_planner:
        split = choose_splitters(all_hist.sum(axis=0)[level], world)
        m = send_matrix(all_hist[:, level, :], split, world)
    # 3b. skewed keys: digit boundaries cannot balance the ranks -> full-resolution splitters
    splitters = t_thr = None
    import time as _time

    plan_t = {}
    if (digit_split_imbalance(m) > balance_tol and world - 1 <= MAX_SPLITTERS and hasattr(eng, "prefix_supported")
            and eng.prefix_supported()):
        n_per_rank = all_hist[:, 0, :].sum(axis=1).astype(np.int64)
        cnt = int(min(n_local, SAMPLES_PER_RANK))
        stride = max(1, n_local // cnt) if cnt else 1
        _t0 = _time.perf_counter()
        smp = np.zeros(SAMPLES_PER_RANK + 1, dtype=np.int64)
        smp[0] = cnt
        smp[1: 1 + cnt] = eng.prefix_sample(local, n_local, stride, cnt).view(np.int64)
        plan_t["sample_ms"] = (_time.perf_counter() - _t0) * 1e3
        _t0 = _time.perf_counter()
        mine_s = eng.to_comm(smp)
        gs = [mine_s.new_empty(mine_s.shape) for _ in range(world)]
        dist.all_gather(gs, mine_s, group=group)
        vals, wts = [], []
        for r, g in enumerate(gs):
            a = g.cpu().numpy()
            c = int(a[0])
            if c:
                vals.append(a[1: 1 + c].view(np.uint64))
                wts.append(np.full(c, n_per_rank[r] / c))
        splitters = (plan_sample_splitters if c_planner else choose_sample_splitters)(np.concatenate(vals), np.concatenate(wts), world)
        plan_t["gather_choose_ms"] = (_time.perf_counter() - _t0) * 1e3
        _t0 = _time.perf_counter()
        less, equal = eng.splitter_count(local, n_local, splitters)
        plan_t["count_ms"] = (_time.perf_counter() - _t0) * 1e3
        _t0 = _time.perf_counter()
        mine_c = eng.to_comm(np.stack([less, equal]).astype(np.int64))
        gc = [mine_c.new_empty(mine_c.shape) for _ in range(world)]
        dist.all_gather(gc, mine_c, group=group)
        allc = np.stack([g.cpu().numpy() for g in gc])  # [P, 2, P-1]
        if c_planner:
            t_thr, m, _masks = plan_splitters(allc[:, 0, :], allc[:, 1, :], n_per_rank, splitters, world,
                                              exact_ties=key.levels <= 8)
        else:
            t_thr, m = plan_splitter_exchange(allc[:, 0, :], allc[:, 1, :], n_per_rank, splitters, world,
                                              exact_ties=key.levels <= 8)
        plan_t["gather_plan_ms"] = (_time.perf_counter() - _t0) * 1e3
    send_counts, recv_counts = m[rank], m[:, rank]
    n_recv = int(recv_counts.sum())
    presplit = presplit_counts = None
    # 4. stable local partition into P contiguous send regions, then the all-to-all
    fused = bool(scratch and scratch.get("peers") is not None)
    part = recv = None
    if not fused:
        part = scratch["part"] if scratch else eng.empty(n_local)
        recv = scratch["recv"] if scratch else eng.empty(int(max(n_recv, 1)))
        if recv.numel() < n_recv * key.elem_size:
            raise _abi.B200SortError(5, f"receive buffer too small for {n_recv} records (skewed splitters)")
    peers = scratch.get("peers") if scratch else None
    # (a rank owns 256 / P digits of the partition level, so its local sort is left with (256 / P) * 256 segments: they
    # fit an SM — 16 384 records, 8192 expected — only up to 512 Mi records in the whole job; a decision every rank
    # takes from the same gathered numbers)
    n_total = int(all_hist[:, 0, :].sum())
    penv = os.environ.get("B200SORT_DIST_PRESPLIT", "")
    presplit = bool(peers is not None and splitters is None and not pipelined and level >= 2 and scratch.get("tmp") is not None
                    and getattr(eng, "presplit_ok", None) and eng.presplit_ok() and penv != "0"
                    and (n_total <= 512 << 20 or penv == "1"))
    if ev:
        ev[1].record()
    if peers is not None:
        # fused: this rank's records for rank g start, in g's buffer, after those of the lower ranks
        if peers.capacity < n_recv * key.elem_size:
            raise _abi.B200SortError(5, f"peer receive buffer too small for {n_recv} records")
        starts = np.concatenate([np.zeros((1, world), dtype=np.int64), np.cumsum(m, axis=0)])[rank]  # [g]
        dst = [peers.ptrs[g] + int(starts[g]) * key.elem_size for g in range(world)]
        peers.fence()  # every rank is done with the previous contents of its receive buffer
        if pipelined and splitters is None:
            _pipelined_exchange(eng, key, local, scratch, peers, bounds, hist_c, level, split, world, rank, starts,
                                send_counts)
        elif splitters is not None:
            eng.partition_splitters(local, None, dst, n_local, splitters, incl_mask_for_rank(t_thr, rank), send_counts)
        elif presplit:
            # the exchange IS the first MSD pass: every digit of the partition level gets its own run in the owner's
            # buffer — buckets in digit order, inside a bucket the senders in rank order — so the receiver's array
            # arrives partitioned on that level and its local sort starts at the second pass
            owner, offset, presplit_counts = presplit_layout(all_hist[:, level, :], split, world, rank)
            dst256 = [peers.ptrs[int(owner[d])] + int(offset[d]) * key.elem_size for d in range(256)]
            pc = eng.partition_peers(local, dst256, n_local, level, hist[level], np.arange(1, 256, dtype=np.uint32), 256)
            assert np.array_equal(pc, hist[level].astype(np.int64)), "partition counts disagree with the histogram"
        else:
            pc = eng.partition_peers(local, dst, n_local, level, hist[level], split, world)
            assert np.array_equal(pc, send_counts), "partition counts disagree with the gathered histograms"
        if ev:
            ev[2].record()
        peers.fence()  # every rank's partition kernel has finished: the received range is complete
        recv = peers.local
        if ev:
            ev[3].record()
    else:
        if splitters is not None:
            eng.partition_splitters(local, part, None, n_local, splitters, incl_mask_for_rank(t_thr, rank), send_counts)
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
    hint = 0
    if (key.levels <= 8 and hasattr(eng, "sort_hinted") and os.environ.get("B200SORT_DIST_HINT", "1") != "0"):
        # which key bits can vary inside this rank's range: the job's varying bits below the highest bit in which the
        # range's bounds differ — the receiver then sorts without counting its records first
        gv = global_vary_bits(all_hist.sum(axis=0))
        full = (1 << (8 * key.levels)) - 1
        if splitters is not None:
            lo_key = int(splitters[rank - 1]) if rank > 0 else 0
            hi_key = int(splitters[rank]) if rank < world - 1 else full
        else:
            # (every level above the partition level is one value for the whole job: it cancels in lo ^ hi)
            owned = np.nonzero((part_of_digit(split, world) == rank) & (all_hist[:, level, :].sum(axis=0) > 0))[0]
            lo_key = (int(owned.min()) if len(owned) else 0) << (8 * level)
            hi_key = ((int(owned.max()) if len(owned) else 0) << (8 * level)) | ((1 << (8 * level)) - 1)
        hint = range_vary_bits(gv, lo_key, hi_key)
    if presplit and scratch["tmp"].numel() >= n_recv * key.elem_size:
        recv, st = eng.sort_presplit(recv, scratch["tmp"], n_recv, level, presplit_counts)
    elif hint:
        st = eng.sort_hinted(recv, n_recv, scratch.get("tmp") if scratch else None, hint)
    else:
        st = eng.sort(recv, n_recv, scratch.get("tmp") if scratch else None)
    if timings is not None:
        timings.update({"level": level, "presplit": bool(presplit), "vary_hint": hex(hint), "split": split.tolist(), "n_recv": n_recv, "local": st,
                        "splitter_mode": "prefix" if splitters is not None else "digit", "splitter_plan_ms": plan_t,
                        "exchange_path": "pipelined" if (pipelined and splitters is None and peers is not None) else
                        ("fused" if peers is not None else "all_to_all"),
                        "sent_bytes": int((send_counts.sum() - send_counts[rank]) * key.elem_size)})
        if ev:
            ev[4].record()
            ev[4].synchronize()
            timings["phase_ms"] = {"hist_allgather_plan": ev[0].elapsed_time(ev[1]), "partition": ev[1].elapsed_time(ev[2]),
                                   "all_to_all": ev[2].elapsed_time(ev[3]), "local_sort": ev[3].elapsed_time(ev[4])}
    return recv, n_recv
