"""Multi-GPU parity worker (run under torchrun by tests/test_gpu_dist.py): every rank sorts its
shard of a seeded array with rdst_b200.dist.sharded_sort (CUDA engine, NCCL exchange); rank 0
gathers the outputs and compares their concatenation with the oracle's stable sort."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_ffi as o
from rdst_b200 import RadixKey
from rdst_b200.dist import sharded_sort


def cases():
    rng = np.random.default_rng(4321)
    yield "u64_uniform", RadixKey.for_dtype("u8"), rng.integers(0, 2**64, 3_000_001, dtype="u8").view(np.uint8)
    yield "u64_small", RadixKey.for_dtype("u8"), rng.integers(0, 70_000, 1_000_000, dtype="u8").view(np.uint8)
    yield "all_equal", RadixKey.for_dtype("u4"), np.full(100_000, 9, dtype="u4").view(np.uint8)
    yield "zipf", RadixKey.for_dtype("u8"), np.minimum(rng.zipf(1.1, 2_000_000), 2**32).astype("u8").view(np.uint8)
    yield "f32", RadixKey.for_dtype("f4"), rng.standard_normal(1_500_003).astype("f4").view(np.uint8)
    zr = np.minimum(rng.zipf(1.1, 2_500_000), 2**32).astype("u8")
    yield "zipf_hashed", RadixKey.for_dtype("u8"), ((zr * np.uint64(0x9E3779B97F4A7C15)) ^ (zr >> np.uint64(7))).view(np.uint8)
    yield "two_values", RadixKey.for_dtype("u8"), np.where(rng.random(1_000_000) < 0.8, 5, 1 << 40).astype("u8").view(np.uint8)
    yield "f64_normal", RadixKey.for_dtype("f8"), rng.standard_normal(1_300_001).view(np.uint8)
    b = np.zeros((400_000, 2), dtype="u8")
    b[:, 0] = rng.integers(0, 2**63, len(b))
    b[:, 1] = np.where(rng.random(len(b)) < 0.9, 7, rng.integers(0, 2**63, len(b)))
    yield "u128_skewed", RadixKey.u128(), b.reshape(-1).view(np.uint8)
    recd = np.dtype([("key", "<u4"), ("idx", "<u8")])
    ad = np.zeros(900_000, dtype=recd)
    ad["key"] = np.where(rng.random(len(ad)) < 0.7, 3, rng.integers(0, 1 << 30, len(ad)))
    ad["idx"] = np.arange(len(ad))
    yield "struct12_dups_stable", RadixKey.field(recd, "key", 4), ad.view(np.uint8)
    rec = np.dtype([("key", "<u4"), ("idx", "<u8")])
    a = np.zeros(1_200_000, dtype=rec)
    a["key"] = rng.integers(0, 1 << 22, len(a))
    a["idx"] = np.arange(len(a))
    yield "struct12_stable", RadixKey.field(rec, "key", 3), a.view(np.uint8)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    ok = True
    from rdst_b200.dist import PeerBuffers

    mode = os.environ.get("DIST_CHECK_MODE", "nccl")
    peers = PeerBuffers(64 << 20) if mode in ("peer", "ce") else None
    part = torch.empty(64 << 20, dtype=torch.uint8, device="cuda") if mode == "ce" else None
    tmp_buf = torch.empty(64 << 20, dtype=torch.uint8, device="cuda") if peers else None
    for name, key, full in cases():
        es = key.elem_size
        n = len(full) // es
        lo, hi = rank * n // world, (rank + 1) * n // world
        shard = torch.from_numpy(full[lo * es: hi * es].copy()).cuda()
        tm = {}
        scratch = None
        if peers:
            # (with a scratch buffer for the local sort the peer mode delivers key-only records split on all 256 digits
            # of the partition level and the local sort starts at its second MSD pass: rdst_b200/dist.py `presplit`)
            scratch = {"peers": peers, "tmp": tmp_buf}
            if part is not None:  # chunked partition pipelined with copy-engine pushes (3 uneven chunks)
                scratch.update({"part": part, "ce_chunks": 3})
        out, n_out = sharded_sort(shard, hi - lo, key, scratch=scratch, timings=tm)
        torch.cuda.synchronize()
        mine = out[: n_out * es].cpu()
        sizes = [torch.zeros(1, dtype=torch.int64, device="cuda") for _ in range(world)]
        dist.all_gather(sizes, torch.tensor([n_out * es], dtype=torch.int64, device="cuda"))
        sizes = [int(s.item()) for s in sizes]
        bufs = [torch.empty(s, dtype=torch.uint8, device="cuda") for s in sizes]
        for r in range(world):  # broadcast-based gather of uneven pieces
            if r == rank:
                bufs[r].copy_(mine)
            dist.broadcast(bufs[r], src=r)
        if rank == 0:
            got = torch.cat(bufs).cpu().numpy().tobytes()
            exp = o.stable_sorted(full, key).tobytes()
            good = got == exp
            print(f"[dist_check] {mode} {name} ({tm.get('exchange_path')}{', presplit' if tm.get('presplit') else ''}): {'ok' if good else 'MISMATCH'} splitters={tm.get('splitter_mode')} sizes={[s // es for s in sizes]}", flush=True)
            ok = ok and good
    if os.environ.get("DIST_CHECK_BIG"):
        ok = big_case(rank, world, mode) and ok
    dist.barrier()
    if peers:
        peers.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


def big_case(rank, world, mode):
    """DIST_CHECK_BIG=<records per rank> (>= 32 Mi so that the local sort takes the hybrid route with bit-window
    digits below the rank's fixed top bits): too large for the CPU oracle, so the GLOBAL result is checked on the
    devices — every rank's range sorted, max(rank r) <= min(rank r + 1), record count and two multiset checksums
    equal to the input's — plus oracle equality on rank 0's first 2 M output records against the oracle's sort of
    the same key range gathered from all inputs."""
    from rdst_b200.dist import PeerBuffers

    n = int(os.environ["DIST_CHECK_BIG"])
    key = RadixKey.for_dtype("u8")
    g = torch.Generator(device="cuda").manual_seed(77 + rank)
    src = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)
    work = src.clone()
    scratch = None
    peers = None
    if mode in ("peer", "ce"):
        peers = PeerBuffers(int(n * 8 * 1.3))
        scratch = {"peers": peers, "tmp": torch.empty(int(n * 8 * 1.3), dtype=torch.uint8, device="cuda")}
        if mode == "ce":
            scratch.update({"part": torch.empty(n * 8, dtype=torch.uint8, device="cuda"), "ce_chunks": 3})
    tm = {}
    out, n_out = sharded_sort(work.view(torch.uint8).view(-1), n, key, scratch=scratch, timings=tm)
    torch.cuda.synchronize()
    res = out[: n_out * 8].view(torch.int64)
    u = res ^ (-2**63)
    ok = bool((u[1:] >= u[:-1]).all())
    fl = torch.tensor([int(u[0]), int(u[-1]), n_out], dtype=torch.int64, device="cuda")
    allfl = [torch.empty_like(fl) for _ in range(world)]
    dist.all_gather(allfl, fl)
    allfl = [[int(v) for v in t.tolist()] for t in allfl]
    ok = ok and all(a[1] <= b[0] for a, b in zip(allfl, allfl[1:]))
    chk = lambda t: [int(t.sum()), int(((t * 0x9E3779B1) ^ (t >> 7)).sum())]
    tot = torch.tensor([n_out] + chk(res) + chk(src), dtype=torch.int64, device="cuda")
    dist.all_reduce(tot)
    tot = [int(v) for v in tot.tolist()]
    ok = ok and tot[0] == n * world and tot[1:3] == tot[3:5]
    # rank 0's first 2 M outputs = the 2 M smallest keys overall: gather every rank's candidates (its 2 M smallest)
    m = min(2_000_000, n)
    cand = torch.sort(src ^ (-2**63)).values[:m] ^ (-2**63)
    allc = [torch.empty_like(cand) for _ in range(world)]
    dist.all_gather(allc, cand)
    if rank == 0:
        pool = torch.cat(allc).cpu().numpy().view("u8")
        exp = o.sorted_copy(pool, key)[:m]
        ok = ok and res[:m].cpu().numpy().view("u8").tobytes() == exp.tobytes()
        print(f"[dist_check] {mode} big u64 x {n} per rank, world {world} (presplit={tm.get('presplit')} local sort hybrid={tm['local'].get('hybrid')} "
              f"unstable={tm['local'].get('unstable')}): {'ok' if ok else 'MISMATCH'} per-rank {[x[2] for x in allfl]}", flush=True)
    if peers:
        del scratch
        peers.close()
    return ok


if __name__ == "__main__":
    main()
