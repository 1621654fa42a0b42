#!/usr/bin/env python3
"""bench.py — Gkeys/s of the sort path on N B200s (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]

Workload at N=1 (default): BASELINE.json configs[1] — u64 keys, 1 B uniform random,
8 levels (uniform data has no trivial level), device resident.  At N>1 (torchrun, one rank
per GPU): configs[4] — the same 1 B u64 keys PER GPU (weak scaling), sharded, top-level
range partition + NCCL all-to-all + local sort.  One step = one full sort of one batch.

Timing: CUDA events on the launching stream around each step's sort; the in-place sort's
input is restored (device-to-device copy) between steps outside the event pairs, so the
timed region starts with the unsorted input resident in HBM.  Inputs (8 GB) are far larger
than L2, no flush needed.  Max over ranks.  `e2e` = same metric through the public host API
with HOST buffers, H2D + sort + D2H inside the timed region: at N=1 radix_sort_builder(pinned numpy
array).sort() -> b200_sort_host (plus `e2e_pageable`: the same call on ordinary pageable memory, what a
Rust Vec is); at N>1 rdst_b200.sort_multi_host -> b200_sort_multi_host, ONE process driving all N GPUs
from one host array of N x n keys (run by rank 0 while the other ranks wait on a CPU barrier).
After the timed loop at N>1 the GLOBAL result is verified on the devices (outside the timed region):
every rank's range sorted under the key's order, max(rank r) <= min(rank r + 1), record count and two
multiset checksums equal to the input's; the outcome is the `verified` object of the JSON line.
--all-configs prints one line per BASELINE.json config that fits this GPU count (kept in profiles/).

--impl reference: the reference's CPU path.  The reference is a pure-Rust crate and this
image has no Rust toolchain, so what is timed is oracle/ (the C restatement of rdst's
default multi-threaded path: StandardTuner -> Scanning/Recombinating/Ska/Lsb/...), with all
host threads, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np

WORKLOADS = {
    # name: (dtype, n per GPU, generator, description)
    "u64_1b_uniform": ("u8", 1_000_000_000, "uniform", "BASELINE configs[1]: u64 keys, 1B uniform random, LSD 8 levels"),
    "u32_100m_uniform": ("u4", 100_000_000, "uniform", "BASELINE configs[0]: u32, 100M uniform random keys"),
    "u64_1b_40bit": ("u8", 1_000_000_000, "low40", "configs[1] variant: keys < 2^40, 3 trivial levels skipped"),
    "u64_1b_61bit": ("u8", 1_000_000_000, "low61",
                     "configs[1] variant: keys < 2^61 (what one rank of an 8-GPU sharded sort holds): MSD digits are bit windows"),
    "f32_500m": ("f4", 500_000_000, "uniform", "BASELINE configs[2]: f32 random bit patterns, 500M"),
    "i64_500m": ("i8", 500_000_000, "uniform", "BASELINE configs[2]: i64, 500M incl. negatives"),
    "u64_1b_lowcard": ("u8", 1_000_000_000, "lowcard", "skew stress: every key byte is 0 or 1 (two buckets per level)"),
    "u64_1b_zipf": ("u8", 1_000_000_000, "zipf", "configs[4] skewed: Zipf(1.1) ranks as u64 keys"),
    "u64_1b_zipf_hashed": ("u8", 1_000_000_000, "zipf_hashed",
                           "configs[4] skewed: hashed Zipf(1.1) ranks (heavy duplicates, every level non-trivial)"),
    "struct12_1b_stable": ("rec12", 1_000_000_000, "struct",
                           "BASELINE configs[3]: {u32 key with LEVELS=3, 8-byte payload} x 1B records, stable order"),
}
# the BASELINE.json configs a 1-GPU run covers (--all-configs)
ALL_CONFIGS_N1 = ["u32_100m_uniform", "u64_1b_uniform", "f32_500m", "i64_500m", "struct12_1b_stable"]
ALL_CONFIGS_NX = ["u64_1b_uniform", "u64_1b_zipf", "u64_1b_zipf_hashed"]
CPU_SAMPLE = 200_000_000  # records of the same generator the CPU baseline sorts per step


def key_of(dtype):
    from rdst_b200 import RadixKey

    if dtype == "rec12":  # benches/struct_sort.rs shape scaled down to BASELINE configs[3]: partial-LEVELS u32 key + payload
        return RadixKey("uint", 12, 3, 0, 4)
    return RadixKey.for_dtype(dtype)


def np_dtype(dtype):
    return np.dtype([("key", "<u4"), ("idx", "<u8")]) if dtype == "rec12" else np.dtype(dtype)


def gen_host(dtype, n, gen, seed):
    if dtype == "rec12":
        rng = np.random.default_rng(seed)
        a = np.zeros(n, dtype=np_dtype(dtype))
        a["key"] = rng.integers(0, 1 << 24, n, dtype=np.uint32)
        a["idx"] = np.arange(n, dtype=np.uint64)
        return a
    rng = np.random.default_rng(seed)
    dt = np.dtype(dtype)
    if gen in ("zipf", "zipf_hashed"):
        r = np.minimum(rng.zipf(1.1, n), 2**32).astype(np.uint64)
        if gen == "zipf_hashed":
            r = (r * np.uint64(0x9E3779B97F4A7C15)) ^ (r >> np.uint64(7))
        return r.astype(dt)
    a = rng.integers(0, 2**64, n, dtype=np.uint64) if dt.itemsize == 8 else rng.integers(0, 2**32, n, dtype=np.uint32)
    if gen == "low40":
        a &= np.uint64((1 << 40) - 1)
    if gen == "low61":
        a &= np.uint64((1 << 61) - 1)
    if gen == "lowcard":
        a &= np.uint64(0x0101010101010101)
    if dt.kind == "f":  # NaN/Inf-free random bit patterns (BASELINE configs[2])
        a = np.where((a >> 23) & 0xFF == 0xFF, a & np.uint32(0x807FFFFF), a).astype(np.uint32)
    return a.view(dt)


def gen_device(torch, dtype, n, gen, seed):
    """Same distributions generated on the device (flat int64/int32 storage of the key bits)."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    if dtype == "rec12":  # [n, 3] int32: key (24 random bits), payload = the record's input index (two words)
        rec = torch.empty((n, 3), dtype=torch.int32, device="cuda")
        step = 1 << 28
        for i in range(0, n, step):
            j = min(n, i + step)
            rec[i:j, 0] = torch.randint(0, 1 << 24, (j - i,), dtype=torch.int32, device="cuda", generator=g)
            rec[i:j, 1] = torch.arange(i, j, dtype=torch.int32, device="cuda")
        rec[:, 2] = 0
        return rec
    dt = np.dtype(dtype)
    if dt.itemsize == 8:
        if gen in ("zipf", "zipf_hashed"):
            u = torch.rand(n, dtype=torch.float64, device="cuda", generator=g)
            # inverse-CDF approximation of Zipf(s=1.1) over 2^32 ranks: r = (1 - u*(1 - R^(1-s)))^(1/(1-s))
            s, R = 1.1, 2.0**32
            r = (1.0 - u * (1.0 - R ** (1.0 - s))) ** (1.0 / (1.0 - s))
            r = r.clamp_(1, R).to(torch.int64)
            if gen == "zipf_hashed":
                r = (r * -7046029254386353131) ^ (r >> 7)  # 0x9E3779B97F4A7C15 as int64; r < 2^33 so >> is logical
            return r
        t = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)
        if gen == "low40":
            t &= (1 << 40) - 1
        if gen.startswith("low61"):  # (low61heavyNN: one rank's range of an 8-way split that owns the job's top key)
            t &= (1 << 61) - 1
            gen = gen[5:] or "low61"
        if gen == "lowcard":
            t &= 0x0101010101010101
        if gen.startswith("heavy"):  # heavyNN: NN % of the records are one key (what the rank owning a job's top key sees)
            m = torch.rand(n, dtype=torch.float32, device="cuda", generator=g) < (int(gen[5:]) / 100.0)
            t[m] = 0x1234567812345678
        return t
    t = torch.randint(-2**31, 2**31, (n,), dtype=torch.int32, device="cuda", generator=g)
    if dt.kind == "f":
        exp_all_ones = ((t >> 23) & 0xFF) == 0xFF
        t = torch.where(exp_all_ones, t & -2139095041, t)  # clear exponent bits 0x7F800000: NaN/Inf-free
    return t


class ClockSampler:
    """SM clocks / throttle reasons sampled DURING the timed region: NVML polled every 10 ms from
    a thread (the timed region is a fraction of a second, too short for an nvidia-smi loop to
    start up), `nvidia-smi -lms` as the fallback."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []
        self.sm, self.mx, self.reasons = [], [], set()
        self.nvml, self.handle, self.stop_flag, self.t = None, None, False, None

    def _nvml_handle(self):
        import pynvml

        pynvml.nvmlInit()
        try:
            import torch

            uuid = str(torch.cuda.get_device_properties(self.index).uuid)
            if not uuid.startswith("GPU-"):
                uuid = "GPU-" + uuid
            h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if hasattr(uuid, "encode") else uuid)
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
        return pynvml, h

    def _poll(self):
        n = self.nvml
        bits = {"hw_slowdown": getattr(n, "nvmlClocksEventReasonHwSlowdown", 0x8),
                "hw_thermal_slowdown": getattr(n, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                "sw_thermal_slowdown": getattr(n, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                "sw_power_cap": getattr(n, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        get_reasons = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag:
            try:
                self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
                r = int(get_reasons(self.handle))
                for name, bit in bits.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.01)

    def start(self):
        try:
            self.nvml, self.handle = self._nvml_handle()
            self.mx.append(float(self.nvml.nvmlDeviceGetMaxClockInfo(self.handle, self.nvml.NVML_CLOCK_SM)))
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.nvml is not None:
            self.stop_flag = True
            self.t.join(timeout=2)
            return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                    "reasons": sorted(self.reasons), "samples": len(self.sm), "how": "NVML polled every 10 ms during the timed region"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])), mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "how": "nvidia-smi -lms 100"}


def cpu_baseline(dtype, gen, steps, warmup, sample=CPU_SAMPLE):
    """oracle/ (C restatement of rdst's default multi-threaded path) timed on the host cores."""
    import oracle_ffi as o

    key = key_of(dtype)
    cores = o.hardware_threads()
    src = gen_host(dtype, sample, gen, 0x5EED0002)
    work = np.empty_like(src)
    times = []
    for i in range(warmup + steps):
        np.copyto(work, src)
        t0 = time.perf_counter()
        if dtype == "rec12":  # rdst's stable order, multi-threaded: a tuner answering MtLsb (src/sorts/mt_lsb_sort.rs)
            o.sort(work, key, tuner="single_algo", algo="MtLsb", multi_threaded=True, threads=cores)
        else:
            o.sort(work, key, tuner="standard", multi_threaded=True, threads=cores)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    assert o.is_sorted(work[:: max(1, sample // 1_000_000)].copy(), key)
    ok = bool(np.all(work[1:] >= work[:-1])) if np_dtype(dtype).kind == "u" else True
    assert ok, "CPU baseline produced an unsorted result"
    return {"value": sample / statistics.median(times) / 1e9, "unit": "Gkeys/s", "cores": cores, "kind": "port",
            "sample": f"{sample} {dtype} keys of the same generator per step (of {WORKLOADS_N[dtype, gen]}), "
                      f"median of {steps} after {warmup} warm-up; oracle/ restates rdst's StandardTuner path "
                      "(the Rust crate cannot be built in this image)",
            "times_s": [round(t, 4) for t in times]}


WORKLOADS_N = {(v[0], v[2]): v[1] for v in WORKLOADS.values()}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path (oracle port)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0  # rank 0 alone runs the CPU arm
    dtype, n, gen, desc = WORKLOADS[args.workload]
    sample = min(args.cpu_sample, n)
    cb = cpu_baseline(dtype, gen, args.steps, args.warmup, sample)
    ms = sample / cb["value"] / 1e6
    line = {
        "impl": "reference", "metric": "Gkeys/s sorted", "value": cb["value"], "unit": "Gkeys/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32 key + 8-byte payload" if dtype == "rec12" else np.dtype(dtype).name.replace("uint", "u").replace("float", "f").replace("int", "i"),
        "data": "synthetic", "config": {"workload": f"{args.workload}: {desc}", "records_per_step": sample,
                                        "note": "bounded sample of the workload, CPU host cores only"},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": "Gkeys/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def key_image(torch, t, dtype):
    """Signed integer image of the stored bit patterns that orders like the key (u: sign bit flipped, i: as is,
    f: IEEE total order) — for on-device checks."""
    kind = np_dtype(dtype).kind
    mn = -2**63 if t.dtype == torch.int64 else -2**31
    if kind == "u":
        return t ^ mn
    if kind == "f":
        return torch.where(t < 0, ~t ^ mn, t)
    return t


def checksums(torch, t):
    """Two wrapping 64-bit multiset checksums of an integer tensor (order-independent)."""
    x = t.to(torch.int64).reshape(-1)
    return int(x.sum()), int(((x * 0x9E3779B1) ^ (x >> 7)).sum())


def run_one(args, workload, torch, dist, world, rank, local_rank, cpu_group):
    from rdst_b200 import _abi, radix_sort_builder, sort_multi_host
    from rdst_b200.dist import CudaEngine, sharded_sort

    dtype, n, gen, desc = WORKLOADS[workload]
    if args.n:
        n = args.n
    key = key_of(dtype)
    es = key.elem_size
    nbytes = n * es
    struct = dtype == "rec12"

    src = gen_device(torch, dtype, n, gen, 0x5EED0002 + rank)           # pristine input, never sorted
    work = torch.empty_like(src)
    in_chk = checksums(torch, src) if world > 1 else None
    eng = CudaEngine(key)
    scratch = {"tmp": torch.empty(nbytes, dtype=torch.uint8, device="cuda")}
    if world > 1:
        slack = 1.0 + args.recv_slack
        cap = int(nbytes * slack) // es * es
        if args.exchange in ("peer", "ce"):
            from rdst_b200.dist import PeerBuffers

            scratch["peers"] = PeerBuffers(cap)
        else:
            scratch["part"] = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
            scratch["recv"] = torch.empty(cap, dtype=torch.uint8, device="cuda")
        scratch["tmp"] = torch.empty(cap, dtype=torch.uint8, device="cuda")
        if args.exchange == "ce":
            # the send regions live in the local sort's scratch buffer: it is idle until the exchange is over
            scratch["part"] = scratch["tmp"]
            scratch["ce_chunks"] = args.ce_chunks

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    wbuf = work.view(torch.uint8).view(-1)
    last = {}

    def one_step(i, timed):
        work.copy_(src)  # restore the unsorted input (untimed)
        if world > 1:
            dist.barrier()
        if timed:
            ev0[i].record()
        out, n_out = sharded_sort(wbuf, n, key, engine=eng, scratch=scratch, timings=last)
        if timed:
            ev1[i].record()
        return out, n_out

    for i in range(args.warmup):
        out, n_out = one_step(i, False)
    barrier()
    launches0 = eng.launches
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        out, n_out = one_step(i, True)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop()
    launches = eng.launches - launches0
    ms_steps = [ev0[i].elapsed_time(ev1[i]) for i in range(args.steps)]
    ms_total = sum(ms_steps)

    # ---- verification of the last step on the device, outside the timed region ----
    # every key type through its order-preserving image; at N>1 the GLOBAL order (boundaries between ranks),
    # the record count and two multiset checksums against the input's
    res = out[: n_out * es].view(torch.int64 if es == 8 else torch.int32)
    verified = {"records_out": int(n_out)}
    if struct:
        r3 = res.view(-1, 3)
        k0, k1, i0, i1 = r3[:-1, 0], r3[1:, 0], r3[:-1, 1], r3[1:, 1]
        verified["sorted"] = bool((k1 >= k0).all()) if n_out > 1 else True
        verified["stable"] = bool(((k1 > k0) | (i1 > i0)).all()) if n_out > 1 else True
        first_last = [int(r3[0, 0]), int(r3[-1, 0])] if n_out else [0, 0]
    else:
        img = key_image(torch, res, dtype)
        step = 1 << 28
        ok = True
        for i in range(0, max(0, n_out - 1), step):
            j = min(n_out - 1, i + step)
            ok = ok and bool((img[i + 1:j + 1] >= img[i:j]).all())
        verified["sorted"] = ok
        first_last = [int(img[0]), int(img[-1])] if n_out else [0, 0]
        del img
    assert verified["sorted"], "result is not sorted"
    if world > 1:
        fl = torch.tensor(first_last + [int(n_out)], dtype=torch.int64, device="cuda")
        allfl = [torch.empty_like(fl) for _ in range(world)]
        dist.all_gather(allfl, fl)
        allfl = [[int(v) for v in t.tolist()] for t in allfl]
        nonempty = [x for x in allfl if x[2] > 0]
        verified["rank_boundaries_ordered"] = all(a[1] <= b[0] for a, b in zip(nonempty, nonempty[1:]))
        tot = torch.tensor([n_out] + list(checksums(torch, res)) + list(in_chk), dtype=torch.int64, device="cuda")
        dist.all_reduce(tot)  # wrapping int64 sums
        tot = [int(v) for v in tot.tolist()]
        verified["records_total"] = tot[0]
        verified["records_expected"] = n * world
        verified["multiset_checksums_match"] = tot[1:3] == tot[3:5]
        verified["records_per_rank"] = [x[2] for x in allfl]
        assert verified["rank_boundaries_ordered"], "ranges of neighbouring ranks overlap"
        assert tot[0] == n * world, "records were lost or duplicated in the exchange"
        assert verified["multiset_checksums_match"], "output is not a permutation of the input"
    elif not struct:
        verified["multiset_checksums_match"] = checksums(torch, res) == checksums(torch, src)
        assert verified["multiset_checksums_match"]
    st_local = last.get("local", {})

    t = torch.tensor([ms_total], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total_max = float(t.item())
    max_n_recv = n
    if world > 1:
        nr = torch.tensor([int(last.get("n_recv", 0))], dtype=torch.int64, device="cuda")
        dist.all_reduce(nr, op=dist.ReduceOp.MAX)
        max_n_recv = int(nr.item())

    # ---- e2e: the public host API on HOST buffers, H2D + sort + D2H per step ----
    e2e = e2e_pageable = None
    if not args.no_e2e and world == 1:
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        npdt = np_dtype(dtype)

        def host_run(pinned):
            hs = torch.empty(n * es, dtype=torch.uint8)
            hw = torch.empty(n * es, dtype=torch.uint8)
            if pinned:
                hs, hw = hs.pin_memory(), hw.pin_memory()
            hs.copy_(src.view(torch.uint8).view(-1))
            harr = hw.numpy().view(npdt)
            times, b = [], None
            for i in range(1 + e2e_steps):
                hw.copy_(hs)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                b = radix_sort_builder(harr, key)
                (b.with_stable(True) if struct else b).sort()  # blocking: returns when the sorted data is back in host memory
                times.append(time.perf_counter() - t0)
            times = times[1:]
            k = harr["key"] if struct else harr
            assert bool(np.all(k[1:] >= k[:-1])) if npdt.kind != "f" else True
            sec = sum(times) / len(times)
            return {"value": n / sec / 1e9, "unit": "Gkeys/s", "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": nbytes,
                    "steps": e2e_steps, "ms_per_step": sec * 1e3,
                    "api": f"rdst_b200.radix_sort_builder({'pinned' if pinned else 'pageable'} numpy array).sort() -> b200_sort_host",
                    "ms_h2d": b.stats["ms_h2d"], "ms_d2h_after_sort": b.stats["ms_d2h"],
                    "ms_sort_device": b.stats["ms_hist"] + b.stats["ms_passes"], "ms_total_device_clock": b.stats["ms_total"],
                    "overlap": "K0/K1 run behind the H2D chunks; D2H starts behind the first slice of the segment sort (pinned)"}

        e2e = host_run(True)
        e2e_pageable = host_run(False)
        e2e_pageable["copy_threads"] = int(os.environ.get("B200SORT_COPY_THREADS", "0")) or "min(8, host threads / 2)"
    elif not args.no_e2e:
        # ONE process drives all N GPUs from one host array (b200_sort_multi_host): rank 0 runs it, the others wait on
        # a CPU barrier.  The per-rank device buffers of the timed loop are released first.
        del work, wbuf, out, res
        scratch.clear()
        _abi.lib().b200_release_scratch(-1)
        torch.cuda.empty_cache()
        dist.barrier(group=cpu_group)
        e2e = {"value": None}
        if rank == 0:
            try:
                import psutil

                avail = psutil.virtual_memory().available
            except Exception:
                avail = 0
            n_e2e = n
            while n_e2e > (1 << 20) and 2.3 * world * n_e2e * es > avail > 0:
                n_e2e //= 2
            total = n_e2e * world
            hs = torch.empty(total * es, dtype=torch.uint8).pin_memory()
            hw = torch.empty(total * es, dtype=torch.uint8).pin_memory()
            for r in range(world):  # the same per-rank generator, shard by shard through this rank's GPU
                shard = gen_device(torch, dtype, n_e2e, gen, 0x5EED0002 + r)
                hs[r * n_e2e * es:(r + 1) * n_e2e * es].copy_(shard.view(torch.uint8).view(-1))
                del shard
            torch.cuda.empty_cache()
            harr = hw.numpy().view(np_dtype(dtype))
            times, st = [], None
            e2e_steps = max(1, min(args.steps, args.e2e_steps, 2))
            for i in range(1 + e2e_steps):
                hw.copy_(hs)
                t0 = time.perf_counter()
                st = sort_multi_host(harr, key, gpus=list(range(world)))
                times.append(time.perf_counter() - t0)
            times = times[1:]
            ok = True
            stepv = 1 << 27
            kk = harr["key"] if struct else harr
            for i in range(0, total - 1, stepv):
                j = min(total - 1, i + stepv)
                ok = ok and bool(np.all(kk[i + 1:j + 1] >= kk[i:j])) if np_dtype(dtype).kind != "f" else ok
            assert ok, "multi-GPU host sort result is not sorted"
            sec = sum(times) / len(times)
            e2e = {"value": total / sec / 1e9, "unit": "Gkeys/s", "h2d_bytes_per_step": total * es, "d2h_bytes_per_step": total * es,
                   "steps": e2e_steps, "ms_per_step": sec * 1e3, "records": total, "records_per_gpu": n_e2e,
                   "api": "rdst_b200.sort_multi_host(pinned numpy array) -> b200_sort_multi_host: one process, all N GPUs, "
                          "one host array of N x n keys sorted in place",
                   "ms_h2d": st["ms_h2d"], "ms_d2h": st["ms_d2h"], "ms_plan": st["ms_plan"], "ms_exchange": st["ms_exchange"],
                   "ms_local_sort": st["ms_local_sort"], "splitter_mode": st["splitter_mode"], "imbalance": st["imbalance"],
                   "verified_sorted": bool(ok)}
            if n_e2e != n:
                e2e["note"] = f"host RAM on this box holds {n_e2e} records per GPU twice over (pinned), not {n}: e2e ran at that size"
            del hs, hw
            _abi.lib().b200_multi_release()
        dist.barrier(group=cpu_group)
    if rank != 0:
        return None

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    passes = int(st_local.get("levels_run", 0))
    ms_passes = float(st_local.get("ms_passes", 0.0))
    n_sorted = int(last.get("n_recv", n)) if world > 1 else n
    roof = None
    traffic_db = {}
    try:  # dram bytes per launch from the committed ncu --set full captures
        traffic_db = json.load(open(os.path.join(ROOT, "profiles", "k2_traffic.json")))
    except Exception:
        pass
    if passes and ms_passes > 0:
        hybrid = int(st_local.get("hybrid", 0)) == 1
        unstable = int(st_local.get("unstable", 0))
        moved_per_launch = 2 * n_sorted * es  # every kernel of the path reads and writes each record once
        whole_bytes = n_sorted * es * (1 + 2 * passes)
        ms_whole = st_local["ms_hist"] + ms_passes
        whole = whole_bytes / (ms_whole * 1e-3) / 1e9
        if hybrid:
            # dominant kernel = the segment sort (K3-U for key-only records, K3 otherwise): ONE read and ONE write of
            # every record through HBM; the levels it covers run in shared memory
            nl = int(st_local["local_levels"])
            ms_k3 = float(st_local["ms_local"])
            ms_k2 = (ms_passes - ms_k3 - float(st_local["ms_joint"])) / 2.0
            kname = ("K3-U local_count_kernel (key-only records: one wide counting-sort level + per-run comparison fix-up)"
                     if unstable & 1 else "K3 local_sort_kernel (stable shared-memory segment sort)")
            ach = moved_per_launch / (ms_k3 * 1e-3) / 1e9
            roof = {"bound": "hbm", "kernel": kname, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                    "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": moved_per_launch,
                    "launches_per_step": 1, "avg_launch_ms": ms_k3, "limiter": "instruction issue and the shared-memory pipe "
                    "(ncu: DRAM ~23 % busy); the kernel moves each record through HBM once and does the work of "
                    f"{nl} LSD levels on chip",
                    "how": "bytes = 2*n*S (one read + one write of every record) / the library's CUDA-event time of the "
                           "segment sort in the last step",
                    "lsd_equivalent_gbps": 2 * n_sorted * es * nl / (ms_k3 * 1e-3) / 1e9,
                    "k2_msd_pass": {"kernel": ("K2-U msd_pass_atomic_kernel" if unstable & 6 else
                                               "lsd_pass_or16_kernel, variant " + str(st_local.get("variant"))),
                                    "algorithmic_bytes_per_launch": moved_per_launch, "launches_per_step": 2,
                                    "avg_launch_ms": ms_k2, "achieved": moved_per_launch / (ms_k2 * 1e-3) / 1e9,
                                    "frac": moved_per_launch / (ms_k2 * 1e-3) / 1e9 / peak},
                    "ms_joint_histogram": float(st_local["ms_joint"]), "slow_segments": int(st_local.get("slow_segments", 0))}
            tr = traffic_db.get(workload + (":k3u" if unstable & 1 else ":k3"))
        else:
            ach = moved_per_launch * passes / (ms_passes * 1e-3) / 1e9
            roof = {"bound": "hbm", "kernel": "K2 LSD pass (lsd_pass_or16_kernel, variant " + str(st_local.get("variant")) + ")",
                    "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": moved_per_launch, "launches_per_step": passes,
                    "avg_launch_ms": ms_passes / passes, "limiter": "the shared-memory pipe (ranking is paid per record)",
                    "how": "CUDA events recorded by the library on the launch stream around the K2 passes of the last step"}
            tr = traffic_db.get(workload)
        if tr and tr.get("records") == n_sorted:
            roof["traffic"] = tr["traffic_bytes_per_launch"]
            roof["traffic_source"] = tr.get("source", "profiles/k2_traffic.json") + " (ncu dram__bytes_read.sum + dram__bytes_write.sum, one launch)"
            roof["achieved_by_traffic"] = roof["traffic"] / (roof["avg_launch_ms"] * 1e-3) / 1e9
            roof["frac_by_traffic"] = roof["achieved_by_traffic"] / peak
        # the contract's whole-sort figure (bytes of an LSD engine, N*S*(1 + 2*levels)) next to what moved
        roof["whole_sort"] = {"algorithmic_bytes": whole_bytes, "achieved": whole, "frac": whole / peak,
                              "frac_of_8TBs_nominal": whole / 8000.0, "ms_hist": st_local["ms_hist"], "ms_passes": ms_passes,
                              "path": ("hybrid (K1, MSD pass x2, joint histogram, segment sort)" if hybrid else "LSD (K1, K2 per level)")}
        moved = (8 if hybrid else 1 + 2 * passes + (2 if st_local.get("copy_back") else 0)) * n_sorted * es
        roof["whole_sort"]["hbm_bytes_moved"] = moved
        roof["whole_sort"]["hbm_moved_frac"] = moved / (ms_whole * 1e-3) / 1e9 / peak
    cb = None
    if world == 1 and not args.no_cpu:
        cb = cpu_baseline(dtype, gen, args.cpu_steps, 1, min(args.cpu_sample, n))
    value = n * world * args.steps / (ms_total_max * 1e-3) / 1e9
    dname = "u32 key + 8-byte payload" if struct else np.dtype(dtype).name.replace("uint", "u").replace("float", "f").replace("int", "i")
    line = {
        "metric": "Gkeys/s sorted", "value": value, "unit": "Gkeys/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total_max / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": dname, "data": "synthetic",
        "config": {"workload": f"{workload}: {desc}", "records_per_gpu": n, "record_bytes": es,
                   "levels_run": passes, "parallelism": f"range-partition x{world}" if world > 1 else "single GPU",
                   "l2": "inputs (>= 0.4 GB) larger than L2; unsorted input restored by a D2D copy between steps, untimed",
                   "timing": "CUDA events per step on the launch stream, summed; max over ranks",
                   "wall_s_timed_loop_incl_restore": t_wall},
        "ms_steps": ms_steps, "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
        "roofline": roof, "cpu_baseline": cb, "verified": verified,
        "route": {k: st_local.get(k) for k in ("hybrid", "unstable", "levels_run", "levels_skipped", "local_levels", "variant",
                                               "copy_back", "slow_segments", "kernel_launches")},
    }
    if e2e_pageable:
        line["e2e_pageable"] = e2e_pageable
    if world > 1:
        line["exchange"] = {"mode": args.exchange}
        line["exchange"].update({k: last.get(k) for k in ("level", "n_recv", "sent_bytes", "phase_ms", "splitter_mode", "splitter_plan_ms", "exchange_path", "presplit", "vary_hint")})
        line["exchange"]["max_n_recv_over_ranks"] = max_n_recv
        line["exchange"]["imbalance"] = max_n_recv / float(n)
        pm = last.get("phase_ms") or {}
        if pm.get("all_to_all"):
            t_x = pm["all_to_all"] + (pm["partition"] if args.exchange in ("peer", "ce") else 0.0)
            line["exchange"]["exchange_ms"] = t_x
            line["exchange"]["all_to_all_GBps_per_direction"] = last["sent_bytes"] / (t_x * 1e-3) / 1e9
            line["exchange"]["nvlink_peak_GBps"] = 900.0
            line["exchange"]["nvlink_measured_peer_copy_GBps"] = 770.0
    return line


def run_b200(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    cpu_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        cpu_group = dist.new_group(backend="gloo")
    names = [args.workload]
    if args.all_configs:
        names = ALL_CONFIGS_N1 if world == 1 else ALL_CONFIGS_NX
    for i, name in enumerate(names):
        if args.all_configs and world > 1 and i > 0:
            args.no_e2e = True  # the host-memory run (N x 8 GB pinned twice over) once per invocation is enough
        line = run_one(args, name, torch, dist, world, rank, local_rank, cpu_group)
        if line is not None:
            print(json.dumps(line), flush=True)
        from rdst_b200 import _abi

        _abi.lib().b200_release_scratch(-1)
        torch.cuda.empty_cache()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="u64_1b_uniform", choices=sorted(WORKLOADS))
    ap.add_argument("--n", "--records", dest="n", type=int, default=0, help="override records per GPU (debug; use --records under torchrun, whose own parser claims --n)")
    ap.add_argument("--cpu-sample", type=int, default=CPU_SAMPLE)
    ap.add_argument("--cpu-steps", type=int, default=3)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--recv-slack", type=float, default=0.3)
    ap.add_argument("--exchange", default="peer", choices=["peer", "ce", "nccl"],
                    help="N>1: fused partition + peer stores over NVLink (CUDA IPC); ce: chunked partition pipelined with "
                         "copy-engine pushes into the same IPC buffers; nccl: partition + NCCL all-to-all")
    ap.add_argument("--ce-chunks", type=int, default=8)
    ap.add_argument("--all-configs", action="store_true",
                    help="one JSON line per BASELINE.json config that fits this GPU count (N=1: configs 0-3; N>1: config 4 uniform and skewed)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = max(args.warmup, 3) if not args.n else args.warmup
    return run_reference(args) if args.impl == "reference" else run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
