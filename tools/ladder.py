#!/usr/bin/env python
"""Latency / throughput ladder of the reference's benches on one B200: the exponential input set
(reference benches/bench_utils.rs:79-100: 200 M records halved down to 1) for u32 and u64 keys
with the generator of benches/basic_sort.rs (uniform random words).  Per size: the median wall
time of one b200_sort_device call on device-resident data (host overhead and the final
synchronisation included — this is what a caller waits for) and the library's own CUDA-event
time.  Also times presorted input (K0's early-out).  Writes one JSON object per line.

    python tools/ladder.py [--max 200000000] [--out gpurun_out/ladder.jsonl]
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from rdst_b200 import RadixKey, _abi  # noqa: E402


def main():
    import torch

    ap = argparse.ArgumentParser()
    ap.add_argument("--max", type=int, default=200_000_000)
    ap.add_argument("--out", default="gpurun_out/ladder.jsonl")
    ap.add_argument("--dtypes", default="u4,u8")
    args = ap.parse_args()
    assert torch.cuda.is_available()
    lib = _abi.lib()
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    out = open(args.out, "w")
    for dt in args.dtypes.split(","):
        item = np.dtype(dt).itemsize
        key = RadixKey.for_dtype(dt)
        desc = key.to_desc()
        opts = _abi.SortOpts()
        opts.stable, opts.variant, opts.device, opts.stream = 0, -1, 0, None
        stats = _abi.Stats()
        g = torch.Generator(device="cuda")
        g.manual_seed(0x5EED0001)
        full = torch.randint(-(1 << 63), (1 << 63) - 1, (args.max * item // 8 + 1,), dtype=torch.int64, device="cuda",
                             generator=g).view(torch.uint8)
        work = torch.empty_like(full)
        n = args.max
        while n >= 1:
            start = ((args.max - n) // 2) * item  # the middle slice, as the reference takes it
            src = full[start:start + n * item]
            dst = work[: n * item]
            reps = 5 if n >= 10_000_000 else (11 if n >= 100_000 else 31)
            for mode in ("random", "presorted"):
                if mode == "presorted":
                    if n < 2:
                        continue
                    dst.copy_(src)
                    _abi.check(lib.b200_sort_device(dst.data_ptr(), None, n, C.byref(desc), C.byref(opts), C.byref(stats)))
                    keep = dst.clone()
                wall, dev = [], []
                for r in range(reps + 2):
                    dst.copy_(keep if mode == "presorted" else src)
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    _abi.check(lib.b200_sort_device(dst.data_ptr(), None, n, C.byref(desc), C.byref(opts), C.byref(stats)))
                    t1 = time.perf_counter()  # the call returns after its own stream synchronisation
                    if r >= 2:
                        wall.append((t1 - t0) * 1e6)
                        dev.append(stats.ms_total * 1e3)
                w = float(np.median(wall))
                line = {"dtype": dt, "n": n, "input": mode, "wall_us": round(w, 2), "device_us": round(float(np.median(dev)), 2),
                        "mkeys_per_s": round(n / w, 2), "launches": int(stats.kernel_launches),
                        "levels_run": int(stats.levels_run), "hybrid": int(stats.hybrid),
                        "already_sorted": int(stats.already_sorted), "reps": reps}
                out.write(json.dumps(line) + "\n")
                out.flush()
                print(json.dumps(line))
            n //= 2
        del full, work
        torch.cuda.empty_cache()
        _abi.check(lib.b200_release_scratch(0))


if __name__ == "__main__":
    main()
