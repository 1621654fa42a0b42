#!/usr/bin/env python3
"""Host entry point timing on one GPU: b200_sort_host on pinned and pageable memory, per copy-thread count.
    python tools/host_path.py [--n 1000000000] [--threads 4,8,16]"""
import argparse, os, subprocess, sys, time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1_000_000_000)
ap.add_argument("--threads", default="")
ap.add_argument("--child", action="store_true")
args = ap.parse_args()
if not args.child and args.threads:
    for t in args.threads.split(","):
        subprocess.run([sys.executable, __file__, "--n", str(args.n), "--child"], env=dict(os.environ, B200SORT_COPY_THREADS=t,
                                                                                          B200SORT_DEBUG_TIMING="1"))
    sys.exit(0)
import numpy as np
import torch
from rdst_b200 import RadixKey, radix_sort_builder

key = RadixKey.for_dtype("u8")
g = torch.Generator(device="cuda").manual_seed(7)
src = torch.randint(-2**63, 2**63 - 1, (args.n,), dtype=torch.int64, device="cuda", generator=g)
print(f"host threads {os.cpu_count()}, B200SORT_COPY_THREADS={os.environ.get('B200SORT_COPY_THREADS')}", flush=True)
for pinned in (True, False):
    hs = torch.empty(args.n, dtype=torch.int64)
    hw = torch.empty(args.n, dtype=torch.int64)
    if pinned:
        hs, hw = hs.pin_memory(), hw.pin_memory()
    hs.copy_(src)
    arr = hw.numpy().view("u8")
    for i in range(3):
        hw.copy_(hs)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        b = radix_sort_builder(arr, key)
        b.sort()
        dt = time.perf_counter() - t0
        print(f"  {'pinned' if pinned else 'pageable'} run {i}: {dt*1e3:.1f} ms  {args.n/dt/1e9:.2f} Gkeys/s", flush=True)
    assert bool(np.all(arr[1:] >= arr[:-1]))
    del hs, hw
