#!/usr/bin/env python3
"""Times the range-partition building blocks of the sharded sort on one GPU (CUDA events):
digit-remap partition, splitter count pass, splitter-mode partition.
  python tools/partition_bench.py [--n 200000000]"""
import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from rdst_b200 import RadixKey, _abi
from rdst_b200.dist import CudaEngine, choose_splitters, part_of_prefix

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=200_000_000)
ap.add_argument("--only", default="")
args = ap.parse_args()
n = args.n
key = RadixKey.for_dtype("u8")
eng = CudaEngine(key)
g = torch.Generator(device="cuda").manual_seed(3)


def timed(fn, reps=3):
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        fn()
        b.record()
        b.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


for name in ("uniform", "zipf"):
    if args.only and name != args.only:
        continue
    if name == "uniform":
        src = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)
    else:
        u = torch.rand(n, dtype=torch.float64, device="cuda", generator=g)
        s, R = 1.1, 2.0**32
        src = ((1.0 - u * (1.0 - R ** (1.0 - s))) ** (1.0 / (1.0 - s))).clamp_(1, R).to(torch.int64)
    buf = src.view(torch.uint8).view(-1)
    dst = torch.empty_like(buf)
    hist = eng.histogram(buf, n)
    t_hist = timed(lambda: eng.histogram(buf, n))
    smp = np.sort(eng.prefix_sample(buf, n, n // 8192, 8192))
    for parts in (2, 8):
        level = max(l for l in range(8) if hist[l].max() != n)
        split = choose_splitters(hist[level], parts)
        t_digit = timed(lambda: eng.partition(buf, dst, n, level, hist[level], split, parts))
        sp = smp[[len(smp) * q // parts for q in range(1, parts)]].astype(np.uint64)
        less, equal = eng.splitter_count(buf, n, sp)
        t_count = timed(lambda: eng.splitter_count(buf, n, sp))
        left = less  # incl_mask all ones: equal keys go right
        pc = np.diff(np.concatenate([[0], left, [n]])).astype(np.uint64)
        t_split = timed(lambda: eng.partition_splitters(buf, dst, None, n, sp, (1 << (parts - 1)) - 1, pc))
        gb = 2 * n * 8 / 1e6
        print(f"{name:8s} parts={parts}: K1 {t_hist:6.2f} ms | digit partition (level {level}) {t_digit:6.2f} ms ({gb / t_digit:6.0f} GB/s)"
              f" | splitter count {t_count:6.2f} ms ({n * 8 / 1e6 / t_count:6.0f} GB/s) | splitter partition {t_split:6.2f} ms"
              f" ({gb / t_split:6.0f} GB/s)  max part {int(pc.max()) / n:.3f}", flush=True)
