#!/usr/bin/env python3
"""Variant sweep on one GPU: per-variant K1 / K2 timings from the library's own CUDA events.
  python tools/sweep.py [--n 200000000] [--dtypes u4,u8] [--reps 3]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from rdst_b200 import RadixKey, _abi, radix_sort_builder

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=200_000_000)
ap.add_argument("--dtypes", default="u4,u8,rec12,rec16")
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--variants", default="")
args = ap.parse_args()
lib = _abi.lib()
nv = lib.b200_variant_count()
variants = [int(v) for v in args.variants.split(",")] if args.variants else list(range(nv))
peak = 6424.4
for name in args.dtypes.split(","):
    if name.startswith("rec"):
        es = int(name[3:])
        key = RadixKey("uint", es, 4, 0, 4)
    else:
        key = RadixKey.for_dtype(name)
        es = key.elem_size
    n = args.n
    g = torch.Generator(device="cuda").manual_seed(1)
    src = torch.randint(-2**31, 2**31, (n * es // 4,), dtype=torch.int32, device="cuda", generator=g)
    work = torch.empty_like(src)
    for v in variants:
        best = None
        for r in range(args.reps):
            work.copy_(src)
            b = radix_sort_builder(work.view(torch.uint8), key).with_variant(v)
            b.sort()
            torch.cuda.synchronize()
            s = b.stats
            if best is None or s["ms_passes"] < best["ms_passes"]:
                best = s
        p = best["levels_run"]
        per = best["ms_passes"] / max(p, 1)
        gbs = 2 * n * es / (per * 1e-3) / 1e9
        print(f"{name:6s} n={n} v{v} {lib.b200_variant_name(v).decode():18s} hist {best['ms_hist']:7.3f} ms ({n*es/best['ms_hist']/1e6:7.1f} GB/s)  "
              f"pass {per:7.3f} ms x{p} ({gbs:7.1f} GB/s, {gbs/peak:5.1%})  total {best['ms_hist']+best['ms_passes']:8.3f} ms  "
              f"{n/(best['ms_hist']+best['ms_passes'])/1e6:7.2f} Gkeys/s", flush=True)
    del src, work
    lib.b200_release_scratch(-1)
