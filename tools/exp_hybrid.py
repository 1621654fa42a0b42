#!/usr/bin/env python3
"""Route experiments on one GPU: for each workload and each environment setting, the best of a few
device-resident sorts (library CUDA events) with the per-phase line of B200SORT_DEBUG_TIMING, and a
sortedness + checksum check of the result.

    python tools/exp_hybrid.py --work u8:1000000000:uniform,u4:100000000:uniform --cfg "base;B200SORT_UNSTABLE=0;B200SORT_MSDU_VARIANT=1"
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("B200SORT_DEBUG_TIMING", "1")
import numpy as np
import torch

import bench
from rdst_b200 import RadixKey, _abi, radix_sort_builder

ap = argparse.ArgumentParser()
ap.add_argument("--work", default="u8:1000000000:uniform")
ap.add_argument("--cfg", default="base")
ap.add_argument("--reps", type=int, default=3)
args = ap.parse_args()
lib = _abi.lib()
for w in args.work.split(","):
    dt, n, gen = w.split(":")
    n = int(n)
    key = RadixKey.for_dtype(dt)
    src = bench.gen_device(torch, dt, n, gen, 0x5EED0002)
    work = torch.empty_like(src)
    chk = int(src.sum())
    for cfg in args.cfg.split(";"):
        sets = [kv.split("=") for kv in cfg.split("+") if "=" in kv]
        for k, v in sets:
            os.environ[k] = v
        best = None
        for r in range(args.reps):
            work.copy_(src)
            torch.cuda.synchronize()
            sys.stderr.flush()
            b = radix_sort_builder(work.view(torch.uint8), key)
            b.sort()
            torch.cuda.synchronize()
            if best is None or b.stats["ms_total"] < best["ms_total"]:
                best = b.stats
        flip = -2**63 if key.elem_size == 8 else -2**31
        if np.dtype(dt).kind == "u":
            u = work ^ flip
        elif np.dtype(dt).kind == "i":
            u = work
        else:  # float total order on the bit patterns
            u = torch.where(work < 0, ~work, work ^ flip) ^ flip
        ok = bool((u[1:] >= u[:-1]).all()) and int(work.sum()) == chk
        del u
        print(f"RESULT {w:28s} cfg={cfg:40s} total {best['ms_total']:8.3f} ms  {n / best['ms_total'] / 1e6:7.2f} Gkeys/s  "
              f"hist {best['ms_hist']:.3f} passes {best['ms_passes']:.3f} joint {best['ms_joint']:.3f} local {best['ms_local']:.3f} "
              f"hybrid={best['hybrid']} unstable={best['unstable']} slow={best['slow_segments']} launches={best['kernel_launches']} ok={ok}",
              flush=True)
        for k, v in sets:
            del os.environ[k]
    del src, work
    lib.b200_release_scratch(-1)
    torch.cuda.empty_cache()
