"""Throughput of b200_sort_segments_device (many independent sorts in one call) against segment size.
usage: python tools/segments_bench.py [records, default 64 Mi]   (one GPU; prints one JSON line per case)"""
import json
import sys

import numpy as np
import torch

from rdst_b200 import RadixKey, radix_sort_builder, sort_segments

REC8 = np.dtype([("idx", "<u4"), ("key", "<i4")])


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 64 << 20
    g = torch.Generator(device="cuda").manual_seed(1)
    for name, key, words in (("u64", RadixKey.for_dtype("u8"), 2), ("u32", RadixKey.for_dtype("u4"), 1),
                             ("rec8_stable", RadixKey.field(REC8, "key", 4), 2)):
        src = torch.randint(-2**31, 2**31 - 1, (n * words,), dtype=torch.int32, device="cuda", generator=g)
        work = torch.empty_like(src)
        for seg in (16, 64, 128, 256, 1024, 4096, 16384):
            off = np.arange(0, n + 1, seg, dtype=np.uint64)
            if off[-1] != n:
                off = np.append(off, np.uint64(n))
            best = None
            for _ in range(4):
                work.copy_(src)
                torch.cuda.synchronize()
                st = sort_segments(work.view(torch.uint8), off, key)
                best = st if best is None or st["ms_total"] < best["ms_total"] else best
            v = work.view(torch.int64) if name == "u64" else None
            print(json.dumps({"case": name, "records": n, "segment": seg, "segments": len(off) - 1, "ms": round(best["ms_total"], 3),
                              "grec_per_s": round(n / best["ms_total"] / 1e6, 2), "gb_per_s_moved": round(2 * n * 4 * words / best["ms_total"] / 1e6, 1),
                              "slow_segments": best["slow_segments"], "unstable": best["unstable"]}), flush=True)
        # the whole array as one sort, for scale
        ms = None
        for _ in range(3):  # (the first call loads the kernels)
            work.copy_(src)
            torch.cuda.synchronize()
            b = radix_sort_builder(work.view(torch.uint8), key)
            b.sort()
            ms = b.stats["ms_total"] if ms is None else min(ms, b.stats["ms_total"])
        print(json.dumps({"case": name, "records": n, "segment": "whole array", "ms": round(ms, 3),
                          "grec_per_s": round(n / ms / 1e6, 2)}), flush=True)


if __name__ == "__main__":
    main()
