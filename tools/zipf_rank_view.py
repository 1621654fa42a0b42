"""What every rank of an 8-GPU sharded sort receives for the hashed-Zipf(1.1) workload (bench.gen_device), computed on the CPU from the
distribution itself: expected per-key counts of the 4 M most frequent keys, digit splitters as rdst_b200/dist.py chooses them, and the
quantities the local sort's planner looks at (rdst_b200/csrc/b200sort.cu sort_hybrid: m1, m2, may_split, oversized segments).
No GPU needed.  Output: profiles/r17_zipf_hashed_rank_view.txt."""
import numpy as np
# Expected per-key counts of the hashed Zipf(1.1) workload at N=8 (8e9 draws), bench.gen_device's inverse-CDF
s, R, NTOT = 1.1, 2.0**32, 8e9
# P(rank <= r) from the inverse CDF: r = (1 - u*(1 - R^(1-s)))^(1/(1-s))  =>  u(r) = (1 - r^(1-s)) / (1 - R^(1-s));  int cast floors r
K = 4_000_000
r = np.arange(1, K + 2, dtype=np.float64)
U = (1 - r ** (1 - s)) / (1 - R ** (1 - s))
p = np.diff(U)                      # P(floor(r) == k) for k = 1..K
cnt = p * NTOT
ranks = np.arange(1, K + 1, dtype=np.int64)
with np.errstate(over='ignore'):
    h = (ranks * np.int64(-7046029254386353131)) ^ (ranks >> 7)
key = h.view(np.uint64)
print("top counts", cnt[:5].astype(np.int64), "sum heavy", cnt.sum() / NTOT)
tail = NTOT - cnt.sum()             # keys beyond K: treat as uniform
lvl7 = (key >> np.uint64(56)).astype(np.int64)
hist7 = np.bincount(lvl7, weights=cnt, minlength=256) + tail / 256
# digit splitters as dist.choose_splitters: boundary g where the running count is closest to g*N/P
cum = np.concatenate([[0], np.cumsum(hist7)])
P = 8
split = []
lo = 0
for g in range(1, P):
    target = NTOT * g / P
    d = np.arange(lo, 257)
    best = d[np.argmin(np.abs(cum[lo:] - target))]
    split.append(int(best)); lo = int(best)
print("split", split)
bounds = [0] + split + [256]
for g in range(P):
    a, b = bounds[g], bounds[g + 1]
    n_g = hist7[a:b].sum()
    m = (lvl7 >= a) & (lvl7 < b)
    kg, cg = key[m], cnt[m]
    tail_g = tail * (b - a) / 256
    # the rank's range: digits a..b-1 of level 7 -> varying bits below msb(lo ^ hi)
    lo_key, hi_key = a << 56, ((b - 1) << 56) | ((1 << 56) - 1)
    top = (lo_key ^ hi_key).bit_length() - 1          # highest varying bit
    w1_lo = top - 7
    w2_lo = w1_lo - 8
    d1 = ((kg >> np.uint64(w1_lo)) & np.uint64(255)).astype(np.int64)
    d2 = ((kg >> np.uint64(w2_lo)) & np.uint64(255)).astype(np.int64)
    # how many (D1) values does the range cover?  uniform tail spreads over the covered D1 values only
    span = ((hi_key >> w1_lo) & 255) - ((lo_key >> w1_lo) & 255) + 1 if (hi_key >> (w1_lo + 8)) == (lo_key >> (w1_lo + 8)) else 256
    h1 = np.bincount(d1, weights=cg, minlength=256)
    h2 = np.bincount(d2, weights=cg, minlength=256) + tail_g / 256
    covered = np.zeros(256); 
    lo_d, hi_d = (lo_key >> w1_lo) & 255, (hi_key >> w1_lo) & 255
    if span == 256: covered[:] = 1
    else: covered[lo_d:hi_d + 1] = 1
    h1 = h1 + tail_g * covered / covered.sum()
    joint = np.zeros((256, 256)); np.add.at(joint, (d1, d2), cg)
    joint += (tail_g / covered.sum() / 256) * covered[:, None]
    m1, m2 = h1.max(), h2.max()
    big = joint > 16384
    print(f"rank {g}: digits {a}..{b-1} n={n_g/1e9:.3f}B top bit {top} D1 bits {w1_lo}..{w1_lo+7} covered {int(covered.sum())} m1/n={m1/n_g:.3f} m2/n={m2/n_g:.3f} "
          f"flat1={m1<=n_g/8} flat2={m2<=n_g/8} may_split={m1<=n_g/4 and m2<=n_g/4} est={m1*m2/n_g:.0f} big segs={int(big.sum())} big_total/n={joint[big].sum()/n_g:.3f} "
          f"max seg={joint.max()/1e6:.1f}M heavy keys>=16384: {int((cg>=16384).sum())}")
