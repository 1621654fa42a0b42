"""CPU-only checks of the exchange planner behind b200_sort_multi_* and rdst_b200.dist (csrc/multi.cu:
b200_multi_plan_digits / b200_multi_choose_sample_splitters / b200_multi_plan_splitters) against the numpy
model of the same rules in rdst_b200/dist.py, which tests/test_dist_gloo.py exercises end to end with a CPU test
double.  The reference has no distributed path; the rules are SURVEY.md §8e steps 2-3."""
import numpy as np
import pytest

from rdst_b200 import dist


def hist_of(keys, levels):
    k = keys.astype(np.uint64)
    return np.stack([np.bincount(((k >> np.uint64(8 * l)) & np.uint64(0xFF)).astype(np.int64), minlength=256)
                     for l in range(levels)]).astype(np.uint64)


def shards(rng, kind, P, n):
    if kind == "uniform":
        return [rng.integers(0, 2**64, n, dtype=np.uint64) for _ in range(P)]
    if kind == "narrow":  # top levels trivial: the partition level is below them
        return [rng.integers(0, 2**37, n, dtype=np.uint64) for _ in range(P)]
    if kind == "zipf":
        return [np.minimum(rng.zipf(1.1, n), 2**32).astype(np.uint64) for _ in range(P)]
    if kind == "ragged":
        return [rng.integers(0, 2**64, max(0, n // (r + 1) - 3 * r), dtype=np.uint64) for r in range(P)]
    if kind == "two_values":
        return [rng.integers(0, 2, n, dtype=np.uint64) << np.uint64(60) for _ in range(P)]
    if kind == "all_equal":
        return [np.full(n, 77, dtype=np.uint64) for _ in range(P)]
    raise AssertionError(kind)


@pytest.mark.parametrize("P", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("kind", ["uniform", "narrow", "zipf", "ragged", "two_values", "all_equal"])
def test_digit_plan_matches_the_numpy_model(P, kind):
    rng = np.random.default_rng(hash((P, kind)) % 2**32)
    all_hist = np.stack([hist_of(s, 8) for s in shards(rng, kind, P, 20_000)])
    level, split, m, imb = dist.plan_digits(all_hist)
    g = all_hist.sum(axis=0)
    assert level == dist.pick_partition_level(g)
    if level < 0:
        assert kind == "all_equal"
        return
    exp_split = dist.choose_splitters(g[level], P)
    assert split.tolist() == exp_split.tolist()
    exp_m = dist.send_matrix(all_hist[:, level, :].astype(np.int64), exp_split, P)
    assert np.array_equal(m, exp_m)
    assert m.sum(axis=1).tolist() == all_hist[:, 0, :].sum(axis=1).tolist()  # every record is sent exactly once
    assert imb == pytest.approx(dist.digit_split_imbalance(exp_m))


@pytest.mark.parametrize("P", [2, 4, 8, 16])
@pytest.mark.parametrize("kind", ["uniform", "zipf", "two_values", "ragged"])
def test_splitter_plan_matches_the_numpy_model(P, kind):
    rng = np.random.default_rng(hash((P, kind, 1)) % 2**32)
    sh = shards(rng, kind, P, 30_000)
    n_per_rank = np.array([len(s) for s in sh], dtype=np.int64)
    vals, wts = [], []
    for s in sh:
        c = min(len(s), 512)
        if c:
            vals.append(s[:: max(1, len(s) // c)][:c])
            wts.append(np.full(len(vals[-1]), len(s) / len(vals[-1])))
    samples, weights = np.concatenate(vals), np.concatenate(wts)
    exp_sp = dist.choose_sample_splitters(samples, weights, P)
    sp = dist.plan_sample_splitters(samples, weights, P)
    assert sp.tolist() == exp_sp.tolist()
    less = np.array([[int((s < x).sum()) for x in sp] for s in sh], dtype=np.int64)
    equal = np.array([[int((s == x).sum()) for x in sp] for s in sh], dtype=np.int64)
    for exact in (True, False):
        exp_t, exp_m = dist.plan_splitter_exchange(less, equal, n_per_rank, sp, P, exact_ties=exact)
        t, m, masks = dist.plan_splitters(less, equal, n_per_rank, sp, P, exact_ties=exact)
        assert t.tolist() == exp_t.tolist() and np.array_equal(m, exp_m)
        assert masks.tolist() == [dist.incl_mask_for_rank(exp_t, r) for r in range(len(sh))]
        assert m.sum(axis=1).tolist() == n_per_rank.tolist()
        # the plan is what the partition rule produces: count each rank's records per part with the numpy rule
        for r, s in enumerate(sh):
            part = dist.part_of_prefix(s, sp, int(masks[r]))
            assert np.bincount(part, minlength=P).tolist() == m[r].tolist()


def test_vary_hints_are_supersets_of_the_truth():
    """rdst_b200/dist.py tells every rank's local sort which key bits can vary inside its range (b200_sort_device_hinted
    skips the counting pass): the job's varying bits below the highest bit in which the range's bounds differ.  A missing
    bit would give a wrong order, so the hint must cover the true varying bits of ANY keys inside the range."""
    from rdst_b200.dist import global_vary_bits, range_vary_bits

    rng = np.random.default_rng(11)
    for bits in (64, 61, 40, 33, 9):
        a = rng.integers(0, 2**bits, 50_000, dtype=np.uint64)
        a |= np.uint64(0x5 << 20) if bits > 24 else np.uint64(0)          # a few constant-one bits
        hist = np.stack([np.bincount(((a >> np.uint64(8 * l)) & np.uint64(255)).astype(np.int64), minlength=256) for l in range(8)])
        gv = global_vary_bits(hist)
        true_vary = int(np.bitwise_or.reduce(a) & ~np.bitwise_and.reduce(a))
        assert gv == true_vary
        s = np.sort(a)
        for parts in (2, 3, 8):
            cuts = [0] + [len(s) * g // parts for g in range(1, parts)] + [len(s)]
            for g in range(parts):
                part = s[cuts[g]:cuts[g + 1]]
                lo, hi = int(part[0]), int(part[-1])
                hint = range_vary_bits(gv, lo, hi)
                part_vary = int(np.bitwise_or.reduce(part) & ~np.bitwise_and.reduce(part))
                assert part_vary & ~hint == 0, (bits, parts, g, hex(part_vary), hex(hint))
                assert hint & ~gv == 0
    assert range_vary_bits(0xFFFF, 7, 7) == 0


def test_presplit_layout_tiles_every_receive_buffer():
    """rdst_b200/dist.py `presplit`: the exchange stores every digit of the partition level as its own run in the owner's
    buffer.  Whatever the histograms, the runs of all senders must tile every receive buffer exactly — bucket by bucket in
    digit order, inside a bucket sender by sender — and the receiver's bucket sizes must be the global counts."""
    from rdst_b200.dist import choose_splitters, presplit_layout, send_matrix

    rng = np.random.default_rng(21)
    for parts in (2, 3, 4, 8):
        for trial in range(6):
            rows = rng.integers(0, 5000, (parts, 256))
            if trial % 2:
                rows[:, rng.integers(0, 256, 180)] = 0                # empty digits
            if trial == 4:
                rows[:, 17] += 400_000                                # a heavy digit
            split = choose_splitters(rows.sum(axis=0), parts)
            m = send_matrix(rows, split, parts)                      # m[r][g] = records r sends to g
            filled = [np.zeros(int(m[:, g].sum()), dtype=np.int64) for g in range(parts)]
            digit_at = [np.full(len(f), -1, dtype=np.int64) for f in filled]
            for r in range(parts):
                owner, offset, counts = presplit_layout(rows, split, parts, r)
                assert int(counts.sum()) == int(m[:, r].sum())
                assert np.array_equal(counts[owner == r], rows.sum(axis=0)[owner == r]) and not counts[owner != r].any()
                for d in range(256):
                    c = int(rows[r, d])
                    g, o_ = int(owner[d]), int(offset[d])
                    assert o_ + c <= len(filled[g])
                    filled[g][o_:o_ + c] += 1
                    digit_at[g][o_:o_ + c] = d
            for g in range(parts):
                assert (filled[g] == 1).all(), (parts, trial, g)      # no gap, no overlap
                assert (np.diff(digit_at[g]) >= 0).all()              # buckets in digit order
