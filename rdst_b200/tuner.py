"""tuner — the reference's Tuner API (reference src/tuner.rs:1-40), re-purposed.

In the reference a Tuner picks one of nine CPU algorithms per chunk and recursion level
(reference src/sorter.rs:67-76).  The device engine does not recurse, so a Tuner is asked ONCE
per sort — with TuningParams for the top level and that level's 256 counts, from the K1
histogram — and its answer selects the engine's route (ALGORITHM_ROUTE): the stable LSB family
forces LSD passes, the in-place family the low-scratch route, everything else leaves the choice
to the planner.  b200_stats (hybrid, levels_run, variant) shows what ran.  The built-in tuners
keep the reference's threshold tables, so the Algorithm reported for a given (len, counts) is
the reference's.
"""
from __future__ import annotations

import enum
from dataclasses import dataclass
from typing import Optional, Sequence


@dataclass
class TuningParams:  # reference src/tuner.rs:1-8
    threads: int
    level: int
    total_levels: int
    input_len: int
    parent_len: Optional[int] = None


class Algorithm(enum.Enum):  # reference src/tuner.rs:10-21 (multi-threaded feature set)
    MtOop = 0
    MtLsb = 1
    Scanning = 2
    Recombinating = 3
    Comparative = 4
    LrLsb = 5
    Lsb = 6
    Regions = 7
    Ska = 8


# What each Algorithm means to the device engine (include/b200sort.h b200_route_hint).  rdst's nine
# algorithms are CPU strategies; on the GPU they collapse to three observable choices:
#   the stable LSB family            -> LSD passes only, stable order   (b200_stats.hybrid == 0)
#   the in-place family (Ska/Regions, what LowMemoryTuner prefers, reference low_memory_tuner.rs:13-43)
#                                    -> LSD passes, small-tile variant, scratch pool released after the call
#   the out-of-place MSB-first ones  -> the planner's own choice (hybrid MSD route when the input qualifies)
HINT_AUTO, HINT_LSD, HINT_LOW_SCRATCH = 0, 1, 2
ALGORITHM_ROUTE = {
    Algorithm.MtOop: HINT_AUTO,
    Algorithm.Scanning: HINT_AUTO,
    Algorithm.Recombinating: HINT_AUTO,
    Algorithm.Comparative: HINT_AUTO,
    Algorithm.MtLsb: HINT_LSD,
    Algorithm.LrLsb: HINT_LSD,
    Algorithm.Lsb: HINT_LSD,
    Algorithm.Regions: HINT_LOW_SCRATCH,
    Algorithm.Ska: HINT_LOW_SCRATCH,
}

STABLE_ALGORITHMS = {Algorithm.Lsb, Algorithm.LrLsb, Algorithm.MtLsb}  # reference lsb_sort.rs:5-7, mt_lsb_sort.rs:12-15


class Tuner:  # reference src/tuner.rs:33-35
    def pick_algorithm(self, p: TuningParams, counts: Sequence[int]) -> Algorithm:
        raise NotImplementedError


def _skewed(p: TuningParams, counts) -> bool:
    """`any bucket >= 2 * (len / 256)` for len >= 5000 (reference standard_tuner.rs:20-24)."""
    if p.input_len < 5000:
        return False
    thr = (p.input_len // 256) * 2
    return any(c >= thr for c in counts)


class StandardTuner(Tuner):
    """Threshold table of reference src/tuners/standard_tuner.rs:13-63 (kept so that the
    Algorithm a user sees for a given (len, level, counts) is the reference's)."""

    def pick_algorithm(self, p, counts):
        if p.input_len <= 128:
            return Algorithm.Comparative
        depth = p.total_levels - p.level - 1
        n = p.input_len
        if _skewed(p, counts):
            if depth == 0:
                if n <= 200_000:
                    return Algorithm.LrLsb
                if n <= 350_000:
                    return Algorithm.Ska
                if n <= 4_000_000:
                    return Algorithm.MtLsb
                return Algorithm.Regions
            if n <= 200_000:
                return Algorithm.LrLsb
            if n <= 800_000:
                return Algorithm.Ska
            if n <= 5_000_000:
                return Algorithm.Recombinating
            return Algorithm.Regions
        if depth > 0:
            if n <= 200_000:
                return Algorithm.Lsb
            if n <= 800_000:
                return Algorithm.Ska
            if n <= 50_000_000:
                return Algorithm.Recombinating
            return Algorithm.Scanning
        if n <= 150_000:
            return Algorithm.Lsb
        if n <= 260_000:
            return Algorithm.Ska
        if n <= 50_000_000:
            return Algorithm.Recombinating
        return Algorithm.Scanning


class LowMemoryTuner(Tuner):
    """reference src/tuners/low_memory_tuner.rs:13-43"""

    def pick_algorithm(self, p, counts):
        n = p.input_len
        if n <= 128:
            return Algorithm.Comparative
        if _skewed(p, counts):
            if n <= 50_000:
                return Algorithm.LrLsb
            if n <= 1_000_000:
                return Algorithm.Ska
            return Algorithm.Regions
        if n <= 50_000:
            return Algorithm.Lsb
        if n <= 1_000_000:
            return Algorithm.Ska
        return Algorithm.Regions


class SingleThreadedTuner(Tuner):
    """reference src/tuners/single_threaded_tuner.rs:13-43"""

    def pick_algorithm(self, p, counts):
        n = p.input_len
        if n <= 128:
            return Algorithm.Comparative
        depth = p.total_levels - p.level - 1
        if _skewed(p, counts):
            return Algorithm.Ska if (n > 100_000 and depth < 2) else Algorithm.LrLsb
        return Algorithm.Ska if (n > 800_000 and depth == 0) else Algorithm.Lsb
