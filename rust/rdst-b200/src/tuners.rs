//! The reference's three tuners (rdst src/tuners/{standard,low_memory,single_threaded}_tuner.rs),
//! restated as threshold tables so that a given `(input_len, level, counts)` yields the Algorithm
//! the reference would report.  (The same tables exist in rdst_b200/tuner.py and, in C, in the test
//! oracle; tests/test_abi_cpu.py cross-checks those two.)

use crate::tuner::{Algorithm, Tuner, TuningParams};

/// "Skewed": some bucket holds at least twice the even share, for inputs of 5000 records and more.
fn skewed(p: &TuningParams, counts: &[usize]) -> bool {
    if p.input_len < 5000 {
        return false;
    }
    let limit = (p.input_len / 256) * 2;
    counts.iter().any(|&c| c >= limit)
}

/// First entry whose bound is `>= len`, else the fallback.
fn by_len(len: usize, table: &[(usize, Algorithm)], otherwise: Algorithm) -> Algorithm {
    table.iter().find(|&&(bound, _)| len <= bound).map(|&(_, a)| a).unwrap_or(otherwise)
}

pub struct StandardTuner;

impl Tuner for StandardTuner {
    fn pick_algorithm(&self, p: &TuningParams, counts: &[usize]) -> Algorithm {
        use Algorithm::*;
        if p.input_len <= 128 {
            return Comparative;
        }
        let depth = p.total_levels - p.level - 1;
        match (skewed(p, counts), depth) {
            (true, 0) => by_len(p.input_len, &[(200_000, LrLsb), (350_000, Ska), (4_000_000, MtLsb)], Regions),
            (true, _) => by_len(p.input_len, &[(200_000, LrLsb), (800_000, Ska), (5_000_000, Recombinating)], Regions),
            (false, 0) => by_len(p.input_len, &[(150_000, Lsb), (260_000, Ska), (50_000_000, Recombinating)], Scanning),
            (false, _) => by_len(p.input_len, &[(200_000, Lsb), (800_000, Ska), (50_000_000, Recombinating)], Scanning),
        }
    }
}

pub struct LowMemoryTuner;

impl Tuner for LowMemoryTuner {
    fn pick_algorithm(&self, p: &TuningParams, counts: &[usize]) -> Algorithm {
        use Algorithm::*;
        if p.input_len <= 128 {
            return Comparative;
        }
        let small = if skewed(p, counts) { LrLsb } else { Lsb };
        by_len(p.input_len, &[(50_000, small), (1_000_000, Ska)], Regions)
    }
}

pub struct SingleThreadedTuner;

impl Tuner for SingleThreadedTuner {
    fn pick_algorithm(&self, p: &TuningParams, counts: &[usize]) -> Algorithm {
        use Algorithm::*;
        if p.input_len <= 128 {
            return Comparative;
        }
        let depth = p.total_levels - p.level - 1;
        if skewed(p, counts) {
            if p.input_len > 100_000 && depth < 2 { Ska } else { LrLsb }
        } else if p.input_len > 800_000 && depth == 0 {
            Ska
        } else {
            Lsb
        }
    }
}
