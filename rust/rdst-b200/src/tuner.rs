//! The policy hook — rdst src/tuner.rs:1-40, same items.
//!
//! On the CPU a `Tuner` is asked per chunk and recursion level (src/sorter.rs:67-76).  The device
//! engine does not recurse, so a `Tuner` is asked ONCE per sort, with `TuningParams` of the top level
//! and that level's 256 counts (from the engine's all-level histogram kernel), and its answer selects
//! the engine's route ([`Algorithm::route_hint`]); `SortStats` reports what ran.

use rdst_b200_sys as sys;

#[derive(Clone)]
pub struct TuningParams {
    pub threads: usize,
    pub level: usize,
    pub total_levels: usize,
    pub input_len: usize,
    pub parent_len: Option<usize>,
}

#[derive(Debug, Clone, Copy, Eq, PartialEq)]
pub enum Algorithm {
    MtOop,
    MtLsb,
    Scanning,
    Recombinating,
    Comparative,
    LrLsb,
    Lsb,
    Regions,
    Ska,
}

impl Algorithm {
    /// What the answer means to the engine (include/b200sort.h `b200_route_hint`):
    /// the stable LSB family forces LSD passes; the in-place family (what `LowMemoryTuner` prefers)
    /// the low-scratch route; the out-of-place MSB-first algorithms leave the choice to the planner.
    pub fn route_hint(self) -> u32 {
        match self {
            Algorithm::Lsb | Algorithm::LrLsb | Algorithm::MtLsb => sys::B200_HINT_LSD,
            Algorithm::Ska | Algorithm::Regions => sys::B200_HINT_LOW_SCRATCH,
            Algorithm::MtOop | Algorithm::Scanning | Algorithm::Recombinating | Algorithm::Comparative => sys::B200_HINT_AUTO,
        }
    }

    /// rdst's stable algorithms (src/sorts/lsb_sort.rs:5-7, src/sorts/mt_lsb_sort.rs:12-15).
    pub fn is_stable(self) -> bool {
        matches!(self, Algorithm::Lsb | Algorithm::LrLsb | Algorithm::MtLsb)
    }
}

pub trait Tuner {
    fn pick_algorithm(&self, p: &TuningParams, counts: &[usize]) -> Algorithm;
}

pub(crate) type TunerRef<'tuner> = &'tuner (dyn Tuner + Send + Sync);
