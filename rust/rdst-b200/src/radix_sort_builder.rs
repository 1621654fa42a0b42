//! `RadixSortBuilder` — rdst src/radix_sort_builder.rs:8-158, same five methods; `sort()` hands the
//! slice to the engine (`b200_sort_host`) instead of `Sorter::route_top_level` (src/sorter.rs:105-119).

use crate::gpu_key::GpuRadixKey;
use crate::tuner::{Tuner, TunerRef, TuningParams};
use crate::tuners::{LowMemoryTuner, SingleThreadedTuner, StandardTuner};
use rdst_b200_sys as sys;
use std::os::raw::c_void;

/// What the engine reports about one sort (`b200_stats`).
pub type SortStats = sys::b200_stats;

enum Policy<'a> {
    Standard,
    LowMemory,
    SingleThreaded,
    Custom(TunerRef<'a>),
}

pub struct RadixSortBuilder<'data, T> {
    data: &'data mut [T],
    multi_threaded: bool,
    policy: Policy<'data>,
    stable: bool,
    device: i32,
    stats: Option<&'data mut SortStats>,
}

impl<'a, T> RadixSortBuilder<'a, T>
where
    T: GpuRadixKey + Send + Sync,
{
    pub(crate) fn new(data: &'a mut [T]) -> Self {
        // the invariant the reference asserts at the same place (src/radix_sort_builder.rs:22)
        assert_ne!(T::LEVELS, 0, "RadixKey must have at least 1 level");
        Self { data, multi_threaded: true, policy: Policy::Standard, stable: false, device: -1, stats: None }
    }

    /// Kept for source compatibility (src/radix_sort_builder.rs:53-57): on the device every CTA runs
    /// in parallel whatever this says.
    pub fn with_parallel(mut self, parallel: bool) -> Self {
        self.multi_threaded = parallel;
        self
    }

    /// `LowMemoryTuner` answers `Ska` / `Regions` for anything large (src/tuners/low_memory_tuner.rs:13-43):
    /// the engine's low-scratch route — LSD passes with the small-tile variant, scratch pool released
    /// when the call returns.  Peak scratch is still one copy of the array.
    pub fn with_low_mem_tuner(mut self) -> Self {
        self.policy = Policy::LowMemory;
        self
    }

    /// Accepted, no effect on the device (src/radix_sort_builder.rs:106-110).
    pub fn with_single_threaded_tuner(mut self) -> Self {
        self.policy = Policy::SingleThreaded;
        self
    }

    /// A user policy (src/radix_sort_builder.rs:128-132): asked once, for the top level, with the
    /// counts of that level; see [`crate::tuner`] for what each `Algorithm` selects.
    pub fn with_tuner(mut self, tuner: &'a (dyn Tuner + Send + Sync)) -> Self {
        self.policy = Policy::Custom(tuner);
        self
    }

    // ---- additions of this crate ----

    /// rdst's stable order (what a tuner answering `Lsb` / `LrLsb` / `MtLsb` yields): records with equal
    /// keys keep their input order.
    pub fn with_stable(mut self, stable: bool) -> Self {
        self.stable = stable;
        self
    }

    /// CUDA device ordinal (default: the current device).
    pub fn with_device(mut self, device: i32) -> Self {
        self.device = device;
        self
    }

    /// Receive the engine's report (route taken, levels run / skipped, phase times).
    pub fn with_stats(mut self, stats: &'a mut SortStats) -> Self {
        self.stats = Some(stats);
        self
    }

    pub fn sort(self) {
        // "By definition, this is already sorted" (src/radix_sort_builder.rs:151)
        if self.data.len() <= 1 {
            return;
        }
        let desc = T::descriptor();
        crate::validate::check_descriptor(&desc, self.data);
        let mut opts = sys::b200_sort_opts { stable: self.stable as u32, device: self.device, ..Default::default() };
        let mut stats = SortStats::default();
        let n = self.data.len();
        let ptr = self.data.as_mut_ptr() as *mut c_void;
        let rc = match self.policy {
            Policy::Standard | Policy::SingleThreaded => unsafe { sys::b200_sort_host(ptr, n, &desc.0, &opts, &mut stats) },
            Policy::LowMemory => {
                // the route LowMemoryTuner's answers (Ska / Regions) map to; asking it would only cost a histogram
                let _ = (&LowMemoryTuner, &StandardTuner, &SingleThreadedTuner);
                opts.route = sys::B200_HINT_LOW_SCRATCH;
                unsafe { sys::b200_sort_host(ptr, n, &desc.0, &opts, &mut stats) }
            }
            Policy::Custom(tuner) => unsafe { sort_with_tuner::<T>(ptr, n, &desc.0, &mut opts, &mut stats, tuner) },
        };
        if rc != 0 {
            crate::engine_panic(rc);
        }
        if let Some(out) = self.stats {
            *out = stats;
        }
    }
}

/// A user `Tuner` sees `TuningParams` of the top level and that level's counts (src/sorter.rs:67-76):
/// the array goes to the device once, the engine's histogram kernel (K1) supplies the counts, the
/// answer becomes the route hint, and the device-resident sort runs on the same copy.
unsafe fn sort_with_tuner<T: GpuRadixKey>(ptr: *mut c_void, n: usize, desc: &sys::b200_key_desc, opts: &mut sys::b200_sort_opts,
                                          stats: &mut SortStats, tuner: TunerRef<'_>) -> i32 {
    let bytes = n * std::mem::size_of::<T>();
    let mut dev: *mut c_void = std::ptr::null_mut();
    let mut handle = [0u8; sys::B200_IPC_HANDLE_BYTES];
    let mut rc = sys::b200_ipc_alloc(&mut dev, bytes, handle.as_mut_ptr() as *mut c_void);
    if rc != 0 {
        return rc;
    }
    rc = sys::b200_copy_h2d(dev, ptr, bytes, opts.device);
    if rc == 0 {
        let levels = T::LEVELS;
        let mut hist = vec![0u64; levels * 256];
        rc = sys::b200_histogram_device(dev, n, desc, opts, hist.as_mut_ptr());
        if rc == 0 {
            let counts: Vec<usize> = hist[(levels - 1) * 256..].iter().map(|&c| c as usize).collect();
            let p = TuningParams { threads: 148, level: levels - 1, total_levels: levels, input_len: n, parent_len: None };
            let algo = tuner.pick_algorithm(&p, &counts);
            opts.route = algo.route_hint();
            if algo.is_stable() {
                opts.stable = 1;
            }
            rc = sys::b200_sort_device(dev, std::ptr::null_mut(), n, desc, opts, stats);
        }
        if rc == 0 {
            rc = sys::b200_copy_d2h(ptr, dev, bytes, opts.device);
        }
    }
    sys::b200_ipc_free(dev);
    rc
}
