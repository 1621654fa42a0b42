//! # rdst-b200
//!
//! The sort path of [rdst](https://github.com/nessex/rdst) — `RadixKey`, `RadixSort`,
//! `RadixSortBuilder`, `tuner::{Tuner, Algorithm, TuningParams}` — running on one or several B200 GPUs
//! through the C ABI of `libb200sort` (include/b200sort.h).  Same names, same argument meaning, same
//! error behaviour as the reference (paths below are into the rdst checkout):
//!
//! ```ignore
//! use rdst_b200::RadixSort;
//! let mut v: Vec<u64> = load();
//! v.radix_sort_unstable();                                   // src/radix_sort.rs:21-27
//! v.radix_sort_builder().with_low_mem_tuner().sort();         // src/radix_sort_builder.rs:86-90
//! ```
//!
//! What changes for a user type: arbitrary `get_level` code cannot run on the device, so a key type
//! also implements [`GpuRadixKey`] — a static description of which record byte each level reads
//! ([`KeyDesc`]).  Every type rdst implements `RadixKey` for has both impls here.  In debug builds
//! (or with the `validate` feature) the descriptor is checked against `get_level` on sampled records
//! before the first launch.
//!
//! There is no CPU fallback: an engine error panics, as the reference panics on a broken invariant
//! (src/radix_sort_builder.rs:22).
//!
//! NOTE: this crate is shipped as source only; the engine's build image has no Rust toolchain.

pub mod device;
mod gpu_key;
pub mod multi;
mod radix_key;
mod radix_sort;
mod radix_sort_builder;
pub mod tuner;
pub mod tuners;
mod validate;

pub use gpu_key::{GpuRadixKey, KeyDesc, KeyKind};
pub use radix_key::RadixKey;
pub use radix_sort::RadixSort;
pub use radix_sort_builder::{RadixSortBuilder, SortStats};

/// Engine error → panic message (the reference API is infallible).
pub(crate) fn engine_panic(code: i32) -> ! {
    use std::ffi::CStr;
    let (what, cuda) = unsafe {
        (
            CStr::from_ptr(rdst_b200_sys::b200_strerror(code)).to_string_lossy().into_owned(),
            CStr::from_ptr(rdst_b200_sys::b200_last_cuda_error()).to_string_lossy().into_owned(),
        )
    };
    if cuda.is_empty() {
        panic!("b200sort: {what} (code {code})");
    }
    panic!("b200sort: {what} (code {code}): {cuda}");
}
