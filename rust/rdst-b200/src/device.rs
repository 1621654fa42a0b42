//! Entry points for callers whose records already live in DEVICE memory (raw device pointers; the caller owns the
//! allocation and orders the work with its own stream).  The reference has no counterpart — it sorts `&mut [T]` in
//! host RAM (src/radix_sort_builder.rs:149-157); these are the engine's additions behind the same C ABI:
//!
//! * [`sort_segments`] — many independent sorts in one call, the tail of rdst's MSB-first recursion
//!   (src/sorter.rs:121-157: every bucket of a split sorted as its own task);
//! * [`sort_static`] — the device sort that allocates nothing (scratch in a static device array; callers under CUDA
//!   graph capture, tight serving loops, harnesses that forbid `cudaMalloc`), limited to
//!   [`STATIC_MAX_RECORDS`] / [`STATIC_MAX_BYTES`];
//! * [`sort_hinted`] — the device sort for a caller that knows which key bits can differ (no counting pass).
//!
//! All three are `unsafe`: the pointers are device addresses the compiler knows nothing about.

use crate::gpu_key::GpuRadixKey;
use rdst_b200_sys as sys;
use std::os::raw::c_void;

pub use sys::{B200_STATIC_MAX_BYTES as STATIC_MAX_BYTES, B200_STATIC_MAX_RECORDS as STATIC_MAX_RECORDS};

fn opts(stream: *mut c_void, device: i32) -> sys::b200_sort_opts {
    let mut o = sys::b200_sort_opts::default();
    o.stream = stream;
    o.device = device;
    o
}

/// Sorts every segment `[seg_off[i], seg_off[i + 1])` of the `n` records at `dev_ptr` on its own, in place.
/// `seg_off`: ascending record offsets, `segments + 1` of them, the last one `<= n`.
///
/// # Safety
/// `dev_ptr` must be a device allocation of at least `n` records of `T` on `device` (-1 = the current one).
pub unsafe fn sort_segments<T: GpuRadixKey>(dev_ptr: *mut T, n: usize, seg_off: &[u64], stream: *mut c_void, device: i32) -> sys::b200_stats {
    assert_ne!(T::LEVELS, 0, "RadixKey must have at least 1 level");
    assert!(!seg_off.is_empty(), "seg_off holds segments + 1 offsets");
    let desc = T::descriptor();
    let o = opts(stream, device);
    let mut stats = sys::b200_stats::default();
    let rc = sys::b200_sort_segments_device(dev_ptr as *mut c_void, n, seg_off.as_ptr(), seg_off.len() - 1, &desc.0, &o, &mut stats);
    if rc != 0 {
        crate::engine_panic(rc);
    }
    stats
}

/// The device sort that allocates nothing.  Only ENQUEUES on `stream` (null = the legacy default stream) and returns;
/// calls that overlap on one GPU race on the shared scratch array.  Panics beyond the static limits.
///
/// # Safety
/// As [`sort_segments`].
pub unsafe fn sort_static<T: GpuRadixKey>(dev_ptr: *mut T, n: usize, stream: *mut c_void, device: i32) {
    assert_ne!(T::LEVELS, 0, "RadixKey must have at least 1 level");
    let desc = T::descriptor();
    let o = opts(stream, device);
    let rc = sys::b200_sort_device_static(dev_ptr as *mut c_void, n, &desc.0, &o);
    if rc != 0 {
        crate::engine_panic(rc);
    }
}

/// `b200_sort_device` for a caller that knows which bits of the key image (level `l` = bits `8l .. 8l + 7`) can differ
/// between two records; a superset of the truth is fine, a missing bit gives a wrong order.
///
/// # Safety
/// As [`sort_segments`]; `dev_tmp` is null or a device scratch of `n` records.
pub unsafe fn sort_hinted<T: GpuRadixKey>(dev_ptr: *mut T, dev_tmp: *mut T, n: usize, vary_bits: u64, stream: *mut c_void,
                                          device: i32) -> sys::b200_stats {
    assert_ne!(T::LEVELS, 0, "RadixKey must have at least 1 level");
    let desc = T::descriptor();
    let o = opts(stream, device);
    let mut stats = sys::b200_stats::default();
    let rc = sys::b200_sort_device_hinted(dev_ptr as *mut c_void, dev_tmp as *mut c_void, n, &desc.0, vary_bits, &o, &mut stats);
    if rc != 0 {
        crate::engine_panic(rc);
    }
    stats
}
