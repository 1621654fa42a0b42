//! One `&mut [T]` sorted over several GPUs of the box in one process (`b200_sort_multi_host`):
//! the slice is cut into shards, every GPU pulls its shard over its own PCIe link, the shards are
//! range-partitioned over NVLink (the partition kernel stores straight into the owners' buffers),
//! sorted locally, and written back in place.  The reference has no counterpart; this is the
//! "inputs larger than one GPU, or 8-GPU mode" of the engine's specification.

use crate::gpu_key::GpuRadixKey;
use rdst_b200_sys as sys;
use std::os::raw::c_void;

pub type MultiStats = sys::b200_multi_stats;

/// `devices`: CUDA ordinals to use (at most 16, all different). Returns the engine's report.
pub fn sort_multi<T: GpuRadixKey + Send + Sync>(data: &mut [T], devices: &[i32]) -> MultiStats {
    assert_ne!(T::LEVELS, 0, "RadixKey must have at least 1 level");
    let mut stats = MultiStats::default();
    if data.len() <= 1 {
        return stats;
    }
    let desc = T::descriptor();
    crate::validate::check_descriptor(&desc, data);
    let rc = unsafe {
        sys::b200_sort_multi_host(data.as_mut_ptr() as *mut c_void, data.len(), devices.len() as i32, devices.as_ptr(), &desc.0,
                                  std::ptr::null(), &mut stats)
    };
    if rc != 0 {
        crate::engine_panic(rc);
    }
    stats
}
