//! Raw bindings of `include/b200sort.h` — one item per declaration of the header, same order.
//! UNCOMPILED in the engine's build image (no Rust toolchain there); layouts are pinned by the
//! `const` assertions at the bottom, which repeat tests/test_abi_cpu.py.
#![allow(non_camel_case_types)]

use std::os::raw::{c_char, c_int, c_void};

pub const B200_MAX_LEVELS: usize = 64;
pub const B200_MAX_GPUS: usize = 16;
pub const B200_IPC_HANDLE_BYTES: usize = 64;

// return codes
pub const B200_OK: c_int = 0;
pub const B200_ERR_LEVELS_ZERO: c_int = 1;
pub const B200_ERR_BAD_DESC: c_int = 2;
pub const B200_ERR_CUDA: c_int = 3;
pub const B200_ERR_NO_DEVICE: c_int = 4;
pub const B200_ERR_ALLOC: c_int = 5;
pub const B200_ERR_BAD_ARG: c_int = 6;
pub const B200_ERR_LEVEL_RANGE: c_int = 7;
pub const B200_ERR_TOO_LARGE: c_int = 9;
pub const B200_STATIC_MAX_RECORDS: usize = 2 << 20;
pub const B200_STATIC_MAX_BYTES: usize = 16 << 20;

// b200_key_kind
pub const B200_KEY_UINT: u32 = 0;
pub const B200_KEY_SINT: u32 = 1;
pub const B200_KEY_FLOAT: u32 = 2;
pub const B200_KEY_BYTES: u32 = 3;

// b200_route_hint
pub const B200_HINT_AUTO: u32 = 0;
pub const B200_HINT_LSD: u32 = 1;
pub const B200_HINT_LOW_SCRATCH: u32 = 2;

// b200_route_kind
pub const B200_ROUTE_LSD: u32 = 0;
pub const B200_ROUTE_HYBRID: u32 = 1;
pub const B200_ROUTE_MSD1: u32 = 2;
pub const B200_ROUTE_SMALL: u32 = 3;
pub const B200_ROUTE_TRIVIAL: u32 = 4;

#[repr(C)]
#[derive(Copy, Clone, Default, Debug, PartialEq, Eq)]
pub struct b200_level {
    pub byte_offset: u16,
    pub xor_mask: u8,
    pub _pad: u8,
}

#[repr(C)]
#[derive(Copy, Clone)]
pub struct b200_key_desc {
    pub elem_size: u32,
    pub elem_align: u32,
    pub levels: u32,
    pub kind: u32,
    pub key_offset: u32,
    pub key_width: u32,
    pub level: [b200_level; B200_MAX_LEVELS],
}

#[repr(C)]
#[derive(Copy, Clone)]
pub struct b200_sort_opts {
    pub stable: u32,
    pub variant: i32,
    pub device: i32,
    pub stream: *mut c_void,
    pub route: u32,
    pub _reserved: u32,
}

impl Default for b200_sort_opts {
    fn default() -> Self {
        Self { stable: 0, variant: -1, device: -1, stream: std::ptr::null_mut(), route: B200_HINT_AUTO, _reserved: 0 }
    }
}

#[repr(C)]
#[derive(Copy, Clone, Default, Debug)]
pub struct b200_stats {
    pub levels_run: u32,
    pub levels_skipped: u32,
    pub variant: u32,
    pub copy_back: u32,
    pub algorithmic_bytes: u64,
    pub ms_hist: f32,
    pub ms_passes: f32,
    pub ms_h2d: f32,
    pub ms_d2h: f32,
    pub ms_total: f32,
    pub tiles_per_pass: u64,
    pub already_sorted: u32,
    pub kernel_launches: u32,
    pub hybrid: u32,
    pub slow_segments: u32,
    pub local_levels: u32,
    pub ms_joint: f32,
    pub ms_local: f32,
    pub unstable: u32,
    pub heavy_segments: u32,
    pub heavy_rounds: u32,
}

#[repr(C)]
#[derive(Copy, Clone, Default, Debug)]
pub struct b200_multi_stats {
    pub ngpus: u32,
    pub level: u32,
    pub splitter_mode: u32,
    pub kernel_launches: u32,
    pub imbalance: f64,
    pub max_sent_bytes: u64,
    pub n_out: [u64; B200_MAX_GPUS],
    pub ms_plan: f32,
    pub ms_exchange: f32,
    pub ms_local_sort: f32,
    pub ms_h2d: f32,
    pub ms_d2h: f32,
    pub ms_total: f32,
    pub local0: b200_stats,
}

#[repr(C)]
#[derive(Copy, Clone, Default, Debug)]
pub struct b200_window_plan {
    pub eligible: u32,
    pub aligned: u32,
    pub vary: u64,
    pub digit_lo: [u32; 2],
    pub digit_width: [u32; 2],
    pub nlocal: u32,
    pub local_lo: [u32; 8],
    pub local_width: [u32; 8],
}

#[repr(C)]
#[derive(Copy, Clone, Default, Debug)]
pub struct b200_route {
    pub route: u32,
    pub active_levels: u32,
    pub levels_run: u32,
    pub msd_level: u32,
    pub max_bucket: u64,
}

extern "C" {
    // ---- the sort ----
    pub fn b200_sort_host(host_ptr: *mut c_void, n: usize, desc: *const b200_key_desc, opts: *const b200_sort_opts,
                          stats: *mut b200_stats) -> c_int;
    pub fn b200_sort_device(dev_ptr: *mut c_void, dev_tmp: *mut c_void, n: usize, desc: *const b200_key_desc,
                            opts: *const b200_sort_opts, stats: *mut b200_stats) -> c_int;
    /// b200_sort_device for a caller that knows which key-image bits can vary (superset allowed): no counting pass.
    pub fn b200_sort_device_hinted(dev_ptr: *mut c_void, dev_tmp: *mut c_void, n: usize, desc: *const b200_key_desc,
                                   vary_bits: u64, opts: *const b200_sort_opts, stats: *mut b200_stats) -> c_int;
    /// Local sort of an array that arrives partitioned on `level` (bucket sizes `counts[256]`); result at *result_out.
    pub fn b200_sort_device_presplit(dev_ptr: *mut c_void, dev_tmp: *mut c_void, n: usize, desc: *const b200_key_desc, level: u32,
                                     counts: *const u64, opts: *const b200_sort_opts, stats: *mut b200_stats,
                                     result_out: *mut *mut c_void) -> c_int;
    /// Every segment [seg_off[i], seg_off[i + 1]) sorted on its own (seg_off: host memory, nseg + 1 offsets).
    pub fn b200_sort_segments_device(dev_ptr: *mut c_void, n: usize, seg_off: *const u64, nseg: usize,
                                     desc: *const b200_key_desc, opts: *const b200_sort_opts, stats: *mut b200_stats) -> c_int;
    /// Allocates nothing (scratch is a static device array); no counterpart in the reference.
    /// Limits: B200_STATIC_MAX_RECORDS / B200_STATIC_MAX_BYTES, else B200_ERR_TOO_LARGE.  Enqueues only.
    pub fn b200_sort_device_static(dev_ptr: *mut c_void, n: usize, desc: *const b200_key_desc,
                                   opts: *const b200_sort_opts) -> c_int;
    // ---- building blocks ----
    pub fn b200_histogram_device(dev_ptr: *const c_void, n: usize, desc: *const b200_key_desc, opts: *const b200_sort_opts,
                                 hist_out: *mut u64) -> c_int;
    pub fn b200_pass_device(dev_src: *const c_void, dev_dst: *mut c_void, n: usize, desc: *const b200_key_desc, level: u32,
                            counts: *const u64, opts: *const b200_sort_opts) -> c_int;
    pub fn b200_partition_device(dev_src: *const c_void, dev_dst: *mut c_void, n: usize, desc: *const b200_key_desc,
                                 level: u32, counts: *const u64, split: *const u32, parts: u32,
                                 part_counts_out: *mut u64, opts: *const b200_sort_opts) -> c_int;
    pub fn b200_partition_peers_device(dev_src: *const c_void, dst_ptrs: *const *mut c_void, n: usize,
                                       desc: *const b200_key_desc, level: u32, counts: *const u64, split: *const u32,
                                       parts: u32, part_counts_out: *mut u64, opts: *const b200_sort_opts) -> c_int;
    pub fn b200_key_prefix_supported(desc: *const b200_key_desc) -> c_int;
    pub fn b200_key_prefix_sample_device(dev_ptr: *const c_void, n: usize, desc: *const b200_key_desc, stride: usize,
                                         count: usize, prefixes_out: *mut u64, opts: *const b200_sort_opts) -> c_int;
    pub fn b200_splitter_count_device(dev_ptr: *const c_void, n: usize, desc: *const b200_key_desc, splitters: *const u64,
                                      nsplit: u32, less_out: *mut u64, equal_out: *mut u64,
                                      opts: *const b200_sort_opts) -> c_int;
    pub fn b200_partition_splitters_device(dev_src: *const c_void, dev_dst: *mut c_void, dst_ptrs: *const *mut c_void,
                                           n: usize, desc: *const b200_key_desc, splitters: *const u64, nsplit: u32,
                                           incl_mask: u32, part_counts: *const u64, opts: *const b200_sort_opts) -> c_int;
    pub fn b200_ipc_alloc(dev_ptr: *mut *mut c_void, bytes: usize, handle_out: *mut c_void) -> c_int;
    pub fn b200_ipc_open(handle: *const c_void, dev_ptr: *mut *mut c_void) -> c_int;
    pub fn b200_ipc_close(dev_ptr: *mut c_void) -> c_int;
    pub fn b200_ipc_free(dev_ptr: *mut c_void) -> c_int;
    // ---- several GPUs, one process ----
    pub fn b200_sort_multi_device(dev_ptrs: *const *mut c_void, n_in: *const usize, ngpus: c_int, devices: *const c_int,
                                  desc: *const b200_key_desc, opts: *const b200_sort_opts, out_ptrs: *mut *mut c_void,
                                  n_out: *mut usize, stats: *mut b200_multi_stats) -> c_int;
    pub fn b200_sort_multi_host(host_ptr: *mut c_void, n: usize, ngpus: c_int, devices: *const c_int,
                                desc: *const b200_key_desc, opts: *const b200_sort_opts,
                                stats: *mut b200_multi_stats) -> c_int;
    pub fn b200_multi_release() -> c_int;
    pub fn b200_multi_plan_digits(all_hist: *const u64, ranks: u32, levels: u32, level_out: *mut i32, split_out: *mut u32,
                                  send_matrix_out: *mut i64, imbalance_out: *mut f64) -> c_int;
    pub fn b200_multi_choose_sample_splitters(samples: *const u64, weights: *const f64, count: usize, parts: u32,
                                              splitters_out: *mut u64) -> c_int;
    pub fn b200_multi_plan_splitters(less: *const i64, equal: *const i64, n_per_rank: *const i64, splitters: *const u64,
                                     ranks: u32, parts: u32, exact_ties: c_int, tie_thresholds_out: *mut i64,
                                     send_matrix_out: *mut i64, incl_masks_out: *mut u32) -> c_int;
    pub fn b200_copy_h2d(dev_dst: *mut c_void, host_src: *const c_void, bytes: usize, device: c_int) -> c_int;
    pub fn b200_copy_d2h(host_dst: *mut c_void, dev_src: *const c_void, bytes: usize, device: c_int) -> c_int;
    // ---- host-only helpers ----
    pub fn b200_get_level(record: *const c_void, desc: *const b200_key_desc, level: u32) -> c_int;
    pub fn b200_host_alloc(out: *mut *mut c_void, bytes: usize) -> c_int;
    pub fn b200_host_free(p: *mut c_void) -> c_int;
    pub fn b200_release_scratch(device: c_int) -> c_int;
    pub fn b200_hybrid_plan(desc: *const b200_key_desc, hist: *const u64, n: usize, out: *mut b200_window_plan) -> c_int;
    pub fn b200_route_plan(desc: *const b200_key_desc, hist: *const u64, n: usize, out: *mut b200_route) -> c_int;
    pub fn b200_variant_count() -> c_int;
    pub fn b200_variant_name(variant: c_int) -> *const c_char;
    pub fn b200_strerror(code: c_int) -> *const c_char;
    pub fn b200_last_cuda_error() -> *const c_char;
    pub fn b200_version() -> *const c_char;
}

// layouts, as tests/test_abi_cpu.py::test_struct_layouts_match_header pins them for the ctypes binding
const _: () = assert!(std::mem::size_of::<b200_level>() == 4);
const _: () = assert!(std::mem::size_of::<b200_key_desc>() == 24 + 4 * B200_MAX_LEVELS);
const _: () = assert!(std::mem::size_of::<b200_sort_opts>() == 32);
const _: () = assert!(std::mem::size_of::<b200_stats>() == 96);
const _: () = assert!(std::mem::size_of::<b200_multi_stats>() == 280);
const _: () = assert!(std::mem::size_of::<b200_route>() == 24);
