//! Debug-build check of a key's lowering: `T::get_level(l)` (host code, the definition) against
//! `b200_get_level(&record, &descriptor, l)` (what the kernels compute) on sampled records, before the
//! first launch.  A wrong `GpuRadixKey::descriptor()` would otherwise sort by the wrong bytes silently.
//! The same cross-check runs against the C oracle in tests/test_abi_cpu.py.

use crate::gpu_key::{GpuRadixKey, KeyDesc};

const SAMPLES: usize = 64;

pub(crate) fn check_descriptor<T: GpuRadixKey>(desc: &KeyDesc, data: &[T]) {
    if !(cfg!(debug_assertions) || cfg!(feature = "validate")) {
        return;
    }
    assert_eq!(desc.levels(), T::LEVELS, "GpuRadixKey::descriptor() declares {} levels, RadixKey::LEVELS is {}", desc.levels(), T::LEVELS);
    let stride = (data.len() / SAMPLES).max(1);
    for rec in data.iter().step_by(stride).take(SAMPLES) {
        for level in 0..T::LEVELS {
            let host = rec.get_level(level);
            let dev = desc.get_level(rec, level);
            assert_eq!(host, dev, "descriptor disagrees with RadixKey::get_level at level {level}: {host:#04x} (get_level) vs {dev:#04x} (descriptor)");
        }
    }
}
