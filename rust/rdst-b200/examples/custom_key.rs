// A struct with a partial-LEVELS key and a payload, sorted in rdst's stable order, and a user Tuner.
use rdst_b200::tuner::{Algorithm, Tuner, TuningParams};
use rdst_b200::{GpuRadixKey, KeyDesc, RadixKey, RadixSort};

#[derive(Copy, Clone, Debug)]
#[repr(C)]
struct Rec {
    key: u32, // values below 2^24: three levels are enough
    payload: u64,
}

impl RadixKey for Rec {
    const LEVELS: usize = 3;
    #[inline]
    fn get_level(&self, level: usize) -> u8 {
        (self.key >> (level * 8)) as u8
    }
}

impl GpuRadixKey for Rec {
    fn descriptor() -> KeyDesc {
        KeyDesc::field::<Rec, u32>(std::mem::offset_of!(Rec, key), 3)
    }
}

struct PreferLsb;

impl Tuner for PreferLsb {
    fn pick_algorithm(&self, p: &TuningParams, _counts: &[usize]) -> Algorithm {
        if p.input_len >= 500_000 { Algorithm::MtLsb } else { Algorithm::Lsb }
    }
}

fn main() {
    let mut v: Vec<Rec> = (0..10_000_000u64).map(|i| Rec { key: (i.wrapping_mul(2654435761) >> 8) as u32 & 0xFF_FFFF, payload: i }).collect();
    v.radix_sort_builder().with_tuner(&PreferLsb).sort(); // Lsb / MtLsb: LSD passes, stable
    assert!(v.windows(2).all(|w| w[0].key < w[1].key || (w[0].key == w[1].key && w[0].payload < w[1].payload)));
}
