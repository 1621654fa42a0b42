//! `RadixKey` — the reference's key trait (rdst src/radix_key.rs:1-5), unchanged: `LEVELS` bytes per
//! key, `get_level(0)` the LEAST significant one.  The impls below give the same bytes as rdst's
//! (src/radix_key_impl.rs:3-185); on the device the engine applies the same rules through the
//! descriptor (`GpuRadixKey`), and the validator compares the two.

pub trait RadixKey: Copy {
    const LEVELS: usize;

    fn get_level(&self, level: usize) -> u8;
}

macro_rules! unsigned_key {
    ($($t:ty),*) => {$(
        impl RadixKey for $t {
            const LEVELS: usize = std::mem::size_of::<$t>();
            #[inline(always)]
            fn get_level(&self, level: usize) -> u8 {
                self.to_le_bytes()[level]
            }
        }
    )*};
}
unsigned_key!(u8, u16, u32, u64, u128, usize);

// signed: two's complement with the sign bit flipped orders like the unsigned image
macro_rules! signed_key {
    ($($t:ty),*) => {$(
        impl RadixKey for $t {
            const LEVELS: usize = std::mem::size_of::<$t>();
            #[inline(always)]
            fn get_level(&self, level: usize) -> u8 {
                (*self ^ <$t>::MIN).to_le_bytes()[level]
            }
        }
    )*};
}
signed_key!(i8, i16, i32, i64, i128, isize);

// floats: IEEE-754 total order — negative values flip every bit, the others only the sign bit
impl RadixKey for f32 {
    const LEVELS: usize = 4;
    #[inline(always)]
    fn get_level(&self, level: usize) -> u8 {
        let b = self.to_bits();
        let image = if b >> 31 == 1 { !b } else { b ^ (1 << 31) };
        image.to_le_bytes()[level]
    }
}

impl RadixKey for f64 {
    const LEVELS: usize = 8;
    #[inline(always)]
    fn get_level(&self, level: usize) -> u8 {
        let b = self.to_bits();
        let image = if b >> 63 == 1 { !b } else { b ^ (1 << 63) };
        image.to_le_bytes()[level]
    }
}

// byte strings compare lexicographically: the FIRST byte is the most significant level
impl<const N: usize> RadixKey for [u8; N] {
    const LEVELS: usize = N;
    #[inline(always)]
    fn get_level(&self, level: usize) -> u8 {
        self[N - 1 - level]
    }
}
