//! The reference's own test shapes (rdst src/radix_sort.rs tests, src/test_utils.rs:97-134) against
//! `slice::sort`; needs a B200 (`cargo test --release -- --ignored`).  Not compiled in the engine's
//! build image; the same cases run through the ctypes binding in tests/test_gpu_parity.py.
use rdst_b200::RadixSort;

fn lcg(seed: &mut u64) -> u64 {
    *seed = seed.wrapping_mul(6364136223846793005).wrapping_add(1442695040888963407);
    *seed
}

macro_rules! int_case {
    ($name:ident, $t:ty) => {
        #[test]
        #[ignore = "needs a B200"]
        fn $name() {
            let mut s = 0x5EEDu64;
            for &n in &[0usize, 1, 2, 127, 128, 129, 16_384, 16_385, 1_000_003, 40_000_000] {
                let mut v: Vec<$t> = (0..n).map(|_| lcg(&mut s) as $t).collect();
                let mut e = v.clone();
                e.sort_unstable();
                v.radix_sort_unstable();
                assert_eq!(v, e, "n = {n}");
                v.radix_sort_builder().with_low_mem_tuner().sort();
                assert_eq!(v, e);
            }
        }
    };
}
int_case!(sorts_u8, u8);
int_case!(sorts_u16, u16);
int_case!(sorts_u32, u32);
int_case!(sorts_u64, u64);
int_case!(sorts_u128, u128);
int_case!(sorts_i8, i8);
int_case!(sorts_i16, i16);
int_case!(sorts_i32, i32);
int_case!(sorts_i64, i64);
int_case!(sorts_i128, i128);

#[test]
#[ignore = "needs a B200"]
fn sorts_floats_in_total_order() {
    let mut s = 7u64;
    let mut v: Vec<f64> = (0..1_000_000).map(|_| f64::from_bits(lcg(&mut s))).collect();
    let mut e = v.clone();
    e.sort_by(|a, b| a.total_cmp(b));
    v.radix_sort_unstable();
    assert!(v.iter().zip(&e).all(|(a, b)| a.to_bits() == b.to_bits()));
}

#[test]
#[ignore = "needs a B200"]
fn sorts_byte_arrays() {
    let mut v: Vec<[u8; 3]> = vec![[3, 3, 3], [1, 2, 3], [3, 2, 1]];
    v.radix_sort_unstable();
    assert_eq!(v, vec![[1, 2, 3], [3, 2, 1], [3, 3, 3]]);
}

#[test]
#[should_panic(expected = "at least 1 level")]
fn zero_levels_panics_at_builder_construction() {
    #[derive(Copy, Clone)]
    struct Z;
    impl rdst_b200::RadixKey for Z {
        const LEVELS: usize = 0;
        fn get_level(&self, _: usize) -> u8 {
            0
        }
    }
    impl rdst_b200::GpuRadixKey for Z {
        fn descriptor() -> rdst_b200::KeyDesc {
            rdst_b200::KeyDesc::bytes::<Z>(&[])
        }
    }
    let mut v: Vec<Z> = vec![];
    let _ = v.radix_sort_builder(); // panics here, even for an empty input (rdst src/radix_sort_builder.rs:22)
}
