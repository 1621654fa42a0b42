// cargo run --release --example basic        (needs a B200 and libb200sort.so)
use rdst_b200::{RadixSort, SortStats};

fn main() {
    let mut v: Vec<u64> = (0..100_000_000u64).map(|i| i.wrapping_mul(0x9E37_79B9_7F4A_7C15) ^ (i >> 7)).collect();
    let mut stats = SortStats::default();
    v.radix_sort_builder().with_stats(&mut stats).sort();
    assert!(v.windows(2).all(|w| w[0] <= w[1]));
    println!("levels run {}, route hybrid={} unstable={}, device sort {:.2} ms, h2d {:.1} ms",
             stats.levels_run, stats.hybrid, stats.unstable, stats.ms_hist + stats.ms_passes, stats.ms_h2d);

    let mut f = vec![3.5f32, -0.0, 0.0, -7.25, f32::INFINITY, 1.0];
    f.radix_sort_unstable();
    println!("{f:?}");
}
