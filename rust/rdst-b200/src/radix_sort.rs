//! `RadixSort` for `Vec<T>` and `[T]` — rdst src/radix_sort.rs:4-45.

use crate::gpu_key::GpuRadixKey;
use crate::radix_sort_builder::RadixSortBuilder;

pub trait RadixSort<T> {
    /// Sorts by the `RadixKey` of `T`, unstable (records with equal keys may be reordered).
    ///
    /// ```ignore
    /// use rdst_b200::RadixSort;
    /// let mut values = [3u32, 1, 2];
    /// values.radix_sort_unstable();
    /// assert_eq!(values, [1, 2, 3]);
    /// ```
    fn radix_sort_unstable(&mut self);

    fn radix_sort_builder(&'_ mut self) -> RadixSortBuilder<'_, T>;
}

impl<T: GpuRadixKey + Send + Sync> RadixSort<T> for Vec<T> {
    fn radix_sort_unstable(&mut self) {
        self.radix_sort_builder().sort();
    }

    fn radix_sort_builder(&'_ mut self) -> RadixSortBuilder<'_, T> {
        RadixSortBuilder::new(self)
    }
}

impl<T: GpuRadixKey + Send + Sync> RadixSort<T> for [T] {
    fn radix_sort_unstable(&mut self) {
        self.radix_sort_builder().sort();
    }

    fn radix_sort_builder(&'_ mut self) -> RadixSortBuilder<'_, T> {
        RadixSortBuilder::new(self)
    }
}
