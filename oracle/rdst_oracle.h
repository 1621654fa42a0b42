/*
 * rdst_oracle.h — CPU restatement of the nessex/rdst sort path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker or the timed CPU baseline — the
 * product (rdst_b200/, libb200sort.so) never links, imports or calls anything here.
 *
 * The reference is pure Rust and there is no Rust toolchain in this image, so the
 * reference itself cannot be compiled (oracle/_ref does not exist).  Every function
 * cites the reference file:line it restates (paths relative to the reference checkout).
 * Parity pinning: the oracle is checked against every known answer the reference's own
 * tests, doctests and examples hold for this path (tests/golden/reference_vectors.json,
 * tests/test_oracle_golden.py) and against an independent model (numpy stable sort on
 * the key bytes).
 */
#ifndef RDST_ORACLE_H
#define RDST_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORACLE_MAX_LEVELS 64

/* Same layout as b200_key_desc (include/b200sort.h) so tests can hand the same
 * bytes to both sides. */
typedef struct {
    uint16_t byte_offset;
    uint8_t xor_mask;
    uint8_t _pad;
} oracle_level;

typedef struct {
    uint32_t elem_size, elem_align, levels, kind, key_offset, key_width;
    oracle_level level[ORACLE_MAX_LEVELS];
} oracle_key_desc;

enum { ORACLE_KEY_UINT = 0, ORACLE_KEY_SINT = 1, ORACLE_KEY_FLOAT = 2, ORACLE_KEY_BYTES = 3 };

/* src/tuner.rs:10-31 (multi-threaded enum order) */
enum {
    ORACLE_ALGO_MT_OOP = 0,
    ORACLE_ALGO_MT_LSB = 1,
    ORACLE_ALGO_SCANNING = 2,
    ORACLE_ALGO_RECOMBINATING = 3,
    ORACLE_ALGO_COMPARATIVE = 4,
    ORACLE_ALGO_LR_LSB = 5,
    ORACLE_ALGO_LSB = 6,
    ORACLE_ALGO_REGIONS = 7,
    ORACLE_ALGO_SKA = 8
};

/* Tuners: src/tuners/{standard,low_memory,single_threaded}_tuner.rs; SINGLE_ALGO is
 * test_utils.rs:40-49 (forces one algorithm at every level). */
enum {
    ORACLE_TUNER_STANDARD = 0,
    ORACLE_TUNER_LOW_MEMORY = 1,
    ORACLE_TUNER_SINGLE_THREADED = 2,
    ORACLE_TUNER_SINGLE_ALGO = 3
};

/* RadixKey::get_level through the descriptor (src/radix_key.rs:1-5). -1 on bad args. */
int oracle_get_level(const void* record, const oracle_key_desc* desc, uint32_t level);

/* The built-in RadixKey impls restated one by one (src/radix_key_impl.rs); `type` is
 * one of "u8","u16","u32","u64","u128","i8","i16","i32","i64","i128","f32","f64". */
int oracle_get_level_builtin(const char* type, const void* value, uint32_t level);
/* [u8; N]: src/radix_key_impl.rs:78-85 */
int oracle_get_level_bytes(const uint8_t* value, uint32_t n, uint32_t level);

/* get_counts_with_ends (src/sort_utils.rs:108-180): counts[256], already_sorted, first, last */
int oracle_get_counts(const void* data, size_t n, const oracle_key_desc* desc, uint32_t level,
                      uint64_t* counts, int* already_sorted, int* first, int* last);
/* get_tile_counts + aggregate (src/sort_utils.rs:192-249): returns already_sorted flag */
int oracle_get_tile_counts_sorted(const void* data, size_t n, const oracle_key_desc* desc,
                                  size_t tile_size, uint32_t level, uint64_t* agg_counts);

/* RadixSortBuilder::sort (src/radix_sort_builder.rs:149-157) -> Sorter::route_top_level
 * (src/sorter.rs:105-119).  threads <= 0: all hardware threads.  multi_threaded = the
 * builder's with_parallel flag.  algo is used by ORACLE_TUNER_SINGLE_ALGO only.
 * Returns 0, or non-zero on bad descriptor / LEVELS == 0. */
int oracle_sort(void* data, size_t n, const oracle_key_desc* desc, int tuner, int algo,
                int multi_threaded, int threads);

/* One stable counting-sort pass of one level, src -> dst: out_of_place_sort /
 * lr_out_of_place_sort (src/sorts/out_of_place_sort.rs:51-108,201-275). */
int oracle_out_of_place_pass(const void* src, void* dst, size_t n, const oracle_key_desc* desc,
                             uint32_t level, int lr);

/* Independent checker (not a restatement): is `data` non-decreasing by key, level
 * LEVELS-1 most significant (README.md:37-38)?  Returns 1/0. */
int oracle_is_sorted(const void* data, size_t n, const oracle_key_desc* desc);

int oracle_hardware_threads(void);
const char* oracle_version(void);

#ifdef __cplusplus
}
#endif
#endif
