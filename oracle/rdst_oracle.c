/*
 * rdst_oracle.c — CPU restatement of the nessex/rdst sort path.  TEST INFRASTRUCTURE ONLY:
 * see the header of rdst_oracle.h for who may load this and how its parity is pinned.
 * The reference (pure Rust, crate rdst 0.20.14 -> unreleased 0.21.0) cannot be built in
 * this image (no rustc/cargo), so this file restates it; citations are file:line into
 * the reference checkout.
 */
#define _GNU_SOURCE
#include "rdst_oracle.h"

#include <stdio.h>
#include <string.h>

#include "rdst_pool.h"

/* ------------------------------------------------------------------------- */
/* context shared by all element-type instantiations                          */
/* ------------------------------------------------------------------------- */
typedef struct {
    uint32_t levels;     /* T::LEVELS */
    int tuner, algo;     /* ORACLE_TUNER_*, ORACLE_ALGO_* (single-algo tuner) */
    int multi_threaded;  /* RadixSortBuilder::with_parallel (src/radix_sort_builder.rs:53-60) */
    int mt_feature;      /* cargo feature "multi-threaded" (Cargo.toml:15-18): 1 unless threads == 1 and
                            the single-threaded tuner is asked for */
    /* descriptor-driven get_level for record types */
    uint32_t elem_size, kind, sign_byte;
    uint16_t off[ORACLE_MAX_LEVELS];
    uint8_t xr[ORACLE_MAX_LEVELS];
} octx;

typedef struct {
    void (*sort_top)(const octx*, void*, size_t);
    void (*counts)(const octx*, const void*, size_t, int, size_t*, int*, uint8_t*, uint8_t*);
    int (*tile_counts)(const octx*, const void*, size_t, size_t, int, size_t*);
    void (*pass)(const octx*, const void*, void*, size_t, int, int);
} otype_vtbl;

static inline size_t div_ceil(size_t a, size_t b) { return (a + b - 1) / b; }

/* src/sort_utils.rs:9-20 */
static inline void get_prefix_sums(const size_t* counts, size_t* sums) {
    size_t running = 0;
    for (int i = 0; i < 256; ++i) {
        sums[i] = running;
        running += counts[i];
    }
}
/* src/sort_utils.rs:22-31 */
static inline void get_end_offsets(const size_t* counts, const size_t* prefix_sums, size_t* end_offsets) {
    for (int i = 0; i < 256; ++i) end_offsets[i] = prefix_sums[i == 255 ? 255 : i + 1];
    end_offsets[255] += counts[255];
}

/* ------------------------------------------------------------------------- */
/* tuners                                                                     */
/* ------------------------------------------------------------------------- */
/* src/tuners/standard_tuner.rs:13-63 */
static int standard_tuner(size_t level, size_t total_levels, size_t input_len, const size_t* counts) {
    if (input_len <= 128) return ORACLE_ALGO_COMPARATIVE;
    const size_t depth = total_levels - level - 1;
    if (input_len >= 5000) {
        const size_t thr = (input_len / 256) * 2;
        for (int i = 0; i < 256; ++i) {
            if (counts[i] >= thr) {
                if (depth == 0) {
                    if (input_len <= 200000) return ORACLE_ALGO_LR_LSB;
                    if (input_len <= 350000) return ORACLE_ALGO_SKA;
                    if (input_len <= 4000000) return ORACLE_ALGO_MT_LSB;
                    return ORACLE_ALGO_REGIONS;
                } else {
                    if (input_len <= 200000) return ORACLE_ALGO_LR_LSB;
                    if (input_len <= 800000) return ORACLE_ALGO_SKA;
                    if (input_len <= 5000000) return ORACLE_ALGO_RECOMBINATING;
                    return ORACLE_ALGO_REGIONS;
                }
            }
        }
    }
    if (depth > 0) {
        if (input_len <= 200000) return ORACLE_ALGO_LSB;
        if (input_len <= 800000) return ORACLE_ALGO_SKA;
        if (input_len <= 50000000) return ORACLE_ALGO_RECOMBINATING;
        return ORACLE_ALGO_SCANNING;
    } else {
        if (input_len <= 150000) return ORACLE_ALGO_LSB;
        if (input_len <= 260000) return ORACLE_ALGO_SKA;
        if (input_len <= 50000000) return ORACLE_ALGO_RECOMBINATING;
        return ORACLE_ALGO_SCANNING;
    }
}
/* src/tuners/low_memory_tuner.rs:13-43 */
static int low_memory_tuner(size_t input_len, const size_t* counts) {
    if (input_len <= 128) return ORACLE_ALGO_COMPARATIVE;
    if (input_len >= 5000) {
        const size_t thr = (input_len / 256) * 2;
        for (int i = 0; i < 256; ++i) {
            if (counts[i] >= thr) {
                if (input_len <= 50000) return ORACLE_ALGO_LR_LSB;
                if (input_len <= 1000000) return ORACLE_ALGO_SKA;
                return ORACLE_ALGO_REGIONS;
            }
        }
    }
    if (input_len <= 50000) return ORACLE_ALGO_LSB;
    if (input_len <= 1000000) return ORACLE_ALGO_SKA;
    return ORACLE_ALGO_REGIONS;
}
/* src/tuners/single_threaded_tuner.rs:13-43 */
static int single_threaded_tuner(size_t level, size_t total_levels, size_t input_len, const size_t* counts) {
    if (input_len <= 128) return ORACLE_ALGO_COMPARATIVE;
    const size_t depth = total_levels - level - 1;
    if (input_len >= 5000) {
        const size_t thr = (input_len / 256) * 2;
        for (int i = 0; i < 256; ++i) {
            if (counts[i] >= thr) return (input_len > 100000 && depth < 2) ? ORACLE_ALGO_SKA : ORACLE_ALGO_LR_LSB;
        }
    }
    return (input_len > 800000 && depth == 0) ? ORACLE_ALGO_SKA : ORACLE_ALGO_LSB;
}

/* Tuner::pick_algorithm (src/tuner.rs:33-35) as called at src/sorter.rs:67-76 */
static int pick_algorithm(const octx* cx, size_t threads, int level, int total_levels, size_t input_len,
                          size_t parent_len, int has_parent, const size_t* counts) {
    (void)threads, (void)parent_len, (void)has_parent;
    switch (cx->tuner) {
        case ORACLE_TUNER_LOW_MEMORY: return low_memory_tuner(input_len, counts);
        case ORACLE_TUNER_SINGLE_THREADED:
            return single_threaded_tuner((size_t)level, (size_t)total_levels, input_len, counts);
        case ORACLE_TUNER_SINGLE_ALGO: return cx->algo; /* test_utils.rs:40-49 */
        default: return standard_tuner((size_t)level, (size_t)total_levels, input_len, counts);
    }
}

/* ------------------------------------------------------------------------- */
/* RadixKey::get_level for the built-in keys (src/radix_key_impl.rs)          */
/* ------------------------------------------------------------------------- */
static inline uint8_t gl_u8(const uint8_t* p, int l) { (void)l; return *p; }                          /* :3-10 */
static inline uint8_t gl_u16(const uint16_t* p, int l) { return (uint8_t)(*p >> (l * 8)); }           /* :12-19 */
static inline uint8_t gl_u32(const uint32_t* p, int l) { return (uint8_t)(*p >> (l * 8)); }           /* :21-28 */
static inline uint8_t gl_u64(const uint64_t* p, int l) { return (uint8_t)(*p >> (l * 8)); }           /* :30-37 */
static inline uint8_t gl_u128(const unsigned __int128* p, int l) { return (uint8_t)(*p >> (l * 8)); } /* :39-46 */
static inline uint8_t gl_i8(const int8_t* p, int l) { (void)l; return (uint8_t)(*p ^ INT8_MIN); }     /* :87-94 */
static inline uint8_t gl_i16(const int16_t* p, int l) {                                               /* :96-103 */
    return (uint8_t)((int16_t)(*p ^ INT16_MIN) >> (l * 8));
}
static inline uint8_t gl_i32(const int32_t* p, int l) { return (uint8_t)((*p ^ INT32_MIN) >> (l * 8)); } /* :105-112 */
static inline uint8_t gl_i64(const int64_t* p, int l) { return (uint8_t)((*p ^ INT64_MIN) >> (l * 8)); } /* :114-121 */
static inline uint8_t gl_i128(const __int128* p, int l) {                                             /* :123-130 */
    const __int128 min = (__int128)((unsigned __int128)1 << 127);
    return (uint8_t)((*p ^ min) >> (l * 8));
}
/* :162-173  s = bits as i32; s ^= (((s >> 31) as u32) >> 1) as i32; ((s ^ MIN) >> 8l) as u8 */
static inline uint8_t gl_f32(const uint32_t* bits, int l) {
    int32_t s = (int32_t)*bits;
    s ^= (int32_t)(((uint32_t)(s >> 31)) >> 1);
    return (uint8_t)((s ^ INT32_MIN) >> (l * 8));
}
/* :175-185 */
static inline uint8_t gl_f64(const uint64_t* bits, int l) {
    int64_t s = (int64_t)*bits;
    s ^= (int64_t)(((uint64_t)(s >> 63)) >> 1);
    return (uint8_t)((s ^ INT64_MIN) >> (l * 8));
}

/* Descriptor form: digit = byte[off[l]] ^ xr[l]; FLOAT keys with the sign bit set flip
 * every byte instead (equivalent to the :162-185 formula, checked in tests). */
static inline uint8_t gl_desc(const octx* cx, const uint8_t* rec, int l) {
    uint8_t b = rec[cx->off[l]];
    if (cx->kind == ORACLE_KEY_FLOAT && (rec[cx->sign_byte] & 0x80u)) return (uint8_t)(b ^ 0xffu);
    return (uint8_t)(b ^ cx->xr[l]);
}

/* ------------------------------------------------------------------------- */
/* element-type instantiations                                                */
/* ------------------------------------------------------------------------- */
#define T uint8_t
#define SFX u8
#define GETL(cx, p, l) gl_u8((p), (l))
#include "rdst_algos.inc"
#define T uint16_t
#define SFX u16
#define GETL(cx, p, l) gl_u16((p), (l))
#include "rdst_algos.inc"
#define T uint32_t
#define SFX u32
#define GETL(cx, p, l) gl_u32((p), (l))
#include "rdst_algos.inc"
#define T uint64_t
#define SFX u64
#define GETL(cx, p, l) gl_u64((p), (l))
#include "rdst_algos.inc"
#define T int8_t
#define SFX i8
#define GETL(cx, p, l) gl_i8((p), (l))
#include "rdst_algos.inc"
#define T int16_t
#define SFX i16
#define GETL(cx, p, l) gl_i16((p), (l))
#include "rdst_algos.inc"
#define T int32_t
#define SFX i32
#define GETL(cx, p, l) gl_i32((p), (l))
#include "rdst_algos.inc"
#define T int64_t
#define SFX i64
#define GETL(cx, p, l) gl_i64((p), (l))
#include "rdst_algos.inc"
#define T uint32_t
#define SFX f32
#define GETL(cx, p, l) gl_f32((p), (l))
#include "rdst_algos.inc"
#define T uint64_t
#define SFX f64
#define GETL(cx, p, l) gl_f64((p), (l))
#include "rdst_algos.inc"
/* records of N bytes whose key is read through the descriptor (structs, [u8; N], partial keys) */
#define ORACLE_REC(N)                                          \
    typedef struct {                                           \
        uint8_t b[N];                                          \
    } rec##N##_t;
ORACLE_REC(1) ORACLE_REC(2) ORACLE_REC(3) ORACLE_REC(4) ORACLE_REC(5) ORACLE_REC(6) ORACLE_REC(7) ORACLE_REC(8)
ORACLE_REC(12) ORACLE_REC(16) ORACLE_REC(20) ORACLE_REC(24) ORACLE_REC(32) ORACLE_REC(40) ORACLE_REC(48) ORACLE_REC(64)
#define T rec1_t
#define SFX r1
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec2_t
#define SFX r2
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec3_t
#define SFX r3
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec4_t
#define SFX r4
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec5_t
#define SFX r5
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec6_t
#define SFX r6
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec7_t
#define SFX r7
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec8_t
#define SFX r8
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec12_t
#define SFX r12
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec16_t
#define SFX r16
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec20_t
#define SFX r20
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec24_t
#define SFX r24
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec32_t
#define SFX r32
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec40_t
#define SFX r40
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec48_t
#define SFX r48
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"
#define T rec64_t
#define SFX r64
#define GETL(cx, p, l) gl_desc((cx), (p)->b, (l))
#include "rdst_algos.inc"

/* ------------------------------------------------------------------------- */
/* dispatch                                                                   */
/* ------------------------------------------------------------------------- */
static int lower(const oracle_key_desc* d, octx* cx) {
    if (!d) return 6;
    if (d->levels == 0) return 1; /* assert_ne!(T::LEVELS, 0), src/radix_sort_builder.rs:22 */
    if (d->levels > ORACLE_MAX_LEVELS || d->elem_size == 0 || d->elem_size > 65535) return 2;
    memset(cx, 0, sizeof(*cx));
    cx->levels = d->levels;
    cx->elem_size = d->elem_size;
    cx->kind = d->kind;
    if (d->kind == ORACLE_KEY_BYTES) {
        for (uint32_t l = 0; l < d->levels; ++l) {
            if (d->level[l].byte_offset >= d->elem_size) return 2;
            cx->off[l] = d->level[l].byte_offset;
            cx->xr[l] = d->level[l].xor_mask;
        }
        return 0;
    }
    if (d->kind > ORACLE_KEY_FLOAT) return 2;
    const uint32_t w = d->key_width;
    if (!(w == 1 || w == 2 || w == 4 || w == 8 || w == 16)) return 2;
    if (d->kind == ORACLE_KEY_FLOAT && !(w == 4 || w == 8)) return 2;
    if ((uint64_t)d->key_offset + w > d->elem_size || d->levels > w) return 2;
    cx->sign_byte = d->key_offset + w - 1;
    for (uint32_t l = 0; l < d->levels; ++l) {
        cx->off[l] = (uint16_t)(d->key_offset + l); /* little-endian byte l of the key word */
        cx->xr[l] = (d->kind != ORACLE_KEY_UINT && l == w - 1) ? 0x80u : 0u;
    }
    return 0;
}

static const otype_vtbl* pick_vtbl(const oracle_key_desc* d) {
    /* key-only primitives with all levels declared: the monomorphised built-in impls */
    if (d->kind != ORACLE_KEY_BYTES && d->key_offset == 0 && d->elem_size == d->key_width &&
        d->levels == d->key_width) {
        switch (d->kind) {
            case ORACLE_KEY_UINT:
                if (d->key_width == 1) return &vtbl_u8;
                if (d->key_width == 2) return &vtbl_u16;
                if (d->key_width == 4) return &vtbl_u32;
                if (d->key_width == 8) return &vtbl_u64;
                break;
            case ORACLE_KEY_SINT:
                if (d->key_width == 1) return &vtbl_i8;
                if (d->key_width == 2) return &vtbl_i16;
                if (d->key_width == 4) return &vtbl_i32;
                if (d->key_width == 8) return &vtbl_i64;
                break;
            case ORACLE_KEY_FLOAT:
                if (d->key_width == 4) return &vtbl_f32;
                if (d->key_width == 8) return &vtbl_f64;
                break;
        }
    }
    switch (d->elem_size) {
        case 1: return &vtbl_r1;
        case 2: return &vtbl_r2;
        case 3: return &vtbl_r3;
        case 4: return &vtbl_r4;
        case 5: return &vtbl_r5;
        case 6: return &vtbl_r6;
        case 7: return &vtbl_r7;
        case 8: return &vtbl_r8;
        case 12: return &vtbl_r12;
        case 16: return &vtbl_r16;
        case 20: return &vtbl_r20;
        case 24: return &vtbl_r24;
        case 32: return &vtbl_r32;
        case 40: return &vtbl_r40;
        case 48: return &vtbl_r48;
        case 64: return &vtbl_r64;
    }
    return NULL;
}

int oracle_hardware_threads(void) {
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int)n : 1;
}

int oracle_get_level(const void* record, const oracle_key_desc* desc, uint32_t level) {
    octx cx;
    if (!record || lower(desc, &cx) != 0 || level >= cx.levels) return -1;
    return gl_desc(&cx, (const uint8_t*)record, (int)level);
}

int oracle_get_level_builtin(const char* type, const void* value, uint32_t level) {
    if (!type || !value) return -1;
#define BI(NAME, CT, FN, LV)                   \
    if (strcmp(type, NAME) == 0) {             \
        CT v;                                  \
        if (level >= LV) return -1;            \
        memcpy(&v, value, sizeof(v));          \
        return FN(&v, (int)level);             \
    }
    BI("u8", uint8_t, gl_u8, 1)
    BI("u16", uint16_t, gl_u16, 2)
    BI("u32", uint32_t, gl_u32, 4)
    BI("u64", uint64_t, gl_u64, 8)
    BI("u128", unsigned __int128, gl_u128, 16)
    BI("usize", uint64_t, gl_u64, 8) /* :68-76, 64-bit target */
    BI("i8", int8_t, gl_i8, 1)
    BI("i16", int16_t, gl_i16, 2)
    BI("i32", int32_t, gl_i32, 4)
    BI("i64", int64_t, gl_i64, 8)
    BI("i128", __int128, gl_i128, 16)
    BI("isize", int64_t, gl_i64, 8) /* :152-160 */
    BI("f32", uint32_t, gl_f32, 4)
    BI("f64", uint64_t, gl_f64, 8)
#undef BI
    return -1;
}

/* src/radix_key_impl.rs:78-85: self[LEVELS - level - 1] */
int oracle_get_level_bytes(const uint8_t* value, uint32_t n, uint32_t level) {
    if (!value || level >= n) return -1;
    return value[n - level - 1];
}

int oracle_get_counts(const void* data, size_t n, const oracle_key_desc* desc, uint32_t level, uint64_t* counts,
                      int* already_sorted, int* first, int* last) {
    octx cx;
    int rc = lower(desc, &cx);
    if (rc) return rc;
    const otype_vtbl* vt = pick_vtbl(desc);
    if (!vt || level >= cx.levels || !counts) return 6;
    size_t c[256];
    int s;
    uint8_t f, l;
    vt->counts(&cx, data, n, (int)level, c, &s, &f, &l);
    for (int k = 0; k < 256; ++k) counts[k] = c[k];
    if (already_sorted) *already_sorted = s;
    if (first) *first = f;
    if (last) *last = l;
    return 0;
}

int oracle_get_tile_counts_sorted(const void* data, size_t n, const oracle_key_desc* desc, size_t tile_size,
                                  uint32_t level, uint64_t* agg_counts) {
    octx cx;
    if (lower(desc, &cx) != 0) return -1;
    const otype_vtbl* vt = pick_vtbl(desc);
    if (!vt || level >= cx.levels || tile_size == 0) return -1;
    size_t c[256];
    int s = vt->tile_counts(&cx, data, n, tile_size, (int)level, c);
    if (agg_counts)
        for (int k = 0; k < 256; ++k) agg_counts[k] = c[k];
    return s;
}

static pthread_mutex_t g_sort_mu = PTHREAD_MUTEX_INITIALIZER; /* the pool is process-global */

int oracle_sort(void* data, size_t n, const oracle_key_desc* desc, int tuner, int algo, int multi_threaded,
                int threads) {
    octx cx;
    int rc = lower(desc, &cx);
    if (rc) return rc; /* LEVELS == 0 is rejected before the len <= 1 early-out, as radix_sort_builder.rs:22 */
    if (n <= 1) return 0; /* radix_sort_builder.rs:151 */
    const otype_vtbl* vt = pick_vtbl(desc);
    if (!vt || !data) return 6;
    if (threads <= 0) threads = oracle_hardware_threads();
    cx.tuner = tuner;
    cx.algo = algo;
    cx.multi_threaded = multi_threaded != 0;
    cx.mt_feature = 1;
    pthread_mutex_lock(&g_sort_mu);
    pool_set_threads(threads);
    vt->sort_top(&cx, data, n);
    pthread_mutex_unlock(&g_sort_mu);
    return 0;
}

int oracle_out_of_place_pass(const void* src, void* dst, size_t n, const oracle_key_desc* desc, uint32_t level,
                             int lr) {
    octx cx;
    int rc = lower(desc, &cx);
    if (rc) return rc;
    const otype_vtbl* vt = pick_vtbl(desc);
    if (!vt || level >= cx.levels) return 6;
    vt->pass(&cx, src, dst, n, (int)level, lr);
    return 0;
}

int oracle_is_sorted(const void* data, size_t n, const oracle_key_desc* desc) {
    octx cx;
    if (lower(desc, &cx) != 0) return -1;
    const uint8_t* p = (const uint8_t*)data;
    for (size_t i = 1; i < n; ++i) {
        const uint8_t* a = p + (i - 1) * cx.elem_size;
        const uint8_t* b = p + i * cx.elem_size;
        for (int l = (int)cx.levels - 1; l >= 0; --l) {
            uint8_t x = gl_desc(&cx, a, l), y = gl_desc(&cx, b, l);
            if (x < y) break;
            if (x > y) return 0;
        }
    }
    return 1;
}

const char* oracle_version(void) { return "rdst-oracle 0.1 (C restatement of rdst 0.20.14+; parity pinned by golden vectors)"; }
