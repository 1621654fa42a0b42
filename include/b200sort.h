/*
 * b200sort.h — C ABI of the B200-native radix sort engine that stands in for
 * the sort path of the Rust crate nessex/rdst.
 *
 * This is the drop-in boundary.  The reference is a pure-Rust library with no
 * FFI of its own; the functions below are what an `extern "C"` block inside the
 * reference's `RadixSortBuilder::sort` would bind (see INTEGRATION.md for the
 * Rust stub).  Each entry point cites the reference interface it replaces
 * (paths relative to the reference checkout).
 *
 * Conventions (mirroring src/radix_sort_builder.rs:149-157):
 *   - the caller owns the data; the sort is in place; scratch is owned by the
 *     library (a per-device cached pool) unless the caller passes its own;
 *   - calls block until the result is in the caller's buffer;
 *   - the reference API is infallible, so every failure is a non-zero return
 *     code that the binding turns into a panic; there is NO CPU fallback;
 *   - thread-safe: concurrent calls serialise on a per-device context.
 */
#ifndef B200SORT_H
#define B200SORT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_MAX_LEVELS 64

/* Return codes. 0 = success. */
enum {
    B200_OK = 0,
    B200_ERR_LEVELS_ZERO = 1,   /* RadixKey::LEVELS == 0 — the reference panics here
                                   (src/radix_sort_builder.rs:22) */
    B200_ERR_BAD_DESC = 2,      /* descriptor inconsistent (offset outside the record, bad width …) */
    B200_ERR_CUDA = 3,          /* a CUDA runtime call failed; see b200_last_cuda_error() */
    B200_ERR_NO_DEVICE = 4,     /* no usable sm_100 device */
    B200_ERR_ALLOC = 5,         /* scratch / staging allocation failed */
    B200_ERR_BAD_ARG = 6,       /* null pointer, unknown variant id … */
    B200_ERR_LEVEL_RANGE = 7,   /* level >= LEVELS (src/radix_key.rs:18 debug_assert) */
    /* 8 is taken by the Python driver (B200_ERR_NCCL, rdst_b200/_abi.py) */
    B200_ERR_TOO_LARGE = 9      /* b200_sort_device_static: more records than the static scratch holds */
};

/*
 * How RadixKey::get_level (src/radix_key.rs:1-5) is lowered for the device.
 * Level 0 is the LEAST significant byte (README.md:37-38).
 *
 *  UINT  — src/radix_key_impl.rs:3-76   (self >> 8*level) as u8
 *  SINT  — src/radix_key_impl.rs:87-160 ((self ^ MIN) >> 8*level) as u8
 *  FLOAT — src/radix_key_impl.rs:162-185 IEEE total-order map, then as SINT
 *  BYTES — src/radix_key_impl.rs:78-85 and user keys (examples/impl_radix_key.rs):
 *          level l reads record byte level[l].byte_offset, xor level[l].xor_mask
 */
typedef enum {
    B200_KEY_UINT = 0,
    B200_KEY_SINT = 1,
    B200_KEY_FLOAT = 2,
    B200_KEY_BYTES = 3
} b200_key_kind;

typedef struct {
    uint16_t byte_offset; /* byte inside the record that holds this level's digit */
    uint8_t xor_mask;     /* applied after the load */
    uint8_t _pad;
} b200_level;

typedef struct {
    uint32_t elem_size;   /* size_of::<T>() — record stride, payload included */
    uint32_t elem_align;  /* align_of::<T>() (informational; 1,2,4,8,16) */
    uint32_t levels;      /* T::LEVELS; may be smaller than key_width (partial keys, src/lib.rs:58-76) */
    uint32_t kind;        /* b200_key_kind */
    uint32_t key_offset;  /* byte offset of the little-endian key word in the record (UINT/SINT/FLOAT) */
    uint32_t key_width;   /* bytes of the key word: 1,2,4,8,16 (FLOAT: 4 or 8) */
    b200_level level[B200_MAX_LEVELS]; /* BYTES kind only: one entry per level, level 0 first */
} b200_key_desc;

/*
 * Options — what RadixSortBuilder carries (src/radix_sort_builder.rs:8-12)
 * re-expressed for the device engine.
 */
/* What a Tuner's answer means on the device (src/tuner.rs:10-35: Algorithm; src/sorter.rs:67-76: where it is
 * asked).  rdst's nine algorithms are CPU strategies; on the GPU they collapse to the choices below, each
 * visible in b200_stats. */
typedef enum {
    B200_HINT_AUTO = 0,        /* the planner decides (Scanning / Recombinating / MtOop / Comparative and no tuner:
                                  hybrid MSD route when the input qualifies, LSD passes otherwise) */
    B200_HINT_LSD = 1,         /* one stable K2 pass per active level, nothing else (Lsb / LrLsb / MtLsb: rdst's
                                  stable LSB family, src/sorts/lsb_sort.rs:5-7): b200_stats.hybrid == 0 */
    B200_HINT_LOW_SCRATCH = 2  /* Ska / Regions and with_low_mem_tuner() (rdst's in-place family,
                                  src/tuners/low_memory_tuner.rs:13-43): LSD passes with the small-tile variant and
                                  the library's cached scratch pool (n * elem_size + look-back rows) is released
                                  before the call returns.  PEAK scratch is still one copy of the array: the
                                  engine has no in-place pass. */
} b200_route_hint;

typedef struct {
    uint32_t stable;   /* 1 = rdst's stable order (Lsb/LrLsb/MtLsb, src/sorts/lsb_sort.rs:5-7): records with a
                          payload keep their input order among equal keys.  0 = any sorted permutation is legal
                          (radix_sort_unstable, src/radix_sort.rs:21-45).  Records with a payload are sorted
                          stably either way; records that are nothing but their key take the unstable kernels
                          (b200_stats.unstable) whose result is bit-identical. */
    int32_t variant;   /* < 0 = let the planner pick the tile / occupancy variant of the pass kernel;
                          else index < b200_variant_count() (forces LSD passes) */
    int32_t device;    /* CUDA device ordinal; -1 = current device */
    void* stream;      /* cudaStream_t to order the work on; NULL = the library's own (blocking) stream,
                          which is ordered against the legacy default stream */
    uint32_t route;    /* b200_route_hint */
    uint32_t _reserved;
} b200_sort_opts;

typedef struct {
    uint32_t levels_run;       /* passes actually executed */
    uint32_t levels_skipped;   /* single-bucket levels (src/sorts/lsb_sort.rs:20-23) */
    uint32_t variant;          /* tile/occupancy variant used by the pass kernel */
    uint32_t copy_back;        /* 1 if a final device copy into the caller's buffer was needed
                                  (src/sorts/lsb_sort.rs:117-126) */
    uint64_t algorithmic_bytes;/* n*elem_size*(1 + 2*levels_run) (+ n*elem_size if copy_back) */
    float ms_hist;             /* K1 + scan, CUDA events */
    float ms_passes;           /* all K2 passes (+ copy back), CUDA events */
    float ms_h2d, ms_d2h;      /* host entry point only */
    float ms_total;            /* first launch to last completion */
    uint64_t tiles_per_pass;
    uint32_t already_sorted;   /* 1 = K0 found the input in key order and nothing else ran (rdst's
                                  already_sorted shortcut, src/sort_utils.rs:125-136, src/sorter.rs:59-65) */
    uint32_t kernel_launches;  /* kernels of this library launched by the call */
    uint32_t hybrid;           /* 1 = MSD x2 + shared-memory segment sort (K3) ran instead of LSD passes;
                                  2 = one MSD pass + K3 on its 256 buckets (mid-size inputs) */
    uint32_t slow_segments;    /* hybrid: segments whose fix-up fell back to a full local LSD */
    uint32_t local_levels;     /* hybrid: levels K3 sorted on chip (levels_run - 2) */
    float ms_joint;            /* hybrid: joint histogram of the two MSD levels + its read-back */
    float ms_local;            /* hybrid: K3 (shared-memory segment sort), CUDA events */
    uint32_t unstable;         /* hybrid, records that are nothing but their key (every byte a level): bit 0 = the
                                  unstable flavour ran (K3-U segment sort; equal keys are identical records, so the
                                  result is bit-identical to the stable engine's — rdst's default entry point is
                                  radix_sort_unstable, src/radix_sort.rs:21-45), bit 1 / bit 2 = MSD pass 1 / 2 ran
                                  the atomic-ranking kernel K2-U (flat digits only) */
    uint32_t heavy_segments;   /* key-only hybrid route: segments too large for one SM (a heavily duplicated key makes its
                                  whole segment long) that were split three ways around a pivot key instead of sorted */
    uint32_t heavy_rounds;     /* pivot-split rounds that took (0 = no oversized segment) */
} b200_stats;

/*
 * Sort `n` records that live in HOST memory, in place.
 * Replaces: RadixSortBuilder::sort (src/radix_sort_builder.rs:149-157) →
 *           Sorter::route_top_level (src/sorter.rs:105-119), i.e. everything
 *           behind RadixSort::radix_sort_unstable (src/radix_sort.rs:21-45).
 * n <= 1 returns immediately (src/radix_sort_builder.rs:151).
 */
int b200_sort_host(void* host_ptr, size_t n, const b200_key_desc* desc,
                   const b200_sort_opts* opts, b200_stats* stats);

/*
 * Sort `n` records that already live in DEVICE memory, in place.
 * `dev_tmp` is an optional caller-owned scratch of n*elem_size bytes
 * (the `tmp_bucket` of src/sorts/lsb_sort.rs:53); NULL = library pool.
 * Replaces: Sorter::lsb_sort_adapter / mt_lsb_sort_adapter and the counting
 *           utilities they call (src/sorts/lsb_sort.rs:38-128,
 *           src/sorts/mt_lsb_sort.rs:136-195, src/sort_utils.rs:33-249).
 */
int b200_sort_device(void* dev_ptr, void* dev_tmp, size_t n, const b200_key_desc* desc,
                     const b200_sort_opts* opts, b200_stats* stats);

/*
 * b200_sort_device for a caller that already knows which bits of the key image (level l = bits 8l .. 8l+7) can differ
 * between two records; a superset of the truth is fine, a missing bit gives a wrong order.  The library then does not
 * count the records first (no K0, no K1): the digits of the MSD route are planned from `vary_bits`.  Inputs the MSD route
 * does not take (records with a payload below 256 Mi records, small arrays ...) are sorted as by b200_sort_device.
 * Used by the sharded sort, whose gathered histograms describe every rank's range (rdst_b200/dist.py).
 * No counterpart in the reference; SURVEY §8(e).
 */
int b200_sort_device_hinted(void* dev_ptr, void* dev_tmp, size_t n, const b200_key_desc* desc, uint64_t vary_bits,
                            const b200_sort_opts* opts, b200_stats* stats);

/*
 * The local sort of the sharded sort when the exchange has already split the records on one level: `dev_ptr` holds n
 * records partitioned on `level` — bucket sizes `counts[256]` (host), buckets in digit order, every level above
 * `level` constant — as b200_partition_peers_device with 256 parts delivers them.  K0, K1 and the first MSD pass are
 * skipped (joint histogram, MSD pass 2 dev_ptr -> dev_tmp, segment sort in dev_tmp); the sorted records are at
 * *result_out, which is dev_ptr or dev_tmp (both n * elem_size bytes, both clobbered).  Records with a payload,
 * level < 2, or segments that do not fit one SM take the general sort from wherever the records are: same result.
 * No counterpart in the reference (it has no distributed path); SURVEY §8(e).
 */
int b200_sort_device_presplit(void* dev_ptr, void* dev_tmp, size_t n, const b200_key_desc* desc, uint32_t level,
                              const uint64_t* counts, const b200_sort_opts* opts, b200_stats* stats, void** result_out);

/*
 * Many independent sorts in one call: every segment [seg_off[i], seg_off[i + 1]) of the device array `dev_ptr`
 * (n records) is sorted on its own, in place.  `seg_off` is HOST memory, nseg + 1 ascending record offsets, the last
 * one <= n; n < 2^32.  Segments of <= 1 record are left alone (src/radix_sort_builder.rs:151).
 * Replaces: the tail of the MSB-first recursion — Sorter::director handing every bucket of a split to
 *           handle_chunk as its own task (src/sorter.rs:121-157), where small buckets end in lsb_sort_adapter
 *           (src/sorts/lsb_sort.rs:38-128) or, below 128 records, comparative_sort
 *           (src/sorts/comparative_sort.rs:29-119).
 * Segments of up to 16 384 records of 4 / 8 bytes (LEVELS <= 8, contiguous key) are sorted in shared memory by the
 * CTAs of one launch (stable; records that are nothing but their key take the unstable kernel, same result);
 * larger segments and other record shapes take the whole-array sort one after the other.
 * stats: tiles_per_pass = segments sorted on chip, heavy_segments = segments that took the whole-array sort,
 * slow_segments = on-chip segments finished by the stable kernel, ms_local = the on-chip launch(es).
 */
int b200_sort_segments_device(void* dev_ptr, size_t n, const uint64_t* seg_off, size_t nseg,
                              const b200_key_desc* desc, const b200_sort_opts* opts, b200_stats* stats);

/*
 * The device sort for callers that cannot allocate: ALLOCATES NOTHING — no cudaMalloc, cudaFree,
 * cudaMallocHost or allocating library routine, not on the first call either.
 * No counterpart in the reference; added (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Its scratch (ping-pong buffer, histograms, tile look-back table, tickets) is a file-scope
 * `static __device__` array of ~20.3 MiB that every GPU loading the library holds until exit, hence the limits:
 *   n <= B200_STATIC_MAX_RECORDS and n * elem_size <= B200_STATIC_MAX_BYTES, else B200_ERR_TOO_LARGE
 *   (b200_last_cuda_error() names the sizes).
 * Sorts `n` records in DEVICE memory in place, result bit-identical to b200_sort_device's.  n <= 16 384
 * records of 4 / 8 bytes: one launch of the shared-memory segment sort; otherwise K1 + one stable pass per
 * declared level (no histogram read-back, so no level is skipped).
 * opts may be NULL; opts->stream = the stream to enqueue on (NULL = the legacy default stream: this entry
 * point creates no stream of its own), opts->device as elsewhere, opts->variant as elsewhere (>= 0 forces
 * the passes), opts->route must be B200_HINT_AUTO.  The call only ENQUEUES and returns without
 * synchronising: the result is there when the stream gets there.
 * Every call on a GPU shares that GPU's array, which keeps the last call's contents; what must start at zero
 * is cleared by each call on its stream.  Calls that overlap on the same GPU (different streams, different
 * threads) RACE on the array: serialise them.
 */
#define B200_STATIC_MAX_RECORDS (2u << 20) /* 2 Mi records */
#define B200_STATIC_MAX_BYTES (16u << 20)  /* 16 MiB of records */
int b200_sort_device_static(void* dev_ptr, size_t n, const b200_key_desc* desc, const b200_sort_opts* opts);

/*
 * K1 on its own: all LEVELS 256-bin histograms of device data in one read.
 * `hist_out` is HOST memory, levels*256 uint64 counts, level 0 first.
 * Replaces: get_counts / get_counts_with_ends / get_tile_counts +
 *           aggregate_tile_counts (src/sort_utils.rs:108-249).
 */
int b200_histogram_device(const void* dev_ptr, size_t n, const b200_key_desc* desc,
                          const b200_sort_opts* opts, uint64_t* hist_out);

/*
 * One stable counting-sort pass over one level, device to device
 * (dev_src and dev_dst must not overlap).  `counts` is the HOST 256-bin
 * histogram of that level (as returned by b200_histogram_device).
 * Replaces: out_of_place_sort / lr_out_of_place_sort and the per-level body of
 *           mt_lsb_sort (src/sorts/out_of_place_sort.rs:51-108,201-275,
 *           src/sorts/mt_lsb_sort.rs:40-133).
 */
int b200_pass_device(const void* dev_src, void* dev_dst, size_t n, const b200_key_desc* desc,
                     uint32_t level, const uint64_t* counts, const b200_sort_opts* opts);

/*
 * Multi-GPU helpers for the range-partitioned sort (one process per GPU; the
 * exchange itself is the caller's collective).  Given this rank's device data
 * and P-1 splitter digits on `level`, partition the records stably into P
 * contiguous destination ranges (dest g = digits in [split[g-1], split[g]))
 * and return the per-destination record counts in host memory.
 * This is b200_pass_device with a 256 -> P bucket remap.
 */
int b200_partition_device(const void* dev_src, void* dev_dst, size_t n, const b200_key_desc* desc,
                          uint32_t level, const uint64_t* counts, const uint32_t* split, uint32_t parts,
                          uint64_t* part_counts_out, const b200_sort_opts* opts);

/*
 * Fused partition + exchange: as b200_partition_device, but part g is written to its own
 * destination pointer dst_ptrs[g] (HOST array of `parts` device pointers, each already offset to
 * where this rank's records start in that destination).  The pointers may be peer GPUs' buffers
 * mapped through CUDA IPC: the partition kernel then stores its staged runs straight over
 * NVLink, which removes the separate all-to-all and one local read + write.
 * Records must be a multiple of 4 bytes.  The caller orders the kernels of different ranks
 * (a stream-ordered collective before and after).
 */
int b200_partition_peers_device(const void* dev_src, void* const* dst_ptrs, size_t n, const b200_key_desc* desc,
                                uint32_t level, const uint64_t* counts, const uint32_t* split, uint32_t parts,
                                uint64_t* part_counts_out, const b200_sort_opts* opts);

/*
 * Full-resolution splitters for skewed keys (sample-sort style).  When one
 * digit of the partition level holds far more than N/P records, digit boundaries cannot balance
 * the ranks.  The caller then samples 64-bit key prefixes (the record's top min(LEVELS, 8) levels
 * as one order-preserving integer, level LEVELS-1 most significant: the bytes
 * RadixKey::get_level returns, reference src/radix_key.rs:1-5), picks P-1 splitter prefixes from
 * the gathered samples, counts every rank's records against them, and partitions with
 *   part(record) = #{ g : prefix > splitters[g]  or  (prefix == splitters[g] and incl_mask bit g) }.
 * incl_mask differs per source rank, which splits a key value heavier than one part between
 * adjacent parts by source rank and keeps the exchange stable.
 * Supported for 4-byte-multiple records up to 16 bytes with a word-aligned UINT / SINT / FLOAT
 * key (b200_key_prefix_supported); other keys keep the digit splitters.
 */
int b200_key_prefix_supported(const b200_key_desc* desc);
/* prefixes_out[i] (HOST) = prefix of record i * stride, i < count; (count - 1) * stride < n */
int b200_key_prefix_sample_device(const void* dev_ptr, size_t n, const b200_key_desc* desc, size_t stride,
                                  size_t count, uint64_t* prefixes_out, const b200_sort_opts* opts);
/* less_out[g] / equal_out[g] (HOST) = records whose prefix is < / == splitters[g]; splitters ascending, nsplit <= 15 */
int b200_splitter_count_device(const void* dev_ptr, size_t n, const b200_key_desc* desc, const uint64_t* splitters,
                               uint32_t nsplit, uint64_t* less_out, uint64_t* equal_out, const b200_sort_opts* opts);
/* Stable partition into nsplit + 1 parts by the rule above.  part_counts (HOST, nsplit + 1 entries,
 * sum n) are this rank's exact per-part counts, derived from b200_splitter_count_device.  Either
 * dev_dst (parts contiguous in one buffer) or dst_ptrs (one device pointer per part, possibly peer
 * GPUs' buffers as in b200_partition_peers_device) is given. */
int b200_partition_splitters_device(const void* dev_src, void* dev_dst, void* const* dst_ptrs, size_t n,
                                    const b200_key_desc* desc, const uint64_t* splitters, uint32_t nsplit,
                                    uint32_t incl_mask, const uint64_t* part_counts, const b200_sort_opts* opts);

/* CUDA IPC helpers for the one-process-per-GPU exchange: allocate a shareable device buffer and
 * its 64-byte handle; open / close a peer's handle; free. */
#define B200_IPC_HANDLE_BYTES 64
int b200_ipc_alloc(void** dev_ptr, size_t bytes, void* handle_out);
int b200_ipc_open(const void* handle, void** dev_ptr);
int b200_ipc_close(void* dev_ptr);
int b200_ipc_free(void* dev_ptr);

/*
 * The sharded sort behind the C ABI: ONE process drives `ngpus` GPUs of one box (one host thread and one stream
 * per GPU, peer access between the devices), SURVEY.md §8e / BASELINE.json configs[4].  The reference has no
 * counterpart (RadixSortBuilder::sort takes one `&mut [T]`, src/radix_sort_builder.rs:149-157); this is the entry a
 * Rust binding calls for inputs larger than one GPU or for 8-GPU mode.  Steps: K1 per shard, histograms merged on
 * the host, partition level = most significant non-trivial level, P-1 splitters (digit boundaries, or sample-sort
 * splitters on the 64-bit key prefix when one digit outweighs a rank's share), fused partition + exchange (every
 * GPU's K2 range-partition pass stores its runs straight into the owners' receive buffers over NVLink), then an
 * independent local sort per GPU.  Stable.  Records of 4, 8, 12 or 16 bytes.
 * `devices` = NULL means devices 0 .. ngpus-1.
 */
#define B200_MAX_GPUS 16
typedef struct {
    uint32_t ngpus;
    uint32_t level;          /* partition level (0xffffffff: every key equal, nothing moved) */
    uint32_t splitter_mode;  /* 0 = digit boundaries, 1 = sample-sort splitters on the key prefix */
    uint32_t kernel_launches;
    double imbalance;        /* max records any GPU received / the even share */
    uint64_t max_sent_bytes; /* largest number of bytes one GPU sent to its peers */
    uint64_t n_out[B200_MAX_GPUS];
    float ms_plan;           /* K1 + host planning (+ sampling / counting in splitter mode), host clock */
    float ms_exchange;       /* fused partition + exchange, host clock between the two fences */
    float ms_local_sort;
    float ms_h2d, ms_d2h;    /* host entry point only */
    float ms_total;
    b200_stats local0;       /* GPU 0's local sort */
} b200_multi_stats;

/* Shard g = n_in[g] records in DEVICE memory of devices[g] at dev_ptrs[g] (read only).  On return GPU g holds the
 * g-th key range, sorted: n_out[g] records at out_ptrs[g], a library-owned buffer on devices[g] that stays valid
 * until the next multi-GPU call or b200_multi_release().  Concatenating the outputs in GPU order gives the sorted
 * array. */
int b200_sort_multi_device(void* const* dev_ptrs, const size_t* n_in, int ngpus, const int* devices,
                           const b200_key_desc* desc, const b200_sort_opts* opts, void** out_ptrs, size_t* n_out,
                           b200_multi_stats* stats);
/* `n` records in HOST memory, sorted in place over `ngpus` GPUs: the array is cut into ngpus contiguous shards, every
 * GPU pulls its shard over its own PCIe link, and GPU g's sorted range is written back behind the ranges of the
 * lower GPUs.  The multi-GPU form of b200_sort_host. */
int b200_sort_multi_host(void* host_ptr, size_t n, int ngpus, const int* devices, const b200_key_desc* desc,
                         const b200_sort_opts* opts, b200_multi_stats* stats);
int b200_multi_release(void);

/* The host-side exchange planner (no GPU), shared by b200_sort_multi_* and the one-process-per-GPU driver
 * (rdst_b200/dist.py).  all_hist = [ranks][levels][256] K1 histograms.  level_out = -1: every key equal.
 * split_out[ranks - 1] digit boundaries (part g = digits in [split[g-1], split[g])), send_matrix_out[ranks][ranks]
 * (m[r][g] = records rank r sends to rank g). */
int b200_multi_plan_digits(const uint64_t* all_hist, uint32_t ranks, uint32_t levels, int32_t* level_out,
                           uint32_t* split_out, int64_t* send_matrix_out, double* imbalance_out);
/* parts - 1 splitter prefixes = weighted quantiles of the gathered key-prefix samples */
int b200_multi_choose_sample_splitters(const uint64_t* samples, const double* weights, size_t count, uint32_t parts,
                                       uint64_t* splitters_out);
/* less / equal = [ranks][parts - 1] counts from b200_splitter_count_device.  tie_thresholds_out[parts - 1]: boundary b
 * sends a record LEFT when prefix < s_b, or prefix == s_b and source rank < t_b; send_matrix_out[ranks][parts];
 * incl_masks_out[ranks] = the incl_mask each rank passes to b200_partition_splitters_device (may be NULL). */
int b200_multi_plan_splitters(const int64_t* less, const int64_t* equal, const int64_t* n_per_rank,
                              const uint64_t* splitters, uint32_t ranks, uint32_t parts, int exact_ties,
                              int64_t* tie_thresholds_out, int64_t* send_matrix_out, uint32_t* incl_masks_out);

/* Host <-> device copies as the host entry points do them: plain DMA for pinned memory; pageable memory (what a
 * Rust Vec is) goes through pinned bounce buffers filled by several host threads while the previous chunk is on
 * the wire.  Blocking. */
int b200_copy_h2d(void* dev_dst, const void* host_src, size_t bytes, int device);
int b200_copy_d2h(void* host_dst, const void* dev_src, size_t bytes, int device);

/*
 * RadixKey::get_level evaluated on the host through the descriptor — lets the
 * binding validate a user key against its lowering before launching.
 * Returns the digit (0..255) or a negative error code.
 */
int b200_get_level(const void* record, const b200_key_desc* desc, uint32_t level);

/* Pinned host memory for the host entry point's fast path (plain DMA, no staging). */
int b200_host_alloc(void** out, size_t bytes);
int b200_host_free(void* p);
/* Release the cached per-device scratch pool. */
int b200_release_scratch(int device);

/*
 * Planner introspection (host only, no GPU): the digits the hybrid route would sort on for an
 * input with these per-level histograms (hist[LEVELS][256], what b200_histogram_device returns).
 * The route is the GPU form of the reference's MSB-first recursion with an LSB finish
 * (src/sorter.rs:24-171, src/tuners/standard_tuner.rs:13-63): two global MSD passes on
 * digit[0] (most significant) and digit[1], then every segment is sorted on chip on the local
 * windows (least significant first).  Digits are bit windows [lo, lo + width) of the 64-bit
 * order-preserving key image (level l = bits 8l .. 8l+7), placed on the bits that vary.
 * eligible = 0: the key shape or the data rule the route out (the LSD passes run).
 */
typedef struct b200_window_plan {
    uint32_t eligible;
    uint32_t aligned;        /* both MSD digits are whole levels (no window histogram pass) */
    uint64_t vary;           /* bits of the key image that differ between some two records */
    uint32_t digit_lo[2], digit_width[2];
    uint32_t nlocal;
    uint32_t local_lo[8], local_width[8];
} b200_window_plan;
int b200_hybrid_plan(const b200_key_desc* desc, const uint64_t* hist, size_t n, b200_window_plan* out);

/*
 * Planner introspection (host only, no GPU): the route b200_sort_device starts on for `n` records
 * with these per-level histograms, given 16-byte aligned buffers and variant = auto — the GPU
 * meaning of Tuner::pick_algorithm (src/tuner.rs:33-35; the reference's own rules:
 * src/tuners/standard_tuner.rs:13-63, size cut-offs src/sorter.rs:33-45).  `hist` may be NULL when
 * the answer does not need it (n <= 1, the small-input path).  The already-sorted early-out (K0) is
 * decided on the data, not here; the hybrid route may still fall back to LSD passes after its joint
 * histogram if a segment does not fit one SM.
 */
typedef enum {
    B200_ROUTE_LSD = 0,     /* one K2 pass per active level */
    B200_ROUTE_HYBRID = 1,  /* two MSD passes + K3 (b200_hybrid_plan gives the digits) */
    B200_ROUTE_MSD1 = 2,    /* one MSD pass on the top active level + K3 on its 256 buckets */
    B200_ROUTE_SMALL = 3,   /* one K3 launch, the whole array as one segment */
    B200_ROUTE_TRIVIAL = 4  /* n <= 1: nothing runs (src/radix_sort_builder.rs:151) */
} b200_route_kind;
typedef struct b200_route {
    uint32_t route;          /* b200_route_kind */
    uint32_t active_levels;  /* levels whose histogram has more than one non-empty bucket */
    uint32_t levels_run;     /* what b200_stats.levels_run reports for this route */
    uint32_t msd_level;      /* MSD1: the level of the MSD pass */
    uint64_t max_bucket;     /* MSD1: the largest bucket of that level (<= 16384) */
} b200_route;
int b200_route_plan(const b200_key_desc* desc, const uint64_t* hist, size_t n, b200_route* out);

/* Tuner introspection: the Tuner now only picks a tile/occupancy variant. */
int b200_variant_count(void);
const char* b200_variant_name(int variant);

const char* b200_strerror(int code);
const char* b200_last_cuda_error(void);
const char* b200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200SORT_H */
