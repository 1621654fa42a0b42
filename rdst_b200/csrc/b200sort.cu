// b200sort.cu — the C ABI of the B200 radix sort engine (include/b200sort.h)
// and the host pass planner behind it.
//
// The planner is what is left of rdst's router and LSB driver on a GPU
//   Sorter::handle_chunk / route*       (reference src/sorter.rs:24-171)
//   Sorter::lsb_sort_adapter            (reference src/sorts/lsb_sort.rs:38-128)
//   mt_lsb_sort_adapter                 (reference src/sorts/mt_lsb_sort.rs:136-195)
// : count every level once (K1), skip levels whose histogram is one bucket
// (lsb_sort.rs:20-23, mt_lsb_sort.rs:170-174, ska_sort.rs:52-54), run one stable
// pass (K2) per remaining level from level 0 upwards while ping-ponging between
// the caller's buffer and a scratch buffer (lsb_sort.rs:86-98), and copy back if
// the result ended in scratch (lsb_sort.rs:117-126).  No recursion, no CPU path.
#include "../../include/b200sort.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "hist.cuh"
#include "splitters.cuh"
#include "local_sort.cuh"
#include "sorted_check.cuh"
#include "launch.h"
#include "onesweep.cuh"
#include "variants.h"
#include "unstable.h"
#include "local_count.cuh"
#include "heavy.cuh"
#include "tiny_sort.cuh"

using namespace b200;

namespace {

thread_local std::string g_last_cuda_error;

#define CUDA_TRY(expr)                                                                          \
    do {                                                                                        \
        cudaError_t _e = (expr);                                                                \
        if (_e != cudaSuccess) {                                                                \
            g_last_cuda_error = std::string(#expr) + ": " + cudaGetErrorString(_e);             \
            cudaGetLastError();                                                                 \
            return B200_ERR_CUDA;                                                               \
        }                                                                                       \
    } while (0)

// ---------------------------------------------------------------------------
// Descriptor lowering: RadixKey::get_level -> (byte offset, xor mask) per level.
// ---------------------------------------------------------------------------
struct Lowered {
    uint32_t elem_size = 0, levels = 0;
    bool is_float = false;
    uint32_t sign_byte = 0;
    LevelLowering lv[MAX_LEVELS];
    bool contiguous = false;  // level l is byte key_offset + l (word kinds)
    uint32_t key_offset = 0, key_width = 0;
};

int lower_desc(const b200_key_desc* d, Lowered* out) {
    if (!d) return B200_ERR_BAD_ARG;
    if (d->levels == 0) return B200_ERR_LEVELS_ZERO;
    if (d->levels > B200_MAX_LEVELS) return B200_ERR_BAD_DESC;
    if (d->elem_size == 0 || d->elem_size > 65535) return B200_ERR_BAD_DESC;
    Lowered lo;
    lo.elem_size = d->elem_size;
    lo.levels = d->levels;
    switch (d->kind) {
        case B200_KEY_UINT:
        case B200_KEY_SINT:
        case B200_KEY_FLOAT: {
            const uint32_t w = d->key_width;
            if (!(w == 1 || w == 2 || w == 4 || w == 8 || w == 16)) return B200_ERR_BAD_DESC;
            if (d->kind == B200_KEY_FLOAT && !(w == 4 || w == 8)) return B200_ERR_BAD_DESC;
            if (uint64_t(d->key_offset) + w > d->elem_size) return B200_ERR_BAD_DESC;
            if (d->levels > w) return B200_ERR_BAD_DESC;
            lo.contiguous = true;
            lo.key_offset = d->key_offset;
            lo.key_width = w;
            lo.is_float = d->kind == B200_KEY_FLOAT;
            lo.sign_byte = d->key_offset + w - 1;
            for (uint32_t l = 0; l < d->levels; ++l) {
                lo.lv[l].byte_offset = d->key_offset + l;
                lo.lv[l].xor_mask = (d->kind != B200_KEY_UINT && l == w - 1) ? 0x80u : 0u;
            }
            break;
        }
        case B200_KEY_BYTES: {
            for (uint32_t l = 0; l < d->levels; ++l) {
                if (d->level[l].byte_offset >= d->elem_size) return B200_ERR_BAD_DESC;
                lo.lv[l].byte_offset = d->level[l].byte_offset;
                lo.lv[l].xor_mask = d->level[l].xor_mask;
            }
            break;
        }
        default:
            return B200_ERR_BAD_DESC;
    }
    *out = lo;
    return B200_OK;
}

uint32_t host_digit(const uint8_t* rec, const Lowered& lo, uint32_t level) {
    uint32_t b = rec[lo.lv[level].byte_offset];
    if (lo.is_float && (rec[lo.sign_byte] & 0x80u)) return b ^ 0xffu;
    return b ^ lo.lv[level].xor_mask;
}

DigitParams digit_params(const Lowered& lo, uint32_t level) {
    DigitParams dp;
    dp.word = lo.lv[level].byte_offset / 4;
    dp.shift = (lo.lv[level].byte_offset % 4) * 8;
    dp.xor_mask = lo.lv[level].xor_mask;
    dp.sign_word = lo.sign_byte / 4;
    dp.sign_shift = (lo.sign_byte % 4) * 8 + 7;
    dp.word2 = dp.word;
    dp.mask = 0xffu;
    return dp;
}

// ---------------------------------------------------------------------------
// Per-device context: stream, events, cached scratch.
// ---------------------------------------------------------------------------
struct DeviceCtx {
    int dev = -1;
    bool ready = false;
    std::mutex mu;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[12] = {};
    int sm_count = 0;
    // small device block: hist [64][256] u64 | base [64][256] u64 | tickets [64] u32 | remap [256] u8 | part base
    uint8_t* d_small = nullptr;
    uint64_t* h_hist = nullptr;  // pinned [64][256]
    uint32_t* h_flag = nullptr;  // pinned: K0's out-of-order flag
    void* d_tmp = nullptr;
    size_t tmp_bytes = 0;
    uint64_t* d_lookback = nullptr;
    size_t lookback_bytes = 0;
    uint8_t* d_sample = nullptr;  // key-prefix samples (splitter choice of the sharded sort)
    size_t sample_bytes = 0;
    // hybrid path: joint histogram [256][256] u64 | segment offsets [65537] u64 | tickets [512] u32
    uint8_t* d_hyb = nullptr;
    size_t hyb_bytes = 0;
    uint64_t* h_joint = nullptr;  // pinned [65537]
    cudaStream_t aux[4] = {};
    cudaEvent_t aux_ev[5] = {};
    void* d_stage = nullptr;  // device copy of host data for b200_sort_host
    size_t stage_bytes = 0;
    void* h_bounce[4] = {};  // pinned bounce buffers for pageable host memory
    size_t bounce_bytes = 0;
    cudaEvent_t bounce_ev[4] = {};
    uint32_t* d_slow = nullptr;  // hybrid key-only route: segments K3-U left to the stable K3 ([0] = count)
    size_t slow_bytes = 0;
    uint8_t* d_heavy = nullptr;  // oversized segments (heavy.cuh): range tables, pivots, tile counts and bases, lists
    size_t heavy_bytes = 0;
    uint8_t* d_batch = nullptr;  // b200_sort_segments_device: {first record, records} of the segments | slow list
    size_t batch_bytes = 0;
};

constexpr size_t SMALL_HIST_OFF = 0;
constexpr size_t SMALL_BASE_OFF = size_t(MAX_LEVELS) * RADIX * 8;
constexpr size_t SMALL_TICKET_OFF = SMALL_BASE_OFF + size_t(MAX_LEVELS) * RADIX * 8;
constexpr size_t SMALL_REMAP_OFF = SMALL_TICKET_OFF + 1024;
constexpr size_t SMALL_PBASE_OFF = SMALL_REMAP_OFF + 256;
constexpr size_t SMALL_DTAB_OFF = SMALL_PBASE_OFF + RADIX * 8;
constexpr size_t SMALL_SPLIT_OFF = SMALL_DTAB_OFF + RADIX * 8;  // less[15] | equal[15] (u64), 256 bytes
constexpr size_t SMALL_FLAG_OFF = SMALL_SPLIT_OFF + 256;  // K0 (sorted check) flag
constexpr size_t SMALL_BYTES = SMALL_FLAG_OFF + 256;

constexpr int MAX_DEVICES = 64;
DeviceCtx g_ctx[MAX_DEVICES];

int get_ctx(int device, DeviceCtx** out) {
    int dev = device;
    if (dev < 0) {
        CUDA_TRY(cudaGetDevice(&dev));
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        g_last_cuda_error = std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e);
        cudaGetLastError();
        return B200_ERR_NO_DEVICE;
    }
    if (dev >= count || dev >= MAX_DEVICES) return B200_ERR_BAD_ARG;
    DeviceCtx& c = g_ctx[dev];
    std::lock_guard<std::mutex> lk(c.mu);
    if (!c.ready) {
        CUDA_TRY(cudaSetDevice(dev));
        cudaDeviceProp prop;
        CUDA_TRY(cudaGetDeviceProperties(&prop, dev));
        if (prop.major < 10) {
            g_last_cuda_error = "device is not sm_100 class";
            return B200_ERR_NO_DEVICE;
        }
        c.dev = dev;
        c.sm_count = prop.multiProcessorCount;
        // a blocking stream: ordered against the legacy default stream, so callers that enqueue work
        // on stream 0 (e.g. a copy into the buffer) and pass stream = NULL are still ordered correctly
        CUDA_TRY(cudaStreamCreateWithFlags(&c.stream, cudaStreamDefault));
        for (auto& ev : c.ev) CUDA_TRY(cudaEventCreate(&ev));
        CUDA_TRY(cudaMalloc(&c.d_small, SMALL_BYTES));
        CUDA_TRY(cudaMallocHost(&c.h_hist, size_t(MAX_LEVELS) * RADIX * 8));
        CUDA_TRY(cudaMallocHost(&c.h_joint, size_t(RADIX * RADIX + 1) * 8));
        CUDA_TRY(cudaMallocHost(&c.h_flag, 64));
        for (auto& a : c.aux) CUDA_TRY(cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking));
        for (auto& ev : c.aux_ev) CUDA_TRY(cudaEventCreate(&ev));
        c.ready = true;
    }
    *out = &c;
    return B200_OK;
}

int ensure_bytes(void** p, size_t* have, size_t need) {
    if (*have >= need && *p) return B200_OK;
    if (*p) {
        cudaFree(*p);
        *p = nullptr;
        *have = 0;
    }
    // grow with a little slack so that repeated slightly larger sorts do not thrash
    size_t want = need + need / 16 + 256;
    cudaError_t e = cudaMalloc(p, want);
    if (e != cudaSuccess) {
        cudaGetLastError();
        want = need;
        e = cudaMalloc(p, want);
    }
    if (e != cudaSuccess) {
        g_last_cuda_error = std::string("cudaMalloc scratch: ") + cudaGetErrorString(e);
        cudaGetLastError();
        *p = nullptr;
        return B200_ERR_ALLOC;
    }
    *have = want;
    return B200_OK;
}

// ---------------------------------------------------------------------------
// Kernel selection.
// ---------------------------------------------------------------------------
bool word_path_ok(const Lowered& lo, const void* a, const void* b) {
    if (lo.elem_size % 4 != 0) return false;
    const uint32_t words = lo.elem_size / 4;
    if (words < 1 || words > 4) return false;
    const uintptr_t pa = reinterpret_cast<uintptr_t>(a), pb = reinterpret_cast<uintptr_t>(b);
    return pa % 16 == 0 && pb % 16 == 0;
}

PassLaunch get_pass_launch(uint32_t words, bool floatk, bool remap, int variant, bool hot, bool seg = false) {
    switch (words) {
        case 1: return get_pass_launch_w1(floatk, remap, variant, hot, seg);
        case 2: return get_pass_launch_w2(floatk, remap, variant, hot, seg);
        case 3: return get_pass_launch_w3(floatk, remap, variant, hot, seg);
        case 4: return get_pass_launch_w4(floatk, remap, variant, hot, seg);
    }
    return PassLaunch{};
}

int default_variant(uint32_t words, uint64_t n, int sm_count) {
    const char* env = getenv("B200SORT_VARIANT");
    if (env) {
        int v = atoi(env);
        if (v >= 0 && v < NUM_VARIANTS) return v;
    }
    // the Tuner's job on the GPU (reference src/tuners/standard_tuner.rs:13-63 picks by input_len):
    // large tiles amortise the per-tile digit scan, small inputs need enough tiles for every SM
    const int v = kDefaultVariant[words - 1];
    const uint64_t big_tile = uint64_t(kVariants[v].threads) * kVariants[v].items[words - 1];
    if (n < big_tile * uint64_t(sm_count) * 2) return kSmallVariant;
    return v;
}

constexpr int BYTE_THREADS = 256, BYTE_ITEMS = 4;

struct PassPlan {
    bool word_path;
    uint32_t words;
    PassLaunch pl;
    uint32_t tile;
    uint32_t num_tiles;
    int grid;
    int variant;
};

int plan_pass(DeviceCtx& c, const Lowered& lo, uint64_t n, const void* a, const void* b, bool remap, int variant,
              PassPlan* out, bool hot = false, bool sub_array = false) {
    PassPlan pp{};
    // sub-arrays of an aligned array start on a record boundary, which is all the word kernels need
    pp.word_path = sub_array ? (lo.elem_size % 4 == 0 && lo.elem_size <= 16) : word_path_ok(lo, a, b);
    if (pp.word_path) {
        pp.words = lo.elem_size / 4;
        if (variant < 0) variant = default_variant(pp.words, n, c.sm_count);
        if (variant >= NUM_VARIANTS) return B200_ERR_BAD_ARG;
        if (remap) variant = kRemapVariant;
        pp.variant = variant;
        pp.pl = get_pass_launch(pp.words, lo.is_float, remap, variant, hot);
        if (!pp.pl.kernel) return B200_ERR_BAD_ARG;
        pp.tile = pp.pl.tile;
        CUDA_TRY(cudaFuncSetAttribute(pp.pl.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(pp.pl.smem)));
        int occ = 0;
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pp.pl.kernel, pp.pl.threads, pp.pl.smem));
        if (occ < 1) {
            g_last_cuda_error = "pass kernel does not fit on an SM";
            return B200_ERR_CUDA;
        }
        const uint64_t tiles = (n + pp.tile - 1) / pp.tile;
        if (tiles > 0xffffffffull) return B200_ERR_BAD_ARG;
        pp.num_tiles = uint32_t(tiles);
        pp.grid = int(std::min<uint64_t>(tiles, uint64_t(c.sm_count) * occ));
    } else {
        pp.variant = 0;
        pp.tile = BYTE_THREADS * BYTE_ITEMS;
        const uint64_t tiles = (n + pp.tile - 1) / pp.tile;
        if (tiles > 0xffffffffull) return B200_ERR_BAD_ARG;
        pp.num_tiles = uint32_t(tiles);
        pp.grid = int(std::min<uint64_t>(tiles, uint64_t(c.sm_count) * 8));
    }
    *out = pp;
    return B200_OK;
}

int launch_pass(DeviceCtx& c, cudaStream_t st, const PassPlan& pp, const Lowered& lo, const void* src, void* dst,
                uint64_t n, uint32_t level, const uint64_t* d_base, uint32_t* d_ticket, uint32_t tag,
                const uint8_t* d_remap, const uint64_t* d_dst_tab = nullptr, const SplitParams* sp = nullptr,
                const uint32_t* hot = nullptr, uint32_t match_bits = 0, uint64_t* lookback = nullptr,
                const uint32_t* seg_tile_start = nullptr, const uint64_t* seg_rec_start = nullptr,
                const uint64_t* seg_bases = nullptr, const DigitParams* window = nullptr, uint32_t nparts = 0) {
    if (pp.word_path) {
        PassParams p;
        memset(&p.sp, 0, sizeof(p.sp));
        if (sp) p.sp = *sp;
        p.src = static_cast<const uint32_t*>(src);
        p.dst = static_cast<uint32_t*>(dst);
        p.n = n;
        p.dp = window ? *window : digit_params(lo, level);
        p.bucket_base = d_base;
        p.lookback = lookback ? lookback : c.d_lookback;
        p.ticket = d_ticket;
        p.tag = tag;
        p.num_tiles = pp.num_tiles;
        p.remap = d_remap;
        p.dst_tab = d_dst_tab;
        p.hot[0] = hot ? hot[0] : 0xffffffffu;
        p.hot[1] = hot ? hot[1] : 0xffffffffu;
        p.match_bits = match_bits;
        p.nparts = nparts;
        p.seg_tile_start = seg_tile_start;
        p.seg_rec_start = seg_rec_start;
        p.seg_bases = seg_bases;
        pp.pl.kernel<<<pp.grid, pp.pl.threads, pp.pl.smem, st>>>(p);
    } else {
        BytePassParams p;
        p.src = static_cast<const uint8_t*>(src);
        p.dst = static_cast<uint8_t*>(dst);
        p.n = n;
        p.elem_size = lo.elem_size;
        p.byte_offset = lo.lv[level].byte_offset;
        p.xor_mask = lo.lv[level].xor_mask;
        p.is_float = lo.is_float ? 1u : 0u;
        p.sign_byte = lo.sign_byte;
        p.bucket_base = d_base;
        p.lookback = c.d_lookback;
        p.ticket = d_ticket;
        p.tag = tag;
        p.num_tiles = pp.num_tiles;
        p.remap = d_remap;
        onesweep_pass_bytes_kernel<BYTE_THREADS, BYTE_ITEMS><<<pp.grid, BYTE_THREADS, 0, st>>>(p);
    }
    CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

// ---- K1 dispatch (instantiations live in hist_w*.cu) ----
// Fills d_hist [levels][256] (u64) on the stream.
int launch_histogram(DeviceCtx& c, cudaStream_t st, const Lowered& lo, const void* src, uint64_t n,
                     unsigned long long* d_hist, const uint32_t* gate = nullptr, bool accumulate = false) {
    if (!accumulate) CUDA_TRY(cudaMemsetAsync(d_hist, 0, size_t(lo.levels) * RADIX * 8, st));
    const bool aligned = reinterpret_cast<uintptr_t>(src) % 16 == 0;
    const uint32_t words = lo.elem_size / 4;
    bool fast = aligned && lo.elem_size % 4 == 0 && words >= 1 && words <= 4 && lo.contiguous &&
                lo.key_offset % 4 == 0 && (!lo.is_float || lo.levels == lo.key_width);
    const uint32_t nkw = (lo.levels + 3) / 4;
    if (fast && lo.key_offset / 4 + nkw > words) fast = false;
    // packed loads assume the key starts the record
    if (fast && words == nkw && lo.key_offset != 0) fast = false;
    const int grid_full = c.sm_count * 4;
    if (fast) {
        HistParams hp;
        hp.src = static_cast<const uint32_t*>(src);
        hp.n = n;
        hp.kw0 = lo.key_offset / 4;
        hp.sign_shift = ((lo.levels - 1) % 4) * 8 + 7;
        for (uint32_t w = 0; w < 4; ++w) hp.xorw[w] = 0u;
        for (uint32_t l = 0; l < lo.levels && l < 16; ++l) hp.xorw[l / 4] |= (lo.lv[l].xor_mask & 0xffu) << (8 * (l % 4));
        hp.ghist = d_hist;
        hp.gate = gate;
        const uint64_t per_thread = 8;
        const int grid =
            int(std::max<uint64_t>(1, std::min<uint64_t>(grid_full, (n + HIST_THREADS * per_thread - 1) /
                                                                         (HIST_THREADS * per_thread))));
        bool ok = false;
        switch (words) {
            case 1: ok = launch_hist_fast_w1(lo.levels, lo.is_float, grid, st, hp); break;
            case 2: ok = launch_hist_fast_w2(lo.levels, lo.is_float, grid, st, hp); break;
            case 3: ok = launch_hist_fast_w3(lo.levels, lo.is_float, grid, st, hp); break;
            case 4: ok = launch_hist_fast_w4(lo.levels, lo.is_float, grid, st, hp); break;
        }
        if (ok) {
            CUDA_TRY(cudaGetLastError());
            return B200_OK;
        }
    }
    HistGenericParams gp;
    gp.src = static_cast<const uint8_t*>(src);
    gp.n = n;
    gp.elem_size = lo.elem_size;
    gp.levels = lo.levels;
    gp.is_float = lo.is_float ? 1u : 0u;
    gp.sign_byte = lo.sign_byte;
    gp.ghist = d_hist;
    gp.gate = gate;
    for (uint32_t l = 0; l < MAX_LEVELS; ++l) gp.lv[l] = l < lo.levels ? lo.lv[l] : LevelLowering{0, 0};
    const size_t smem = size_t(lo.levels) * RADIX * 4;
    // (per device, so set on every call like the other kernels' attributes)
    CUDA_TRY(cudaFuncSetAttribute(hist_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_LEVELS * RADIX * 4));
    const int grid = int(std::max<uint64_t>(1, std::min<uint64_t>(grid_full, (n + HIST_THREADS - 1) / HIST_THREADS)));
    hist_generic_kernel<<<grid, HIST_THREADS, smem, st>>>(gp);
    CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

// Word indices / xor / mask of the 64-bit key prefix (top min(LEVELS, 8) levels).  Only word-path
// records with a contiguous key are supported; the caller falls back to digit splitters otherwise.
int lower_prefix(const Lowered& lo, SplitParams* sp) {
    memset(sp, 0, sizeof(*sp));
    if (!lo.contiguous || lo.elem_size % 4 != 0 || lo.elem_size > 16 || lo.key_offset % 4 != 0) return B200_ERR_BAD_DESC;
    if (lo.is_float && lo.levels != lo.key_width) return B200_ERR_BAD_DESC;
    const uint32_t L = lo.levels, words = lo.elem_size / 4, kw0 = lo.key_offset / 4;
    if (L > 16 || (L > 8 && (L - 8) % 4 != 0)) return B200_ERR_BAD_DESC;
    const uint32_t first = L > 8 ? L - 8 : 0;  // lowest level inside the prefix
    const uint32_t nl = L - first;              // 1..8 levels
    sp->lo_word = kw0 + first / 4;
    sp->hi_word = nl > 4 ? sp->lo_word + 1 : 0xffffffffu;
    if (sp->lo_word >= words || (nl > 4 && sp->hi_word >= words)) return B200_ERR_BAD_DESC;
    sp->mask = nl >= 8 ? ~0ull : ((1ull << (8 * nl)) - 1);
    for (uint32_t i = 0; i < nl; ++i) sp->xorv |= uint64_t(lo.lv[first + i].xor_mask & 0xffu) << (8 * i);
    sp->sign_word = kw0 + (L - 1) / 4;
    sp->sign_shift = ((L - 1) % 4) * 8 + 7;
    return B200_OK;
}

// Buckets holding more than 1/8 of the records are "hot": the half-warp pass kernel matches them by
// ballot instead of a same-address shared-memory RED.OR (which serialises).  Measured gain is small
// (+3 % on two-bucket levels, profiles/r1s_skew.txt), so only clearly skewed levels take this path.
static void pick_hot_buckets(const uint64_t* hist, uint64_t n, uint32_t hot[2]) {
    static const bool off = getenv("B200SORT_NO_HOT") != nullptr;
    hot[0] = hot[1] = 0xffffffffu;
    if (off) return;
    static const uint64_t div = getenv("B200SORT_HOT_DIV") ? std::max(2, atoi(getenv("B200SORT_HOT_DIV"))) : 8;
    uint64_t best0 = n / div, best1 = n / div;
    for (uint32_t b = 0; b < uint32_t(RADIX); ++b) {
        const uint64_t cnt = hist[b];
        if (cnt > best0) {
            best1 = best0, hot[1] = hot[0];
            best0 = cnt, hot[0] = b;
        } else if (cnt > best1) {
            best1 = cnt, hot[1] = b;
        }
    }
    if (hot[0] == 0xffffffffu) hot[1] = 0xffffffffu;
}

static uint32_t match_bits_for(uint32_t parts) {
    static const bool off = getenv("B200SORT_NO_BALLOT_PARTS") != nullptr;
    if (off || parts > 4) return 0;  // measured: ballots win at 2 parts, lose to OR-matching at 8
    uint32_t b = 1;
    while ((1u << b) < parts) ++b;
    return b;
}

// ---------------------------------------------------------------------------
// Hybrid path: two global MSD passes (K2 on the two most significant active levels), then K3 sorts
// every one of the 65 536 segments inside one SM's shared memory (local_sort.cuh).  The GPU form
// of rdst's own MSB-then-LSB route (reference src/sorter.rs:24-171); results are identical to
// the pure LSD path (both are THE stable order by the key).
// ---------------------------------------------------------------------------
constexpr size_t HYB_JOINT_OFF = 0;
constexpr size_t HYB_SEG_OFF = size_t(RADIX) * RADIX * 8;
constexpr size_t HYB_TICKET_OFF = HYB_SEG_OFF + (size_t(RADIX) * RADIX + 1) * 8 + 8;
constexpr size_t HYB_TILESTART_OFF = HYB_TICKET_OFF + 4096;  // [257] u32
constexpr size_t HYB_RECSTART_OFF = HYB_TILESTART_OFF + 2048;  // [257] u64
constexpr size_t HYB_WHIST_OFF = HYB_RECSTART_OFF + 4096;     // [2][256] u64 window histograms
constexpr size_t HYB_P1_TILESTART_OFF = HYB_WHIST_OFF + 4096;  // pass 1 as one trivial segment: [257] u32
constexpr size_t HYB_P1_RECSTART_OFF = HYB_P1_TILESTART_OFF + 2048;  // [257] u64
constexpr size_t HYB_P1_BASES_OFF = HYB_P1_RECSTART_OFF + 4096;      // [256] u64
constexpr size_t HYB_BYTES = HYB_P1_BASES_OFF + 2048;

// bits [bit_lo, bit_lo + width) of the 64-bit key image (lower_prefix) as a digit
DigitParams window_params(const Lowered& lo, const SplitParams& key, uint32_t bit_lo, uint32_t width) {
    DigitParams dp = digit_params(lo, 0);  // sign word / shift
    const bool in_hi = bit_lo >= 32;
    dp.word = in_hi ? key.hi_word : key.lo_word;
    dp.shift = bit_lo % 32;
    dp.word2 = (!in_hi && dp.shift + width > 32) ? key.hi_word : dp.word;
    dp.mask = (1u << width) - 1u;
    dp.xor_mask = uint32_t(key.xorv >> bit_lo) & dp.mask;
    return dp;
}

int hybrid_mode() {  // 0 = never, 1 = whenever the key shape allows (tests), 2 = auto (default)
    const char* env = getenv("B200SORT_HYBRID");
    if (!env) return 2;
    return atoi(env) == 0 ? 0 : (atoi(env) == 1 ? 1 : 2);
}

template <int WORDS>
int launch_local_sort_w(bool floatk, int grid, cudaStream_t st, const LocalSortParams& lp) {
    const size_t smem = LocalSmem<WORDS>::TOTAL;
    if (floatk) {
        CUDA_TRY(cudaFuncSetAttribute(local_sort_kernel<WORDS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        local_sort_kernel<WORDS, true><<<grid, LS_THREADS, smem, st>>>(lp);
    } else {
        CUDA_TRY(cudaFuncSetAttribute(local_sort_kernel<WORDS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        local_sort_kernel<WORDS, false><<<grid, LS_THREADS, smem, st>>>(lp);
    }
    CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

template <int WORDS>
int launch_hist2_w(bool floatk, int grid, cudaStream_t st, const Hist2Params& hp) {
    if (floatk)
        hist2_kernel<WORDS, true><<<grid, H2_THREADS, 0, st>>>(hp);
    else
        hist2_kernel<WORDS, false><<<grid, H2_THREADS, 0, st>>>(hp);
    CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

// The digits of the hybrid path, from K1's per-level histograms (host only; b200_hybrid_plan exposes it).
// D1 = the top 8 varying bits of the key image, D2 = the 8 bits below D1 starting at the next varying
// bit, local windows = byte-wide windows of what is left, anchored under D2 (least significant first;
// windows without a varying bit are dropped: sorting on constant bits is an exact no-op).
struct WindowPlan {
    uint64_t vary;            // bits of the key image that differ between some two records
    uint32_t w1_lo, w2_lo, w2_width;
    uint32_t nlocal, local_lo[LS_MAXLEV], local_width[LS_MAXLEV];
    bool aligned;             // D1 and D2 are whole levels: K1 already has their histograms
};
bool plan_windows_from_vary(uint64_t vary, WindowPlan* wp);
bool plan_windows(const Lowered& lo, const uint64_t* hist, uint64_t n, WindowPlan* wp) {
    memset(wp, 0, sizeof(*wp));
    if (lo.levels > 8) return false;
    uint64_t vary = 0;
    for (uint32_t l = 0; l < lo.levels; ++l) {
        const uint64_t* h = hist + size_t(l) * RADIX;
        for (uint32_t bit = 0; bit < 8; ++bit) {
            uint64_t ones = 0;
            for (uint32_t d = 0; d < uint32_t(RADIX); ++d)
                if (d >> bit & 1u) ones += h[d];
            if (ones != 0 && ones != n) vary |= 1ull << (8 * l + bit);
        }
    }
    return plan_windows_from_vary(vary, wp);
}
// the windows from the set of varying bits alone (a superset is fine: windows over constant bits sort nothing)
bool plan_windows_from_vary(uint64_t vary, WindowPlan* wp) {
    memset(wp, 0, sizeof(*wp));
    wp->vary = vary;
    if (!vary) return false;
    const int T1 = 63 - __builtin_clzll(vary);
    if (T1 < 8) return false;
    wp->w1_lo = uint32_t(T1 - 7);
    const uint64_t below1 = vary & ((1ull << wp->w1_lo) - 1);
    if (!below1) return false;
    const int T2 = 63 - __builtin_clzll(below1);
    wp->w2_lo = T2 >= 7 ? uint32_t(T2 - 7) : 0u;
    wp->w2_width = uint32_t(T2) + 1 - wp->w2_lo;
    uint32_t los[8], widths[8], cnt = 0;
    for (uint32_t hi = wp->w2_lo; hi > 0;) {
        const uint32_t width = std::min<uint32_t>(8, hi);
        hi -= width;
        if ((vary >> hi) & ((1ull << width) - 1)) los[cnt] = hi, widths[cnt++] = width;
    }
    if (cnt == 0 || cnt > uint32_t(LS_MAXLEV)) return false;
    for (uint32_t i = 0; i < cnt; ++i) wp->local_lo[i] = los[cnt - 1 - i], wp->local_width[i] = widths[cnt - 1 - i];
    wp->nlocal = cnt;
    wp->aligned = wp->w1_lo % 8 == 0 && wp->w2_lo % 8 == 0 && wp->w2_width == 8;
    return true;
}

template <int WORDS>
int launch_hist_win_w(bool floatk, int grid, cudaStream_t st, const HistWinParams& hp) {
    if (floatk)
        hist_win_kernel<WORDS, true><<<grid, H2_THREADS, 0, st>>>(hp);
    else
        hist_win_kernel<WORDS, false><<<grid, H2_THREADS, 0, st>>>(hp);
    CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

// ---- the unstable (key-only) flavour of the hybrid route: K2-U + K3-U (pass_atomic.cuh, local_count.cuh) ----
// A record that is nothing but its key (every byte is a level) may be permuted freely among equal
// MSD digits — rdst's default entry point is radix_sort_UNSTABLE and its MSB algorithms do exactly
// that (reference src/radix_sort.rs:21-45, src/sorts/ska_sort.rs, regions_sort.rs) — and equal keys
// are identical records, so the result is still bit-identical to the stable engine's.
bool key_only_record(const Lowered& lo) {
    const char* env = getenv("B200SORT_UNSTABLE");  // read per call: tests flip it
    const bool off = env && atoi(env) == 0;
    return !off && lo.contiguous && lo.key_offset == 0 && lo.levels == lo.elem_size && lo.key_width == lo.elem_size;
}
// same-address shared-memory atomics serialise: digits with a bucket above 1/8 of the records stay on the
// stable pass kernel (ballot path for hot buckets)
bool digit_is_flat(const uint64_t* hist, uint64_t n) {
    uint64_t m = 0;
    for (int b = 0; b < RADIX; ++b) m = std::max(m, hist[b]);
    return m <= n / 8;
}
bool heavy_split_enabled() {
    const char* env = getenv("B200SORT_HEAVY");  // read per call: tests flip it
    return !(env && atoi(env) == 0);
}
uint64_t hybrid_min_records(bool key_only, uint32_t words) {
    const char* env = getenv("B200SORT_HYBRID_MIN");
    if (env) return strtoull(env, nullptr, 10);
    // stable route: K3 holds one segment per SM at a time, so small segments leave it idle.  Key-only route:
    // K3-U runs several CTAs per SM on small segments, but every segment still costs a few microseconds of
    // claim / scan / barrier latency, so below ~500 records per segment the LSD passes win (measured, uniform
    // keys, profiles/r4_k3u_variants.txt: u64 25 M 1.33 against 1.21 ms, 50 M 1.72 against 2.21 ms; u32 100 M
    // 1.79 against 1.74 ms, 500 M 6.27 against 8.12 ms)
    if (!key_only) return 256ull << 20;
    return words == 1 ? (128ull << 20) : (32ull << 20);
}
MsdULaunch get_msdu_launch(uint32_t words, bool floatk) {
    const int v = getenv("B200SORT_MSDU_VARIANT") ? atoi(getenv("B200SORT_MSDU_VARIANT")) : 0;  // read per call
    const int vv = (v >= 0 && v < NUM_MSDU_VARIANTS) ? v : 0;
    return words == 1 ? get_msdu_launch_w1(floatk, vv) : get_msdu_launch_w2(floatk, vv);
}

// ---------------------------------------------------------------------------
// Oversized segments of the key-only hybrid route (heavy.cuh): split three ways around a pivot key, round
// by round; equal parts are final, small parts go to K3-U, parts that are still oversized take the next round.
// `ranges` = {first record, records} of the oversized segments, all living in `data`.  On return *ok says
// whether everything was finished (else the caller falls back to LSD passes over `data`, which is a
// permutation of the input at any time).
// ---------------------------------------------------------------------------
// A part of distinct keys shrinks like a quicksort partition, not by a fixed factor: with the median-of-five pivot
// 120 k distinct keys need 4-6 rounds to get under 16 384, and in 6.5 % of the runs more than 6 (the positions the
// split hands out are atomic, so the samples differ from run to run).  Later rounds touch only what is still
// oversized; 12 covers parts of several million distinct keys.
constexpr uint32_t HEAVY_MAX_ROUNDS = 12;
constexpr size_t HEAVY_MAX_RANGES = 1 << 16;

template <int WORDS>
int launch_heavy_w(int which, bool floatk, int grid, int threads, cudaStream_t st, const HeavyRanges& hr) {
    if (which == 0) {
        if (floatk) pivot_kernel<WORDS, true><<<grid, threads, 0, st>>>(hr);
        else pivot_kernel<WORDS, false><<<grid, threads, 0, st>>>(hr);
    } else {
        if (floatk) split_ends_kernel<WORDS, true><<<grid, HV_THREADS, 0, st>>>(hr);
        else split_ends_kernel<WORDS, false><<<grid, HV_THREADS, 0, st>>>(hr);
    }
    CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

int sort_heavy(DeviceCtx& c, cudaStream_t st, void* data, void* tmp, const Lowered& lo, uint32_t words, std::vector<uint2> ranges,
               LocalCountParams cp, LocalSortParams lp, uint32_t* launches, uint32_t* rounds_out, uint32_t* slow_out, bool* ok) {
    *ok = false;
    *rounds_out = 0;
    const uint32_t cap = uint32_t(LS_CAP);
    void* src = data;
    void* dst = tmp;
    uint32_t slow_total = 0;
    const char* renv = getenv("B200SORT_HEAVY_ROUNDS");  // read per call: tests flip it
    const uint32_t max_rounds = renv ? uint32_t(atoi(renv)) : HEAVY_MAX_ROUNDS;
    for (uint32_t round = 0; round < max_rounds && !ranges.empty(); ++round) {
        const size_t nr = ranges.size();
        if (nr > HEAVY_MAX_RANGES) break;
        // ---- tiles of the ranges ----
        std::vector<uint32_t> h_off(nr), h_cnt(nr), h_ts(nr + 1);
        uint64_t T = 0;
        for (size_t i = 0; i < nr; ++i) {
            h_off[i] = ranges[i].x, h_cnt[i] = ranges[i].y, h_ts[i] = uint32_t(T);
            T += (uint64_t(ranges[i].y) + HV_TILE - 1) / HV_TILE;
        }
        h_ts[nr] = uint32_t(T);
        // device scratch: off | cnt | tile_start | pivot | ends | pairs | chunks | chunk owners | slow list
        uint64_t recs = 0;
        for (size_t i = 0; i < nr; ++i) recs += ranges[i].y;
        const size_t max_chunks = size_t(recs / 65536) + nr + 1;  // equal parts cut into pieces of <= 65 536 records
        const size_t o_off = 0, o_cnt = o_off + nr * 4, o_ts = o_cnt + nr * 4, o_piv = (o_ts + (nr + 1) * 4 + 7) / 8 * 8,
                     o_ends = o_piv + nr * 8, o_pairs = o_ends + nr * 8, o_chunks = o_pairs + 2 * nr * 8,
                     o_owner = o_chunks + max_chunks * 8, o_slow = o_owner + max_chunks * 4, total = o_slow + (2 * nr + 1) * 4;
        int rc = ensure_bytes(reinterpret_cast<void**>(&c.d_heavy), &c.heavy_bytes, total);
        if (rc) return rc;
        uint8_t* d = c.d_heavy;
        CUDA_TRY(cudaMemcpyAsync(d + o_off, h_off.data(), nr * 4, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d + o_cnt, h_cnt.data(), nr * 4, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d + o_ts, h_ts.data(), (nr + 1) * 4, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemsetAsync(d + o_ends, 0, nr * 8, st));
        HeavyRanges hr;
        memset(&hr, 0, sizeof(hr));
        hr.src = static_cast<const uint32_t*>(src);
        hr.dst = static_cast<uint32_t*>(dst);
        hr.range_off = reinterpret_cast<const uint32_t*>(d + o_off);
        hr.range_cnt = reinterpret_cast<const uint32_t*>(d + o_cnt);
        hr.range_tile_start = reinterpret_cast<const uint32_t*>(d + o_ts);
        hr.nranges = uint32_t(nr);
        hr.ntiles = uint32_t(T);
        hr.pivot = reinterpret_cast<unsigned long long*>(d + o_piv);
        hr.ends = reinterpret_cast<uint32_t*>(d + o_ends);
        hr.xorv = cp.key.xorv;
        const int tgrid = int(std::min<uint64_t>(T, uint64_t(c.sm_count) * 8));
        // ---- a pivot per range; one read of the ranges moves the smaller / larger records to the other buffer ----
        rc = words == 1 ? launch_heavy_w<1>(0, lo.is_float, int((nr + 127) / 128), 128, st, hr)
                        : launch_heavy_w<2>(0, lo.is_float, int((nr + 127) / 128), 128, st, hr);
        if (rc) return rc;
        rc = words == 1 ? launch_heavy_w<1>(1, lo.is_float, tgrid, HV_THREADS, st, hr) : launch_heavy_w<2>(1, lo.is_float, tgrid, HV_THREADS, st, hr);
        if (rc) return rc;
        *launches += 2;
        std::vector<uint32_t> h_ends(2 * nr);
        CUDA_TRY(cudaMemcpyAsync(h_ends.data(), d + o_ends, nr * 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        // ---- what happens to every part: equal part = the pivot record written into `data`, small parts to K3-U
        // (read from `dst`, written sorted into `data`), oversized parts to the next round (they live in `dst`) ----
        std::vector<uint2> next, pairs, chunks;
        std::vector<uint32_t> owner;
        uint32_t max_small = 0;
        for (size_t i = 0; i < nr; ++i) {
            const uint32_t off = h_off[i], cnt = h_cnt[i], LT = h_ends[2 * i], GT = h_ends[2 * i + 1], EQ = cnt - LT - GT;
            // the equal records were never moved: where the range lives in `data` and nothing smaller or larger was found
            // they are in place already; otherwise the middle of the range is written
            if (EQ && !(src == data && LT == 0 && GT == 0))
                for (uint32_t at = 0; at < EQ; at += 65536)
                    chunks.push_back(make_uint2(off + LT + at, std::min<uint32_t>(65536, EQ - at))), owner.push_back(uint32_t(i));
            const uint2 parts[2] = {make_uint2(off, LT), make_uint2(off + cnt - GT, GT)};
            for (const uint2& pt : parts) {
                if (pt.y == 0) continue;
                if (pt.y > cap) next.push_back(pt);
                else pairs.push_back(pt), max_small = std::max(max_small, pt.y);
            }
        }
        if (chunks.size() > max_chunks || pairs.size() > 2 * nr) {
            g_last_cuda_error = "heavy segments: list sizes exceed their bounds";
            return B200_ERR_CUDA;
        }
        if (!chunks.empty()) {
            CUDA_TRY(cudaMemcpyAsync(d + o_chunks, chunks.data(), chunks.size() * 8, cudaMemcpyHostToDevice, st));
            CUDA_TRY(cudaMemcpyAsync(d + o_owner, owner.data(), owner.size() * 4, cudaMemcpyHostToDevice, st));
            const int cgrid = int(std::min<uint64_t>(chunks.size(), uint64_t(c.sm_count) * 8));
            if (words == 1)
                fill_kernel<1><<<cgrid, 256, 0, st>>>(static_cast<uint32_t*>(data), reinterpret_cast<const uint2*>(d + o_chunks),
                                                      reinterpret_cast<const uint32_t*>(d + o_owner), uint32_t(chunks.size()), hr.pivot);
            else
                fill_kernel<2><<<cgrid, 256, 0, st>>>(static_cast<uint32_t*>(data), reinterpret_cast<const uint2*>(d + o_chunks),
                                                      reinterpret_cast<const uint32_t*>(d + o_owner), uint32_t(chunks.size()), hr.pivot);
            CUDA_TRY(cudaGetLastError());
            *launches += 1;
        }
        // ---- small parts: K3-U reads them where the split put them and writes them, sorted, into `data` ----
        if (!pairs.empty()) {
            CUDA_TRY(cudaMemcpyAsync(d + o_pairs, pairs.data(), pairs.size() * 8, cudaMemcpyHostToDevice, st));
            uint32_t* d_slow2 = reinterpret_cast<uint32_t*>(d + o_slow);
            CUDA_TRY(cudaMemsetAsync(d_slow2, 0, 4, st));
            CUDA_TRY(cudaMemsetAsync(cp.ticket, 0, 4, st));
            const int size = max_small <= 4096 ? 0 : (max_small <= 8192 ? 1 : 2);
            const LocalCountLaunch ll = words == 1 ? get_local_count_launch_w1(lo.is_float, size, true) : get_local_count_launch_w2(lo.is_float, size, false);
            cp.data = static_cast<uint32_t*>(data);
            cp.src = dst != data ? static_cast<const uint32_t*>(dst) : nullptr;
            cp.seg_off = nullptr;
            cp.seg_pairs = reinterpret_cast<const uint2*>(d + o_pairs);
            cp.seg_first = 0;
            cp.nseg = uint32_t(pairs.size());
            cp.bmax = std::min<uint32_t>(cp.bmax, uint32_t(ll.cbits));
            cp.slow_list = d_slow2;
            cp.prof = nullptr;
            CUDA_TRY(cudaFuncSetAttribute(ll.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(ll.smem)));
            int occ = 1;
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ll.kernel, ll.threads, ll.smem));
            const int ugrid = int(std::min<uint64_t>(uint64_t(c.sm_count) * std::max(occ, 1), pairs.size()));
            ll.kernel<<<ugrid, ll.threads, ll.smem, st>>>(cp);
            CUDA_TRY(cudaGetLastError());
            // what K3-U could not finish (now in `data`, as a permutation): the stable segment sort, from the list
            CUDA_TRY(cudaMemsetAsync(lp.ticket, 0, 4, st));
            lp.data = static_cast<uint32_t*>(data);
            lp.seg_off = nullptr;
            lp.seg_pairs = cp.seg_pairs;
            lp.nseg = cp.nseg;
            lp.list = d_slow2;
            const int lgrid = int(std::min<uint64_t>(uint64_t(c.sm_count), pairs.size()));
            rc = words == 1 ? launch_local_sort_w<1>(lo.is_float, lgrid, st, lp) : launch_local_sort_w<2>(lo.is_float, lgrid, st, lp);
            if (rc) return rc;
            *launches += 2;
            uint32_t slow = 0;
            CUDA_TRY(cudaMemcpyAsync(&slow, d_slow2, 4, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
            slow_total += slow;
        }
        ranges.swap(next);
        std::swap(src, dst);  // the parts of this round live in the buffer the split wrote into
        *rounds_out = round + 1;
    }
    if (slow_out) *slow_out = slow_total;
    if (!ranges.empty()) {
        // still oversized after the last round: bring those ranges home and let the caller run the LSD passes
        if (src != data) {
            std::vector<uint2> chunks;
            for (const uint2& r : ranges)
                for (uint32_t at = 0; at < r.y; at += 65536) chunks.push_back(make_uint2(r.x + at, std::min<uint32_t>(65536, r.y - at)));
            int rc = ensure_bytes(reinterpret_cast<void**>(&c.d_heavy), &c.heavy_bytes, chunks.size() * 8);
            if (rc) return rc;
            CUDA_TRY(cudaMemcpyAsync(c.d_heavy, chunks.data(), chunks.size() * 8, cudaMemcpyHostToDevice, st));
            const int cgrid = int(std::min<uint64_t>(chunks.size(), uint64_t(c.sm_count) * 8));
            if (words == 1)
                copy_ranges_kernel<1><<<cgrid, 256, 0, st>>>(static_cast<const uint32_t*>(tmp), static_cast<uint32_t*>(data),
                                                             reinterpret_cast<const uint2*>(c.d_heavy), uint32_t(chunks.size()));
            else
                copy_ranges_kernel<2><<<cgrid, 256, 0, st>>>(static_cast<const uint32_t*>(tmp), static_cast<uint32_t*>(data),
                                                             reinterpret_cast<const uint2*>(c.d_heavy), uint32_t(chunks.size()));
            CUDA_TRY(cudaGetLastError());
            CUDA_TRY(cudaStreamSynchronize(st));
        }
        return B200_OK;
    }
    *ok = true;
    return B200_OK;
}

// How one sort call is driven (what b200_sort_opts carries, plus what the host entry point adds).
struct SortCtl {
    int variant = -1;             // >= 0: that tile variant, LSD passes only
    uint32_t route = B200_HINT_AUTO;
    bool result_anywhere = false; // the caller takes the result from wherever it ends up (no copy back)
    // host entry point: K0 and K1 already ran, chunk by chunk behind the H2D copies; the device histogram
    // and K0's flag are complete when the stream gets here
    bool pre_counted = false;
    // host entry point, pinned memory: D2H of the sorted array starts behind the first finished slice of
    // the segment sort instead of after the whole sort
    void* drain_host = nullptr;
    cudaStream_t drain_stream = nullptr;
    cudaEvent_t drain_event = nullptr;
    bool drained = false;         // out: the result is already on its way to drain_host (all of it)
    uint32_t heavy_segments = 0, heavy_rounds = 0;  // out: oversized segments split around pivots (heavy.cuh)
    // the array arrives already partitioned on one level (the sharded sort's exchange did MSD pass 1 on the way,
    // b200_sort_device_presplit): sort_hybrid skips its first pass, runs the second one data -> tmp and leaves the
    // result in `tmp`
    const uint64_t* presplit_counts = nullptr;  // [256] bucket sizes on presplit_level, in order
    uint32_t presplit_level = 0;
    // the caller knows which bits of the key image can differ between two records (the sharded sort: the gathered
    // histograms say so for every rank's range): K0 and K1 are skipped, the digits are planned from the hint and their
    // histograms come from the window histogram pass.  A superset of the truth is fine.  0 = no hint.
    uint64_t vary_hint = 0;
};

// Returns B200_OK with *done = true when the array was sorted here (result in `data`);
// *done = false means "not eligible / segments too large": the caller runs the LSD passes.
// *src_out tells the caller where the (possibly re-partitioned) records are.
int sort_hybrid(DeviceCtx& c, cudaStream_t st, void* data, void* tmp, uint64_t n, const Lowered& lo,
                const std::vector<uint32_t>& active, const unsigned long long* d_base, uint32_t* d_ticket, bool* done,
                void** src_out, uint32_t* launches, uint32_t* slow_segments, uint32_t* local_levels,
                uint32_t* unstable_out, SortCtl* ctl) {
    *done = false;
    *src_out = data;
    const bool pre = ctl && ctl->presplit_counts;
    const int mode = pre ? 1 : hybrid_mode();
    if (mode == 0 || (!pre && active.size() < 3)) return B200_OK;
    const uint32_t words = lo.elem_size / 4;
    if (lo.elem_size % 4 != 0 || (words != 1 && words != 2) || lo.levels > 8) return B200_OK;
    if (!word_path_ok(lo, data, tmp) || n >= (1ull << 32)) return B200_OK;
    SplitParams key;
    if (lower_prefix(lo, &key) != B200_OK) return B200_OK;
    const bool key_only = key_only_record(lo);
    // stable route: worth it from 256 M records and five active levels on (it costs about as much as 4.5
    // LSD passes whatever the key width, and K3 handles one segment per SM at a time).  Unstable route:
    // cheaper passes and a segment sort that keeps the SMs busy on small segments, so it pays from four
    // active levels and a few million records on.
    if (mode == 2 && (active.size() < (key_only ? 4u : 5u) || n < hybrid_min_records(key_only, words))) return B200_OK;

    // ---- the digits: bit windows over the varying bits of the key image (plan_windows) ----
    // keys that do not fill their type — or one rank's range of a sharded sort — still split 65 536 ways
    WindowPlan wp;
    if (pre) {
        // D1 = the level the array is partitioned on, D2 = the level below, local windows = every level below that.
        // Which of the lower bits vary is not known (no K1 ran on this array); windows over constant bits sort
        // nothing and cost little, and K3-U finds every segment's top bit by itself (AUTOTOP instantiations).
        const uint32_t L = ctl->presplit_level;
        if (!key_only || L < 2 || L >= lo.levels) return B200_OK;
        memset(&wp, 0, sizeof(wp));
        wp.w1_lo = 8 * L, wp.w2_lo = 8 * (L - 1), wp.w2_width = 8;
        wp.nlocal = L - 1;
        for (uint32_t i = 0; i + 1 < L; ++i) wp.local_lo[i] = 8 * i, wp.local_width[i] = 8;
        wp.vary = (L + 1 >= 8) ? ~0ull : ((1ull << (8 * (L + 1))) - 1);
        wp.aligned = true;
    } else if (ctl && ctl->vary_hint) {
        if (!plan_windows_from_vary(ctl->vary_hint, &wp)) return B200_OK;
        wp.aligned = false;  // no K1 ran: the digits' histograms come from the window histogram pass
    } else if (!plan_windows(lo, c.h_hist, n, &wp)) {
        return B200_OK;
    }
    const uint32_t w1_lo = wp.w1_lo, w2_lo = wp.w2_lo, w2_width = wp.w2_width, nlocal = wp.nlocal;
    const DigitParams W1 = window_params(lo, key, w1_lo, 8), W2 = window_params(lo, key, w2_lo, w2_width);
    DigitParams local[LS_MAXLEV];
    uint32_t local_lo[LS_MAXLEV];
    for (uint32_t i = 0; i < nlocal; ++i) {
        local[i] = window_params(lo, key, wp.local_lo[i], wp.local_width[i]);
        local_lo[i] = wp.local_lo[i];
    }
    {   // the kernels take the upper bits of a window that crosses words from the record's last word
        auto crossing_ok = [&](const DigitParams& w) { return w.word2 == w.word || w.word2 == words - 1; };
        bool ok = crossing_ok(W1) && crossing_ok(W2);
        for (uint32_t i = 0; i < nlocal; ++i) ok = ok && crossing_ok(local[i]);
        if (!ok) return B200_OK;
    }
    // byte-aligned windows are levels: K1 already has their histograms and bucket bases
    const bool aligned = wp.aligned;
    const uint32_t A1 = w1_lo / 8, A2 = w2_lo / 8;

    int rc = ensure_bytes(reinterpret_cast<void**>(&c.d_hyb), &c.hyb_bytes, HYB_BYTES);
    if (rc) return rc;
    unsigned long long* d_joint = reinterpret_cast<unsigned long long*>(c.d_hyb + HYB_JOINT_OFF);
    uint64_t* d_seg = reinterpret_cast<uint64_t*>(c.d_hyb + HYB_SEG_OFF);
    uint32_t* d_tk = reinterpret_cast<uint32_t*>(c.d_hyb + HYB_TICKET_OFF);
    unsigned long long* d_whist = reinterpret_cast<unsigned long long*>(c.d_hyb + HYB_WHIST_OFF);

    static const bool dbg = getenv("B200SORT_DEBUG_TIMING") != nullptr;
    if (dbg) cudaEventRecord(c.ev[3], st);
    uint64_t hw[2 * RADIX];  // histograms of the two MSD digits
    void* const A = pre ? data : tmp;  // where MSD pass 2 reads (the array partitioned on D1)
    void* const B = pre ? tmp : data;  // where it writes: the segments' home
    if (pre) {
        uint64_t s1 = 0;
        for (int b = 0; b < RADIX; ++b) hw[b] = ctl->presplit_counts[b], hw[RADIX + b] = 0, s1 += hw[b];
        if (s1 != n) return B200_ERR_BAD_ARG;
    } else if (aligned) {
        memcpy(hw, c.h_hist + size_t(A1) * RADIX, RADIX * 8);
        memcpy(hw + RADIX, c.h_hist + size_t(A2) * RADIX, RADIX * 8);
    } else {
        HistWinParams hp;
        hp.src = static_cast<const uint32_t*>(data);
        hp.n = n;
        hp.l1 = W1;
        hp.l2 = W2;
        hp.hist = d_whist;
        CUDA_TRY(cudaMemsetAsync(d_whist, 0, 2 * RADIX * 8, st));
        const int grid = int(std::min<uint64_t>(uint64_t(c.sm_count) * 8, (n + H2_THREADS * 4 - 1) / (H2_THREADS * 4)));
        rc = words == 1 ? launch_hist_win_w<1>(lo.is_float, grid, st, hp) : launch_hist_win_w<2>(lo.is_float, grid, st, hp);
        if (rc) return rc;
        *launches += 1;
        CUDA_TRY(cudaMemcpyAsync(c.h_joint, d_whist, 2 * RADIX * 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        memcpy(hw, c.h_joint, 2 * RADIX * 8);
        uint64_t s1 = 0, s2 = 0;
        for (int b = 0; b < RADIX; ++b) s1 += hw[b], s2 += hw[RADIX + b];
        if (s1 != n || s2 != n) {
            g_last_cuda_error = "hybrid: window histograms do not sum to n";
            return B200_ERR_CUDA;
        }
    }
    const uint64_t* h1 = hw;
    if (mode == 2) {
        // only when the segments can be expected to fit one SM (independence estimate; the exact
        // check follows the joint histogram)
        uint64_t m1 = 0, m2 = 0;
        for (int b = 0; b < RADIX; ++b) m1 = std::max(m1, hw[b]), m2 = std::max(m2, hw[RADIX + b]);
        // key-only records may still take the route when oversized segments are to be expected: those of heavily
        // duplicated keys are split around a pivot instead of sorted (heavy.cuh).  Marginals with a bucket above a
        // quarter of the records are clustered keys, not duplicates: passes.
        const bool may_split = key_only && heavy_split_enabled() && m1 <= n / 4 && m2 <= n / 4;
        if (double(m1) * double(m2) / double(n) > 0.97 * LS_CAP && !may_split) return B200_OK;
    }

    // ---- the pass kernels: stable K2 (SEG instantiation) or, for key-only records, K2-U ----
    PassPlan pp{};
    rc = plan_pass(c, lo, n, data, tmp, false, -1, &pp);
    if (rc) return rc;
    PassPlan ps = pp;  // stable, segmented (also the instantiation that reads bit windows)
    ps.pl = get_pass_launch(words, lo.is_float, false, pp.variant, false, true);
    if (!ps.pl.kernel || !ps.pl.seg || ps.pl.tile != int(pp.tile)) {
        g_last_cuda_error = "hybrid: no segmented instantiation of the pass kernel for this variant";
        return B200_ERR_CUDA;
    }
    PassPlan pu = pp;  // unstable
    if (key_only) {
        const MsdULaunch ul = get_msdu_launch(words, lo.is_float);
        pu.pl = PassLaunch{ul.kernel, ul.smem, ul.tile, ul.threads, false, true};
        pu.tile = uint32_t(ul.tile);
    }
    auto occupancy = [&](const PassPlan& q, int* occ) -> int {
        CUDA_TRY(cudaFuncSetAttribute(q.pl.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(q.pl.smem)));
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, q.pl.kernel, q.pl.threads, q.pl.smem));
        *occ = std::max(*occ, 1);
        return B200_OK;
    };
    // (presplit: D2's histogram is not known yet; uniform-looking keys are what takes this route, and a skewed
    // digit costs K2-U time, not correctness)
    const bool u1 = key_only && digit_is_flat(hw, n), u2 = key_only && (pre || digit_is_flat(hw + RADIX, n));
    const PassPlan& k1 = u1 ? pu : ps;
    const PassPlan& k2 = u2 ? pu : ps;
    int occ1 = 1, occ2 = 1;
    rc = occupancy(k1, &occ1);
    if (rc) return rc;
    rc = occupancy(k2, &occ2);
    if (rc) return rc;
    const uint64_t tiles1 = (n + k1.tile - 1) / k1.tile;
    // look-back rows: pass 1 needs tiles1; pass 2 at most n / tile + 256
    const uint64_t min_tile = std::min<uint64_t>(std::min(k1.tile, k2.tile), 4096);
    const size_t lb_rows = size_t(std::max<uint64_t>(tiles1, n / min_tile + 2 * RADIX + 1));
    rc = ensure_bytes(reinterpret_cast<void**>(&c.d_lookback), &c.lookback_bytes, lb_rows * RADIX * 8);
    if (rc) return rc;
    CUDA_TRY(cudaMemsetAsync(c.d_lookback, 0, size_t(tiles1) * RADIX * 8, st));
    CUDA_TRY(cudaMemsetAsync(d_ticket, 0, 1024, st));
    CUDA_TRY(cudaMemsetAsync(c.d_hyb, 0, HYB_SEG_OFF, st));
    CUDA_TRY(cudaMemsetAsync(d_tk, 0, 4096, st));

    // ---- MSD pass 1: data -> tmp on the first digit ----
    if (pre) {
        rc = B200_OK;  // done by whoever partitioned the array
    } else if (aligned && !u1) {
        rc = launch_pass(c, st, pp, lo, data, tmp, n, A1, reinterpret_cast<const uint64_t*>(d_base + size_t(A1) * RADIX),
                         d_ticket, 1, nullptr);
    } else {
        // the whole array as one segment (bucket 0) of the segmented kernel
        uint32_t h_ts[RADIX + 1];
        uint64_t h_rs[RADIX + 1], h_b1[RADIX];
        h_ts[0] = 0, h_rs[0] = 0;
        for (int b = 1; b <= RADIX; ++b) h_ts[b] = uint32_t(tiles1), h_rs[b] = n;
        uint64_t run = 0;
        for (int b = 0; b < RADIX; ++b) h_b1[b] = run, run += hw[b];
        uint32_t* d_ts = reinterpret_cast<uint32_t*>(c.d_hyb + HYB_P1_TILESTART_OFF);
        uint64_t* d_rs = reinterpret_cast<uint64_t*>(c.d_hyb + HYB_P1_RECSTART_OFF);
        uint64_t* d_b1 = reinterpret_cast<uint64_t*>(c.d_hyb + HYB_P1_BASES_OFF);
        CUDA_TRY(cudaMemcpyAsync(d_ts, h_ts, sizeof(h_ts), cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d_rs, h_rs, sizeof(h_rs), cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemcpyAsync(d_b1, h_b1, sizeof(h_b1), cudaMemcpyHostToDevice, st));
        PassPlan p1 = k1;
        p1.num_tiles = uint32_t(tiles1);
        p1.grid = int(std::min<uint64_t>(tiles1, uint64_t(c.sm_count) * occ1));
        rc = launch_pass(c, st, p1, lo, data, tmp, n, 0, d_b1, d_ticket, 1, nullptr, nullptr, nullptr, nullptr, 0, nullptr,
                         d_ts, d_rs, d_b1, &W1);
    }
    if (rc) return rc;
    if (!pre) *launches += 1;
    *src_out = A;
    cudaEventRecord(c.ev[4], st);
    // ---- joint histogram of the two digits on the partitioned array = the 65 536 segment sizes ----
    Hist2Params hp;
    hp.src = static_cast<const uint32_t*>(A);
    hp.n = n;
    hp.l1 = W1;
    hp.l2 = W2;
    hp.joint = d_joint;
    const int hgrid = int(std::min<uint64_t>(uint64_t(c.sm_count) * 8, (n + H2_CHUNK - 1) / H2_CHUNK));
    rc = words == 1 ? launch_hist2_w<1>(lo.is_float, hgrid, st, hp) : launch_hist2_w<2>(lo.is_float, hgrid, st, hp);
    if (rc) return rc;
    *launches += 1;
    CUDA_TRY(cudaMemcpyAsync(c.h_joint, d_joint, size_t(RADIX) * RADIX * 8, cudaMemcpyDeviceToHost, st));
    cudaEventRecord(c.ev[5], st);
    CUDA_TRY(cudaStreamSynchronize(st));
    // ---- segment offsets; give up (LSD from `tmp`) if a segment does not fit one SM ----
    uint64_t run = 0, max_seg = 0, nonempty = 0;
    for (int b1 = 0; b1 < RADIX; ++b1) {
        uint64_t in_bucket = 0;
        for (int b2 = 0; b2 < RADIX; ++b2) {
            const uint64_t cnt = c.h_joint[b1 * RADIX + b2];
            c.h_joint[b1 * RADIX + b2] = run;
            run += cnt;
            in_bucket += cnt;
            max_seg = std::max(max_seg, cnt);
            nonempty += cnt ? 1 : 0;
        }
        if (in_bucket != h1[b1]) {
            g_last_cuda_error = "hybrid: joint histogram disagrees with K1";
            return B200_ERR_CUDA;
        }
    }
    c.h_joint[RADIX * RADIX] = run;
    if (run != n) {
        g_last_cuda_error = "hybrid: joint histogram does not sum to n";
        return B200_ERR_CUDA;
    }
    std::vector<uint2> big;  // oversized segments {first record, records}
    if (max_seg > uint64_t(LS_CAP)) {
        // records with a payload, or too much of the array in oversized segments: *done stays false, records are in tmp
        if (!key_only || !heavy_split_enabled()) return B200_OK;
        uint64_t big_total = 0;
        for (uint32_t i = 0; i < uint32_t(RADIX * RADIX); ++i) {
            const uint64_t cnt = c.h_joint[i + 1] - c.h_joint[i];
            if (cnt > uint64_t(LS_CAP)) big.push_back(make_uint2(uint32_t(c.h_joint[i]), uint32_t(cnt))), big_total += cnt;
        }
        if (big.size() > 16384 || big_total > n / 10 * 9) return B200_OK;
    }
    CUDA_TRY(cudaMemcpyAsync(d_seg, c.h_joint, size_t(RADIX * RADIX + 1) * 8, cudaMemcpyHostToDevice, st));
    // ---- MSD pass 2: tmp -> data on the second digit, every bucket of the first partitioned on its own ----
    uint32_t h_tile_start[RADIX + 1];
    uint64_t h_rec_start[RADIX + 1];
    {
        uint64_t t = 0, rsum = 0;
        for (int b1 = 0; b1 < RADIX; ++b1) {
            h_tile_start[b1] = uint32_t(t);
            h_rec_start[b1] = rsum;
            t += (h1[b1] + k2.tile - 1) / k2.tile;
            rsum += h1[b1];
        }
        h_tile_start[RADIX] = uint32_t(t);
        h_rec_start[RADIX] = rsum;
        if (t > lb_rows) {
            g_last_cuda_error = "hybrid: look-back rows exhausted";
            return B200_ERR_CUDA;
        }
    }
    uint32_t* d_tile_start = reinterpret_cast<uint32_t*>(c.d_hyb + HYB_TILESTART_OFF);
    uint64_t* d_rec_start = reinterpret_cast<uint64_t*>(c.d_hyb + HYB_RECSTART_OFF);
    CUDA_TRY(cudaMemcpyAsync(d_tile_start, h_tile_start, sizeof(h_tile_start), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(d_rec_start, h_rec_start, sizeof(h_rec_start), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemsetAsync(c.d_lookback, 0, size_t(h_tile_start[RADIX]) * RADIX * 8, st));
    PassPlan p2 = k2;
    p2.num_tiles = h_tile_start[RADIX];
    p2.grid = int(std::min<uint64_t>(p2.num_tiles, uint64_t(c.sm_count) * occ2));
    rc = launch_pass(c, st, p2, lo, A, B, n, 0, d_seg, d_tk + 1, 1, nullptr, nullptr, nullptr, nullptr, 0, nullptr,
                     d_tile_start, d_rec_start, d_seg, &W2);
    if (rc) return rc;
    *launches += 1;
    cudaEventRecord(c.ev[6], st);
    // ---- K3: every segment sorted in shared memory, in place in `data` ----
    LocalSortParams lp;
    memset(&lp, 0, sizeof(lp));
    lp.data = static_cast<uint32_t*>(B);
    lp.seg_off = d_seg;
    lp.nseg = RADIX * RADIX;
    lp.ticket = d_tk;
    lp.nlev = nlocal;
    for (uint32_t i = 0; i < lp.nlev; ++i) lp.lv[i] = local[i];
    lp.key = key;
    lp.slow_segments = d_tk + 300;
    uint32_t nfast = 0;
    if (key_only) {
        // K3-U: one wide counting-sort level + comparison fix-up fused into the write-out; segments it cannot
        // finish (a long run of unequal keys inside one counter) go to the stable K3 in list mode
        const uint32_t slow_bytes = (RADIX * RADIX + 1) * 4;
        rc = ensure_bytes(reinterpret_cast<void**>(&c.d_slow), &c.slow_bytes, slow_bytes);
        if (rc) return rc;
        CUDA_TRY(cudaMemsetAsync(c.d_slow, 0, 4, st));
        const int force_size = getenv("B200SORT_K3U_SIZE") ? atoi(getenv("B200SORT_K3U_SIZE")) : -1;  // read per call: tests flip it
        int size = max_seg <= 4096 ? 0 : (max_seg <= 8192 ? 1 : 2);  // (oversized segments are skipped by K3-U: size 2 then)
        if (force_size >= 0 && force_size <= 2) size = std::max(size, force_size);
        // twice the threads with half the records each: pays for 4-byte keys (0.69 against 0.81 ms at 100 M u32,
        // 3.59 against 3.88 ms at 500 M f32), loses for 8-byte keys whose 64-register budget spills
        // (profiles/r4_k3u_variants.txt)
        const bool wide = getenv("B200SORT_K3U_WIDE") ? atoi(getenv("B200SORT_K3U_WIDE")) != 0 : words == 1;
        const LocalCountLaunch ll = words == 1 ? get_local_count_launch_w1(lo.is_float, size, wide, pre)
                                               : get_local_count_launch_w2(lo.is_float, size, wide, pre);
        LocalCountParams cp;
        memset(&cp, 0, sizeof(cp));
        cp.data = static_cast<uint32_t*>(B);
        cp.seg_off = d_seg;
        cp.key = key;
        // the digit starts at the highest bit below D2 that varies at all (plan_windows guarantees there is one)
        cp.top_bit = 64u - uint32_t(__builtin_clzll(wp.vary & ((1ull << w2_lo) - 1ull)));
        const int bmax_env = getenv("B200SORT_K3U_BMAX") ? atoi(getenv("B200SORT_K3U_BMAX")) : 0;
        cp.bmax = uint32_t(bmax_env > 0 ? std::min(bmax_env, ll.cbits) : ll.cbits);
        cp.slow_list = c.d_slow;
        static const bool k3uprof = getenv("B200SORT_K3_PROFILE") != nullptr;
        cp.prof = k3uprof ? reinterpret_cast<unsigned long long*>(d_tk + 320) : nullptr;
        CUDA_TRY(cudaFuncSetAttribute(ll.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(ll.smem)));
        int occ = 1;
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ll.kernel, ll.threads, ll.smem));
        occ = std::max(occ, 1);
        const int ugrid = int(std::min<uint64_t>(uint64_t(c.sm_count) * occ, nonempty ? nonempty : 1));
        // Segments are claimed in memory order, so the sort can run in slices of consecutive segments; the host
        // entry point uses that to start the D2H copy of a finished slice while the next one is sorted.
        const bool drain = ctl && ctl->drain_host && n * lo.elem_size >= (size_t(256) << 20) && big.empty();
        const uint32_t slices = drain ? 8u : 1u;
        for (uint32_t sl = 0; sl < slices; ++sl) {
            cp.seg_first = uint32_t(uint64_t(RADIX * RADIX) * sl / slices);
            cp.nseg = uint32_t(uint64_t(RADIX * RADIX) * (sl + 1) / slices);
            cp.ticket = d_tk + 2 + sl;
            ll.kernel<<<ugrid, ll.threads, ll.smem, st>>>(cp);
            CUDA_TRY(cudaGetLastError());
            *launches += 1;
            if (drain) {
                const size_t lo_b = size_t(c.h_joint[cp.seg_first]) * lo.elem_size, hi_b = size_t(c.h_joint[cp.nseg]) * lo.elem_size;
                CUDA_TRY(cudaEventRecord(ctl->drain_event, st));
                CUDA_TRY(cudaStreamWaitEvent(ctl->drain_stream, ctl->drain_event, 0));
                if (hi_b > lo_b)
                    CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t*>(ctl->drain_host) + lo_b, static_cast<const uint8_t*>(B) + lo_b,
                                             hi_b - lo_b, cudaMemcpyDeviceToHost, ctl->drain_stream));
            }
        }
        if (drain) ctl->drained = true;  // (the caller re-copies everything if a segment went to the slow list)
        if (k3uprof) {
            unsigned long long h[8];
            cudaMemcpyAsync(h, d_tk + 320, sizeof(h), cudaMemcpyDeviceToHost, st);
            cudaStreamSynchronize(st);
            double tot = 0;
            for (int i = 0; i < 8; ++i) tot += double(h[i]);
            fprintf(stderr, "[b200sort] K3-U cycles of thread 0, share per phase: wait/claim %.1f%% count %.1f%% scan %.1f%% scatter %.1f%% "
                            "fetch-issue %.1f%% fix-up %.1f%% write-out %.1f%% zero %.1f%% (%.0f cycles per segment)\n",
                    100 * h[0] / tot, 100 * h[1] / tot, 100 * h[2] / tot, 100 * h[3] / tot, 100 * h[4] / tot, 100 * h[5] / tot,
                    100 * h[6] / tot, 100 * h[7] / tot, tot / double(std::max<uint64_t>(1, nonempty)));
        }
        // the leftovers: full stable LSD over every local level, segments from the list
        lp.list = c.d_slow;
        lp.nfast = 0;
        lp.fix_shift = 0;
        lp.prof = nullptr;
        const int lgrid = int(std::min<uint64_t>(uint64_t(c.sm_count), nonempty ? nonempty : 1));
        rc = words == 1 ? launch_local_sort_w<1>(lo.is_float, lgrid, st, lp) : launch_local_sort_w<2>(lo.is_float, lgrid, st, lp);
        if (rc) return rc;
        *launches += 1;
        if (unstable_out) *unstable_out = 1u | (u1 ? 2u : 0u) | (u2 ? 4u : 0u);
        if (!big.empty()) {
            bool finished = false;
            uint32_t rounds = 0, slow2 = 0;
            cp.ticket = d_tk + 12;
            rc = sort_heavy(c, st, B, A, lo, words, big, cp, lp, launches, &rounds, &slow2, &finished);
            if (rc) return rc;
            if (ctl) ctl->heavy_segments = uint32_t(big.size()), ctl->heavy_rounds = rounds;
            if (dbg) fprintf(stderr, "[b200sort] hybrid: %zu oversized segments, %u pivot-split rounds, %s\n", big.size(), rounds,
                             finished ? "finished" : "NOT finished: LSD passes over the array");
            if (!finished) {
                *src_out = B;  // a permutation of the input; the caller runs the LSD passes on it
                return B200_OK;
            }
        }
    } else {
        // fast levels: as many of the top local levels as make runs of equal prefixes rare
        const double avg = double(n) / double(std::max<uint64_t>(1, nonempty));
        nfast = avg <= 64.0 ? 1 : 2;
        if (max_seg > 4096 * 4) nfast = 2;
        if (nfast >= lp.nlev) nfast = 0;  // nothing would be left for the fix-up: plain LSD over all local levels
        const char* force_slow = getenv("B200SORT_HYBRID_NFAST");
        if (force_slow) nfast = uint32_t(atoi(force_slow)) < lp.nlev ? uint32_t(atoi(force_slow)) : 0;
        lp.nfast = nfast;
        lp.fix_shift = nfast ? local_lo[lp.nlev - nfast] : 0;
        static const bool k3prof = getenv("B200SORT_K3_PROFILE") != nullptr;
        lp.prof = k3prof ? reinterpret_cast<unsigned long long*>(d_tk + 320) : nullptr;
        const int lgrid = int(std::min<uint64_t>(uint64_t(c.sm_count), nonempty ? nonempty : 1));
        rc = words == 1 ? launch_local_sort_w<1>(lo.is_float, lgrid, st, lp) : launch_local_sort_w<2>(lo.is_float, lgrid, st, lp);
        if (rc) return rc;
        *launches += 1;
        if (slow_segments) {
            CUDA_TRY(cudaMemcpyAsync(slow_segments, d_tk + 300, 4, cudaMemcpyDeviceToHost, st));
        }
        if (k3prof) {
            unsigned long long h[8];
            cudaMemcpyAsync(h, d_tk + 320, sizeof(h), cudaMemcpyDeviceToHost, st);
            cudaStreamSynchronize(st);
            double tot = 0;
            for (int i = 0; i < 8; ++i) tot += double(h[i]);
            fprintf(stderr, "[b200sort] K3 cycles of thread 0, share per phase: wait/claim %.1f%% rank %.1f%% digit scan %.1f%% scatter %.1f%% "
                            "read-back %.1f%% fix-up heads %.1f%% fix-up runs %.1f%% prefetch+write-out %.1f%% (%.0f cycles per segment)\n",
                    100 * h[0] / tot, 100 * h[1] / tot, 100 * h[2] / tot, 100 * h[3] / tot, 100 * h[4] / tot, 100 * h[5] / tot,
                    100 * h[6] / tot, 100 * h[7] / tot, tot / double(std::max<uint64_t>(1, nonempty)));
        }
    }
    cudaEventRecord(c.ev[7], st);
    // (after the event: with the host entry point's drain copies queued on the D2H engine this small copy waits
    // behind them, and must not count as sort time)
    if (key_only && slow_segments) CUDA_TRY(cudaMemcpyAsync(slow_segments, c.d_slow, 4, cudaMemcpyDeviceToHost, st));
    if (dbg) {
        cudaEventSynchronize(c.ev[7]);
        float a = 0, b = 0, d = 0, e = 0;
        cudaEventElapsedTime(&a, c.ev[3], c.ev[4]);
        cudaEventElapsedTime(&b, c.ev[4], c.ev[5]);
        cudaEventElapsedTime(&d, c.ev[5], c.ev[6]);
        cudaEventElapsedTime(&e, c.ev[6], c.ev[7]);
        fprintf(stderr, "[b200sort] hybrid%s: MSD pass 1%s %.3f ms, joint hist + D2H %.3f ms, MSD pass 2%s %.3f ms, K3%s %.3f ms"
                        " (max segment %llu, nfast %u of %u local levels; digits: bits %u..%u, %u..%u%s)\n",
                key_only ? " (key-only)" : "", u1 ? " (K2-U)" : "", a, b, u2 ? " (K2-U)" : "", d, key_only ? "-U" : "", e,
                (unsigned long long)max_seg, nfast, lp.nlev, w1_lo, w1_lo + 7, w2_lo, w2_lo + w2_width - 1,
                aligned ? "" : ", window histogram pass");
    }
    *done = true;
    *src_out = B;
    if (local_levels) *local_levels = nlocal;
    return B200_OK;
}

// ---------------------------------------------------------------------------
// Small inputs: the whole array is one segment of K3.  rdst hands chunks of up to 128 records to a
// comparison sort and runs everything else below its thresholds single-threaded in cache
// (reference src/sorter.rs:33-45, src/sorts/comparative_sort.rs:29-119); on the GPU an array that
// fits one SM's shared memory is sorted by ONE launch — no histogram, no read-back, no passes
// through HBM, no scratch — which is what bounds the latency of small sorts.
// ---------------------------------------------------------------------------
bool small_shape_ok(const Lowered& lo, uint64_t n, SplitParams* key) {  // host-only part of the rule (b200_route_plan)
    const char* env = getenv("B200SORT_SMALL");  // read per call: tests flip it
    if ((env && atoi(env) == 0) || hybrid_mode() == 0 || n > uint64_t(LS_CAP)) return false;
    if (lo.elem_size != 4 && lo.elem_size != 8) return false;
    if (lo.levels > uint32_t(LS_MAXLEV)) return false;
    return lower_prefix(lo, key) == B200_OK;
}
bool small_sort_ok(const Lowered& lo, const void* data, uint64_t n, SplitParams* key) {
    return small_shape_ok(lo, n, key) && word_path_ok(lo, data, data);
}

int sort_small(DeviceCtx& c, cudaStream_t st, void* data, uint64_t n, const Lowered& lo, const SplitParams& key) {
    int rc = ensure_bytes(reinterpret_cast<void**>(&c.d_hyb), &c.hyb_bytes, HYB_BYTES);
    if (rc) return rc;
    uint64_t* d_seg = reinterpret_cast<uint64_t*>(c.d_hyb + HYB_SEG_OFF);
    uint32_t* d_tk = reinterpret_cast<uint32_t*>(c.d_hyb + HYB_TICKET_OFF);
    c.h_joint[0] = 0;
    c.h_joint[1] = n;
    CUDA_TRY(cudaMemcpyAsync(d_seg, c.h_joint, 16, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemsetAsync(d_tk, 0, 4096, st));
    LocalSortParams lp;
    memset(&lp, 0, sizeof(lp));
    lp.data = static_cast<uint32_t*>(data);
    lp.seg_off = d_seg;
    lp.nseg = 1;
    lp.ticket = d_tk;
    lp.nlev = lo.levels;
    for (uint32_t l = 0; l < lo.levels; ++l) lp.lv[l] = window_params(lo, key, 8 * l, 8);
    lp.nfast = 0;  // a plain stable LSD over every level, all in shared memory
    lp.fix_shift = 0;
    lp.key = key;
    lp.slow_segments = d_tk + 300;
    lp.prof = nullptr;
    return lo.elem_size == 4 ? launch_local_sort_w<1>(lo.is_float, 1, st, lp) : launch_local_sort_w<2>(lo.is_float, 1, st, lp);
}

// ---------------------------------------------------------------------------
// Mid-size inputs: ONE MSD pass on the top active level, then K3 on its 256 buckets.  This is rdst's
// own default shape — split on the most significant byte, finish every bucket with an LSB sort that
// runs in cache (reference src/sorter.rs:105-157, src/tuners/standard_tuner.rs:13-63,
// src/sorts/lsb_sort.rs:38-128) — for inputs whose top-level buckets each fit one SM's shared
// memory (<= 16 384 records, i.e. up to ~4 M uniform keys): K0, K1, scan, one K2 pass and one K3
// launch instead of one K2 pass per level, each of which costs ~15 us of launch and look-back
// latency at these sizes whatever the data.  The records end up sorted in `tmp`.
// ---------------------------------------------------------------------------
// The host-only part of the rule: key shape, active levels and the bucket sizes of the MSD level.
bool msd1_shape_ok(const Lowered& lo, uint64_t n, const uint64_t* hist, const std::vector<uint32_t>& active,
                   SplitParams* key, uint64_t* max_seg_out, uint64_t* nonempty_out) {
    const char* env = getenv("B200SORT_MSD1");  // read per call: tests flip it
    if ((env && atoi(env) == 0) || hybrid_mode() == 0) return false;
    // measured (profiles/r3_k0_small_ladder.md): the MSD pass + K3 + copy back cost about as much as
    // 3.7 LSD passes at latency-bound sizes, so five active levels always pay, four only from ~1 M records
    if (active.size() < 4 || (active.size() == 4 && n < (1ull << 20))) return false;
    if ((lo.elem_size != 4 && lo.elem_size != 8) || lo.levels > uint32_t(LS_MAXLEV)) return false;
    if (lower_prefix(lo, key) != B200_OK) return false;
    const uint64_t* h = hist + size_t(active.back()) * RADIX;
    uint64_t max_seg = 0, nonempty = 0, total = 0;
    for (int b = 0; b < RADIX; ++b) {
        total += h[b];
        max_seg = std::max(max_seg, h[b]);
        nonempty += h[b] ? 1 : 0;
    }
    *max_seg_out = max_seg;
    *nonempty_out = nonempty;
    return total == n && max_seg <= uint64_t(LS_CAP);
}

int sort_msd1(DeviceCtx& c, cudaStream_t st, void* data, void* tmp, uint64_t n, const Lowered& lo,
              const std::vector<uint32_t>& active, const unsigned long long* d_base, uint32_t* d_ticket,
              const PassPlan& pp, bool* done, uint32_t* local_levels) {
    *done = false;
    SplitParams key;
    uint64_t max_seg = 0, nonempty = 0;
    if (!msd1_shape_ok(lo, n, c.h_hist, active, &key, &max_seg, &nonempty)) return B200_OK;
    if (!pp.word_path || !word_path_ok(lo, data, tmp)) return B200_OK;
    const uint32_t top = active.back();
    {
        const uint64_t* h = c.h_hist + size_t(top) * RADIX;
        uint64_t run = 0;
        for (int b = 0; b < RADIX; ++b) {
            c.h_joint[b] = run;
            run += h[b];
        }
        c.h_joint[RADIX] = run;
    }

    int rc = ensure_bytes(reinterpret_cast<void**>(&c.d_hyb), &c.hyb_bytes, HYB_BYTES);
    if (rc) return rc;
    uint64_t* d_seg = reinterpret_cast<uint64_t*>(c.d_hyb + HYB_SEG_OFF);
    uint32_t* d_tk = reinterpret_cast<uint32_t*>(c.d_hyb + HYB_TICKET_OFF);
    const size_t lb_bytes = size_t(pp.num_tiles) * RADIX * 8;
    rc = ensure_bytes(reinterpret_cast<void**>(&c.d_lookback), &c.lookback_bytes, lb_bytes);
    if (rc) return rc;
    CUDA_TRY(cudaMemsetAsync(c.d_lookback, 0, lb_bytes, st));
    CUDA_TRY(cudaMemsetAsync(d_ticket, 0, 1024, st));
    CUDA_TRY(cudaMemsetAsync(d_tk, 0, 4096, st));
    CUDA_TRY(cudaMemcpyAsync(d_seg, c.h_joint, size_t(RADIX + 1) * 8, cudaMemcpyHostToDevice, st));
    // ---- the MSD pass: data -> tmp, stable, on the top active level ----
    rc = launch_pass(c, st, pp, lo, data, tmp, n, top, reinterpret_cast<const uint64_t*>(d_base + size_t(top) * RADIX),
                     d_ticket, 1, nullptr);
    if (rc) return rc;
    // ---- K3: every bucket sorted on the remaining active levels, in place in tmp ----
    LocalSortParams lp;
    memset(&lp, 0, sizeof(lp));
    lp.data = static_cast<uint32_t*>(tmp);
    lp.seg_off = d_seg;
    lp.nseg = RADIX;
    lp.ticket = d_tk;
    lp.nlev = uint32_t(active.size()) - 1;
    for (uint32_t i = 0; i < lp.nlev; ++i) lp.lv[i] = window_params(lo, key, 8 * active[i], 8);
    // fast levels as on the hybrid path: radix-sort the top one or two local levels, insertion-sort what
    // is left of the runs; long runs (duplicates) switch a segment to the full local LSD inside K3
    const double avg = double(n) / double(std::max<uint64_t>(1, nonempty));
    uint32_t nfast = avg <= 64.0 ? 1 : 2;
    if (nfast >= lp.nlev) nfast = 0;
    const char* force = getenv("B200SORT_HYBRID_NFAST");
    if (force) nfast = uint32_t(atoi(force)) < lp.nlev ? uint32_t(atoi(force)) : 0;
    lp.nfast = nfast;
    lp.fix_shift = nfast ? 8 * active[lp.nlev - nfast] : 0;
    lp.key = key;
    lp.slow_segments = d_tk + 300;
    lp.prof = nullptr;
    const int grid = int(std::min<uint64_t>(uint64_t(c.sm_count), std::max<uint64_t>(1, nonempty)));
    rc = lo.elem_size == 4 ? launch_local_sort_w<1>(lo.is_float, grid, st, lp) : launch_local_sort_w<2>(lo.is_float, grid, st, lp);
    if (rc) return rc;
    *done = true;
    *local_levels = lp.nlev;
    return B200_OK;
}

// ---------------------------------------------------------------------------
// K0: is the input already in key order?  (rdst's already_sorted shortcut,
// reference src/sort_utils.rs:125-136 and src/sorter.rs:59-65.)
// ---------------------------------------------------------------------------
int sorted_check_mode() {  // B200SORT_SORTED_CHECK=0 switches the early-out off
    const char* env = getenv("B200SORT_SORTED_CHECK");  // read per call: tests flip it
    return env ? atoi(env) : 1;
}

// Enqueues K0 over records [0, n) and the read-back of its flag into c.h_flag (no synchronisation).
int launch_sorted_check(DeviceCtx& c, cudaStream_t st, const Lowered& lo, const void* data, uint64_t n, bool first = true,
                        bool last = true) {
    uint32_t* d_flag = reinterpret_cast<uint32_t*>(c.d_small + SMALL_FLAG_OFF);
    SortedCheckParams p;
    memset(&p, 0, sizeof(p));
    p.data = static_cast<const uint8_t*>(data);
    p.n_pairs = n - 1;
    p.flag = d_flag;
    p.elem_size = lo.elem_size;
    p.levels = lo.levels;
    p.key_offset = lo.key_offset;
    p.is_float = lo.is_float ? 1u : 0u;
    p.sign_byte = lo.sign_byte;
    for (uint32_t l = 0; l < lo.levels; ++l) p.lv[l] = lo.lv[l];
    const uint32_t w = lo.key_width;
    const bool word = lo.contiguous && (w == 1 || w == 2 || w == 4 || w == 8) && lo.levels <= w && lo.key_offset % w == 0 &&
                      lo.elem_size % w == 0 && reinterpret_cast<uintptr_t>(data) % w == 0;
    if (word) {
        p.keep = lo.levels >= 8 ? ~0ull : ((1ull << (8 * lo.levels)) - 1ull);
        for (uint32_t l = 0; l < lo.levels; ++l) p.flip |= uint64_t(lo.lv[l].xor_mask & 0xffu) << (8 * l);
    }
    if (first) CUDA_TRY(cudaMemsetAsync(d_flag, 0, 4, st));
    const uint64_t per_cta = uint64_t(SC_THREADS) * SC_ITEMS;
    const int grid = int(std::max<uint64_t>(1, std::min<uint64_t>(uint64_t(c.sm_count) * 8, (p.n_pairs + per_cta - 1) / per_cta)));
    // key-only arrays: the vectorised kernel (B200SORT_SORTED_CHECK=2 keeps the scalar one, for tests)
    const bool vec = word && lo.elem_size == w && lo.key_offset == 0 && reinterpret_cast<uintptr_t>(data) % 16 == 0 &&
                     sorted_check_mode() != 2;
    if (vec) {
        const uint64_t nvec = n * w / 16;
        const uint64_t per_vcta = uint64_t(SC_THREADS) * SC_VEC_ITEMS;
        const int vgrid = int(std::max<uint64_t>(1, std::min<uint64_t>(uint64_t(c.sm_count) * 8, (nvec + per_vcta - 1) / per_vcta)));
        if (w == 8) sorted_check_vec_kernel<8><<<vgrid, SC_THREADS, 0, st>>>(p, nvec);
        else if (w == 4) sorted_check_vec_kernel<4><<<vgrid, SC_THREADS, 0, st>>>(p, nvec);
        else if (w == 2) sorted_check_vec_kernel<2><<<vgrid, SC_THREADS, 0, st>>>(p, nvec);
        else sorted_check_vec_kernel<1><<<vgrid, SC_THREADS, 0, st>>>(p, nvec);
    } else if (!word) sorted_check_kernel<0><<<grid, SC_THREADS, 0, st>>>(p);
    else if (w == 8) sorted_check_kernel<8><<<grid, SC_THREADS, 0, st>>>(p);
    else if (w == 4) sorted_check_kernel<4><<<grid, SC_THREADS, 0, st>>>(p);
    else if (w == 2) sorted_check_kernel<2><<<grid, SC_THREADS, 0, st>>>(p);
    else sorted_check_kernel<1><<<grid, SC_THREADS, 0, st>>>(p);
    CUDA_TRY(cudaGetLastError());
    if (last) CUDA_TRY(cudaMemcpyAsync(c.h_flag, d_flag, 4, cudaMemcpyDeviceToHost, st));
    return B200_OK;
}

// ---------------------------------------------------------------------------
// The sort proper, device resident.
// ---------------------------------------------------------------------------
int sort_device_impl(DeviceCtx& c, cudaStream_t st, void* data, void* tmp, uint64_t n, const Lowered& lo,
                     SortCtl& ctl, void** result_ptr, b200_stats* stats) {
    // what the route hint means (include/b200sort.h b200_route_hint): LSD = passes only, like an explicit variant;
    // LOW_SCRATCH = passes with the small-tile variant (the caller releases the scratch pool afterwards)
    int variant = ctl.variant;
    if (variant < 0 && ctl.route == B200_HINT_LOW_SCRATCH) variant = kSmallVariant;
    const bool auto_route = ctl.variant < 0 && ctl.route == B200_HINT_AUTO;
    const bool result_anywhere = ctl.result_anywhere;
    unsigned long long* d_hist = reinterpret_cast<unsigned long long*>(c.d_small + SMALL_HIST_OFF);
    unsigned long long* d_base = reinterpret_cast<unsigned long long*>(c.d_small + SMALL_BASE_OFF);
    uint32_t* d_ticket = reinterpret_cast<uint32_t*>(c.d_small + SMALL_TICKET_OFF);

    CUDA_TRY(cudaEventRecord(c.ev[0], st));
    int rc = B200_OK;
    SplitParams small_key;
    if (auto_route && !ctl.pre_counted && small_sort_ok(lo, data, n, &small_key)) {
        rc = sort_small(c, st, data, n, lo, small_key);
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(c.ev[1], st));
        CUDA_TRY(cudaEventRecord(c.ev[2], st));
        CUDA_TRY(cudaStreamSynchronize(st));
        if (result_ptr) *result_ptr = data;
        if (stats) {
            stats->levels_run = lo.levels;  // no histogram: every declared level is sorted on chip
            stats->levels_skipped = 0;
            stats->local_levels = lo.levels;
            stats->algorithmic_bytes = n * lo.elem_size * (1ull + 2ull * lo.levels);
            stats->kernel_launches = 1;
            stats->tiles_per_pass = 1;
            CUDA_TRY(cudaEventElapsedTime(&stats->ms_local, c.ev[0], c.ev[1]));
            stats->ms_passes = stats->ms_total = stats->ms_local;
        }
        return B200_OK;
    }
    // The caller says which key bits can vary (SortCtl::vary_hint): no K0, no K1 — the hybrid route plans its digits
    // from the hint and counts them itself.  If it declines or gives up, the general path below runs on wherever the
    // records are, without the hint.
    if (ctl.vary_hint && auto_route && !ctl.pre_counted) {
        std::vector<uint32_t> hinted;
        for (uint32_t l = 0; l < lo.levels && l < 8; ++l)
            if ((ctl.vary_hint >> (8 * l)) & 0xffull) hinted.push_back(l);
        bool hdone = false;
        void* where = data;
        uint32_t hl = 0, hslow = 0, hlocal = 0, hunst = 0;
        CUDA_TRY(cudaEventRecord(c.ev[1], st));
        rc = sort_hybrid(c, st, data, tmp, n, lo, hinted, d_base, d_ticket, &hdone, &where, &hl, &hslow, &hlocal, &hunst, &ctl);
        if (rc) return rc;
        if (hdone) {
            CUDA_TRY(cudaEventRecord(c.ev[2], st));
            CUDA_TRY(cudaStreamSynchronize(st));
            if (result_ptr) *result_ptr = where;
            if (stats) {
                stats->levels_run = uint32_t(hinted.size());
                stats->levels_skipped = lo.levels - uint32_t(hinted.size());
                stats->hybrid = 1;
                stats->unstable = hunst;
                stats->heavy_segments = ctl.heavy_segments;
                stats->heavy_rounds = ctl.heavy_rounds;
                stats->slow_segments = hslow;
                stats->local_levels = hlocal;
                stats->kernel_launches = hl;
                stats->algorithmic_bytes = n * lo.elem_size * (1ull + 2ull * hinted.size());
                CUDA_TRY(cudaEventElapsedTime(&stats->ms_joint, c.ev[4], c.ev[5]));
                CUDA_TRY(cudaEventElapsedTime(&stats->ms_local, c.ev[6], c.ev[7]));
                CUDA_TRY(cudaEventElapsedTime(&stats->ms_passes, c.ev[1], c.ev[7]));
                CUDA_TRY(cudaEventElapsedTime(&stats->ms_total, c.ev[0], c.ev[2]));
            }
            return B200_OK;
        }
        ctl.vary_hint = 0;
        ctl.heavy_segments = ctl.heavy_rounds = 0;
        if (where != data) {  // MSD pass 1 ran before the route gave up: the records live in `tmp` now
            SortCtl plain;
            plain.variant = ctl.variant;
            plain.route = ctl.route;
            plain.result_anywhere = true;
            void* res = nullptr;
            rc = sort_device_impl(c, st, tmp, data, n, lo, plain, &res, stats);
            if (rc) return rc;
            if (res != data && !result_anywhere) {
                CUDA_TRY(cudaMemcpyAsync(data, res, n * lo.elem_size, cudaMemcpyDeviceToDevice, st));
                CUDA_TRY(cudaStreamSynchronize(st));
                res = data;
            }
            if (result_ptr) *result_ptr = res;
            return B200_OK;
        }
    }
    // K0, the already-sorted early-out, rides in front of K1 on the same stream.  K1 is gated on K0's
    // flag on the device (its CTAs leave at once when no pair was out of order), so a presorted input
    // costs one read of the array and a random one pays no extra synchronisation: the flag comes back
    // with the histograms.
    const bool check = sorted_check_mode() != 0 && n >= 2;
    static const bool dbg = getenv("B200SORT_DEBUG_TIMING") != nullptr;
    if (ctl.pre_counted) {
        // the host entry point ran K0 and K1 behind its H2D chunks; only the flag has to come back
        if (check) {
            *c.h_flag = 1;
            CUDA_TRY(cudaMemcpyAsync(c.h_flag, c.d_small + SMALL_FLAG_OFF, 4, cudaMemcpyDeviceToHost, st));
        }
    } else {
        if (check) {
            *c.h_flag = 1;
            rc = launch_sorted_check(c, st, lo, data, n);
            if (rc) return rc;
            if (dbg) cudaEventRecord(c.ev[8], st);
        }
        rc = launch_histogram(c, st, lo, data, n, d_hist,
                              check ? reinterpret_cast<const uint32_t*>(c.d_small + SMALL_FLAG_OFF) : nullptr);
        if (rc) return rc;
        if (dbg && check) {
            cudaEventRecord(c.ev[9], st);
            cudaEventSynchronize(c.ev[9]);
            float k0 = 0, k1 = 0;
            cudaEventElapsedTime(&k0, c.ev[0], c.ev[8]);
            cudaEventElapsedTime(&k1, c.ev[8], c.ev[9]);
            fprintf(stderr, "[b200sort] K0 (memset + kernel + flag D2H) %.3f ms, K1 %.3f ms, sorted = %d\n", k0, k1, *c.h_flag == 0);
        }
    }
    if (dbg) cudaEventRecord(c.ev[8], st);
    scan_hist_kernel<<<lo.levels, RADIX, 0, st>>>(d_hist, d_base);
    CUDA_TRY(cudaGetLastError());
    if (dbg) cudaEventRecord(c.ev[9], st);
    CUDA_TRY(cudaMemcpyAsync(c.h_hist, d_hist, size_t(lo.levels) * RADIX * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaEventRecord(c.ev[1], st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (dbg) {
        float a = 0, b = 0, d = 0;
        cudaEventElapsedTime(&a, c.ev[0], c.ev[8]);
        cudaEventElapsedTime(&b, c.ev[8], c.ev[9]);
        cudaEventElapsedTime(&d, c.ev[9], c.ev[1]);
        fprintf(stderr, "[b200sort] memset+K1 %.3f ms, scan %.3f ms, D2H hist %.3f ms\n", a, b, d);
    }

    if (check && *c.h_flag == 0) {  // K0: no pair out of order; K1's CTAs left at once and nothing was counted
        CUDA_TRY(cudaEventRecord(c.ev[2], st));
        CUDA_TRY(cudaStreamSynchronize(st));
        if (result_ptr) *result_ptr = data;
        if (stats) {
            stats->levels_run = 0;
            stats->levels_skipped = lo.levels;
            stats->already_sorted = 1;
            stats->algorithmic_bytes = n * lo.elem_size;
            stats->kernel_launches = 3;  // K0, the gated K1, scan
            CUDA_TRY(cudaEventElapsedTime(&stats->ms_hist, c.ev[0], c.ev[1]));
            stats->ms_total = stats->ms_hist;
        }
        return B200_OK;
    }

    // levels with a single non-empty bucket are exact no-ops for a stable LSD sort
    std::vector<uint32_t> active;
    for (uint32_t l = 0; l < lo.levels; ++l) {
        bool single = false;
        for (int b = 0; b < RADIX; ++b)
            if (c.h_hist[l * RADIX + b] == n) single = true;
        if (!single) active.push_back(l);
    }

    PassPlan pp{};
    rc = plan_pass(c, lo, n, data, tmp, false, variant, &pp);
    if (rc) return rc;
    PassPlan pp_hot{};  // same tiles, the instantiation with the hot-bucket ballot path
    rc = plan_pass(c, lo, n, data, tmp, false, pp.variant, &pp_hot, true);
    if (rc) return rc;
    void* src = data;
    void* dst = tmp;
    // the Tuner's algorithm choice on the GPU: MSD x2 + on-chip segment sort for large inputs whose
    // segments fit an SM, LSD passes otherwise (an explicit variant request always means LSD passes)
    bool hybrid_done = false;
    uint32_t hybrid_launches = 0, slow_segments = 0, local_levels = 0, unstable = 0;
    if (auto_route) {
        void* where = data;
        rc = sort_hybrid(c, st, data, tmp, n, lo, active, d_base, d_ticket, &hybrid_done, &where, &hybrid_launches,
                         &slow_segments, &local_levels, &unstable, &ctl);
        if (rc) return rc;
        if (!hybrid_done && where == tmp) std::swap(src, dst);  // MSD pass 1 ran: the records now live in tmp
    }
    bool msd1_done = false;
    uint32_t msd1_local = 0;
    if (auto_route && !hybrid_done && src == data) {
        rc = sort_msd1(c, st, data, tmp, n, lo, active, d_base, d_ticket, pp, &msd1_done, &msd1_local);
        if (rc) return rc;
        if (msd1_done) {
            std::swap(src, dst);  // the sorted records are in tmp
            hybrid_launches += 2;
        }
    }
    const uint32_t n_active = uint32_t(active.size());
    if (hybrid_done || msd1_done) active.clear();  // nothing left for the LSD loop
    if (!active.empty()) {
        const size_t lb_bytes = size_t(pp.num_tiles) * RADIX * 8;
        rc = ensure_bytes(reinterpret_cast<void**>(&c.d_lookback), &c.lookback_bytes, lb_bytes);
        if (rc) return rc;
        CUDA_TRY(cudaMemsetAsync(c.d_lookback, 0, lb_bytes, st));
        CUDA_TRY(cudaMemsetAsync(d_ticket, 0, 1024, st));
    }
    uint32_t pass = 0;
    for (uint32_t l : active) {
        uint32_t hot[2];
        pick_hot_buckets(c.h_hist + size_t(l) * RADIX, n, hot);
        const bool use_hot = hot[0] != 0xffffffffu && pp_hot.word_path && pp_hot.pl.hot;
        rc = launch_pass(c, st, use_hot ? pp_hot : pp, lo, src, dst, n, l,
                         reinterpret_cast<const uint64_t*>(d_base + size_t(l) * RADIX), d_ticket + pass, pass + 1, nullptr,
                         nullptr, nullptr, use_hot ? hot : nullptr);
        if (rc) return rc;
        std::swap(src, dst);
        ++pass;
    }
    uint32_t copy_back = 0;
    if (src != data && !result_anywhere) {
        CUDA_TRY(cudaMemcpyAsync(data, src, n * lo.elem_size, cudaMemcpyDeviceToDevice, st));
        copy_back = 1;
        src = data;
    }
    if (result_ptr) *result_ptr = src;
    CUDA_TRY(cudaEventRecord(c.ev[2], st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (ctl.drained && slow_segments) ctl.drained = false;  // some segments were finished after their slice was copied
    const uint32_t launches = (ctl.pre_counted ? 1 : 2 + (check ? 1 : 0)) + pass + hybrid_launches;  // (K0 +) K1 + scan + passes
    if (hybrid_done || msd1_done) pass = n_active;  // levels the result is sorted on
    if (stats) {
        stats->levels_run = pass;
        stats->levels_skipped = lo.levels - pass;
        stats->hybrid = hybrid_done ? 1u : (msd1_done ? 2u : 0u);
        stats->unstable = hybrid_done ? unstable : 0u;
        stats->heavy_segments = ctl.heavy_segments;
        stats->heavy_rounds = ctl.heavy_rounds;
        stats->slow_segments = slow_segments;
        stats->local_levels = hybrid_done ? local_levels : (msd1_done ? msd1_local : 0);
        stats->ms_joint = stats->ms_local = 0.0f;
        if (hybrid_done) {
            CUDA_TRY(cudaEventElapsedTime(&stats->ms_joint, c.ev[4], c.ev[5]));
            CUDA_TRY(cudaEventElapsedTime(&stats->ms_local, c.ev[6], c.ev[7]));
        }
        stats->variant = uint32_t(pp.variant);
        stats->copy_back = copy_back;
        stats->algorithmic_bytes = n * lo.elem_size * (1ull + 2ull * pass) + (copy_back ? 2ull * n * lo.elem_size : 0);
        CUDA_TRY(cudaEventElapsedTime(&stats->ms_hist, c.ev[0], c.ev[1]));
        CUDA_TRY(cudaEventElapsedTime(&stats->ms_passes, c.ev[1], hybrid_done ? c.ev[7] : c.ev[2]));
        CUDA_TRY(cudaEventElapsedTime(&stats->ms_total, c.ev[0], c.ev[2]));
        stats->tiles_per_pass = pp.num_tiles;
        stats->kernel_launches = launches;
    }
    return B200_OK;
}

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

}  // namespace

// ===========================================================================
// C ABI
// ===========================================================================
extern "C" {

int b200_sort_device(void* dev_ptr, void* dev_tmp, size_t n, const b200_key_desc* desc, const b200_sort_opts* opts,
                     b200_stats* stats) {
    Lowered lo;
    int rc = lower_desc(desc, &lo);
    if (rc) return rc;
    if (stats) memset(stats, 0, sizeof(*stats));
    if (opts && opts->route > B200_HINT_LOW_SCRATCH) return B200_ERR_BAD_ARG;
    if (n <= 1) return B200_OK;  // reference src/radix_sort_builder.rs:151
    if (!dev_ptr) return B200_ERR_BAD_ARG;
    DeviceCtx* c = nullptr;
    rc = get_ctx(opts ? opts->device : -1, &c);
    if (rc) return rc;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        DeviceGuard guard(c->dev);
        cudaStream_t st = (opts && opts->stream) ? static_cast<cudaStream_t>(opts->stream) : c->stream;
        void* tmp = dev_tmp;
        if (!tmp) {
            rc = ensure_bytes(&c->d_tmp, &c->tmp_bytes, n * lo.elem_size);
            if (rc) return rc;
            tmp = c->d_tmp;
        }
        SortCtl ctl;
        ctl.variant = opts ? opts->variant : -1;
        ctl.route = opts ? opts->route : uint32_t(B200_HINT_AUTO);
        rc = sort_device_impl(*c, st, dev_ptr, tmp, n, lo, ctl, nullptr, stats);
    }
    if (rc == B200_OK && opts && opts->route == B200_HINT_LOW_SCRATCH) rc = b200_release_scratch(c->dev);
    return rc;
}

}  // extern "C"

namespace {

// ---- host <-> device copies -------------------------------------------------------------------
// Pinned memory: plain DMA.  Pageable memory (what a Rust Vec is): the DMA engine cannot read it, so the
// bytes go through pinned bounce buffers; several host threads fill (drain) a buffer while the previous
// one is on the wire, so the copy runs at PCIe speed instead of one core's memcpy speed.
constexpr size_t BOUNCE_CHUNK = size_t(64) << 20;
constexpr int NBOUNCE = 4;

bool host_is_pinned(const void* p) {
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, p) == cudaSuccess) return attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged;
    cudaGetLastError();
    return false;
}

int copy_threads() {
    static const int t = [] {
        const char* env = getenv("B200SORT_COPY_THREADS");
        if (env && atoi(env) > 0) return std::min(atoi(env), 64);
        const unsigned hc = std::thread::hardware_concurrency();
        return int(std::max(1u, std::min(16u, hc ? hc : 4u)));  // measured on a 16-thread host: 8 -> 436 ms, 16 -> 417 ms per 8 GB round trip
    }();
    return t;
}

// A small persistent pool: spawning threads per 32 MB chunk cost ~60 ms on an 8 GB array.
class CopyPool {
public:
    explicit CopyPool(int workers) {
        for (int i = 0; i < workers; ++i) th_.emplace_back([this] { loop(); });
    }
    ~CopyPool() {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& t : th_) t.join();
    }
    // copies [src, src + len) to dst with the workers and the calling thread; returns when done
    void run(void* dst, const void* src, size_t len) {
        const size_t parts = th_.size() + 1;
        const size_t per = ((len + parts - 1) / parts + 4095) / 4096 * 4096;
        {
            std::lock_guard<std::mutex> lk(m_);
            dst_ = static_cast<uint8_t*>(dst), src_ = static_cast<const uint8_t*>(src), len_ = len, per_ = per;
            next_ = 1;
            pending_ = 0;
            for (size_t i = 1; i < parts; ++i)
                if (per * i < len) ++pending_;
            ++gen_;
        }
        cv_.notify_all();
        memcpy(dst, src, std::min(per, len));
        std::unique_lock<std::mutex> lk(m_);
        done_.wait(lk, [&] { return pending_ == 0; });
    }

private:
    void loop() {
        uint64_t seen = 0;
        std::unique_lock<std::mutex> lk(m_);
        while (true) {
            cv_.wait(lk, [&] { return stop_ || (gen_ != seen && next_ * per_ < len_); });
            if (stop_) return;
            while (next_ * per_ < len_) {
                const size_t off = next_++ * per_;
                const size_t n = std::min(per_, len_ - off);
                uint8_t* d = dst_ + off;
                const uint8_t* sp = src_ + off;
                lk.unlock();
                memcpy(d, sp, n);
                lk.lock();
                if (--pending_ == 0) done_.notify_all();
            }
            seen = gen_;
        }
    }
    std::vector<std::thread> th_;
    std::mutex m_;
    std::condition_variable cv_, done_;
    uint8_t* dst_ = nullptr;
    const uint8_t* src_ = nullptr;
    size_t len_ = 0, per_ = 1, next_ = 0, pending_ = 0;
    uint64_t gen_ = 0;
    bool stop_ = false;
};

std::mutex g_pool_mu;  // one bulk copy at a time uses the pool (calls for different devices take turns chunk by chunk)

void par_memcpy(void* dst, const void* src, size_t len) {
    const int T = copy_threads();
    if (T <= 1 || len < (size_t(4) << 20)) {
        memcpy(dst, src, len);
        return;
    }
    static CopyPool* pool = new CopyPool(T - 1);  // lives for the process
    std::lock_guard<std::mutex> lk(g_pool_mu);
    pool->run(dst, src, len);
}

int ensure_bounce(DeviceCtx& c) {
    if (c.h_bounce[0]) return B200_OK;
    for (int i = 0; i < NBOUNCE; ++i) {
        CUDA_TRY(cudaMallocHost(&c.h_bounce[i], BOUNCE_CHUNK));
        CUDA_TRY(cudaEventCreateWithFlags(&c.bounce_ev[i], cudaEventDisableTiming));
    }
    c.bounce_bytes = BOUNCE_CHUNK;
    return B200_OK;
}

// Enqueues (pinned) or performs (pageable: returns when the last chunk is queued) the copy of `bytes` bytes to
// the device on stream `st`.  `chunk_done(lo, hi)` is called after the bytes [lo, hi) have been queued, so the
// caller can queue work on them behind the copy.
template <class F>
int copy_in(DeviceCtx& c, cudaStream_t st, void* dev_dst, const void* host_src, size_t bytes, size_t chunk, size_t align,
            F chunk_done) {
    const bool pinned = host_is_pinned(host_src);
    if (!pinned) {
        int rc = ensure_bounce(c);
        if (rc) return rc;
        chunk = std::max<size_t>(align, std::min(chunk, BOUNCE_CHUNK) / align * align);  // chunks end on `align` boundaries
    }
    int k = 0;
    for (size_t off = 0; off < bytes; off += chunk) {
        const size_t len = std::min(chunk, bytes - off);
        const uint8_t* src = static_cast<const uint8_t*>(host_src) + off;
        if (pinned) {
            CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t*>(dev_dst) + off, src, len, cudaMemcpyHostToDevice, st));
        } else {
            CUDA_TRY(cudaEventSynchronize(c.bounce_ev[k]));  // the DMA that last used this buffer is done
            par_memcpy(c.h_bounce[k], src, len);
            CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t*>(dev_dst) + off, c.h_bounce[k], len, cudaMemcpyHostToDevice, st));
            CUDA_TRY(cudaEventRecord(c.bounce_ev[k], st));
            k = (k + 1) % NBOUNCE;
        }
        int rc = chunk_done(off, off + len);
        if (rc) return rc;
    }
    return B200_OK;
}

// Copies `bytes` bytes from the device to host memory on stream `st`; returns when they are in host memory
// (pageable) or queued (pinned: the caller synchronises the stream).
int copy_out(DeviceCtx& c, cudaStream_t st, void* host_dst, const void* dev_src, size_t bytes) {
    if (host_is_pinned(host_dst)) {
        CUDA_TRY(cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, st));
        return B200_OK;
    }
    int rc = ensure_bounce(c);
    if (rc) return rc;
    size_t pend_off[NBOUNCE] = {}, pend_len[NBOUNCE] = {};
    size_t off = 0;
    int k = 0;
    bool any = true;
    while (off < bytes || any) {
        if (pend_len[k]) {  // the oldest DMA: wait for it, then spread its bytes with all copy threads
            CUDA_TRY(cudaEventSynchronize(c.bounce_ev[k]));
            par_memcpy(static_cast<uint8_t*>(host_dst) + pend_off[k], c.h_bounce[k], pend_len[k]);
            pend_len[k] = 0;
        }
        if (off < bytes) {
            const size_t len = std::min(BOUNCE_CHUNK, bytes - off);
            CUDA_TRY(cudaMemcpyAsync(c.h_bounce[k], static_cast<const uint8_t*>(dev_src) + off, len, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaEventRecord(c.bounce_ev[k], st));
            pend_off[k] = off, pend_len[k] = len;
            off += len;
        }
        k = (k + 1) % NBOUNCE;
        any = false;
        for (int i = 0; i < NBOUNCE; ++i) any = any || pend_len[i] != 0;
    }
    return B200_OK;
}

}  // namespace

extern "C" {

int b200_copy_h2d(void* dev_dst, const void* host_src, size_t bytes, int device) {
    if (bytes == 0) return B200_OK;
    if (!dev_dst || !host_src) return B200_ERR_BAD_ARG;
    DeviceCtx* c = nullptr;
    int rc = get_ctx(device, &c);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard guard(c->dev);
    rc = copy_in(*c, c->stream, dev_dst, host_src, bytes, size_t(256) << 20, 4096, [](size_t, size_t) { return int(B200_OK); });
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return B200_OK;
}

int b200_copy_d2h(void* host_dst, const void* dev_src, size_t bytes, int device) {
    if (bytes == 0) return B200_OK;
    if (!host_dst || !dev_src) return B200_ERR_BAD_ARG;
    DeviceCtx* c = nullptr;
    int rc = get_ctx(device, &c);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard guard(c->dev);
    rc = copy_out(*c, c->stream, host_dst, dev_src, bytes);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return B200_OK;
}

// The `&mut [T]`-in-host-RAM drop-in.  Timeline on a large input: the H2D copy runs in chunks on a copy stream;
// K0 (sorted check) and K1 (all-level histogram) of chunk i run on the compute stream while chunk i + 1 is on
// the wire, so the sort proper starts the moment the last byte lands; on the key-only hybrid route the D2H copy
// starts behind the first finished slice of the segment sort.
int b200_sort_host(void* host_ptr, size_t n, const b200_key_desc* desc, const b200_sort_opts* opts,
                   b200_stats* stats) {
    Lowered lo;
    int rc = lower_desc(desc, &lo);
    if (rc) return rc;
    if (stats) memset(stats, 0, sizeof(*stats));
    if (opts && opts->route > B200_HINT_LOW_SCRATCH) return B200_ERR_BAD_ARG;
    if (n <= 1) return B200_OK;
    if (!host_ptr) return B200_ERR_BAD_ARG;
    DeviceCtx* c = nullptr;
    rc = get_ctx(opts ? opts->device : -1, &c);
    if (rc) return rc;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        DeviceGuard guard(c->dev);
        cudaStream_t st = (opts && opts->stream) ? static_cast<cudaStream_t>(opts->stream) : c->stream;
        cudaStream_t cs = c->aux[0];  // copies
        const size_t bytes = n * lo.elem_size;
        rc = ensure_bytes(&c->d_stage, &c->stage_bytes, bytes);
        if (rc) return rc;
        rc = ensure_bytes(&c->d_tmp, &c->tmp_bytes, bytes);
        if (rc) return rc;
        const bool pinned = host_is_pinned(host_ptr);
        SortCtl ctl;
        ctl.variant = opts ? opts->variant : -1;
        ctl.route = opts ? opts->route : uint32_t(B200_HINT_AUTO);
        ctl.result_anywhere = true;
        // count behind the copy when the input is large enough for the chunks to matter
        static const bool overlap_off = getenv("B200SORT_HOST_OVERLAP") && atoi(getenv("B200SORT_HOST_OVERLAP")) == 0;
        const bool pre = !overlap_off && bytes >= (size_t(64) << 20);
        const bool check = sorted_check_mode() != 0;
        unsigned long long* d_hist = reinterpret_cast<unsigned long long*>(c->d_small + SMALL_HIST_OFF);
        const size_t rec_align = size_t(lo.elem_size) * 16;  // chunk boundaries keep 16-byte alignment for every record size
        size_t chunk = std::max<size_t>(size_t(128) << 20, bytes / 16) / rec_align * rec_align;
        if (!pre) chunk = (bytes + rec_align - 1) / rec_align * rec_align;

        CUDA_TRY(cudaEventRecord(c->ev[10], st));
        CUDA_TRY(cudaStreamWaitEvent(cs, c->ev[10], 0));  // the copies start after whatever the caller queued before
        if (pre) CUDA_TRY(cudaMemsetAsync(d_hist, 0, size_t(lo.levels) * RADIX * 8, st));
        uint8_t* stage = static_cast<uint8_t*>(c->d_stage);
        rc = copy_in(*c, cs, stage, host_ptr, bytes, chunk, rec_align, [&](size_t lo_b, size_t hi_b) -> int {
            if (!pre) return B200_OK;
            CUDA_TRY(cudaEventRecord(c->aux_ev[0], cs));
            CUDA_TRY(cudaStreamWaitEvent(st, c->aux_ev[0], 0));
            const uint64_t r0 = lo_b / lo.elem_size, r1 = hi_b / lo.elem_size;
            if (check) {  // pairs (r0 - 1, r0) .. (r1 - 2, r1 - 1): the pair that straddles two chunks belongs to the later one
                const uint64_t first = r0 ? r0 - 1 : 0;
                if (r1 - first >= 2) {
                    int e = launch_sorted_check(*c, st, lo, stage + first * lo.elem_size, r1 - first, r0 == 0, false);
                    if (e) return e;
                }
            }
            return launch_histogram(*c, st, lo, stage + lo_b, r1 - r0, d_hist, nullptr, true);
        });
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(c->aux_ev[1], cs));
        CUDA_TRY(cudaStreamWaitEvent(st, c->aux_ev[1], 0));
        CUDA_TRY(cudaEventRecord(c->ev[11], st));
        ctl.pre_counted = pre;
        if (pinned) {
            ctl.drain_host = host_ptr;
            ctl.drain_stream = cs;
            ctl.drain_event = c->aux_ev[2];
        }
        void* result = nullptr;
        b200_stats local{};
        rc = sort_device_impl(*c, st, c->d_stage, c->d_tmp, n, lo, ctl, &result, &local);
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(c->aux_ev[3], st));  // (ev[5] / ev[6] belong to the hybrid route's phase times)
        if (local.already_sorted) {
            // K0 found the input in key order: the caller's buffer already holds the result
        } else if (ctl.drained) {
            CUDA_TRY(cudaStreamSynchronize(cs));
        } else {
            if (ctl.drain_host) CUDA_TRY(cudaStreamSynchronize(cs));  // partial drains of a sort that then took the slow list
            rc = copy_out(*c, st, host_ptr, result, bytes);
            if (rc) return rc;
        }
        CUDA_TRY(cudaEventRecord(c->aux_ev[4], st));
        CUDA_TRY(cudaStreamSynchronize(st));
        static const bool hdbg = getenv("B200SORT_DEBUG_TIMING") != nullptr;
        if (hdbg) {
            float a = 0, b = 0, d = 0;
            cudaEventElapsedTime(&a, c->ev[10], c->ev[11]);
            cudaEventElapsedTime(&b, c->ev[11], c->aux_ev[3]);
            cudaEventElapsedTime(&d, c->aux_ev[3], c->aux_ev[4]);
            fprintf(stderr, "[b200sort] host entry (%s memory, %d copy threads): H2D%s %.1f ms, sort%s %.1f ms, D2H after the sort %.1f ms\n",
                    pinned ? "pinned" : "pageable", copy_threads(), pre ? " + K0/K1 behind the chunks" : "", a,
                    ctl.drained ? " + D2H of finished slices" : "", b, d);
        }
        if (stats) {
            *stats = local;
            if (pre) stats->kernel_launches += uint32_t((bytes + chunk - 1) / chunk) * (check ? 2u : 1u);
            CUDA_TRY(cudaEventElapsedTime(&stats->ms_h2d, c->ev[10], c->ev[11]));
            CUDA_TRY(cudaEventElapsedTime(&stats->ms_d2h, c->aux_ev[3], c->aux_ev[4]));
            CUDA_TRY(cudaEventElapsedTime(&stats->ms_total, c->ev[10], c->aux_ev[4]));
        }
    }
    if (opts && opts->route == B200_HINT_LOW_SCRATCH) return b200_release_scratch(c->dev);
    return B200_OK;
}

int b200_histogram_device(const void* dev_ptr, size_t n, const b200_key_desc* desc, const b200_sort_opts* opts,
                          uint64_t* hist_out) {
    Lowered lo;
    int rc = lower_desc(desc, &lo);
    if (rc) return rc;
    if (!hist_out) return B200_ERR_BAD_ARG;
    memset(hist_out, 0, size_t(lo.levels) * RADIX * 8);
    if (n == 0) return B200_OK;
    if (!dev_ptr) return B200_ERR_BAD_ARG;
    DeviceCtx* c = nullptr;
    rc = get_ctx(opts ? opts->device : -1, &c);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard guard(c->dev);
    cudaStream_t st = (opts && opts->stream) ? static_cast<cudaStream_t>(opts->stream) : c->stream;
    unsigned long long* d_hist = reinterpret_cast<unsigned long long*>(c->d_small + SMALL_HIST_OFF);
    rc = launch_histogram(*c, st, lo, dev_ptr, n, d_hist);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(c->h_hist, d_hist, size_t(lo.levels) * RADIX * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    memcpy(hist_out, c->h_hist, size_t(lo.levels) * RADIX * 8);
    return B200_OK;
}

static int pass_common(const void* dev_src, void* dev_dst, size_t n, const b200_key_desc* desc, uint32_t level,
                       const uint64_t* counts, const uint32_t* split, uint32_t parts, uint64_t* part_counts_out,
                       const b200_sort_opts* opts, void* const* dst_ptrs = nullptr) {
    Lowered lo;
    int rc = lower_desc(desc, &lo);
    if (rc) return rc;
    if (level >= lo.levels) return B200_ERR_LEVEL_RANGE;
    if (!counts) return B200_ERR_BAD_ARG;
    const bool remap = split != nullptr;
    if (remap && (parts == 0 || parts > RADIX)) return B200_ERR_BAD_ARG;
    uint64_t total = 0;
    for (int b = 0; b < RADIX; ++b) total += counts[b];
    if (total != n) return B200_ERR_BAD_ARG;
    // bucket bases; with a split table digits are first mapped to their part
    uint8_t h_remap[RADIX];
    uint64_t h_base[RADIX];
    if (remap) {
        uint64_t pc[RADIX] = {0};
        for (int b = 0; b < RADIX; ++b) {
            uint32_t part = 0;
            while (part + 1 < parts && uint32_t(b) >= split[part]) ++part;
            h_remap[b] = uint8_t(part);
            pc[part] += counts[b];
        }
        uint64_t run = 0;
        for (int q = 0; q < RADIX; ++q) {
            h_base[q] = dst_ptrs ? 0 : run;  // with per-part destinations every part starts at its own pointer
            run += pc[q];
        }
        if (part_counts_out)
            for (uint32_t q = 0; q < parts; ++q) part_counts_out[q] = pc[q];
    } else {
        uint64_t run = 0;
        for (int b = 0; b < RADIX; ++b) {
            h_base[b] = run;
            run += counts[b];
        }
    }
    if (n == 0) return B200_OK;
    if (!dev_src || (!dev_dst && !dst_ptrs)) return B200_ERR_BAD_ARG;
    uint64_t h_dtab[RADIX] = {0};
    if (dst_ptrs) {
        for (uint32_t q = 0; q < parts; ++q) {
            if (!dst_ptrs[q]) return B200_ERR_BAD_ARG;
            h_dtab[q] = reinterpret_cast<uint64_t>(dst_ptrs[q]);
            if (h_dtab[q] % 16 != 0 && lo.elem_size % 16 == 0) return B200_ERR_BAD_ARG;
        }
        dev_dst = dst_ptrs[0];
    }
    DeviceCtx* c = nullptr;
    rc = get_ctx(opts ? opts->device : -1, &c);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard guard(c->dev);
    cudaStream_t st = (opts && opts->stream) ? static_cast<cudaStream_t>(opts->stream) : c->stream;
    PassPlan pp{};
    uint32_t hot[2];
    pick_hot_buckets(counts, n, hot);
    rc = plan_pass(*c, lo, n, dev_src, dst_ptrs ? dev_src : dev_dst, remap, opts ? opts->variant : -1, &pp,
                   !remap && hot[0] != 0xffffffffu);
    if (rc) return rc;
    if (dst_ptrs && !pp.word_path) return B200_ERR_BAD_DESC;  // per-part destinations need 4-byte-multiple records
    const size_t lb_bytes = size_t(pp.num_tiles) * RADIX * 8;
    rc = ensure_bytes(reinterpret_cast<void**>(&c->d_lookback), &c->lookback_bytes, lb_bytes);
    if (rc) return rc;
    uint64_t* d_pbase = reinterpret_cast<uint64_t*>(c->d_small + SMALL_PBASE_OFF);
    uint8_t* d_remap = c->d_small + SMALL_REMAP_OFF;
    uint32_t* d_ticket = reinterpret_cast<uint32_t*>(c->d_small + SMALL_TICKET_OFF);
    CUDA_TRY(cudaMemsetAsync(c->d_lookback, 0, lb_bytes, st));
    CUDA_TRY(cudaMemsetAsync(d_ticket, 0, 1024, st));
    CUDA_TRY(cudaMemcpyAsync(d_pbase, h_base, sizeof(h_base), cudaMemcpyHostToDevice, st));
    if (remap) CUDA_TRY(cudaMemcpyAsync(d_remap, h_remap, sizeof(h_remap), cudaMemcpyHostToDevice, st));
    uint64_t* d_dtab = reinterpret_cast<uint64_t*>(c->d_small + SMALL_DTAB_OFF);
    if (dst_ptrs) CUDA_TRY(cudaMemcpyAsync(d_dtab, h_dtab, sizeof(h_dtab), cudaMemcpyHostToDevice, st));
    rc = launch_pass(*c, st, pp, lo, dev_src, dev_dst, n, level, d_pbase, d_ticket, 1, remap ? d_remap : nullptr,
                     dst_ptrs ? d_dtab : nullptr, nullptr, (pp.word_path && pp.pl.hot) ? hot : nullptr,
                     remap ? match_bits_for(parts) : 0, nullptr, nullptr, nullptr, nullptr, nullptr, remap ? parts : 0);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(st));
    return B200_OK;
}

int b200_pass_device(const void* dev_src, void* dev_dst, size_t n, const b200_key_desc* desc, uint32_t level,
                     const uint64_t* counts, const b200_sort_opts* opts) {
    return pass_common(dev_src, dev_dst, n, desc, level, counts, nullptr, 0, nullptr, opts);
}

int b200_partition_device(const void* dev_src, void* dev_dst, size_t n, const b200_key_desc* desc, uint32_t level,
                          const uint64_t* counts, const uint32_t* split, uint32_t parts, uint64_t* part_counts_out,
                          const b200_sort_opts* opts) {
    if (!split && parts > 1) return B200_ERR_BAD_ARG;
    static const uint32_t none[1] = {256};
    return pass_common(dev_src, dev_dst, n, desc, level, counts, split ? split : none, parts ? parts : 1,
                       part_counts_out, opts);
}

int b200_partition_peers_device(const void* dev_src, void* const* dst_ptrs, size_t n, const b200_key_desc* desc,
                                uint32_t level, const uint64_t* counts, const uint32_t* split, uint32_t parts,
                                uint64_t* part_counts_out, const b200_sort_opts* opts) {
    if (!dst_ptrs || parts == 0) return B200_ERR_BAD_ARG;
    if (!split && parts > 1) return B200_ERR_BAD_ARG;
    static const uint32_t none[1] = {256};
    return pass_common(dev_src, nullptr, n, desc, level, counts, split ? split : none, parts, part_counts_out, opts,
                       dst_ptrs);
}

// ---- full-resolution splitters (sample-sort style) for the sharded sort ----
static int prefix_src(const void* dev_ptr, size_t n, const b200_key_desc* desc, Lowered* lo, PrefixSrc* ps) {
    int rc = lower_desc(desc, lo);
    if (rc) return rc;
    rc = lower_prefix(*lo, &ps->sp);
    if (rc) return rc;
    if (reinterpret_cast<uintptr_t>(dev_ptr) % 4 != 0) return B200_ERR_BAD_DESC;
    ps->src = static_cast<const uint32_t*>(dev_ptr);
    ps->n = n;
    ps->words = lo->elem_size / 4;
    ps->is_float = lo->is_float ? 1u : 0u;
    return B200_OK;
}

int b200_key_prefix_supported(const b200_key_desc* desc) {
    Lowered lo;
    if (lower_desc(desc, &lo)) return 0;
    SplitParams sp;
    return lower_prefix(lo, &sp) == B200_OK ? 1 : 0;
}

int b200_key_prefix_sample_device(const void* dev_ptr, size_t n, const b200_key_desc* desc, size_t stride,
                                  size_t count, uint64_t* prefixes_out, const b200_sort_opts* opts) {
    Lowered lo;
    PrefixSrc ps;
    int rc = prefix_src(dev_ptr, n, desc, &lo, &ps);
    if (rc) return rc;
    if (count == 0) return B200_OK;
    if (!dev_ptr || !prefixes_out || stride == 0 || (count - 1) * stride >= n) return B200_ERR_BAD_ARG;
    DeviceCtx* c = nullptr;
    rc = get_ctx(opts ? opts->device : -1, &c);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard guard(c->dev);
    cudaStream_t st = (opts && opts->stream) ? static_cast<cudaStream_t>(opts->stream) : c->stream;
    rc = ensure_bytes(reinterpret_cast<void**>(&c->d_sample), &c->sample_bytes, count * 8);
    if (rc) return rc;
    unsigned long long* d_out = reinterpret_cast<unsigned long long*>(c->d_sample);
    prefix_sample_kernel<<<unsigned((count + 255) / 256), 256, 0, st>>>(ps, stride, count, d_out);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(prefixes_out, d_out, count * 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return B200_OK;
}

int b200_splitter_count_device(const void* dev_ptr, size_t n, const b200_key_desc* desc, const uint64_t* splitters,
                               uint32_t nsplit, uint64_t* less_out, uint64_t* equal_out, const b200_sort_opts* opts) {
    Lowered lo;
    PrefixSrc ps;
    int rc = prefix_src(dev_ptr, n, desc, &lo, &ps);
    if (rc) return rc;
    if (nsplit == 0 || nsplit > uint32_t(MAX_SPLITTERS) || !splitters || !less_out || !equal_out) return B200_ERR_BAD_ARG;
    for (uint32_t g = 0; g < nsplit; ++g) {
        if (g && splitters[g] < splitters[g - 1]) return B200_ERR_BAD_ARG;
        ps.sp.split[g] = splitters[g];
        less_out[g] = equal_out[g] = 0;
    }
    ps.sp.nsplit = nsplit;
    if (n == 0) return B200_OK;
    if (!dev_ptr) return B200_ERR_BAD_ARG;
    DeviceCtx* c = nullptr;
    rc = get_ctx(opts ? opts->device : -1, &c);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard guard(c->dev);
    cudaStream_t st = (opts && opts->stream) ? static_cast<cudaStream_t>(opts->stream) : c->stream;
    unsigned long long* d_cnt = reinterpret_cast<unsigned long long*>(c->d_small + SMALL_SPLIT_OFF);
    CUDA_TRY(cudaMemsetAsync(d_cnt, 0, 256, st));
    const int grid = int(std::max<uint64_t>(1, std::min<uint64_t>(uint64_t(c->sm_count) * 8, (n + COUNT_THREADS - 1) / COUNT_THREADS)));
    splitter_count_kernel<<<grid, COUNT_THREADS, 0, st>>>(ps, d_cnt, d_cnt + 16);
    CUDA_TRY(cudaGetLastError());
    uint64_t h[32];
    CUDA_TRY(cudaMemcpyAsync(h, d_cnt, 256, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    for (uint32_t g = 0; g < nsplit; ++g) {
        less_out[g] = h[g];
        equal_out[g] = h[16 + g];
    }
    return B200_OK;
}

int b200_partition_splitters_device(const void* dev_src, void* dev_dst, void* const* dst_ptrs, size_t n,
                                    const b200_key_desc* desc, const uint64_t* splitters, uint32_t nsplit,
                                    uint32_t incl_mask, const uint64_t* part_counts, const b200_sort_opts* opts) {
    Lowered lo;
    int rc = lower_desc(desc, &lo);
    if (rc) return rc;
    SplitParams sp;
    rc = lower_prefix(lo, &sp);
    if (rc) return rc;
    if (nsplit == 0 || nsplit > uint32_t(MAX_SPLITTERS) || !splitters || !part_counts) return B200_ERR_BAD_ARG;
    const uint32_t parts = nsplit + 1;
    uint64_t total = 0;
    uint64_t h_base[RADIX] = {0};
    for (uint32_t q = 0; q < parts; ++q) {
        h_base[q] = dst_ptrs ? 0 : total;  // with per-part destinations every part starts at its own pointer
        total += part_counts[q];
    }
    if (total != n) return B200_ERR_BAD_ARG;
    // fold the tie rule into strict thresholds: right of boundary g  <=>  prefix > s_g - incl_g.
    // (s_g = 0 with incl: always right -> part0; thresholds must ascend, as the parts are ordered)
    uint32_t part0 = 0, nth = 0;
    uint64_t prev = 0;
    for (uint32_t g = 0; g < nsplit; ++g) {
        if (g && splitters[g] < splitters[g - 1]) return B200_ERR_BAD_ARG;
        const bool incl = (incl_mask >> g) & 1u;
        if (incl && splitters[g] == 0) {
            if (nth) return B200_ERR_BAD_ARG;
            ++part0;
            continue;
        }
        const uint64_t th = splitters[g] - (incl ? 1 : 0);
        if (nth && th < prev) return B200_ERR_BAD_ARG;  // incl flags of equal splitters must not increase
        sp.split[nth++] = prev = th;
    }
    for (uint32_t g = nth; g < uint32_t(MAX_SPLITTERS); ++g) sp.split[g] = ~0ull;
    sp.nsplit = nsplit;
    sp.incl_mask = incl_mask;
    sp.part0 = part0;
    sp.ncmp = nth <= 1 ? 1 : nth <= 3 ? 3 : nth <= 7 ? 7 : uint32_t(MAX_SPLITTERS);
    if (n == 0) return B200_OK;
    if (!dev_src || (!dev_dst && !dst_ptrs)) return B200_ERR_BAD_ARG;
    uint64_t h_dtab[RADIX] = {0};
    if (dst_ptrs) {
        for (uint32_t q = 0; q < parts; ++q) {
            if (!dst_ptrs[q]) return B200_ERR_BAD_ARG;
            h_dtab[q] = reinterpret_cast<uint64_t>(dst_ptrs[q]);
            if (h_dtab[q] % 16 != 0 && lo.elem_size % 16 == 0) return B200_ERR_BAD_ARG;
        }
        dev_dst = dst_ptrs[0];
    }
    DeviceCtx* c = nullptr;
    rc = get_ctx(opts ? opts->device : -1, &c);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard guard(c->dev);
    cudaStream_t st = (opts && opts->stream) ? static_cast<cudaStream_t>(opts->stream) : c->stream;
    PassPlan pp{};
    rc = plan_pass(*c, lo, n, dev_src, dst_ptrs ? dev_src : dev_dst, true, -1, &pp);
    if (rc) return rc;
    if (!pp.word_path) return B200_ERR_BAD_DESC;
    const size_t lb_bytes = size_t(pp.num_tiles) * RADIX * 8;
    rc = ensure_bytes(reinterpret_cast<void**>(&c->d_lookback), &c->lookback_bytes, lb_bytes);
    if (rc) return rc;
    uint64_t* d_pbase = reinterpret_cast<uint64_t*>(c->d_small + SMALL_PBASE_OFF);
    uint8_t* d_remap = c->d_small + SMALL_REMAP_OFF;
    uint32_t* d_ticket = reinterpret_cast<uint32_t*>(c->d_small + SMALL_TICKET_OFF);
    uint64_t* d_dtab = reinterpret_cast<uint64_t*>(c->d_small + SMALL_DTAB_OFF);
    CUDA_TRY(cudaMemsetAsync(c->d_lookback, 0, lb_bytes, st));
    CUDA_TRY(cudaMemsetAsync(d_ticket, 0, 1024, st));
    CUDA_TRY(cudaMemcpyAsync(d_pbase, h_base, sizeof(h_base), cudaMemcpyHostToDevice, st));
    if (dst_ptrs) CUDA_TRY(cudaMemcpyAsync(d_dtab, h_dtab, sizeof(h_dtab), cudaMemcpyHostToDevice, st));
    rc = launch_pass(*c, st, pp, lo, dev_src, dev_dst, n, 0, d_pbase, d_ticket, 1, d_remap, dst_ptrs ? d_dtab : nullptr, &sp,
                     nullptr, match_bits_for(parts), nullptr, nullptr, nullptr, nullptr, nullptr, parts);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(st));
    return B200_OK;
}

// ---- CUDA IPC helpers: one process per GPU shares its receive buffer with its peers ----
int b200_ipc_alloc(void** dev_ptr, size_t bytes, void* handle_out) {
    if (!dev_ptr || !handle_out) return B200_ERR_BAD_ARG;
    CUDA_TRY(cudaMalloc(dev_ptr, bytes ? bytes : 256));
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, *dev_ptr));
    memcpy(handle_out, &h, sizeof(h));
    return B200_OK;
}
int b200_ipc_open(const void* handle, void** dev_ptr) {
    if (!handle || !dev_ptr) return B200_ERR_BAD_ARG;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    CUDA_TRY(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return B200_OK;
}
int b200_ipc_close(void* dev_ptr) {
    if (dev_ptr) CUDA_TRY(cudaIpcCloseMemHandle(dev_ptr));
    return B200_OK;
}
int b200_ipc_free(void* dev_ptr) {
    if (dev_ptr) CUDA_TRY(cudaFree(dev_ptr));
    return B200_OK;
}

int b200_get_level(const void* record, const b200_key_desc* desc, uint32_t level) {
    Lowered lo;
    int rc = lower_desc(desc, &lo);
    if (rc) return -rc;
    if (!record) return -B200_ERR_BAD_ARG;
    if (level >= lo.levels) return -B200_ERR_LEVEL_RANGE;  // reference src/radix_key.rs:18
    return int(host_digit(static_cast<const uint8_t*>(record), lo, level));
}

int b200_host_alloc(void** out, size_t bytes) {
    if (!out) return B200_ERR_BAD_ARG;
    CUDA_TRY(cudaMallocHost(out, bytes ? bytes : 1));
    return B200_OK;
}
int b200_host_free(void* p) {
    if (p) CUDA_TRY(cudaFreeHost(p));
    return B200_OK;
}

int b200_release_scratch(int device) {
    DeviceCtx* c = nullptr;
    int rc = get_ctx(device, &c);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard guard(c->dev);
    void** ptrs[] = {&c->d_tmp, reinterpret_cast<void**>(&c->d_lookback), &c->d_stage,
                     reinterpret_cast<void**>(&c->d_sample), reinterpret_cast<void**>(&c->d_slow),
                     reinterpret_cast<void**>(&c->d_batch)};
    size_t* sizes[] = {&c->tmp_bytes, &c->lookback_bytes, &c->stage_bytes, &c->sample_bytes, &c->slow_bytes, &c->batch_bytes};
    for (int i = 0; i < 6; ++i) {
        if (*ptrs[i]) cudaFree(*ptrs[i]);
        *ptrs[i] = nullptr;
        *sizes[i] = 0;
    }
    return B200_OK;
}

int b200_hybrid_plan(const b200_key_desc* desc, const uint64_t* hist, size_t n, b200_window_plan* out) {
    Lowered lo;
    int rc = lower_desc(desc, &lo);
    if (rc) return rc;
    if (!hist || !out) return B200_ERR_BAD_ARG;
    memset(out, 0, sizeof(*out));
    SplitParams key;
    const uint32_t words = lo.elem_size / 4;
    if (lo.elem_size % 4 != 0 || (words != 1 && words != 2) || lo.levels > 8 || lower_prefix(lo, &key) != B200_OK) return B200_OK;
    WindowPlan wp;
    if (!plan_windows(lo, hist, n, &wp)) return B200_OK;
    out->eligible = 1;
    out->aligned = wp.aligned ? 1u : 0u;
    out->vary = wp.vary;
    out->digit_lo[0] = wp.w1_lo, out->digit_width[0] = 8;
    out->digit_lo[1] = wp.w2_lo, out->digit_width[1] = wp.w2_width;
    out->nlocal = wp.nlocal;
    for (uint32_t i = 0; i < wp.nlocal; ++i) out->local_lo[i] = wp.local_lo[i], out->local_width[i] = wp.local_width[i];
    return B200_OK;
}

int b200_route_plan(const b200_key_desc* desc, const uint64_t* hist, size_t n, b200_route* out) {
    Lowered lo;
    int rc = lower_desc(desc, &lo);
    if (rc) return rc;
    if (!out) return B200_ERR_BAD_ARG;
    memset(out, 0, sizeof(*out));
    if (n <= 1) {  // reference src/radix_sort_builder.rs:151
        out->route = B200_ROUTE_TRIVIAL;
        return B200_OK;
    }
    SplitParams key;
    if (small_shape_ok(lo, n, &key)) {
        out->route = B200_ROUTE_SMALL;
        out->active_levels = out->levels_run = lo.levels;  // no histogram is taken: every declared level is sorted
        return B200_OK;
    }
    if (!hist) return B200_ERR_BAD_ARG;
    std::vector<uint32_t> active;
    for (uint32_t l = 0; l < lo.levels; ++l) {
        bool single = false;
        for (int b = 0; b < RADIX; ++b)
            if (hist[size_t(l) * RADIX + b] == n) single = true;
        if (!single) active.push_back(l);
    }
    out->active_levels = out->levels_run = uint32_t(active.size());
    out->route = B200_ROUTE_LSD;
    if (active.empty()) return B200_OK;
    // hybrid: the checks sort_hybrid makes before its first launch
    const int mode = hybrid_mode();
    const uint32_t words = lo.elem_size / 4;
    if (mode != 0 && active.size() >= 3 && lo.elem_size % 4 == 0 && (words == 1 || words == 2) && lo.levels <= 8 &&
        uint64_t(n) < (1ull << 32) && lower_prefix(lo, &key) == B200_OK &&
        !(mode == 2 && (active.size() < (key_only_record(lo) ? 4u : 5u) || uint64_t(n) < hybrid_min_records(key_only_record(lo), words)))) {
        WindowPlan wp;
        if (plan_windows(lo, hist, n, &wp)) {
            bool ok = true;
            if (wp.aligned && mode == 2) {  // segments expected to fit an SM (independence estimate)
                uint64_t m1 = 0, m2 = 0;
                for (int b = 0; b < RADIX; ++b) {
                    m1 = std::max(m1, hist[size_t(wp.w1_lo / 8) * RADIX + b]);
                    m2 = std::max(m2, hist[size_t(wp.w2_lo / 8) * RADIX + b]);
                }
                ok = double(m1) * double(m2) / double(n) <= 0.97 * LS_CAP ||
                     (key_only_record(lo) && heavy_split_enabled() && m1 <= n / 4 && m2 <= n / 4);  // pivot splits (heavy.cuh)
            }
            if (ok) {
                out->route = B200_ROUTE_HYBRID;
                return B200_OK;
            }
        }
    }
    uint64_t max_seg = 0, nonempty = 0;
    if (msd1_shape_ok(lo, n, hist, active, &key, &max_seg, &nonempty)) {
        out->route = B200_ROUTE_MSD1;
        out->msd_level = active.back();
        out->max_bucket = max_seg;
    }
    return B200_OK;
}

int b200_variant_count(void) { return NUM_VARIANTS; }
const char* b200_variant_name(int variant) {
    if (variant < 0 || variant >= NUM_VARIANTS) return nullptr;
    return kVariants[variant].name;
}

const char* b200_strerror(int code) {
    switch (code) {
        case B200_OK: return "ok";
        case B200_ERR_LEVELS_ZERO: return "RadixKey::LEVELS is zero";
        case B200_ERR_BAD_DESC: return "inconsistent key descriptor";
        case B200_ERR_CUDA: return "CUDA runtime error";
        case B200_ERR_NO_DEVICE: return "no usable sm_100 device";
        case B200_ERR_ALLOC: return "device scratch allocation failed";
        case B200_ERR_BAD_ARG: return "bad argument";
        case B200_ERR_LEVEL_RANGE: return "level >= LEVELS";
        case B200_ERR_TOO_LARGE: return "more records than the static scratch holds";
    }
    return "unknown error";
}
const char* b200_last_cuda_error(void) { return g_last_cuda_error.c_str(); }
const char* b200_version(void) { return "b200sort 0.1.0 (sm_100a)"; }

}  // extern "C"

// b200_sort_device_hinted — b200_sort_device for a caller that knows which bits of the key image can differ between
// two records (a superset is fine).  The local sort of the sharded sort: the gathered histograms say which bits vary
// in every rank's range, so the receiver need not count its records again (K0 and K1 are skipped).
extern "C" int b200_sort_device_hinted(void* dev_ptr, void* dev_tmp, size_t n, const b200_key_desc* desc, uint64_t vary_bits,
                                       const b200_sort_opts* opts, b200_stats* stats) {
    Lowered lo;
    int rc = lower_desc(desc, &lo);
    if (rc) return rc;
    if (stats) memset(stats, 0, sizeof(*stats));
    if (opts && opts->route > B200_HINT_LOW_SCRATCH) return B200_ERR_BAD_ARG;
    if (n <= 1) return B200_OK;
    if (!dev_ptr) return B200_ERR_BAD_ARG;
    DeviceCtx* c = nullptr;
    rc = get_ctx(opts ? opts->device : -1, &c);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard guard(c->dev);
    cudaStream_t st = (opts && opts->stream) ? static_cast<cudaStream_t>(opts->stream) : c->stream;
    void* tmp = dev_tmp;
    if (!tmp) {
        rc = ensure_bytes(&c->d_tmp, &c->tmp_bytes, n * lo.elem_size);
        if (rc) return rc;
        tmp = c->d_tmp;
    }
    SortCtl ctl;
    ctl.variant = opts ? opts->variant : -1;
    ctl.route = opts ? opts->route : uint32_t(B200_HINT_AUTO);
    if (lo.levels <= 8) ctl.vary_hint = lo.levels == 8 ? vary_bits : (vary_bits & ((1ull << (8 * lo.levels)) - 1));
    return sort_device_impl(*c, st, dev_ptr, tmp, n, lo, ctl, nullptr, stats);
}

// ===========================================================================
// b200_sort_device_presplit — the local sort of the sharded sort when the exchange has already done MSD pass 1:
// the array arrives partitioned on `level` (bucket sizes `counts`, buckets in digit order; every level above is
// constant), so K0, K1 and the first MSD pass are skipped: joint histogram -> MSD pass 2 (dev_ptr -> dev_tmp) ->
// K3-U in dev_tmp.  The result is wherever it ends up (*result_out = dev_ptr or dev_tmp): a copy home would cost
// a quarter of what was saved.  Records that are not key-only, or segments that do not fit: the general sort runs
// from where the records are — same result, none of the saving.
// ===========================================================================
extern "C" int b200_sort_device_presplit(void* dev_ptr, void* dev_tmp, size_t n, const b200_key_desc* desc, uint32_t level,
                                         const uint64_t* counts, const b200_sort_opts* opts, b200_stats* stats, void** result_out) {
    Lowered lo;
    int rc = lower_desc(desc, &lo);
    if (rc) return rc;
    if (stats) memset(stats, 0, sizeof(*stats));
    if (level >= lo.levels) return B200_ERR_LEVEL_RANGE;
    if (!counts || !result_out) return B200_ERR_BAD_ARG;
    *result_out = dev_ptr;
    if (n <= 1) return B200_OK;
    if (!dev_ptr || !dev_tmp) return B200_ERR_BAD_ARG;
    DeviceCtx* cp_ = nullptr;
    rc = get_ctx(opts ? opts->device : -1, &cp_);
    if (rc) return rc;
    DeviceCtx& c = *cp_;
    std::lock_guard<std::mutex> lk(c.mu);
    DeviceGuard guard(c.dev);
    cudaStream_t st = (opts && opts->stream) ? static_cast<cudaStream_t>(opts->stream) : c.stream;
    SortCtl ctl;
    ctl.presplit_counts = counts;
    ctl.presplit_level = level;
    ctl.result_anywhere = true;
    bool done = false;
    void* where = dev_ptr;
    uint32_t launches = 0, slow = 0, local_levels = 0, unstable = 0;
    const std::vector<uint32_t> no_levels;
    CUDA_TRY(cudaEventRecord(c.ev[0], st));
    rc = sort_hybrid(c, st, dev_ptr, dev_tmp, n, lo, no_levels, reinterpret_cast<unsigned long long*>(c.d_small + SMALL_BASE_OFF),
                     reinterpret_cast<uint32_t*>(c.d_small + SMALL_TICKET_OFF), &done, &where, &launches, &slow, &local_levels,
                     &unstable, &ctl);
    if (rc) return rc;
    if (!done) {  // not this route's shape, or it gave up: `where` holds a permutation of the input
        SortCtl plain;
        plain.variant = opts ? opts->variant : -1;
        plain.result_anywhere = true;
        return sort_device_impl(c, st, where, where == dev_ptr ? dev_tmp : dev_ptr, n, lo, plain, result_out, stats);
    }
    CUDA_TRY(cudaEventRecord(c.ev[2], st));
    CUDA_TRY(cudaStreamSynchronize(st));
    *result_out = where;
    if (stats) {
        stats->levels_run = level + 1;
        stats->levels_skipped = lo.levels - (level + 1);
        stats->hybrid = 1;
        stats->unstable = unstable;
        stats->slow_segments = slow;
        stats->local_levels = local_levels;
        stats->heavy_segments = ctl.heavy_segments;
        stats->heavy_rounds = ctl.heavy_rounds;
        stats->kernel_launches = launches;
        stats->algorithmic_bytes = n * lo.elem_size * (1ull + 2ull * (level + 1));
        CUDA_TRY(cudaEventElapsedTime(&stats->ms_joint, c.ev[4], c.ev[5]));
        CUDA_TRY(cudaEventElapsedTime(&stats->ms_local, c.ev[6], c.ev[7]));
        CUDA_TRY(cudaEventElapsedTime(&stats->ms_total, c.ev[0], c.ev[2]));
        stats->ms_passes = stats->ms_total;
    }
    return B200_OK;
}

// ===========================================================================
// b200_sort_segments_device — many independent sorts in one call: every segment [seg_off[i], seg_off[i + 1]) of one
// device array is sorted on its own.  rdst's MSB-first recursion ends in exactly this: thousands of small buckets,
// each finished by an in-cache LSB sort or, below 128 records, a comparison sort (reference src/sorter.rs:121-157,
// src/sorts/lsb_sort.rs:38-128, src/sorts/comparative_sort.rs:29-119), one rayon task per bucket.  Here the segments
// that fit one SM (<= 16 384 records of 4 / 8 bytes) are claimed by the CTAs of ONE launch of the segment sort —
// K3-U for records that are nothing but their key (one wide counting-sort level + comparison fix-up; every segment
// cuts its digit from the highest bit in which ITS keys differ, the AUTOTOP instantiations, because the buckets of an
// MSD split share their leading bytes), the stable K3 over every declared level otherwise — and segments that do not
// fit take the whole-array sort, one after the other.
// ===========================================================================
extern "C" int b200_sort_segments_device(void* dev_ptr, size_t n, const uint64_t* seg_off, size_t nseg, const b200_key_desc* desc,
                                         const b200_sort_opts* opts, b200_stats* stats) {
    Lowered lo;
    int rc = lower_desc(desc, &lo);
    if (rc) return rc;
    if (stats) memset(stats, 0, sizeof(*stats));
    if (opts && opts->route > B200_HINT_LOW_SCRATCH) return B200_ERR_BAD_ARG;
    if (nseg == 0 || n == 0) return B200_OK;
    if (!seg_off || nseg > 0xfffffffeull || n > 0xffffffffull) return B200_ERR_BAD_ARG;
    // the segments, checked: offsets ascend and stay inside the array
    const char* tenv = getenv("B200SORT_TINY");  // read per call: tests flip it
    // a warp per segment (K3-T) up to 128 records for key-only records, 256 for records with a payload (whose
    // CTA-wide stable sort runs every declared level): measured, profiles/r9_segments_bench.jsonl
    const uint32_t tiny_cap = (tenv && atoi(tenv) == 0) ? 0u : (key_only_record(lo) ? 128u : uint32_t(TS_CAP));
    std::vector<uint2> tiny;   // 2 .. tiny_cap records: one warp each
    std::vector<uint2> pairs;  // .. 16 384 records: one CTA each
    std::vector<uint2> big;    // more: the whole-array sort, one by one
    uint32_t max_small = 0;
    for (size_t i = 0; i < nseg; ++i) {
        if (seg_off[i + 1] < seg_off[i] || seg_off[i + 1] > n) return B200_ERR_BAD_ARG;
        const uint64_t cnt = seg_off[i + 1] - seg_off[i];
        if (cnt <= 1) continue;  // reference src/radix_sort_builder.rs:151
        if (cnt <= tiny_cap) {
            tiny.push_back(make_uint2(uint32_t(seg_off[i]), uint32_t(cnt)));
        } else if (cnt <= uint64_t(LS_CAP)) {
            pairs.push_back(make_uint2(uint32_t(seg_off[i]), uint32_t(cnt)));
            max_small = std::max(max_small, uint32_t(cnt));
        } else {
            big.push_back(make_uint2(uint32_t(seg_off[i]), uint32_t(cnt)));
        }
    }
    if (tiny.empty() && pairs.empty() && big.empty()) return B200_OK;
    if (!dev_ptr) return B200_ERR_BAD_ARG;
    DeviceCtx* cp_ = nullptr;
    rc = get_ctx(opts ? opts->device : -1, &cp_);
    if (rc) return rc;
    DeviceCtx& c = *cp_;
    std::lock_guard<std::mutex> lk(c.mu);
    DeviceGuard guard(c.dev);
    cudaStream_t st = (opts && opts->stream) ? static_cast<cudaStream_t>(opts->stream) : c.stream;
    SplitParams key;
    const bool on_chip = (lo.elem_size == 4 || lo.elem_size == 8) && lo.levels <= uint32_t(LS_MAXLEV) && hybrid_mode() != 0 &&
                         (!opts || opts->variant < 0) && lower_prefix(lo, &key) == B200_OK && word_path_ok(lo, dev_ptr, dev_ptr);
    if (!on_chip) {  // shapes the segment sort does not take (2-byte keys, 12-byte records ...): every segment on its own
        big.insert(big.end(), pairs.begin(), pairs.end());
        big.insert(big.end(), tiny.begin(), tiny.end());
        pairs.clear();
        tiny.clear();
    }
    CUDA_TRY(cudaEventRecord(c.ev[10], st));
    uint32_t launches = 0, slow = 0, unstable = 0;
    const uint32_t words = lo.elem_size / 4;
    const size_t np = pairs.size(), nt = tiny.size();
    const size_t o_pairs = 0, o_tiny = np * 8, o_slow = o_tiny + nt * 8, o_tk = o_slow + (np + 1) * 4, total = (o_tk + 4096 + 15) / 16 * 16;
    if (np + nt) {
        rc = ensure_bytes(reinterpret_cast<void**>(&c.d_batch), &c.batch_bytes, total);
        if (rc) return rc;
    }
    if (nt) {  // K3-T: a warp per segment
        CUDA_TRY(cudaMemcpyAsync(c.d_batch + o_tiny, tiny.data(), nt * 8, cudaMemcpyHostToDevice, st));
        TinySortParams tp;
        memset(&tp, 0, sizeof(tp));
        tp.data = static_cast<uint32_t*>(dev_ptr);
        tp.seg_pairs = reinterpret_cast<const uint2*>(c.d_batch + o_tiny);
        tp.nseg = uint32_t(nt);
        tp.key = key;
        const int tgrid = int(std::min<uint64_t>((nt + TS_WARPS - 1) / TS_WARPS, uint64_t(c.sm_count) * 8));
        if (words == 1) {
            if (lo.is_float) tiny_sort_kernel<1, true><<<tgrid, TS_THREADS, 0, st>>>(tp);
            else tiny_sort_kernel<1, false><<<tgrid, TS_THREADS, 0, st>>>(tp);
        } else {
            if (lo.is_float) tiny_sort_kernel<2, true><<<tgrid, TS_THREADS, 0, st>>>(tp);
            else tiny_sort_kernel<2, false><<<tgrid, TS_THREADS, 0, st>>>(tp);
        }
        CUDA_TRY(cudaGetLastError());
        ++launches;
    }
    if (np) {
        uint8_t* d = c.d_batch;
        const size_t tk_at = (o_tk + 15) / 16 * 16;
        uint32_t* d_slow = reinterpret_cast<uint32_t*>(d + o_slow);
        uint32_t* d_tk = reinterpret_cast<uint32_t*>(d + tk_at);
        CUDA_TRY(cudaMemcpyAsync(d + o_pairs, pairs.data(), np * 8, cudaMemcpyHostToDevice, st));
        CUDA_TRY(cudaMemsetAsync(d_slow, 0, 4, st));
        CUDA_TRY(cudaMemsetAsync(d_tk, 0, 2048, st));
        LocalSortParams lp;
        memset(&lp, 0, sizeof(lp));
        lp.data = static_cast<uint32_t*>(dev_ptr);
        lp.seg_pairs = reinterpret_cast<const uint2*>(d + o_pairs);
        lp.nseg = uint32_t(np);
        lp.ticket = d_tk;
        lp.nlev = lo.levels;
        for (uint32_t l = 0; l < lo.levels; ++l) lp.lv[l] = window_params(lo, key, 8 * l, 8);
        lp.nfast = 0;  // a plain stable LSD over every declared level, all in shared memory
        lp.key = key;
        lp.slow_segments = d_tk + 300;
        if (key_only_record(lo)) {
            const int size = max_small <= 4096 ? 0 : (max_small <= 8192 ? 1 : 2);
            const LocalCountLaunch ll = words == 1 ? get_local_count_launch_w1(lo.is_float, size, true, true)
                                                   : get_local_count_launch_w2(lo.is_float, size, false, true);
            LocalCountParams cp;
            memset(&cp, 0, sizeof(cp));
            cp.data = static_cast<uint32_t*>(dev_ptr);
            cp.seg_pairs = lp.seg_pairs;
            cp.nseg = uint32_t(np);
            cp.ticket = d_tk + 16;
            cp.key = key;
            cp.top_bit = 8 * lo.levels;
            cp.bmax = uint32_t(ll.cbits);
            cp.slow_list = d_slow;
            CUDA_TRY(cudaFuncSetAttribute(ll.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(ll.smem)));
            int occ = 1;
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ll.kernel, ll.threads, ll.smem));
            const int ugrid = int(std::min<uint64_t>(uint64_t(c.sm_count) * std::max(occ, 1), np));
            ll.kernel<<<ugrid, ll.threads, ll.smem, st>>>(cp);
            CUDA_TRY(cudaGetLastError());
            lp.list = d_slow;  // what K3-U left behind (a long run of unequal keys in one counter): the stable K3, from the list
            unstable = 1;
            ++launches;
        }
        const int lgrid = int(std::min<uint64_t>(uint64_t(c.sm_count), np));
        rc = words == 1 ? launch_local_sort_w<1>(lo.is_float, lgrid, st, lp) : launch_local_sort_w<2>(lo.is_float, lgrid, st, lp);
        if (rc) return rc;
        ++launches;
        if (unstable) CUDA_TRY(cudaMemcpyAsync(&slow, d_slow, 4, cudaMemcpyDeviceToHost, st));
    }
    CUDA_TRY(cudaEventRecord(c.ev[11], st));
    // ---- segments that do not fit an SM: the whole-array sort on each (through an aligned staging copy when the
    // segment does not start on a 16-byte boundary, which the vector loads of the pass kernels need) ----
    if (!big.empty()) {
        uint32_t mx = 0;
        for (const uint2& b : big) mx = std::max(mx, b.y);
        rc = ensure_bytes(&c.d_tmp, &c.tmp_bytes, size_t(mx) * lo.elem_size);
        if (rc) return rc;
        for (const uint2& b : big) {
            uint8_t* at = static_cast<uint8_t*>(dev_ptr) + size_t(b.x) * lo.elem_size;
            const size_t bytes = size_t(b.y) * lo.elem_size;
            const bool staged = reinterpret_cast<uintptr_t>(at) % 16 != 0;
            void* where = at;
            if (staged) {
                rc = ensure_bytes(&c.d_stage, &c.stage_bytes, bytes);
                if (rc) return rc;
                CUDA_TRY(cudaMemcpyAsync(c.d_stage, at, bytes, cudaMemcpyDeviceToDevice, st));
                where = c.d_stage;
            }
            SortCtl ctl;
            ctl.variant = opts ? opts->variant : -1;
            ctl.route = opts ? opts->route : uint32_t(B200_HINT_AUTO);
            b200_stats sub;
            memset(&sub, 0, sizeof(sub));
            rc = sort_device_impl(c, st, where, c.d_tmp, b.y, lo, ctl, nullptr, &sub);
            if (rc) return rc;
            if (staged) CUDA_TRY(cudaMemcpyAsync(at, c.d_stage, bytes, cudaMemcpyDeviceToDevice, st));
            launches += sub.kernel_launches;
        }
    }
    CUDA_TRY(cudaEventRecord(c.ev[2], st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (stats) {
        stats->levels_run = lo.levels;
        stats->local_levels = (pairs.empty() && tiny.empty()) ? 0 : lo.levels;
        stats->kernel_launches = launches;
        stats->slow_segments = slow;
        stats->unstable = unstable;
        stats->tiles_per_pass = pairs.size() + tiny.size();   // segments sorted on chip
        stats->heavy_segments = uint32_t(big.size());  // segments that took the whole-array sort
        uint64_t recs = 0;
        for (const uint2& p : pairs) recs += p.y;
        for (const uint2& p : tiny) recs += p.y;
        for (const uint2& b : big) recs += b.y;
        stats->algorithmic_bytes = 2ull * recs * lo.elem_size;
        CUDA_TRY(cudaEventElapsedTime(&stats->ms_local, c.ev[10], c.ev[11]));
        CUDA_TRY(cudaEventElapsedTime(&stats->ms_total, c.ev[10], c.ev[2]));
        stats->ms_passes = stats->ms_total;
    }
    return B200_OK;
}

// ===========================================================================
// b200_sort_device_static — the sort with its scratch in a file-scope device array: no cudaMalloc, no
// cudaFree, no allocating library routine, not on the first call either.  No counterpart in the reference;
// added for callers that cannot allocate (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  It shares nothing with the entry points
// above except the kernels: no DeviceCtx pool, no library stream, no events, no pinned host buffers.
//
//   n <= 16 384 records of 4 / 8 bytes (the `small` route's shape): ONE launch of K3, the shared-memory segment
//     sort; its segment table and ticket live in the static array and are written by a one-CTA set-up kernel
//   otherwise: K1 (all level histograms) -> scan -> one stable K2 pass per level, ping-ponging between the
//     caller's buffer and the static array, with a device copy back when LEVELS is odd.  No histogram comes
//     back to the host, so nothing is skipped: every declared level is a pass.
//
// Everything is enqueued on the caller's stream (NULL = the legacy default stream) and the call returns
// without synchronising.  Every call on a device shares that device's copy of the array, which keeps the last
// call's contents: what has to start at zero (histograms, tile look-back table, tickets) is cleared on the
// stream by every call, and two calls that overlap on the same GPU race.
// ===========================================================================
namespace {

static_assert(B200_STATIC_MAX_RECORDS % 1024 == 0, "look-back table is sized in tiles of >= 1024 records");
constexpr size_t ST_HIST_OFF = 0;                                                  // [64][256] u64
constexpr size_t ST_BASE_OFF = ST_HIST_OFF + size_t(MAX_LEVELS) * RADIX * 8;       // [64][256] u64
constexpr size_t ST_TICKET_OFF = ST_BASE_OFF + size_t(MAX_LEVELS) * RADIX * 8;     // one u32 per pass, 1024 bytes
constexpr size_t ST_SEG_OFF = ST_TICKET_OFF + 1024;                                // K3: segment offsets [2] u64
constexpr size_t ST_K3TICKET_OFF = ST_SEG_OFF + 256;                               // K3: ticket + telemetry, 4096 bytes
constexpr size_t ST_LOOKBACK_OFF = ST_K3TICKET_OFF + 4096;                         // [tiles][256] u64, tiles of >= 1024 records
constexpr size_t ST_LOOKBACK_BYTES = size_t(B200_STATIC_MAX_RECORDS / 1024) * RADIX * 8;
constexpr size_t ST_TMP_OFF = ST_LOOKBACK_OFF + ST_LOOKBACK_BYTES;                 // the ping-pong buffer
constexpr size_t ST_BYTES = ST_TMP_OFF + size_t(B200_STATIC_MAX_BYTES);
static_assert(ST_TMP_OFF % 256 == 0 && ST_LOOKBACK_OFF % 256 == 0 && ST_SEG_OFF % 256 == 0, "sub-buffers stay 256-byte aligned");

}  // namespace

// ~20.3 MiB in every GPU that loads this module, until the process exits.  In HOST code this name is a
// stand-in, not a device address: static_scratch_base() below asks the runtime for the address.
static __device__ __align__(256) unsigned char g_static_scratch[ST_BYTES];

namespace {

// Device code may name the array directly: the set-up of the one-launch route.
__global__ void static_small_setup_kernel(unsigned long long n) {
    unsigned long long* seg = reinterpret_cast<unsigned long long*>(g_static_scratch + ST_SEG_OFF);
    uint32_t* tk = reinterpret_cast<uint32_t*>(g_static_scratch + ST_K3TICKET_OFF);
    for (uint32_t i = threadIdx.x; i < 1024u; i += blockDim.x) tk[i] = 0u;
    if (threadIdx.x == 0) {
        seg[0] = 0ull;
        seg[1] = n;
    }
}

struct StaticDev {
    bool ready = false;
    unsigned char* base = nullptr;  // this device's copy of g_static_scratch
    int sm_count = 0;
};
std::mutex g_static_mu;
StaticDev g_static_dev[MAX_DEVICES];

// Once per device, with that device current (the caller holds a DeviceGuard).
int static_scratch_base(int dev, StaticDev** out) {
    std::lock_guard<std::mutex> lk(g_static_mu);
    StaticDev& s = g_static_dev[dev];
    if (!s.ready) {
        int major = 0;
        CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        if (major < 10) {
            g_last_cuda_error = "device is not sm_100 class";
            return B200_ERR_NO_DEVICE;
        }
        CUDA_TRY(cudaDeviceGetAttribute(&s.sm_count, cudaDevAttrMultiProcessorCount, dev));
        void* p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p || reinterpret_cast<uintptr_t>(p) % 256 != 0) {
            g_last_cuda_error = "cudaGetSymbolAddress(g_static_scratch) returned an unusable address";
            return B200_ERR_CUDA;
        }
        s.base = static_cast<unsigned char*>(p);
        s.ready = true;
    }
    *out = &s;
    return B200_OK;
}

}  // namespace

extern "C" int b200_sort_device_static(void* dev_ptr, size_t n, const b200_key_desc* desc, const b200_sort_opts* opts) {
    Lowered lo;
    int rc = lower_desc(desc, &lo);
    if (rc) return rc;
    if (opts && opts->route != B200_HINT_AUTO) return B200_ERR_BAD_ARG;  // one route per size, nothing to hint
    if (opts && opts->variant >= NUM_VARIANTS) return B200_ERR_BAD_ARG;
    if (n <= 1) return B200_OK;  // reference src/radix_sort_builder.rs:151
    if (!dev_ptr) return B200_ERR_BAD_ARG;
    if (n > size_t(B200_STATIC_MAX_RECORDS) || n * size_t(lo.elem_size) > size_t(B200_STATIC_MAX_BYTES)) {
        char msg[256];
        snprintf(msg, sizeof(msg),
                 "b200_sort_device_static: %zu records of %u bytes exceed the static scratch (at most %u records and %u bytes); "
                 "use b200_sort_device",
                 n, lo.elem_size, unsigned(B200_STATIC_MAX_RECORDS), unsigned(B200_STATIC_MAX_BYTES));
        g_last_cuda_error = msg;
        return B200_ERR_TOO_LARGE;
    }
    int dev = opts ? opts->device : -1;
    if (dev < 0) CUDA_TRY(cudaGetDevice(&dev));
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        g_last_cuda_error = std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e);
        cudaGetLastError();
        return B200_ERR_NO_DEVICE;
    }
    if (dev >= count || dev >= MAX_DEVICES) return B200_ERR_BAD_ARG;
    DeviceGuard guard(dev);
    StaticDev* sd = nullptr;
    rc = static_scratch_base(dev, &sd);
    if (rc) return rc;
    unsigned char* base = sd->base;
    cudaStream_t st = opts ? static_cast<cudaStream_t>(opts->stream) : nullptr;

    // ---- n fits one SM: ONE launch of K3 over the whole array ----
    SplitParams key;
    if ((!opts || opts->variant < 0) && small_sort_ok(lo, dev_ptr, n, &key)) {
        static_small_setup_kernel<<<1, 256, 0, st>>>(n);
        CUDA_TRY(cudaGetLastError());
        uint32_t* d_tk = reinterpret_cast<uint32_t*>(base + ST_K3TICKET_OFF);
        LocalSortParams lp;
        memset(&lp, 0, sizeof(lp));
        lp.data = static_cast<uint32_t*>(dev_ptr);
        lp.seg_off = reinterpret_cast<const uint64_t*>(base + ST_SEG_OFF);
        lp.nseg = 1;
        lp.ticket = d_tk;
        lp.nlev = lo.levels;
        for (uint32_t l = 0; l < lo.levels; ++l) lp.lv[l] = window_params(lo, key, 8 * l, 8);
        lp.nfast = 0;  // a plain stable LSD over every level, all in shared memory
        lp.key = key;
        lp.slow_segments = d_tk + 300;
        return lo.elem_size == 4 ? launch_local_sort_w<1>(lo.is_float, 1, st, lp) : launch_local_sort_w<2>(lo.is_float, 1, st, lp);
    }

    // ---- K1 -> scan -> one stable pass per level ----
    DeviceCtx c;  // a view for the launch helpers, on the stack: SM count and the look-back table; owns nothing
    c.dev = dev;
    c.sm_count = sd->sm_count;
    c.d_lookback = reinterpret_cast<uint64_t*>(base + ST_LOOKBACK_OFF);
    c.lookback_bytes = ST_LOOKBACK_BYTES;
    unsigned long long* d_hist = reinterpret_cast<unsigned long long*>(base + ST_HIST_OFF);
    unsigned long long* d_base = reinterpret_cast<unsigned long long*>(base + ST_BASE_OFF);
    uint32_t* d_ticket = reinterpret_cast<uint32_t*>(base + ST_TICKET_OFF);
    void* tmp = base + ST_TMP_OFF;
    PassPlan pp{};
    rc = plan_pass(c, lo, n, dev_ptr, tmp, false, opts ? opts->variant : -1, &pp);
    if (rc) return rc;
    const size_t lb_bytes = size_t(pp.num_tiles) * RADIX * 8;
    if (lb_bytes > ST_LOOKBACK_BYTES) {
        g_last_cuda_error = "b200_sort_device_static: tile look-back table does not fit the static scratch";
        return B200_ERR_TOO_LARGE;
    }
    rc = launch_histogram(c, st, lo, dev_ptr, n, d_hist);  // clears the histograms first
    if (rc) return rc;
    scan_hist_kernel<<<lo.levels, RADIX, 0, st>>>(d_hist, d_base);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemsetAsync(c.d_lookback, 0, lb_bytes, st));
    CUDA_TRY(cudaMemsetAsync(d_ticket, 0, 1024, st));
    void* src = dev_ptr;
    void* dst = tmp;
    for (uint32_t l = 0; l < lo.levels; ++l) {
        rc = launch_pass(c, st, pp, lo, src, dst, n, l, reinterpret_cast<const uint64_t*>(d_base + size_t(l) * RADIX),
                         d_ticket + l, l + 1, nullptr);
        if (rc) return rc;
        std::swap(src, dst);
    }
    if (src != dev_ptr) CUDA_TRY(cudaMemcpyAsync(dev_ptr, src, n * lo.elem_size, cudaMemcpyDeviceToDevice, st));
    return B200_OK;
}
