// multi.cu — the sharded sort across the GPUs of one box behind the C ABI: b200_sort_multi_device /
// b200_sort_multi_host (include/b200sort.h), plus the host-side exchange planner they share with the
// one-process-per-GPU driver (rdst_b200/dist.py calls the same b200_multi_plan_* entry points).
//
// The reference has no distributed path: RadixSortBuilder::sort takes one `&mut [T]` in host RAM
// (reference src/radix_sort_builder.rs:149-157).  This is the widening BASELINE.json configs[4]
// names, in the form a Rust binding can call: ONE process, one host thread and one stream per GPU,
// peer access between the devices (cudaDeviceEnablePeerAccess), and the same fused partition +
// exchange kernel as the multi-process path — every GPU's K2 range-partition pass stores its runs
// straight into the owners' receive buffers over NVLink — then an independent local sort per GPU.
// Steps: K1 per shard -> histograms merged on the host -> partition level and P-1
// splitters (digit boundaries; sample-sort splitters when one digit is heavier than a rank's share)
// -> fused partition/exchange -> local sort.  Stable: records travel in (source rank, source
// position) order and every step is stable.
#include "../../include/b200sort.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <numeric>
#include <string>
#include <thread>
#include <vector>

namespace {

constexpr int RADIX = 256;
constexpr int MAX_GPUS = 16;
constexpr uint32_t SAMPLES_PER_RANK = 8192;
constexpr uint32_t MAX_SPLIT = 15;

struct Barrier {
    std::mutex m;
    std::condition_variable cv;
    int n, waiting = 0;
    uint64_t gen = 0;
    explicit Barrier(int n_) : n(n_) {}
    void wait() {
        std::unique_lock<std::mutex> lk(m);
        const uint64_t g = gen;
        if (++waiting == n) {
            waiting = 0;
            ++gen;
            cv.notify_all();
        } else {
            cv.wait(lk, [&] { return gen != g; });
        }
    }
};

double now_ms() {
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

// ---- the planner (mirrors the numpy model in rdst_b200/dist.py; tests/test_multi_plan_cpu.py pins both) ----
// Most significant level whose GLOBAL histogram is not a single bucket (rdst's trivial-level skip,
// reference src/sorts/ska_sort.rs:52-54, applied to the splitters); -1 = every key equal.
int pick_partition_level(const uint64_t* gh, uint32_t levels) {
    uint64_t n = 0;
    for (int b = 0; b < RADIX; ++b) n += gh[b];
    for (int l = int(levels) - 1; l >= 0; --l) {
        uint64_t mx = 0;
        for (int b = 0; b < RADIX; ++b) mx = std::max(mx, gh[size_t(l) * RADIX + b]);
        if (mx != n) return l;
    }
    return -1;
}

// P-1 digit boundaries on one 256-bin row: part g = digits in [split[g-1], split[g]).  Boundary g sits
// where the running count is closest to g * N / P (first minimiser at or after the previous boundary).
void choose_splitters(const uint64_t* row, uint32_t parts, uint32_t* split) {
    uint64_t cum[RADIX + 1];
    cum[0] = 0;
    for (int d = 0; d < RADIX; ++d) cum[d + 1] = cum[d] + row[d];
    const uint64_t total = cum[RADIX];
    uint32_t lo = 0;
    for (uint32_t g = 1; g < parts; ++g) {
        const double target = double(total * g) / double(parts);
        uint32_t best = lo;
        double bd = std::abs(double(cum[lo]) - target);
        for (uint32_t d = lo + 1; d <= uint32_t(RADIX); ++d) {
            const double dd = std::abs(double(cum[d]) - target);
            if (dd < bd) bd = dd, best = d;
        }
        split[g - 1] = best;
        lo = best;
    }
}

void part_of_digit(const uint32_t* split, uint32_t parts, uint32_t* pod) {
    for (uint32_t d = 0; d < uint32_t(RADIX); ++d) {
        uint32_t part = 0;
        while (part + 1 < parts && d >= split[part]) ++part;
        pod[d] = part;
    }
}

// rows[r] = rank r's 256-bin row on the partition level -> m[r][g] = records rank r sends to rank g
void send_matrix(const uint64_t* rows, uint32_t ranks, const uint32_t* split, uint32_t parts, int64_t* m) {
    uint32_t pod[RADIX];
    part_of_digit(split, parts, pod);
    std::fill(m, m + size_t(ranks) * parts, int64_t(0));
    for (uint32_t r = 0; r < ranks; ++r)
        for (int d = 0; d < RADIX; ++d) m[size_t(r) * parts + pod[d]] += int64_t(rows[size_t(r) * RADIX + d]);
}

double matrix_imbalance(const int64_t* m, uint32_t ranks, uint32_t parts) {
    int64_t total = 0, mx = 0;
    for (uint32_t g = 0; g < parts; ++g) {
        int64_t col = 0;
        for (uint32_t r = 0; r < ranks; ++r) col += m[size_t(r) * parts + g];
        total += col;
        mx = std::max(mx, col);
    }
    return total ? double(mx) * parts / double(total) : 1.0;
}

// P-1 splitter prefixes = weighted quantiles g/P of the gathered samples (weights = records a sample
// stands for).  Duplicates are kept: a value heavier than one part yields equal splitters.
void choose_sample_splitters(const uint64_t* samples, const double* weights, size_t count, uint32_t parts, uint64_t* out) {
    std::vector<size_t> order(count);
    std::iota(order.begin(), order.end(), size_t(0));
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return samples[a] < samples[b]; });
    std::vector<double> cum(count);
    double run = 0;
    for (size_t i = 0; i < count; ++i) run += weights[order[i]], cum[i] = run;
    const double total = run;
    for (uint32_t g = 1; g < parts; ++g) {
        const double x = total * g / parts;
        size_t idx = size_t(std::lower_bound(cum.begin(), cum.end(), x) - cum.begin());
        idx = std::min(idx, count - 1);
        out[g - 1] = samples[order[idx]];
    }
}

// Tie thresholds t[S] and the send matrix m[R][P] for the splitter partition: boundary b puts a record
// LEFT when prefix < s_b, or prefix == s_b and source rank < t_b; t_b (0..R) is the threshold whose
// left count is closest to (b + 1) * N / P, non-decreasing over equal splitters.  exact_ties = 0 (keys
// longer than the 8-level prefix): equal prefixes are not equal keys, a tie group is never split.
int plan_splitter_exchange(const int64_t* less, const int64_t* equal, const int64_t* n_per_rank, const uint64_t* splitters,
                           uint32_t ranks, uint32_t parts, int exact_ties, int64_t* t, int64_t* m) {
    const uint32_t S = parts - 1;
    int64_t total = 0;
    for (uint32_t r = 0; r < ranks; ++r) total += n_per_rank[r];
    for (uint32_t b = 0; b < S; ++b) {
        const double target = double(total * int64_t(b + 1)) / double(parts);
        int64_t base = 0;
        for (uint32_t r = 0; r < ranks; ++r) base += less[size_t(r) * S + b];
        int64_t best_t = 0;
        double bd = 0;
        int64_t cand = base;
        for (uint32_t tt = 0; tt <= ranks; ++tt) {  // left count for t = tt
            if (tt) cand += equal[size_t(tt - 1) * S + b];
            if (!exact_ties && tt != 0 && tt != ranks) continue;
            const double dd = std::abs(double(cand) - target);
            if (tt == 0 || dd < bd) bd = dd, best_t = tt;
        }
        if (b && splitters[b] == splitters[b - 1]) best_t = std::max(best_t, t[b - 1]);
        t[b] = best_t;
    }
    for (uint32_t r = 0; r < ranks; ++r) {
        int64_t prev = 0;
        for (uint32_t b = 0; b < S; ++b) {
            const int64_t left = less[size_t(r) * S + b] + (int64_t(r) < t[b] ? equal[size_t(r) * S + b] : 0);
            if (left < prev) return B200_ERR_BAD_ARG;  // splitter boundaries out of order
            m[size_t(r) * parts + b] = left - prev;
            prev = left;
        }
        if (n_per_rank[r] < prev) return B200_ERR_BAD_ARG;
        m[size_t(r) * parts + S] = n_per_rank[r] - prev;
    }
    return B200_OK;
}

uint32_t incl_mask_for_rank(const int64_t* t, uint32_t S, uint32_t rank) {
    uint32_t mask = 0;
    for (uint32_t b = 0; b < S; ++b)
        if (int64_t(rank) >= t[b]) mask |= 1u << b;
    return mask;
}

// ---- per-process multi-GPU state ----
struct MultiState {
    std::mutex mu;
    void* recv[MAX_GPUS] = {};
    size_t recv_bytes[MAX_GPUS] = {};
    void* in[MAX_GPUS] = {};   // host entry point: device copy of a shard
    size_t in_bytes[MAX_GPUS] = {};
    void* tmp[MAX_GPUS] = {};  // scratch of the local sort
    size_t tmp_bytes[MAX_GPUS] = {};
    int dev_of[MAX_GPUS] = {-1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1};
    bool peers_on[64][64] = {};
};
MultiState g_multi;
thread_local std::string g_multi_err;

bool grow(void** p, size_t* have, size_t need) {
    if (*p && *have >= need) return true;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *have = 0;
    const size_t want = need + need / 16 + 256;
    if (cudaMalloc(p, want) != cudaSuccess) {
        cudaGetLastError();
        if (cudaMalloc(p, need) != cudaSuccess) {
            cudaGetLastError();
            *p = nullptr;
            return false;
        }
        *have = need;
        return true;
    }
    *have = want;
    return true;
}

struct Shared {  // what the per-GPU threads of one call share
    int P = 0;
    const int* devices = nullptr;
    const b200_key_desc* desc = nullptr;
    uint32_t L = 0, es = 0;
    std::vector<uint64_t> hist;       // [P][L][256]
    std::vector<uint64_t> samples;    // [P][SAMPLES_PER_RANK]
    std::vector<uint32_t> nsamp;      // [P]
    std::vector<int64_t> less, equal; // [P][S]
    std::vector<int64_t> m;           // [P][P]
    std::vector<size_t> n_in;
    int level = -1;
    uint32_t split[MAX_GPUS] = {};
    bool prefix_mode = false;
    uint64_t splitters[MAX_SPLIT] = {};
    int64_t tthr[MAX_SPLIT] = {};
    double imbalance = 1.0;
    int rc[MAX_GPUS] = {};
    double t_plan = 0, t_part = 0, t_sort = 0;
    b200_stats local[MAX_GPUS];
};

int first_error(const Shared& sh) {
    for (int g = 0; g < sh.P; ++g)
        if (sh.rc[g]) return sh.rc[g];
    return B200_OK;
}

// One GPU's share of the sharded sort.  `src` = this GPU's shard (device memory), n records.
void worker(Shared& sh, Barrier& bar, int g, const void* src, double balance_tol) {
    const int P = sh.P, dev = sh.devices[g];
    const uint32_t L = sh.L, es = sh.es;
    const size_t n = sh.n_in[g];
    b200_sort_opts o{};
    o.stable = 1, o.variant = -1, o.device = dev, o.stream = nullptr;
    auto fail = [&](int rc) { sh.rc[g] = rc; };
    cudaSetDevice(dev);
    const double t0 = now_ms();
    // 1. K1 on the shard
    int rc = b200_histogram_device(src, n, sh.desc, &o, sh.hist.data() + size_t(g) * L * RADIX);
    if (rc) fail(rc);
    bar.wait();
    if (first_error(sh)) return;
    // 2. global histogram, partition level, digit splitters (thread 0 plans, everyone reads)
    if (g == 0) {
        std::vector<uint64_t> gh(size_t(L) * RADIX, 0);
        for (int r = 0; r < P; ++r)
            for (size_t i = 0; i < gh.size(); ++i) gh[i] += sh.hist[size_t(r) * L * RADIX + i];
        sh.level = pick_partition_level(gh.data(), L);
        if (sh.level >= 0) {
            choose_splitters(gh.data() + size_t(sh.level) * RADIX, uint32_t(P), sh.split);
            std::vector<uint64_t> rows(size_t(P) * RADIX);
            for (int r = 0; r < P; ++r)
                memcpy(rows.data() + size_t(r) * RADIX, sh.hist.data() + (size_t(r) * L + sh.level) * RADIX, RADIX * 8);
            send_matrix(rows.data(), uint32_t(P), sh.split, uint32_t(P), sh.m.data());
            sh.imbalance = matrix_imbalance(sh.m.data(), uint32_t(P), uint32_t(P));
            sh.prefix_mode = sh.imbalance > balance_tol && uint32_t(P - 1) <= MAX_SPLIT && b200_key_prefix_supported(sh.desc);
        }
    }
    bar.wait();
    if (sh.level < 0) return;  // every key equal: the shards are the result
    // 2b. skewed keys: sample-sort splitters on the 64-bit key prefix
    if (sh.prefix_mode) {
        const uint32_t cnt = uint32_t(std::min<size_t>(n, SAMPLES_PER_RANK));
        const size_t stride = cnt ? std::max<size_t>(1, n / cnt) : 1;
        sh.nsamp[g] = cnt;
        if (cnt) {
            rc = b200_key_prefix_sample_device(src, n, sh.desc, stride, cnt, sh.samples.data() + size_t(g) * SAMPLES_PER_RANK, &o);
            if (rc) fail(rc);
        }
        bar.wait();
        if (first_error(sh)) return;
        if (g == 0) {
            std::vector<uint64_t> v;
            std::vector<double> w;
            for (int r = 0; r < P; ++r)
                for (uint32_t i = 0; i < sh.nsamp[r]; ++i) {
                    v.push_back(sh.samples[size_t(r) * SAMPLES_PER_RANK + i]);
                    w.push_back(double(sh.n_in[r]) / double(sh.nsamp[r]));
                }
            if (v.empty()) sh.prefix_mode = false;
            else choose_sample_splitters(v.data(), w.data(), v.size(), uint32_t(P), sh.splitters);
        }
        bar.wait();
    }
    if (sh.prefix_mode) {
        const uint32_t S = uint32_t(P - 1);
        std::vector<uint64_t> le(S), eq(S);
        rc = b200_splitter_count_device(src, n, sh.desc, sh.splitters, S, le.data(), eq.data(), &o);
        if (rc) fail(rc);
        for (uint32_t b = 0; b < S; ++b) sh.less[size_t(g) * S + b] = int64_t(le[b]), sh.equal[size_t(g) * S + b] = int64_t(eq[b]);
        bar.wait();
        if (first_error(sh)) return;
        if (g == 0) {
            std::vector<int64_t> npr(P);
            for (int r = 0; r < P; ++r) npr[r] = int64_t(sh.n_in[r]);
            const int prc = plan_splitter_exchange(sh.less.data(), sh.equal.data(), npr.data(), sh.splitters, uint32_t(P), uint32_t(P),
                                                   sh.desc->levels <= 8, sh.tthr, sh.m.data());
            if (prc) fail(prc);
            sh.imbalance = matrix_imbalance(sh.m.data(), uint32_t(P), uint32_t(P));
        }
        bar.wait();
        if (first_error(sh)) return;
    }
    // 3. receive buffers (this GPU's own), then every rank's destination pointers
    size_t n_recv = 0;
    for (int r = 0; r < P; ++r) n_recv += size_t(sh.m[size_t(r) * P + g]);
    if (!grow(&g_multi.recv[g], &g_multi.recv_bytes[g], std::max<size_t>(n_recv, 1) * es) ||
        !grow(&g_multi.tmp[g], &g_multi.tmp_bytes[g], std::max<size_t>(n_recv, 1) * es))
        fail(B200_ERR_ALLOC);
    if (g == 0) sh.t_plan = now_ms() - t0;
    bar.wait();  // all receive buffers exist; nobody still reads the previous contents
    if (first_error(sh)) return;
    const double t1 = now_ms();
    void* dst[MAX_GPUS];
    std::vector<uint64_t> pc(P);
    for (int q = 0; q < P; ++q) {
        size_t start = 0;  // this rank's records for q begin after those of the lower ranks
        for (int r = 0; r < g; ++r) start += size_t(sh.m[size_t(r) * P + q]);
        dst[q] = static_cast<uint8_t*>(g_multi.recv[q]) + start * es;
        pc[q] = uint64_t(sh.m[size_t(g) * P + q]);
    }
    // 4. fused partition + exchange: the pass kernel stores its runs straight into the owners' buffers
    if (n) {
        if (sh.prefix_mode) {
            rc = b200_partition_splitters_device(src, nullptr, dst, n, sh.desc, sh.splitters, uint32_t(P - 1),
                                                 incl_mask_for_rank(sh.tthr, uint32_t(P - 1), uint32_t(g)), pc.data(), &o);
        } else {
            std::vector<uint64_t> got(P);
            rc = b200_partition_peers_device(src, dst, n, sh.desc, uint32_t(sh.level),
                                             sh.hist.data() + (size_t(g) * L + sh.level) * RADIX, sh.split, uint32_t(P), got.data(), &o);
            if (!rc && got != pc) rc = B200_ERR_BAD_ARG;
        }
        if (rc) fail(rc);
    }
    bar.wait();  // every rank's partition kernel has finished: the received range is complete
    if (first_error(sh)) return;
    if (g == 0) sh.t_part = now_ms() - t1;
    const double t2 = now_ms();
    // 5. independent local sort of the received range
    if (n_recv > 1) {
        rc = b200_sort_device(g_multi.recv[g], g_multi.tmp[g], n_recv, sh.desc, &o, &sh.local[g]);
        if (rc) fail(rc);
    }
    bar.wait();
    if (g == 0) sh.t_sort = now_ms() - t2;
}

int enable_peers(const int* devices, int P) {
    for (int a = 0; a < P; ++a) {
        if (cudaSetDevice(devices[a]) != cudaSuccess) return B200_ERR_NO_DEVICE;
        for (int b = 0; b < P; ++b) {
            if (a == b || g_multi.peers_on[devices[a]][devices[b]]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[a], devices[b]);
            if (!can) {
                g_multi_err = "no peer access between devices " + std::to_string(devices[a]) + " and " + std::to_string(devices[b]);
                return B200_ERR_NO_DEVICE;
            }
            const cudaError_t e = cudaDeviceEnablePeerAccess(devices[b], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                g_multi_err = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e);
                cudaGetLastError();
                return B200_ERR_CUDA;
            }
            cudaGetLastError();
            g_multi.peers_on[devices[a]][devices[b]] = true;
        }
    }
    return B200_OK;
}

int check_args(int ngpus, const int* devices, const b200_key_desc* desc, int* devs_out) {
    if (ngpus < 1 || ngpus > MAX_GPUS || !desc) return B200_ERR_BAD_ARG;
    if (desc->levels == 0) return B200_ERR_LEVELS_ZERO;
    if (desc->elem_size % 4 != 0 || desc->elem_size > 16) return B200_ERR_BAD_DESC;  // fused exchange: 4..16-byte records
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        cudaGetLastError();
        return B200_ERR_NO_DEVICE;
    }
    for (int g = 0; g < ngpus; ++g) {
        devs_out[g] = devices ? devices[g] : g;
        if (devs_out[g] < 0 || devs_out[g] >= count || devs_out[g] >= 64) return B200_ERR_BAD_ARG;
        for (int h = 0; h < g; ++h)
            if (devs_out[h] == devs_out[g]) return B200_ERR_BAD_ARG;
    }
    return B200_OK;
}

int run_sharded(int P, const int* devs, const b200_key_desc* desc, const void* const* src, const size_t* n_in, double balance_tol,
                Shared& sh) {
    sh.P = P;
    sh.devices = devs;
    sh.desc = desc;
    sh.L = desc->levels;
    sh.es = desc->elem_size;
    sh.hist.assign(size_t(P) * sh.L * RADIX, 0);
    sh.samples.assign(size_t(P) * SAMPLES_PER_RANK, 0);
    sh.nsamp.assign(P, 0);
    sh.less.assign(size_t(P) * std::max(1, P - 1), 0);
    sh.equal.assign(size_t(P) * std::max(1, P - 1), 0);
    sh.m.assign(size_t(P) * P, 0);
    sh.n_in.assign(n_in, n_in + P);
    memset(sh.local, 0, sizeof(sh.local));
    for (int g = 0; g < P; ++g) {  // a slot's cached buffers live on the device that used the slot last
        if (g_multi.dev_of[g] != devs[g]) {
            void** ps[] = {&g_multi.recv[g], &g_multi.tmp[g], &g_multi.in[g]};
            size_t* ss[] = {&g_multi.recv_bytes[g], &g_multi.tmp_bytes[g], &g_multi.in_bytes[g]};
            for (int i = 0; i < 3; ++i) {
                if (*ps[i]) cudaFree(*ps[i]);
                *ps[i] = nullptr;
                *ss[i] = 0;
            }
            g_multi.dev_of[g] = devs[g];
        }
    }
    Barrier bar(P);
    std::vector<std::thread> th;
    for (int g = 0; g < P; ++g) th.emplace_back(worker, std::ref(sh), std::ref(bar), g, src[g], balance_tol);
    for (auto& t : th) t.join();
    return first_error(sh);
}

void fill_stats(const Shared& sh, b200_multi_stats* st, const size_t* n_out) {
    if (!st) return;
    st->ngpus = uint32_t(sh.P);
    st->level = sh.level < 0 ? 0xffffffffu : uint32_t(sh.level);
    st->splitter_mode = sh.prefix_mode ? 1u : 0u;
    st->imbalance = sh.imbalance;
    st->ms_plan = float(sh.t_plan);
    st->ms_exchange = float(sh.t_part);
    st->ms_local_sort = float(sh.t_sort);
    uint64_t sent = 0;
    for (int g = 0; g < sh.P; ++g) {
        uint64_t s = 0;
        for (int q = 0; q < sh.P; ++q)
            if (q != g) s += uint64_t(sh.m[size_t(g) * sh.P + q]);
        sent = std::max(sent, s * sh.es);
        st->n_out[g] = n_out[g];
        st->kernel_launches += sh.local[g].kernel_launches + 2;  // + K1 and the partition pass
    }
    st->max_sent_bytes = sent;
    st->local0 = sh.local[0];
}

}  // namespace

extern "C" {

int b200_multi_plan_digits(const uint64_t* all_hist, uint32_t ranks, uint32_t levels, int32_t* level_out, uint32_t* split_out,
                           int64_t* send_matrix_out, double* imbalance_out) {
    if (!all_hist || !level_out || !split_out || !send_matrix_out || ranks < 1 || ranks > uint32_t(MAX_GPUS) || levels < 1)
        return B200_ERR_BAD_ARG;
    std::vector<uint64_t> gh(size_t(levels) * RADIX, 0);
    for (uint32_t r = 0; r < ranks; ++r)
        for (size_t i = 0; i < gh.size(); ++i) gh[i] += all_hist[size_t(r) * levels * RADIX + i];
    const int level = pick_partition_level(gh.data(), levels);
    *level_out = level;
    if (level < 0) return B200_OK;
    choose_splitters(gh.data() + size_t(level) * RADIX, ranks, split_out);
    std::vector<uint64_t> rows(size_t(ranks) * RADIX);
    for (uint32_t r = 0; r < ranks; ++r)
        memcpy(rows.data() + size_t(r) * RADIX, all_hist + (size_t(r) * levels + level) * RADIX, RADIX * 8);
    send_matrix(rows.data(), ranks, split_out, ranks, send_matrix_out);
    if (imbalance_out) *imbalance_out = matrix_imbalance(send_matrix_out, ranks, ranks);
    return B200_OK;
}

int b200_multi_choose_sample_splitters(const uint64_t* samples, const double* weights, size_t count, uint32_t parts,
                                       uint64_t* splitters_out) {
    if (!samples || !weights || !splitters_out || count == 0 || parts < 2 || parts - 1 > MAX_SPLIT) return B200_ERR_BAD_ARG;
    choose_sample_splitters(samples, weights, count, parts, splitters_out);
    return B200_OK;
}

int b200_multi_plan_splitters(const int64_t* less, const int64_t* equal, const int64_t* n_per_rank, const uint64_t* splitters,
                              uint32_t ranks, uint32_t parts, int exact_ties, int64_t* tie_thresholds_out,
                              int64_t* send_matrix_out, uint32_t* incl_masks_out) {
    if (!less || !equal || !n_per_rank || !splitters || !tie_thresholds_out || !send_matrix_out || parts < 2 ||
        parts - 1 > MAX_SPLIT || ranks < 1 || ranks > uint32_t(MAX_GPUS))
        return B200_ERR_BAD_ARG;
    const int rc = plan_splitter_exchange(less, equal, n_per_rank, splitters, ranks, parts, exact_ties, tie_thresholds_out,
                                          send_matrix_out);
    if (rc) return rc;
    if (incl_masks_out)
        for (uint32_t r = 0; r < ranks; ++r) incl_masks_out[r] = incl_mask_for_rank(tie_thresholds_out, parts - 1, r);
    return B200_OK;
}

int b200_sort_multi_device(void* const* dev_ptrs, const size_t* n_in, int ngpus, const int* devices, const b200_key_desc* desc,
                           const b200_sort_opts* opts, void** out_ptrs, size_t* n_out, b200_multi_stats* stats) {
    (void)opts;
    int devs[MAX_GPUS];
    int rc = check_args(ngpus, devices, desc, devs);
    if (rc) return rc;
    if (!dev_ptrs || !n_in || !out_ptrs || !n_out) return B200_ERR_BAD_ARG;
    if (stats) memset(stats, 0, sizeof(*stats));
    std::lock_guard<std::mutex> lk(g_multi.mu);
    int prev = -1;
    cudaGetDevice(&prev);
    rc = enable_peers(devs, ngpus);
    if (rc == B200_OK) {
        Shared sh;
        const double t0 = now_ms();
        rc = run_sharded(ngpus, devs, desc, reinterpret_cast<const void* const*>(dev_ptrs), n_in, 1.25, sh);
        if (rc == B200_OK) {
            for (int g = 0; g < ngpus; ++g) {
                if (sh.level < 0) {  // every key equal: the shards are the result
                    out_ptrs[g] = dev_ptrs[g];
                    n_out[g] = n_in[g];
                } else {
                    size_t nr = 0;
                    for (int r = 0; r < ngpus; ++r) nr += size_t(sh.m[size_t(r) * ngpus + g]);
                    out_ptrs[g] = g_multi.recv[g];
                    n_out[g] = nr;
                }
            }
            fill_stats(sh, stats, n_out);
            if (stats) stats->ms_total = float(now_ms() - t0);
        }
    }
    if (prev >= 0) cudaSetDevice(prev);
    return rc;
}

int b200_sort_multi_host(void* host_ptr, size_t n, int ngpus, const int* devices, const b200_key_desc* desc,
                         const b200_sort_opts* opts, b200_multi_stats* stats) {
    (void)opts;
    int devs[MAX_GPUS];
    int rc = check_args(ngpus, devices, desc, devs);
    if (rc) return rc;
    if (stats) memset(stats, 0, sizeof(*stats));
    if (n <= 1) return B200_OK;  // reference src/radix_sort_builder.rs:151
    if (!host_ptr) return B200_ERR_BAD_ARG;
    std::lock_guard<std::mutex> lk(g_multi.mu);
    int prev = -1;
    cudaGetDevice(&prev);
    rc = enable_peers(devs, ngpus);
    const size_t es = desc->elem_size;
    const int P = ngpus;
    std::vector<size_t> n_in(P), off(P + 1, 0);
    for (int g = 0; g < P; ++g) {
        n_in[g] = n / P + (size_t(g) < n % P ? 1 : 0);
        off[g + 1] = off[g] + n_in[g];
    }
    for (int g = 0; g < P; ++g) {
        if (g_multi.dev_of[g] != devs[g]) {
            void** ps[] = {&g_multi.recv[g], &g_multi.tmp[g], &g_multi.in[g]};
            size_t* ss[] = {&g_multi.recv_bytes[g], &g_multi.tmp_bytes[g], &g_multi.in_bytes[g]};
            for (int i = 0; i < 3; ++i) {
                if (*ps[i]) cudaFree(*ps[i]);
                *ps[i] = nullptr;
                *ss[i] = 0;
            }
            g_multi.dev_of[g] = devs[g];
        }
    }
    const double t0 = now_ms();
    std::vector<int> rcs(P, 0);
    // H2D: every GPU pulls its shard over its own PCIe link, one host thread each
    if (rc == B200_OK) {
        std::vector<std::thread> th;
        for (int g = 0; g < P; ++g)
            th.emplace_back([&, g] {
                cudaSetDevice(devs[g]);
                if (!grow(&g_multi.in[g], &g_multi.in_bytes[g], std::max<size_t>(n_in[g], 1) * es)) {
                    rcs[g] = B200_ERR_ALLOC;
                    return;
                }
                rcs[g] = b200_copy_h2d(g_multi.in[g], static_cast<const uint8_t*>(host_ptr) + off[g] * es, n_in[g] * es, devs[g]);
            });
        for (auto& t : th) t.join();
        for (int g = 0; g < P; ++g)
            if (rcs[g]) rc = rcs[g];
    }
    const double t1 = now_ms();
    Shared sh;
    std::vector<size_t> n_out(P, 0);
    if (rc == B200_OK) {
        std::vector<const void*> src(P);
        for (int g = 0; g < P; ++g) src[g] = g_multi.in[g];
        rc = run_sharded(P, devs, desc, src.data(), n_in.data(), 1.25, sh);
    }
    const double t2 = now_ms();
    if (rc == B200_OK) {
        // D2H: GPU g's sorted range lands behind the ranges of the lower ranks
        std::vector<size_t> ooff(P + 1, 0);
        for (int g = 0; g < P; ++g) {
            size_t nr = n_in[g];
            if (sh.level >= 0) {
                nr = 0;
                for (int r = 0; r < P; ++r) nr += size_t(sh.m[size_t(r) * P + g]);
            }
            n_out[g] = nr;
            ooff[g + 1] = ooff[g] + nr;
        }
        if (ooff[P] != n) rc = B200_ERR_CUDA;
        if (rc == B200_OK) {
            std::vector<std::thread> th;
            for (int g = 0; g < P; ++g)
                th.emplace_back([&, g] {
                    cudaSetDevice(devs[g]);
                    const void* from = sh.level >= 0 ? g_multi.recv[g] : g_multi.in[g];
                    rcs[g] = b200_copy_d2h(static_cast<uint8_t*>(host_ptr) + ooff[g] * es, from, n_out[g] * es, devs[g]);
                });
            for (auto& t : th) t.join();
            for (int g = 0; g < P; ++g)
                if (rcs[g]) rc = rcs[g];
        }
    }
    if (rc == B200_OK && stats) {
        fill_stats(sh, stats, n_out.data());
        stats->ms_h2d = float(t1 - t0);
        stats->ms_d2h = float(now_ms() - t2);
        stats->ms_total = float(now_ms() - t0);
    }
    if (prev >= 0) cudaSetDevice(prev);
    return rc;
}

int b200_multi_release(void) {
    std::lock_guard<std::mutex> lk(g_multi.mu);
    int prev = -1;
    cudaGetDevice(&prev);
    int count = 0;
    cudaGetDeviceCount(&count);
    for (int g = 0; g < MAX_GPUS; ++g) {
        void** ps[] = {&g_multi.recv[g], &g_multi.in[g], &g_multi.tmp[g]};
        size_t* ss[] = {&g_multi.recv_bytes[g], &g_multi.in_bytes[g], &g_multi.tmp_bytes[g]};
        for (int i = 0; i < 3; ++i) {
            if (*ps[i]) {
                cudaPointerAttributes a;
                if (cudaPointerGetAttributes(&a, *ps[i]) == cudaSuccess && a.device >= 0 && a.device < count) cudaSetDevice(a.device);
                cudaFree(*ps[i]);
            }
            *ps[i] = nullptr;
            *ss[i] = 0;
        }
    }
    cudaGetLastError();
    if (prev >= 0) cudaSetDevice(prev);
    return B200_OK;
}

}  // extern "C"
