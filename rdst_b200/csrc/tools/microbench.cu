// microbench.cu — instruction-throughput probes that size the K1/K2 design on
// B200: MATCH.ANY vs eight ballots, shared-memory atomics (spread, same
// address, with and without return), and plain HBM copy / read bandwidth.
// Prints warp-instructions per cycle per SM (64 resident warps per SM).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../common.cuh"
using namespace b200;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

constexpr int ITERS = 2048;

__device__ __forceinline__ uint32_t lcg(uint32_t x) { return x * 1664525u + 1013904223u; }

template <int MODE>
__global__ void __launch_bounds__(1024) probe(uint32_t* out, long long* cycles, int skew) {
    __shared__ uint32_t sh[32 * 256];
    for (int i = threadIdx.x; i < 32 * 256; i += 1024) sh[i] = 0;
    __syncthreads();
    uint32_t x = (blockIdx.x * 1024 + threadIdx.x) * 2654435761u + 12345u;
    uint32_t acc = 0;
    uint32_t* my = sh + (threadIdx.x >> 5) * 256;
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < ITERS; ++i) {
        x = lcg(x);
        uint32_t d = skew == 0 ? (x >> 24) : skew == 1 ? 7u : skew == 2 ? ((x >> 24) & (x >> 16) & (x >> 8) & 0xffu)
                     : skew == 3 ? ((x & 0x100u) ? 7u : (x >> 24))
                     // skew 4: Zipf-like, one digit holds 1/8 of the records, another 1/16, the rest uniform
                     : ((x & 0x700u) == 0 ? 7u : (x & 0xf00u) == 0x100u ? 9u : (x >> 24));
        if (MODE == 0) acc += __match_any_sync(0xffffffffu, d);
        if (MODE == 1) acc += match_digit<false>(d);
        if (MODE == 2) atomicAdd(&my[d], 1u);                 // warp-private, no return
        if (MODE == 3) acc += atomicAdd(&my[d], 1u);          // warp-private, with return
        if (MODE == 4) atomicAdd(&sh[d], 1u);                 // CTA-shared, no return
        if (MODE == 5) { uint32_t v = my[d]; __syncwarp(); my[d] = v + 1; __syncwarp(); acc += v; }  // LDS+STS
        if (MODE == 9) atomicOr(&my[d], 1u << (threadIdx.x & 31));  // RED.OR, warp-private
        if (MODE == 10) {  // the half-warp OR-matching rank step of pass_or16.cuh
            uint32_t* slot = sh + ((threadIdx.x >> 4) & 31) * 256 + d;
            const uint32_t bit = 1u << (threadIdx.x & 15);
            atomicOr(slot, bit); __syncwarp();
            const uint32_t v = *(volatile uint32_t*)slot; __syncwarp();
            const uint32_t m = v & 0xffffu, c = v >> 16, lt = m & (bit - 1u);
            if (lt == 0) *(volatile uint32_t*)slot = (c + __popc(m)) << 16;
            __syncwarp(); acc += c + __popc(lt);
        }
        if (MODE == 11) atomicAdd(&my[d], 1u << (threadIdx.x & 31));  // RED.ADD of distinct bits == OR
        if (MODE == 12) {  // the half-warp rank step with ADD instead of OR
            uint32_t* slot = sh + ((threadIdx.x >> 4) & 31) * 256 + d;
            const uint32_t bit = 1u << (threadIdx.x & 15);
            atomicAdd(slot, bit); __syncwarp();
            const uint32_t v = *(volatile uint32_t*)slot; __syncwarp();
            const uint32_t m = v & 0xffffu, c = v >> 16, lt = m & (bit - 1u);
            if (lt == 0) *(volatile uint32_t*)slot = (c + __popc(m)) << 16;
            __syncwarp(); acc += c + __popc(lt);
        }
        if (MODE == 6) acc += d;                                // loop overhead only
        if (MODE == 7) {                                         // full rank step, ballots
            uint32_t m = match_digit<false>(d); uint32_t lt = m & lanemask_lt(); uint32_t b = my[d]; __syncwarp();
            if (lt == 0) my[d] = b + __popc(m); __syncwarp(); acc += b + __popc(lt);
        }
        if (MODE == 8) {                                         // full rank step, MATCH.ANY
            uint32_t m = __match_any_sync(0xffffffffu, d); uint32_t lt = m & lanemask_lt(); uint32_t b = my[d]; __syncwarp();
            if (lt == 0) my[d] = b + __popc(m); __syncwarp(); acc += b + __popc(lt);
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    out[blockIdx.x * 1024 + threadIdx.x] = acc + sh[threadIdx.x];
}

template <int MODE>
void run_probe(const char* name, int sms, uint32_t* d_out, long long* d_cyc, int skew) {
    int grid = sms * 2;
    probe<MODE><<<grid, 1024>>>(d_out, d_cyc, skew);
    probe<MODE><<<grid, 1024>>>(d_out, d_cyc, skew);
    CK(cudaDeviceSynchronize());
    std::vector<long long> h(grid);
    CK(cudaMemcpy(h.data(), d_cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost));
    double avg = 0; for (auto c : h) avg += double(c); avg /= grid;
    double warp_ops = 64.0 * ITERS;  // per SM
    printf("%-34s skew=%d  cycles/iter/SM(64 warps)=%8.2f  warp-instr-groups/cycle/SM=%.4f  cycles per warp-op=%.3f\n",
           name, skew, avg / ITERS, warp_ops / avg, avg / warp_ops);
}

__global__ void copy_kernel(const uint4* __restrict__ a, uint4* __restrict__ b, size_t n) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x, s = size_t(gridDim.x) * blockDim.x;
    for (; i < n; i += s) b[i] = ldg_stream(a + i);
}
__global__ void read_kernel(const uint4* __restrict__ a, uint32_t* out, size_t n) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x, s = size_t(gridDim.x) * blockDim.x;
    uint32_t acc = 0;
    for (; i < n; i += s) { uint4 v = ldg_stream(a + i); acc += v.x ^ v.y ^ v.z ^ v.w; }
    if (acc == 0x12345678u) out[0] = acc;
}

int main() {
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    int sms = prop.multiProcessorCount;
    printf("device %s, %d SMs, clock %d kHz\n", prop.name, sms, prop.clockRate);
    uint32_t* d_out; long long* d_cyc;
    CK(cudaMalloc(&d_out, size_t(sms) * 2 * 1024 * 4)); CK(cudaMalloc(&d_cyc, sms * 2 * 8));
    for (int skew = 0; skew < 5; ++skew) {
        run_probe<2>("ATOMS.ADD warp-private no-ret", sms, d_out, d_cyc, skew);
        run_probe<9>("RED.OR warp-private", sms, d_out, d_cyc, skew);
        run_probe<11>("RED.ADD lane bit warp-private", sms, d_out, d_cyc, skew);
        run_probe<10>("half-warp OR rank step", sms, d_out, d_cyc, skew);
        run_probe<12>("half-warp ADD rank step", sms, d_out, d_cyc, skew);
    }
    if (getenv("MICROBENCH_ATOMICS_ONLY")) return 0;
    for (int skew = 0; skew < 2; ++skew) {
        run_probe<6>("loop overhead", sms, d_out, d_cyc, skew);
        run_probe<0>("MATCH.ANY", sms, d_out, d_cyc, skew);
        run_probe<1>("8x ballot match", sms, d_out, d_cyc, skew);
        run_probe<2>("ATOMS warp-private no-ret", sms, d_out, d_cyc, skew);
        run_probe<3>("ATOMS warp-private ret", sms, d_out, d_cyc, skew);
        run_probe<4>("ATOMS cta-shared no-ret", sms, d_out, d_cyc, skew);
        run_probe<5>("LDS+STS counter", sms, d_out, d_cyc, skew);
        run_probe<7>("rank step (ballots)", sms, d_out, d_cyc, skew);
        run_probe<8>("rank step (MATCH.ANY)", sms, d_out, d_cyc, skew);
    }
    // HBM bandwidth
    size_t bytes = size_t(4) << 30; uint4 *a, *b;
    CK(cudaMalloc(&a, bytes)); CK(cudaMalloc(&b, bytes)); CK(cudaMemset(a, 1, bytes)); CK(cudaMemset(b, 2, bytes));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int rep = 0; rep < 3; ++rep) {
        CK(cudaEventRecord(e0)); copy_kernel<<<sms * 16, 512>>>(a, b, bytes / 16); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("copy 4 GiB: %.3f ms  %.1f GB/s (read+write)\n", ms, 2.0 * bytes / ms / 1e6);
        CK(cudaEventRecord(e0)); read_kernel<<<sms * 16, 512>>>(a, d_out, bytes / 16); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("read 4 GiB: %.3f ms  %.1f GB/s\n", ms, 1.0 * bytes / ms / 1e6);
        CK(cudaEventRecord(e0)); CK(cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice)); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("cudaMemcpy D2D 4 GiB: %.3f ms  %.1f GB/s (read+write)\n", ms, 2.0 * bytes / ms / 1e6);
    }
    return 0;
}
