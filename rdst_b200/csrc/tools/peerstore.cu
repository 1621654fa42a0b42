// peerstore.cu — how fast can SM-issued stores push data to a peer GPU over NVLink?
// Single process, two GPUs, peer access: a grid on GPU 0 copies a local buffer into GPU 1's memory
// with 8- or 16-byte stores per lane, in contiguous chunks of a given size per warp (the partition
// kernel's write-out writes runs of a few KB per destination).  Reference: cudaMemcpyPeerAsync.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

// every warp copies chunks of `chunk` bytes; consecutive chunks of a warp are `nwarps * chunk` apart
template <typename V>
__global__ void push(const V* __restrict__ src, V* __restrict__ dst, size_t bytes, size_t chunk) {
    const size_t warp = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) / 32, nwarps = size_t(gridDim.x) * blockDim.x / 32;
    const int lane = threadIdx.x & 31;
    const size_t per = chunk / sizeof(V);
    for (size_t c = warp; c * chunk < bytes; c += nwarps) {
        const V* s = src + c * per;
        V* d = dst + c * per;
        for (size_t i = lane; i < per; i += 32) d[i] = s[i];
    }
}

int main() {
    int n = 0; CK(cudaGetDeviceCount(&n));
    if (n < 2) { printf("needs 2 GPUs\n"); return 0; }
    const size_t bytes = size_t(2) << 30;
    void *a, *b;
    CK(cudaSetDevice(1)); CK(cudaMalloc(&b, bytes)); CK(cudaMemset(b, 0, bytes));
    CK(cudaSetDevice(0)); CK(cudaMalloc(&a, bytes)); CK(cudaMemset(a, 1, bytes));
    CK(cudaDeviceEnablePeerAccess(1, 0));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float ms;
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaEventRecord(e0)); CK(cudaMemcpyPeerAsync(b, 1, a, 0, bytes)); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("cudaMemcpyPeerAsync              %8.3f ms  %7.1f GB/s\n", ms, bytes / ms / 1e6);
    }
    const size_t chunks[] = {256, 1024, 8192, 65536};
    const int ctas[] = {148, 296, 592, 1184};
    for (int w = 0; w < 2; ++w)
        for (size_t chunk : chunks)
            for (int g : ctas) {
                for (int rep = 0; rep < 2; ++rep) {
                    CK(cudaEventRecord(e0));
                    if (w == 0) push<uint2><<<g, 256>>>((const uint2*)a, (uint2*)b, bytes, chunk);
                    else push<uint4><<<g, 256>>>((const uint4*)a, (uint4*)b, bytes, chunk);
                    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
                    CK(cudaEventElapsedTime(&ms, e0, e1));
                }
                printf("SM stores %2d B/lane chunk %6zu B grid %4d x 256   %8.3f ms  %7.1f GB/s\n", w ? 16 : 8, chunk, g, ms,
                       bytes / ms / 1e6);
            }
    // both directions at once (what the exchange of a sharded sort does: every GPU sends and receives): GPU 0 pushes into
    // GPU 1 while GPU 1 pushes into GPU 0; per-direction rate = bytes / the slower of the two kernels
    {
        void *a2, *b2;
        CK(cudaSetDevice(0)); CK(cudaMalloc(&a2, bytes));
        CK(cudaSetDevice(1)); CK(cudaMalloc(&b2, bytes)); CK(cudaMemset(b2, 2, bytes)); CK(cudaDeviceEnablePeerAccess(0, 0));
        cudaEvent_t f0, f1; CK(cudaEventCreate(&f0)); CK(cudaEventCreate(&f1));
        for (int g : {148, 296, 592}) {
            float m0 = 0, m1 = 0;
            for (int rep = 0; rep < 3; ++rep) {
                CK(cudaSetDevice(0)); CK(cudaDeviceSynchronize());
                CK(cudaSetDevice(1)); CK(cudaDeviceSynchronize());
                CK(cudaSetDevice(0)); CK(cudaEventRecord(e0)); push<uint2><<<g, 256>>>((const uint2*)a, (uint2*)b, bytes, 8192); CK(cudaEventRecord(e1));
                CK(cudaSetDevice(1)); CK(cudaEventRecord(f0)); push<uint2><<<g, 256>>>((const uint2*)b2, (uint2*)a2, bytes, 8192); CK(cudaEventRecord(f1));
                CK(cudaEventSynchronize(f1)); CK(cudaEventElapsedTime(&m1, f0, f1));
                CK(cudaSetDevice(0)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&m0, e0, e1));
            }
            printf("SM stores  8 B/lane chunk   8192 B grid %4d x 256  BOTH DIRECTIONS  %8.3f / %8.3f ms  %7.1f GB/s per direction\n", g, m0, m1,
                   bytes / (m0 > m1 ? m0 : m1) / 1e6);
        }
    }
    // destination runs that do not start on a 128-byte line (the partition kernel's runs start wherever the previous
    // run ended): every 256-byte warp store then touches three lines, two of them partially
    for (size_t off : {size_t(0), size_t(8), size_t(40), size_t(64)}) {
        for (int rep = 0; rep < 2; ++rep) {
            CK(cudaSetDevice(0));
            CK(cudaEventRecord(e0));
            push<uint2><<<296, 256>>>((const uint2*)a, (uint2*)((char*)b + off), bytes - 8192, 8192);
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
            CK(cudaEventElapsedTime(&ms, e0, e1));
        }
        printf("SM stores  8 B/lane chunk   8192 B grid  296 x 256  destination offset %3zu B  %8.3f ms  %7.1f GB/s\n", off, ms, (bytes - 8192) / ms / 1e6);
    }
    return 0;
}
