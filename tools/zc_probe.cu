// zc_probe.cu -- how fast can a kernel pull SELECTED 32-byte sectors of 288-byte records out of
// pinned host memory over PCIe (zero-copy), compared with copying the whole batch first?
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/zc_probe tools/zc_probe.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

// mode bit0: sector 0, bit1: sector 8, bit2: all 9 sectors; frac_full: every k-th record is read whole
__global__ void pull(const uint4 *recs, uint64_t n, int mode, int every_full, uint32_t *sink) {
    uint32_t acc = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint4 *r = recs + i * 18;
        if (mode & 1) { uint4 a = r[0], b = r[1]; acc ^= a.x ^ a.w ^ b.x ^ b.w; }
        if (mode & 2) { uint4 a = r[16], b = r[17]; acc ^= a.x ^ a.w ^ b.x ^ b.w; }
        if ((mode & 4) || (every_full && (i % every_full) == 0)) {
#pragma unroll
            for (int k = 2; k < 16; k++) { uint4 a = r[k]; acc ^= a.x ^ a.w; }
        }
    }
    if (acc == 0x12345678u) *sink = acc;
}
// warp-cooperative: a warp reads the chosen sectors of 32 records with 2 lanes per 32-byte sector
__global__ void pull_coop(const uint4 *recs, uint64_t n, uint32_t *sink) {
    uint32_t acc = 0;
    const int lane = threadIdx.x & 31;
    const uint64_t warp = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5, nw = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t base = warp * 16; base < n; base += nw * 16) {   // 16 records per warp step: lanes (rec, half)
        const uint64_t i = base + (lane >> 1);
        if (i < n) {
            const uint4 *r = recs + i * 18;
            uint4 a = r[lane & 1], b = r[16 + (lane & 1)];
            acc ^= a.x ^ a.w ^ b.x ^ b.w;
        }
    }
    if (acc == 0x12345678u) *sink = acc;
}
// whole batch streamed by the kernel (coalesced 16 B per lane)
__global__ void pull_stream(const uint4 *recs, uint64_t n16, uint32_t *sink) {
    uint32_t acc = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x) {
        uint4 a = recs[i]; acc ^= a.x ^ a.w;
    }
    if (acc == 0x12345678u) *sink = acc;
}

int main(int argc, char **argv) {
    const uint64_t n = argc > 1 ? strtoull(argv[1], 0, 10) : 10000000ull;
    const size_t bytes = n * 288;
    void *h; CK(cudaHostAlloc(&h, bytes, cudaHostAllocMapped));
    memset(h, 1, bytes);
    void *dview; CK(cudaHostGetDevicePointer(&dview, h, 0));
    void *d; CK(cudaMalloc(&d, bytes));
    uint32_t *sink; CK(cudaMalloc(&sink, 4));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms;
    for (int rep = 0; rep < 2; rep++) {
        cudaEventRecord(e0); CK(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice)); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        cudaEventElapsedTime(&ms, e0, e1);
        printf("memcpy H2D whole batch: %.2f ms  %.1f GB/s\n", ms, bytes / ms / 1e6);
    }
    struct { const char *name; int mode, every; double payload; } cases[] = {
        {"sector 0 only (32 B/rec)", 1, 0, 32}, {"sectors 0+8 (64 B/rec)", 3, 0, 64},
        {"sectors 0+8, every 8th record whole", 3, 8, 64 + 224.0 / 8}, {"all 9 sectors, thread per record", 7, 0, 288}};
    for (int grid : {148 * 4, 148 * 16}) {
        for (auto &c : cases) {
            for (int rep = 0; rep < 2; rep++) {
                cudaEventRecord(e0); pull<<<grid, 256>>>((const uint4 *)dview, n, c.mode, c.every, sink); cudaEventRecord(e1);
                CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1);
            }
            printf("grid %5d  %-40s %.2f ms  payload %.1f GB/s  (%.0f M rec/s)\n", grid, c.name, ms, n * c.payload / ms / 1e6, n / ms / 1e3);
        }
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0); pull_coop<<<grid, 256>>>((const uint4 *)dview, n, sink); cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1);
        }
        printf("grid %5d  %-40s %.2f ms  payload %.1f GB/s\n", grid, "sectors 0+8, 2 lanes per sector", ms, n * 64.0 / ms / 1e6);
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0); pull_stream<<<grid, 256>>>((const uint4 *)dview, n * 18, sink); cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1);
        }
        printf("grid %5d  %-40s %.2f ms  payload %.1f GB/s\n", grid, "whole batch streamed by the kernel", ms, bytes / ms / 1e6);
    }
    return 0;
}
