// sector_probe.cu -- what does HBM deliver when a kernel reads only SOME 32-byte sectors of
// every 288-byte record (device memory)? Decides whether lazy sector loading in the filter pays.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/sector_probe tools/sector_probe.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

// thread per record: chunks 0,1 (+17) (+ chunks 2..5 for every `every2`-th record) (+ rest for every `every3`-th)
__global__ void pull(const uint4 *recs, uint64_t n, int tail, int every2, int every3, uint32_t *sink) {
    uint32_t acc = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint4 *r = recs + i * 18;
        uint4 a = r[0], b = r[1];
        acc ^= a.x ^ a.w ^ b.x ^ b.w;
        if (tail) { uint4 c = r[17]; acc ^= c.x ^ c.w; }
        if (every2 && (i % every2) == 0) { uint4 c = r[2], d = r[3], e = r[4], f = r[5]; acc ^= c.x ^ d.x ^ e.x ^ f.w; }
        if (every3 && (i % every3) == 0) {
#pragma unroll
            for (int k = 6; k < 17; k++) { uint4 c = r[k]; acc ^= c.x ^ c.w; }
        }
    }
    if (acc == 0x12345678u) *sink = acc;
}
// warp per 32 records, single-warp CTAs like the filter: lane i <-> record i, same access pattern
__global__ void __launch_bounds__(32) pull_tiles(const uint4 *recs, uint64_t n_tiles, int every2, uint32_t *ticket, uint32_t *sink) {
    uint32_t acc = 0;
    const int lane = threadIdx.x;
    while (true) {
        uint32_t t = 0;
        if (lane == 0) t = atomicAdd(ticket, 1u);
        t = __shfl_sync(0xffffffffu, t, 0);
        if (t >= n_tiles) break;
        const uint4 *r = recs + ((uint64_t)t * 32 + lane) * 18;
        uint4 a = r[0], b = r[1], c = r[17];
        acc ^= a.x ^ a.w ^ b.x ^ b.w ^ c.x ^ c.w;
        if (every2 && ((t * 32 + lane) % every2) == 0) { uint4 c2 = r[2], d = r[3], e = r[4], f = r[5]; acc ^= c2.x ^ d.x ^ e.x ^ f.w; }
    }
    if (acc == 0x12345678u) *sink = acc;
}
__global__ void pull_stream(const uint4 *recs, uint64_t n16, uint32_t *sink) {
    uint32_t acc = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x) { uint4 a = recs[i]; acc ^= a.x ^ a.w; }
    if (acc == 0x12345678u) *sink = acc;
}
int main(int argc, char **argv) {
    const uint64_t n = argc > 1 ? strtoull(argv[1], 0, 10) : 10000000ull;
    const size_t bytes = n * 288;
    void *d; CK(cudaMalloc(&d, bytes)); CK(cudaMemset(d, 1, bytes));
    uint32_t *sink; CK(cudaMalloc(&sink, 8));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms;
    auto report = [&](const char *name, double payload) { printf("%-58s %.3f ms  %5.0f M rec/s  payload %.0f GB/s\n", name, ms, n / ms / 1e3, n * payload / ms / 1e6); };
    for (int rep = 0; rep < 3; rep++) { cudaEventRecord(e0); pull_stream<<<148 * 16, 256>>>((const uint4 *)d, n * 18, sink); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1); }
    report("whole records, streamed (288 B/rec)", 288);
    struct { const char *name; int tail, e2, e3; double payload; } cs[] = {
        {"sector 0 only (32 B/rec)", 0, 0, 0, 32}, {"sectors 0+8 (64 B/rec)", 1, 0, 0, 64},
        {"sectors 0+8, every 3rd +sectors 1,2", 1, 3, 0, 64 + 64.0 / 3}, {"sectors 0+8, 1/3 +s1,2, 1/10 whole", 1, 3, 10, 64 + 64.0 / 3 + 17.6}};
    for (int grid : {148 * 8, 148 * 32}) for (auto &c : cs) {
        for (int rep = 0; rep < 3; rep++) { cudaEventRecord(e0); pull<<<grid, 256>>>((const uint4 *)d, n, c.tail, c.e2, c.e3, sink); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1); }
        char nm[96]; snprintf(nm, sizeof nm, "grid %5d  %s", grid, c.name); report(nm, c.payload);
    }
    for (int e2 : {0, 3}) {
        for (int rep = 0; rep < 3; rep++) { cudaMemset(sink + 1, 0, 4); cudaEventRecord(e0); pull_tiles<<<148 * 19, 32>>>((const uint4 *)d, n / 32, e2, sink + 1, sink); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1); }
        report(e2 ? "single-warp CTAs (19/SM), sectors 0+8, 1/3 +s1,2" : "single-warp CTAs (19/SM), sectors 0+8", e2 ? 64 + 64.0 / 3 : 64);
    }
    return 0;
}
