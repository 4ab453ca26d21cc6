// nvlink_probe.cu -- how fast can a kernel on GPU 0 WRITE into GPU 1's memory, depending on how
// the stores are shaped? (a) fully coalesced 16 B per lane, (b) runs of `run` 16-byte slots at
// scattered places (the exchange writes about 5 slots = 86 B per survivor), (c) the same runs
// staged so that consecutive lanes write consecutive runs of one destination region.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/nvlink_probe tools/nvlink_probe.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void w_coalesced(uint4 *dst, size_t n16) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
        dst[i] = make_uint4((uint32_t)i, 1u, 2u, 3u);
}
// entry e owns `run` consecutive slots; entries are placed at pseudo-random run-aligned positions;
// lane l of a warp writes slot (k % run) of entry (k / run): neighbouring lanes mostly hit one entry
__global__ void w_runs(uint4 *dst, size_t n_entries, int run, size_t span_entries, int scatter) {
    const size_t total = n_entries * run;
    for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < total; k += (size_t)gridDim.x * blockDim.x) {
        const size_t e = k / run, s = k - e * run;
        size_t pos = e;
        if (scatter) pos = (e * 2654435761ull) % span_entries;  // scattered entry placement
        dst[pos * run + s] = make_uint4((uint32_t)k, 1u, 2u, 3u);
    }
}
int main() {
    int n = 0; CK(cudaGetDeviceCount(&n));
    if (n < 2) { printf("needs 2 GPUs\n"); return 0; }
    const size_t bytes = 512ull << 20;
    void *remote, *local;
    CK(cudaSetDevice(1)); CK(cudaMalloc(&remote, bytes)); CK(cudaMemset(remote, 0, bytes));
    CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0)); CK(cudaMalloc(&local, bytes));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms;
    for (int target = 0; target < 2; target++) {
        uint4 *d = (uint4 *)(target ? remote : local);
        const char *tn = target ? "peer " : "local";
        const size_t payload = 100ull << 20;  // what one rank sends per step, roughly
        for (int rep = 0; rep < 3; rep++) { cudaEventRecord(e0); w_coalesced<<<148 * 4, 256>>>(d, payload / 16); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1); }
        printf("%s coalesced 16 B per lane               %.3f ms  %.0f GB/s\n", tn, ms, payload / ms / 1e6);
        for (int run : {1, 3, 5, 8, 32}) for (int scatter : {0, 1}) {
            const size_t ne = payload / 16 / run;
            for (int rep = 0; rep < 3; rep++) { cudaEventRecord(e0); w_runs<<<148 * 4, 256>>>(d, ne, run, bytes / 16 / run, scatter); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms, e0, e1); }
            printf("%s runs of %2d slots (%3d B), %s   %.3f ms  %.0f GB/s\n", tn, run, run * 16, scatter ? "scattered " : "contiguous", ms, payload / ms / 1e6);
        }
    }
    return 0;
}
