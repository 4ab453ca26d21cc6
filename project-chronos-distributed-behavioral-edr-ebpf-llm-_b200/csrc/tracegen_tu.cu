// tracegen_tu.cu -- the synthetic trace generator as a translation unit (and therefore a CUDA
// module) of its own.
//
// gen_record inlines to about 218 KB of machine code, as much as all other kernels together; it
// is benchmark scaffolding, not the hot path, and compiles separately so that a change to the
// kernels does not pay for it. (While the filter's bistable timing was hunted it was suspected of
// pushing the filter's module over a code-size threshold; that turned out to be chance, see
// profiles/README.md.)
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>

#include "chronos.h"
#include "common.cuh"
#include "tracegen.cuh"

namespace {

int sm_count_of(int device) {
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || sms <= 0) {
        cudaGetLastError();
        sms = 148;
    }
    return sms;
}

__global__ void tracegen_kernel(chronos_tracegen_cfg cfg, const uint64_t *cdf, uint64_t first, uint64_t n,
                                uint8_t *dst) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint8_t rec[CH_REC_BYTES];
        chronos_gen::gen_record(cfg, first + i, cdf, rec);
        uint32_t *d = reinterpret_cast<uint32_t *>(dst + i * CH_REC_BYTES);
        const uint32_t *s = reinterpret_cast<const uint32_t *>(rec);
        for (int k = 0; k < CH_REC_WORDS; k++) d[k] = s[k];
    }
}

}  // namespace

extern "C" {

// ---- trace generator ------------------------------------------------------------------------
int chronos_tracegen_host(const chronos_tracegen_cfg *cfg, const uint64_t *cdf, uint64_t first, uint64_t n, void *dst) {
    if (!cfg || (n && !dst) || cfg->n_pids == 0 || (cfg->use_cdf && !cdf)) return CHRONOS_E_ARG;
    uint8_t *d = static_cast<uint8_t *>(dst);
    for (uint64_t i = 0; i < n; i++) chronos_gen::gen_record(*cfg, first + i, cdf, d + i * CH_REC_BYTES);
    return CHRONOS_OK;
}

int chronos_tracegen_device(int device, const chronos_tracegen_cfg *cfg, const uint64_t *cdf_dev, uint64_t first,
                            uint64_t n, void *dst_dev) {
    if (!cfg || (n && !dst_dev) || cfg->n_pids == 0 || (cfg->use_cdf && !cdf_dev)) return CHRONOS_E_ARG;
    if (cudaSetDevice(device) != cudaSuccess) return CHRONOS_E_CUDA;
    if (!n) return CHRONOS_OK;
    tracegen_kernel<<<(unsigned)std::min<uint64_t>((n + 127) / 128, (uint64_t)sm_count_of(device) * 16), 128>>>(
        *cfg, cdf_dev, first, n, static_cast<uint8_t *>(dst_dev));
    if (cudaGetLastError() != cudaSuccess) return CHRONOS_E_CUDA;
    return cudaDeviceSynchronize() == cudaSuccess ? CHRONOS_OK : CHRONOS_E_CUDA;
}

}  // extern "C"
