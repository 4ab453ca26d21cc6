// Host-only build of the synthetic trace generator (csrc/tracegen.cuh compiled by g++, no CUDA).
// bench.py's CPU reference arm generates its sample with this library, so that the arm never
// maps the product's CUDA library. Same source as the CUDA twin: the bytes are identical
// (tests/test_abi.py::test_host_only_generator_matches_the_library).
#include <stdint.h>

#include "tracegen.cuh"

extern "C" int chronos_hostgen(const chronos_tracegen_cfg *cfg, const uint64_t *cdf, uint64_t first, uint64_t n, void *dst) {
    if (!cfg || (n && !dst) || cfg->n_pids == 0 || (cfg->use_cdf && !cdf)) return -1;
    uint8_t *d = static_cast<uint8_t *>(dst);
    for (uint64_t i = 0; i < n; i++) chronos_gen::gen_record(*cfg, first + i, cdf, d + i * 288);
    return 0;
}
