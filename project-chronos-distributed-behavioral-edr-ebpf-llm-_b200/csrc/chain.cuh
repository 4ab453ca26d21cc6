// chain.cuh -- the persistent per-PID table.
//
// It holds (1) the key set of process_memory (chronos_sensor.py:105): every PID that ever
// had a surviving event, including those whose list is currently [] (:157), and (2) the
// state of the kill-chain DFA per PID (the extension described in window.cuh), which must
// survive window flushes and batches.
//
// Open addressing, u64 slots: bit 63 valid | bits 33:32 DFA state | low 32 bits pid. The host
// keeps it at most half full (ensure_pid_table in chronos.cu).
#pragma once
#include "common.cuh"

namespace chronos {

constexpr uint64_t PT_VALID = 1ull << 63;

// returns the slot index of pid, inserting {pid, state 0} if absent (count++ on insert)
__device__ __forceinline__ uint32_t pt_find_or_insert(uint64_t *tab, uint32_t mask, uint32_t pid, uint32_t *count) {
    uint32_t i = hash32(pid) & mask;
    while (true) {
        uint64_t v = ld_relaxed_u64(tab + i);
        if (v == 0) {
            const uint64_t mine = PT_VALID | pid;
            const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long *>(tab + i), 0ull, (unsigned long long)mine);
            if (old == 0) {
                atomicAdd(count, 1u);
                return i;
            }
            v = old;
        }
        if ((uint32_t)v == pid) return i;
        i = (i + 1) & mask;
    }
}

__global__ void pt_rehash_kernel(const uint64_t *old_tab, uint32_t old_cap, uint64_t *new_tab, uint32_t new_mask) {
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < old_cap; k += gridDim.x * blockDim.x) {
        const uint64_t v = old_tab[k];
        if (!(v & PT_VALID)) continue;
        uint32_t i = hash32((uint32_t)v) & new_mask;
        while (atomicCAS(reinterpret_cast<unsigned long long *>(new_tab + i), 0ull, (unsigned long long)v) != 0ull)
            i = (i + 1) & new_mask;
    }
}

__global__ void pt_dump_kernel(const uint64_t *tab, uint32_t cap, uint32_t *out, uint32_t *n_out) {
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < cap; k += gridDim.x * blockDim.x) {
        const uint64_t v = tab[k];
        if (v & PT_VALID) out[atomicAdd(n_out, 1u)] = (uint32_t)v;
    }
}

}  // namespace chronos
