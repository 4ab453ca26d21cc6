// chain.cuh -- the persistent per-PID table.
//
// It holds (1) the key set of process_memory (chronos_sensor.py:105): every PID that ever
// had a surviving event, including those whose list is currently [] (:157), and (2) the
// state of the kill-chain DFA per PID (the extension described in window.cuh), which must
// survive window flushes and batches.
//
// Open addressing, u64 slots: bit 63 valid | bits 47:40 epoch | bits 33:32 DFA state | low 32
// bits pid. The host keeps it at most half full (ensure_pid_table in chronos.cu).
#pragma once
#include "common.cuh"

namespace chronos {

constexpr uint64_t PT_VALID = 1ull << 63;
__host__ __device__ __forceinline__ uint64_t pt_make(uint32_t epoch, uint32_t dfa, uint32_t pid) {
    return PT_VALID | ((uint64_t)(epoch & 0xFFu) << 40) | ((uint64_t)(dfa & 3u) << 32) | pid;
}
__device__ __forceinline__ bool pt_live(uint64_t v, uint32_t epoch) {
    return (v & PT_VALID) && (uint32_t)((v >> 40) & 0xFFu) == epoch;
}

// Home slot of a pid. group_forward probes in ASCENDING pid order (the entries are sorted), so
// the map keeps that order: pids below 2^22 (Linux pid_max) map to themselves and neighbouring
// heads touch neighbouring slots -- a stream through the table instead of one DRAM + TLB miss
// per head (measured on the 1 M-PID workload: 187 -> see profiles/). Larger pids get their high
// bits folded in so that keys differing only there do not pile up on one slot.
__host__ __device__ __forceinline__ uint32_t pt_home(uint32_t pid, uint32_t mask) {
    return (pid ^ ((pid >> 22) * 0x9E3779B1u)) & mask;
}

// Slot index of pid, inserting {pid, state 0} if absent (inserted++ then; the caller adds its
// total to the table's count ONCE per warp: one atomic per new PID on a single address was the
// largest cost of group_forward on the 1 M-PID workload). Slots written under an older epoch
// count as empty: chronos_reset only bumps the epoch.
// *dfa_out (optional) receives the chain state stored for the pid (0 for a new one), so that
// the caller does not have to read the slot again.
__device__ __forceinline__ uint32_t pt_find_or_insert(uint64_t *tab, uint32_t mask, uint32_t epoch, uint32_t pid,
                                                      uint32_t &inserted, uint32_t *dfa_out = nullptr) {
    uint32_t i = pt_home(pid, mask);
    while (true) {
        uint64_t v = ld_relaxed_u64(tab + i);
        while (!pt_live(v, epoch)) {
            const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long *>(tab + i), (unsigned long long)v,
                                           (unsigned long long)pt_make(epoch, 0, pid));
            if (old == v) {
                inserted++;
                if (dfa_out) *dfa_out = 0;
                return i;
            }
            v = old;  // somebody else took the slot: look at what is there now
        }
        if ((uint32_t)v == pid) {
            if (dfa_out) *dfa_out = (uint32_t)(v >> 32) & 3u;
            return i;
        }
        i = (i + 1) & mask;
    }
}

__global__ void pt_rehash_kernel(const uint64_t *old_tab, uint32_t old_cap, uint64_t *new_tab, uint32_t new_mask,
                                 uint32_t epoch) {
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < old_cap; k += gridDim.x * blockDim.x) {
        const uint64_t v = old_tab[k];
        if (!pt_live(v, epoch)) continue;
        uint32_t i = pt_home((uint32_t)v, new_mask);
        while (atomicCAS(reinterpret_cast<unsigned long long *>(new_tab + i), 0ull, (unsigned long long)v) != 0ull)
            i = (i + 1) & new_mask;
    }
}

__global__ void pt_dump_kernel(const uint64_t *tab, uint32_t cap, uint32_t epoch, uint32_t *out, uint32_t *n_out) {
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < cap; k += gridDim.x * blockDim.x) {
        const uint64_t v = tab[k];
        if (pt_live(v, epoch)) out[atomicAdd(n_out, 1u)] = (uint32_t)v;
    }
}

// checkpoint: every live entry as (dfa << 32 | pid), any order
__global__ void pt_dump64_kernel(const uint64_t *tab, uint32_t cap, uint32_t epoch, uint64_t *out, uint32_t *n_out) {
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < cap; k += gridDim.x * blockDim.x) {
        const uint64_t v = tab[k];
        if (pt_live(v, epoch)) out[atomicAdd(n_out, 1u)] = v & 0x3FFFFFFFFull;
    }
}

// checkpoint: entries (dfa << 32 | pid) back into an empty table
__global__ void pt_load_kernel(uint64_t *tab, uint32_t mask, uint32_t epoch, const uint64_t *in, uint32_t n) {
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const uint64_t e = in[k];
        uint32_t inserted = 0;
        const uint32_t slot = pt_find_or_insert(tab, mask, epoch, (uint32_t)e, inserted);
        st_relaxed_u64(tab + slot, pt_make(epoch, (uint32_t)(e >> 32) & 3u, (uint32_t)e));
    }
}

}  // namespace chronos
