// sort.cuh -- kernel 2: stable group-by on pid.
//
// process_memory (chronos_sensor.py:105) is a dict keyed by pid whose lists keep arrival
// order (:138); on the GPU that is a STABLE sort of (pid -> entry id) pairs. The schema
// has no timestamp (:18-23), so stability on pid alone is what preserves arrival order.
//
// LSD radix sort, <=8-bit digits, only over the bit range in which the keys actually
// differ (planned on the host from the OR/AND the filter kernel folded). One pass =
//   radix_hist    per-tile digit histograms            (digit-major table)
//   radix_scan    exclusive scan of the table          (single CTA)
//   radix_scatter stable scatter: per-warp sub-ranges, __match_any_sync ranking
// Values are entry ids; pass 0 synthesises them (iota) instead of reading them.
#pragma once
#include "common.cuh"

namespace chronos {

constexpr int RADIX_BINS = 256;
constexpr int RADIX_THREADS = 256;
constexpr int RADIX_WARPS = RADIX_THREADS / 32;

struct RadixParams {
    const uint32_t *keys_in;
    const uint32_t *vals_in;  // nullptr => value = index
    uint32_t *keys_out;
    uint32_t *vals_out;
    uint32_t *table;          // RADIX_BINS * n_tiles counters, digit-major
    uint32_t n;
    uint32_t tile_keys;       // keys per tile (multiple of RADIX_WARPS*32)
    uint32_t n_tiles;
    uint32_t shift;
    uint32_t mask;            // (1<<bits)-1
};

__global__ void __launch_bounds__(RADIX_THREADS) radix_hist(const RadixParams p) {
    __shared__ uint32_t h[RADIX_BINS];
    const uint32_t tile = blockIdx.x;
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t lo = tile * p.tile_keys;
    const uint32_t hi = min(p.n, lo + p.tile_keys);
    for (uint32_t i = lo + threadIdx.x; i < hi; i += RADIX_THREADS)
        atomicAdd(&h[(p.keys_in[i] >> p.shift) & p.mask], 1u);
    __syncthreads();
    p.table[(size_t)threadIdx.x * p.n_tiles + tile] = h[threadIdx.x];
}

// Exclusive scan of `len` counters in place, one CTA of 1024 threads.
__global__ void __launch_bounds__(1024) radix_scan(uint32_t *table, uint32_t len) {
    __shared__ uint32_t warp_sums[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t per = (len + 1023u) / 1024u;
    const uint32_t lo = min(len, (uint32_t)tid * per), hi = min(len, lo + per);
    uint32_t sum = 0;
    for (uint32_t i = lo; i < hi; i++) sum += table[i];
    uint32_t inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = warp_sums[lane], winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += t;
        }
        warp_sums[lane] = winc - w;
    }
    __syncthreads();
    uint32_t run = warp_sums[warp] + inc - sum;
    for (uint32_t i = lo; i < hi; i++) {
        uint32_t v = table[i];
        table[i] = run;
        run += v;
    }
}

__global__ void __launch_bounds__(RADIX_THREADS) radix_scatter(const RadixParams p) {
    __shared__ uint32_t whist[RADIX_WARPS][RADIX_BINS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tile = blockIdx.x;
    for (int i = tid; i < RADIX_WARPS * RADIX_BINS; i += RADIX_THREADS) (&whist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t t_lo = tile * p.tile_keys;
    const uint32_t t_hi = min(p.n, t_lo + p.tile_keys);
    const uint32_t sub = p.tile_keys / RADIX_WARPS;  // multiple of 32
    const uint32_t w_lo = min(t_hi, t_lo + warp * sub);
    const uint32_t w_hi = min(t_hi, w_lo + sub);
    // 1) per-warp digit counts over the warp's contiguous sub-range
    for (uint32_t i = w_lo + lane; i < w_hi; i += 32) atomicAdd(&whist[warp][(p.keys_in[i] >> p.shift) & p.mask], 1u);
    __syncthreads();
    // 2) digit d: running base over warps, seeded with the global offset of (d, tile)
    {
        uint32_t run = p.table[(size_t)tid * p.n_tiles + tile];
#pragma unroll
        for (int w = 0; w < RADIX_WARPS; w++) {
            uint32_t c = whist[w][tid];
            whist[w][tid] = run;
            run += c;
        }
    }
    __syncthreads();
    // 3) stable placement, 32 keys per step, in index order
    for (uint32_t i0 = w_lo; i0 < w_hi; i0 += 32) {
        const uint32_t i = i0 + lane;
        const bool valid = i < w_hi;
        uint32_t key = 0, d = RADIX_BINS;  // invalid lanes form their own group
        if (valid) {
            key = p.keys_in[i];
            d = (key >> p.shift) & p.mask;
        }
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const uint32_t before = __popc(peers & lanemask_lt());
        uint32_t base = 0;
        if (valid) base = whist[warp][d];
        __syncwarp();
        if (valid && before == 0) whist[warp][d] = base + __popc(peers);
        __syncwarp();
        if (valid) {
            const uint32_t dst = base + before;
            p.keys_out[dst] = key;
            p.vals_out[dst] = p.vals_in ? p.vals_in[i] : i;
        }
    }
}

}  // namespace chronos
