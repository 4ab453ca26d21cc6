// sort.cuh -- kernel 2: stable group-by on pid.
//
// process_memory (chronos_sensor.py:105) is a dict keyed by pid whose lists keep arrival
// order (:138); on the GPU that is a STABLE sort of (pid -> entry id) pairs. The schema
// has no timestamp (:18-23), so stability on pid alone is what preserves arrival order.
//
// LSD radix sort, <=8-bit digits, only over the bit range in which the keys actually
// differ (planned on the host from the OR/AND the filter kernel folded). One pass =
//   radix_hist    per-tile digit histograms            (digit-major table)
//   scan_exclusive_u32  exclusive scan of the table     (look-back, multi-CTA)
//   radix_scatter stable scatter: per-warp sub-ranges, __match_any_sync ranking
// Values are entry ids; pass 0 synthesises them (iota) instead of reading them.
#pragma once
#include "common.cuh"

namespace chronos {

constexpr int RADIX_BINS = 256;
constexpr int RADIX_THREADS = 256;
constexpr int RADIX_WARPS = RADIX_THREADS / 32;

struct RadixParams {
    // plan mode (plan != nullptr): pass `pass` of the batch's sort; reads set A when the pass
    // index is even and set B when it is odd, writes the other; extents come from the plan
    const Plan *plan;
    int pass;
    uint32_t *keyA, *keyB, *valA, *valB;
    // explicit mode (plan == nullptr)
    const uint32_t *keys_in;
    const uint32_t *vals_in;  // nullptr => value = index | class bits << 27 (see synth_value)
    const uint8_t *meta_lo;   // class bits of entries [0, split) (carried-in pending), may be nullptr
    const uint8_t *meta_hi;   // class bits of entries [split, n) (this batch's survivors), may be nullptr
    uint32_t split;
    uint32_t *keys_out;
    uint32_t *vals_out;
    uint32_t *table;          // RADIX_BINS * n_tiles counters, digit-major
    uint32_t n;
    uint32_t tile_keys;       // keys per tile (multiple of RADIX_WARPS*32)
    uint32_t n_tiles;
    uint32_t shift;
    uint32_t mask;            // (1<<bits)-1
};

// Pass 0 synthesises the sort values instead of reading them: the entry id in the low 27
// bits and the event's class bits (CHRONOS_META_*) above, so that the kernels after the
// sort never gather them again. Carried-in entries lose the trigger bit (a pending event
// cannot trigger again, chronos_sensor.py:141 looks at the current event only).
__device__ __forceinline__ uint32_t synth_value(const uint8_t *meta_lo, const uint8_t *meta_hi, uint32_t split, uint32_t i) {
    uint32_t m = 0;
    if (i < split) m = meta_lo ? (uint32_t)(meta_lo[i] & 0x1Eu) : 0u;
    else m = meta_hi ? (uint32_t)(meta_hi[i - split] & 0x1Fu) : 0u;
    return i | (m << 27);
}

__global__ void synth_values_kernel(uint32_t *v, uint32_t n, const uint8_t *meta_lo, const uint8_t *meta_hi, uint32_t split,
                                    const Plan *plan) {
    if (plan) n = plan->M;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        v[i] = synth_value(meta_lo, meta_hi, split, i);
}

// One radix pass resolved to plain values (plan mode reads the device plan; explicit mode the
// parameter struct). active == false: this pass or this tile has nothing to do.
struct RadixView {
    const uint32_t *keys_in, *vals_in;
    uint32_t *keys_out, *vals_out;
    uint32_t n, tile_keys, n_tiles, shift, mask;
    bool active;
};
__device__ __forceinline__ RadixView radix_view(const RadixParams &p) {
    RadixView v;
    if (!p.plan) {
        v.keys_in = p.keys_in; v.vals_in = p.vals_in; v.keys_out = p.keys_out; v.vals_out = p.vals_out;
        v.n = p.n; v.tile_keys = p.tile_keys; v.n_tiles = p.n_tiles; v.shift = p.shift; v.mask = p.mask;
        v.active = blockIdx.x < p.n_tiles;
        return v;
    }
    const Plan &pl = *p.plan;
    const bool odd = p.pass & 1;
    v.keys_in = odd ? p.keyB : p.keyA;
    v.vals_in = odd ? p.valB : p.valA;
    v.keys_out = odd ? p.keyA : p.keyB;
    v.vals_out = odd ? p.valA : p.valB;
    v.n = pl.M;
    v.tile_keys = pl.tile_keys;
    v.n_tiles = pl.n_tiles;
    v.shift = pl.shift[p.pass & 3];
    v.mask = pl.mask[p.pass & 3];
    v.active = (uint32_t)p.pass < pl.n_passes && blockIdx.x < pl.n_tiles;
    return v;
}

__global__ void __launch_bounds__(RADIX_THREADS) radix_hist(const RadixParams p) {
    __shared__ uint32_t h[RADIX_BINS];
    const RadixView v = radix_view(p);
    if (!v.active) return;
    const uint32_t tile = blockIdx.x;
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t lo = tile * v.tile_keys;
    const uint32_t hi = min(v.n, lo + v.tile_keys);
    for (uint32_t i = lo + threadIdx.x; i < hi; i += RADIX_THREADS)
        atomicAdd(&h[(v.keys_in[i] >> v.shift) & v.mask], 1u);
    __syncthreads();
    p.table[(size_t)threadIdx.x * v.n_tiles + tile] = h[threadIdx.x];
}

// Exclusive scan of n counters in place: persistent CTAs, 4096 counters per tile, chained
// with decoupled look-back. Status words carry an epoch tag ([63:34] epoch, [33:32] state,
// [31:0] value) so the same status buffer serves every scan without being cleared.
constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

struct ScanParams {
    uint32_t *data;
    uint32_t n;
    const uint32_t *n_ptr;  // optional: the count lives on the device (overrides n)
    uint32_t *ticket;   // zeroed counter
    uint64_t *status;   // ceil(n / SCAN_TILE) words, any content from older epochs
    uint32_t epoch;     // > 0, unique per launch sharing `status`
};

__device__ __forceinline__ uint32_t lookback_count_epoch(uint64_t *status, uint32_t tile, uint32_t agg, int lane,
                                                         uint32_t epoch) {
    const uint64_t tag = (uint64_t)epoch << 34;
    if (tile == 0) {
        if (lane == 0) st_relaxed_u64(status, tag | (ST_PREFIX << 32) | agg);
        __syncwarp();
        return 0;
    }
    if (lane == 0) st_relaxed_u64(status + tile, tag | (ST_AGG << 32) | agg);
    uint32_t excl = 0;
    int64_t look = (int64_t)tile - 1;
    while (true) {
        int64_t idx = look - lane;
        uint64_t v = tag | (ST_PREFIX << 32);  // virtual tiles below 0: prefix 0
        if (idx >= 0) {
            do {
                v = ld_relaxed_u64(status + idx);
            } while ((v >> 34) != epoch || ((v >> 32) & 3u) == ST_NONE);
        }
        const uint32_t is_p = __ballot_sync(0xffffffffu, ((v >> 32) & 3u) == ST_PREFIX);
        if (is_p) {
            const int first = __ffs((int)is_p) - 1;
            excl += __reduce_add_sync(0xffffffffu, lane <= first ? (uint32_t)v : 0u);
            break;
        }
        excl += __reduce_add_sync(0xffffffffu, (uint32_t)v);
        look -= 32;
    }
    if (lane == 0) st_relaxed_u64(status + tile, tag | (ST_PREFIX << 32) | (uint64_t)(excl + agg));
    __syncwarp();  // reconverge before the caller's __syncthreads()
    return excl;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_exclusive_u32(const ScanParams p) {
    __shared__ uint32_t warp_sums[SCAN_THREADS / 32];
    __shared__ uint32_t s_tile, s_base;
    const uint32_t n = p.n_ptr ? *p.n_ptr : p.n;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    while (true) {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
        const uint32_t i0 = tile * SCAN_TILE + tid * SCAN_ITEMS;
        uint32_t v[SCAN_ITEMS];
        if (i0 + SCAN_ITEMS <= n) {
            const uint4 a = *reinterpret_cast<const uint4 *>(p.data + i0);
            const uint4 b = *reinterpret_cast<const uint4 *>(p.data + i0 + 4);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        } else {
#pragma unroll
            for (int k = 0; k < SCAN_ITEMS; k++) v[k] = (i0 + k < n) ? p.data[i0 + k] : 0u;
        }
        uint32_t sum = 0;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) sum += v[k];
        uint32_t inc = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += t;
        }
        if (lane == 31) warp_sums[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = lane < SCAN_THREADS / 32 ? warp_sums[lane] : 0u, winc = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                uint32_t t = __shfl_up_sync(0xffffffffu, winc, d);
                if (lane >= d) winc += t;
            }
            if (lane < SCAN_THREADS / 32) warp_sums[lane] = winc - w;
            const uint32_t tot = __shfl_sync(0xffffffffu, winc, 31);
            const uint32_t excl = lookback_count_epoch(p.status, tile, tot, lane, p.epoch);
            __syncwarp();
            if (lane == 0) s_base = excl;
        }
        __syncthreads();
        uint32_t run = s_base + warp_sums[warp] + inc - sum;
        uint32_t o[SCAN_ITEMS];
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            o[k] = run;
            run += v[k];
        }
        if (i0 + SCAN_ITEMS <= n) {
            *reinterpret_cast<uint4 *>(p.data + i0) = make_uint4(o[0], o[1], o[2], o[3]);
            *reinterpret_cast<uint4 *>(p.data + i0 + 4) = make_uint4(o[4], o[5], o[6], o[7]);
        } else {
#pragma unroll
            for (int k = 0; k < SCAN_ITEMS; k++)
                if (i0 + k < n) p.data[i0 + k] = o[k];
        }
    }
}

__global__ void __launch_bounds__(RADIX_THREADS) radix_scatter(const RadixParams p) {
    __shared__ uint32_t whist[RADIX_WARPS][RADIX_BINS];
    const RadixView v = radix_view(p);
    if (!v.active) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tile = blockIdx.x;
    for (int i = tid; i < RADIX_WARPS * RADIX_BINS; i += RADIX_THREADS) (&whist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t t_lo = tile * v.tile_keys;
    const uint32_t t_hi = min(v.n, t_lo + v.tile_keys);
    const uint32_t sub = v.tile_keys / RADIX_WARPS;  // multiple of 32
    const uint32_t w_lo = min(t_hi, t_lo + warp * sub);
    const uint32_t w_hi = min(t_hi, w_lo + sub);
    // 1) per-warp digit counts over the warp's contiguous sub-range
    for (uint32_t i = w_lo + lane; i < w_hi; i += 32) atomicAdd(&whist[warp][(v.keys_in[i] >> v.shift) & v.mask], 1u);
    __syncthreads();
    // 2) digit d: running base over warps, seeded with the global offset of (d, tile)
    {
        uint32_t run = p.table[(size_t)tid * v.n_tiles + tile];
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
            key = v.keys_in[i];
            d = (key >> v.shift) & v.mask;
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
            v.keys_out[dst] = key;
            v.vals_out[dst] = v.vals_in ? v.vals_in[i] : synth_value(p.meta_lo, p.meta_hi, p.split, i);
        }
    }
}

}  // namespace chronos
