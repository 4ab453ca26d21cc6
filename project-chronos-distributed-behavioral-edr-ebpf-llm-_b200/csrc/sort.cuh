// sort.cuh -- kernel 2: stable group-by on pid.
//
// process_memory (chronos_sensor.py:105) is a dict keyed by pid whose lists keep arrival
// order (:138); on the GPU that is a STABLE sort of (pid -> entry id) pairs. The schema
// has no timestamp (:18-23), so stability on pid alone is what preserves arrival order.
//
// LSD radix sort, <=8-bit digits, only over the bit range in which the keys actually
// differ (planned ON THE DEVICE from the OR/AND the filter kernel folded):
//   sort_prepare   the plan + the global digit histogram of every pass, one read of the keys
//   onesweep_pass  one kernel per pass: per-tile digit counts, per-digit decoupled look-back
//                  across tiles, stable scatter (per-warp sub-ranges, __match_any_sync ranking)
// Values are entry ids; pass 0 synthesises them (iota | class bits) instead of reading them.
// scan_exclusive_u32 (bottom of this file) is the stand-alone exclusive scan used by the
// compaction and the exchange.
#pragma once
#include "common.cuh"

namespace chronos {

constexpr int RADIX_BINS = 256;
constexpr int RADIX_THREADS = 256;
constexpr int RADIX_WARPS = RADIX_THREADS / 32;

#ifndef CHRONOS_RADIX_TILE
#define CHRONOS_RADIX_TILE 2048  // (measured together with CHRONOS_WIN_ITEMS, window.cuh)
#endif
#ifndef CHRONOS_RADIX_LB
#define CHRONOS_RADIX_LB 8     // predecessors' status words loaded per look-back round
#endif
constexpr uint32_t RADIX_TILE_KEYS = CHRONOS_RADIX_TILE;  // keys per tile (multiple of RADIX_WARPS*32)
constexpr int RADIX_MAX_PASSES = 4;

// What the filter stage and the previous batch leave on the device for the planner.
struct SortInputs {
    const uint32_t *surv_count, *key_or, *key_and;  // scalars of the batch in flight
    Carry *carry;                                   // pending count and pid bit range carried in
    uint32_t m_max, win_tile;
    uint32_t s_max, w_max;  // capacities of the survivor / window arrays (checked builds assert against them)
};

struct SortParams {
    Plan *plan;               // written by sort_prepare (block 0), read by everything after it
    SortInputs in;
    uint32_t *keyA, *keyB, *valA, *valB;   // pass i reads set A when i is even, set B when odd
    // The unsorted entries are never materialised: entry i < P is the i-th pending entry of the
    // previous batch (arrays right-aligned in a layout of pend_cap entries), entry i >= P the
    // (i-P)-th survivor of this batch. Pass 0 reads its keys from these two sources.
    const uint32_t *pend_key;
    const uint8_t *pend_meta;
    uint32_t pend_cap;
    const uint32_t *surv_pid;
    const uint8_t *surv_meta;
    uint32_t *ghist;          // RADIX_MAX_PASSES x RADIX_BINS global digit counts, zeroed per batch
    int pass;                 // onesweep_pass only
    uint32_t *ticket;         // zeroed per batch, one per pass
    uint64_t *status;         // look-back words: one per (tile, digit); epoch-tagged, never cleared
    uint32_t epoch;           // unique per onesweep_pass launch
};

__device__ __forceinline__ uint32_t entry_key(const SortParams &p, uint32_t P, uint32_t i) {
    return i < P ? p.pend_key[p.pend_cap - P + i] : p.surv_pid[i - P];
}
// Pass 0 synthesises the sort values instead of reading them: the entry id in the low 27
// bits and the event's class bits (CHRONOS_META_*) above, so that the kernels after the
// sort never gather them again. Carried-in entries lose the trigger bit (a pending event
// cannot trigger again, chronos_sensor.py:141 looks at the current event only).
__device__ __forceinline__ uint32_t synth_value(const SortParams &p, uint32_t P, uint32_t i) {
    const uint32_t m = i < P ? (uint32_t)(p.pend_meta[p.pend_cap - P + i] & 0x1Eu) : (uint32_t)(p.surv_meta[i - P] & 0x1Fu);
    return i | (m << 27);
}

// Sizes of the batch, decided on the device once the survivor count is there (no host round
// trip): radix passes over the bits in which the pids differ, tilings, and where the sorted
// arrays end up after n_passes ping-pong passes.
__device__ __forceinline__ void make_plan(const SortParams &p, Plan &pl) {
    const uint32_t S = *p.in.surv_count, P = p.in.carry->P;
    uint32_t M = P + S;
    pl.cap_m = p.in.m_max;
    pl.cap_s = p.in.s_max;
    pl.cap_w = p.in.w_max;
    pl.overflow = M > p.in.m_max ? 1u : 0u;
    if (M > p.in.m_max) M = p.in.m_max;  // never run past the buffers; the host reports the overflow
    pl.S = M > P ? M - P : 0u;
    pl.P = P;
    pl.M = M;
    // bit range of every pid seen since the last reset: a superset of what this batch needs,
    // and known without reading the pending keys again
    pl.key_or = p.in.carry->key_or | *p.in.key_or;
    pl.key_and = p.in.carry->key_and & *p.in.key_and;
    const uint32_t diff = pl.key_or ^ pl.key_and;
    uint32_t passes = 0;
    for (int i = 0; i < RADIX_MAX_PASSES; i++) { pl.shift[i] = 0; pl.mask[i] = 0; }
    if (M >= 2 && diff != 0) {
        const int lo = __ffs((int)diff) - 1, hi = 32 - __clz((int)diff);
        const int bits = hi - lo;
        passes = (uint32_t)(bits + 7) / 8;
        const int per = (bits + (int)passes - 1) / (int)passes;
        int sft = lo;
        for (uint32_t i = 0; i < passes; i++) {
            const int b = min(per, hi - sft);
            pl.shift[i] = (uint32_t)sft;
            pl.mask[i] = (1u << b) - 1u;
            sft += b;
        }
    }
    pl.n_passes = passes;
    pl.tile_keys = RADIX_TILE_KEYS;
    pl.n_tiles = (M + RADIX_TILE_KEYS - 1) / RADIX_TILE_KEYS;
    pl.table_len = 0;
    pl.tiles_m = (M + p.in.win_tile - 1) / p.in.win_tile;
    pl.tiles_s = (pl.S + p.in.win_tile - 1) / p.in.win_tile;
    pl.skey = (passes & 1u) ? p.keyB : p.keyA;
    pl.sval = (passes & 1u) ? p.valB : p.valA;
}

// Kernel 1 of the sort: the plan, and the GLOBAL digit histogram of every pass in one read of
// the keys (the passes then need no histogram kernel and no table scan of their own).
__device__ __forceinline__ void sort_prepare_body(const SortParams &p) {
    __shared__ Plan pl;
    __shared__ uint32_t h[RADIX_MAX_PASSES][RADIX_BINS];
    if (threadIdx.x == 0) {
        make_plan(p, pl);
        if (blockIdx.x == 0) {
            *p.plan = pl;
            // (other blocks may read the carry before or after this: OR / AND are idempotent)
            p.in.carry->key_or = pl.key_or;
            p.in.carry->key_and = pl.key_and;
        }
    }
    for (int i = threadIdx.x; i < RADIX_MAX_PASSES * RADIX_BINS; i += RADIX_THREADS) (&h[0][0])[i] = 0;
    __syncthreads();
    const uint32_t M = pl.M, np = pl.n_passes;
    if (np == 0) {  // nothing to sort: the entries themselves, in place
        for (uint32_t i = blockIdx.x * RADIX_THREADS + threadIdx.x; i < M; i += gridDim.x * RADIX_THREADS) {
            p.keyA[i] = entry_key(p, pl.P, i);
            p.valA[i] = synth_value(p, pl.P, i);
        }
        return;
    }
    for (uint32_t i = blockIdx.x * RADIX_THREADS + threadIdx.x; i < M; i += gridDim.x * RADIX_THREADS) {
        const uint32_t k = entry_key(p, pl.P, i);
        for (uint32_t q = 0; q < np; q++) atomicAdd(&h[q][(k >> pl.shift[q]) & pl.mask[q]], 1u);
    }
    __syncthreads();
    for (uint32_t q = 0; q < np; q++) {
        const uint32_t c = h[q][threadIdx.x];
        if (c) atomicAdd(&p.ghist[q * RADIX_BINS + threadIdx.x], c);
    }
}

__global__ void __launch_bounds__(RADIX_THREADS) sort_prepare(const SortParams p) { sort_prepare_body(p); }

// Kernel 2..: one radix pass in ONE kernel. Tiles are taken in ticket order; a tile counts its
// digits, publishes the counts, learns how many keys with the same digit precede it from its
// predecessors' status words (decoupled look-back, one chain per digit = one per thread) and
// scatters: stable, because a tile's keys go behind those of every earlier tile and the
// placement inside the tile is in index order.
// status word: [63:34] epoch | [33:32] ST_* | [31:0] count
// (pass, ticket and epoch are explicit arguments: the fused tail kernel runs all passes of one
//  parameter block)
__device__ __forceinline__ void onesweep_pass_body(const SortParams &p, const int pass, uint32_t *const ticket,
                                                   const uint32_t epoch) {
    __shared__ uint32_t whist[RADIX_WARPS][RADIX_BINS];
    __shared__ uint32_t gbase[RADIX_BINS];
    __shared__ uint32_t warp_tot[RADIX_WARPS];
    __shared__ uint32_t s_tile;
    const Plan &pl = *p.plan;
    if ((uint32_t)pass >= pl.n_passes) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool odd = pass & 1;
    const bool first = pass == 0;
    const uint32_t *keys_in = odd ? p.keyB : p.keyA;   // (pass 0 reads the entries instead)
    const uint32_t *vals_in = first ? nullptr : (odd ? p.valB : p.valA);
    uint32_t *keys_out = odd ? p.keyA : p.keyB;
    uint32_t *vals_out = odd ? p.valA : p.valB;
    const uint32_t n = pl.M, n_tiles = pl.n_tiles, shift = pl.shift[pass], mask = pl.mask[pass], split = pl.P;
    const uint64_t tag = (uint64_t)epoch << 34;
    {   // where digit d starts in the output: exclusive scan of the global digit counts
        const uint32_t c = p.ghist[pass * RADIX_BINS + tid];
        uint32_t inc = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += t;
        }
        if (lane == 31) warp_tot[warp] = inc;
        __syncthreads();
        uint32_t before = 0;
#pragma unroll
        for (int w = 0; w < RADIX_WARPS; w++)
            if (w < warp) before += warp_tot[w];
        gbase[tid] = before + inc - c;
    }
    while (true) {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(ticket, 1u);
        for (int i = tid; i < RADIX_WARPS * RADIX_BINS; i += RADIX_THREADS) (&whist[0][0])[i] = 0;
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
        const uint32_t t_lo = tile * RADIX_TILE_KEYS;
        const uint32_t t_hi = min(n, t_lo + RADIX_TILE_KEYS);
        constexpr uint32_t sub = RADIX_TILE_KEYS / RADIX_WARPS;  // multiple of 32
        const uint32_t w_lo = min(t_hi, t_lo + warp * sub);
        const uint32_t w_hi = min(t_hi, w_lo + sub);
        // 1) per-warp digit counts over the warp's contiguous sub-range. The warp's keys and values
        //    stay in registers for the placement in step 3: read there, step by step, every one of
        //    its rounds began with a dependent trip to L2.
        constexpr int PER_LANE = (int)(sub / 32);
        uint32_t kreg[PER_LANE], vreg[PER_LANE];
#pragma unroll
        for (int it = 0; it < PER_LANE; it++) {
            const uint32_t i = w_lo + it * 32 + lane;
            kreg[it] = 0;
            vreg[it] = 0;
            if (i < w_hi) {
                kreg[it] = first ? entry_key(p, split, i) : keys_in[i];
                vreg[it] = first ? synth_value(p, split, i) : vals_in[i];
            }
        }
#pragma unroll
        for (int it = 0; it < PER_LANE; it++)
            if (w_lo + it * 32 + lane < w_hi) atomicAdd(&whist[warp][(kreg[it] >> shift) & mask], 1u);
        __syncthreads();
        // 2) digit d = tid: tile total, published; keys of digit d in earlier tiles by look-back
        uint32_t total = 0;
#pragma unroll
        for (int w = 0; w < RADIX_WARPS; w++) total += whist[w][tid];
        uint64_t *st = p.status + (size_t)tile * RADIX_BINS + tid;
        uint32_t excl = 0;
        if (tile == 0) {
            st_relaxed_u64(st, tag | (ST_PREFIX << 32) | total);
        } else {
            st_relaxed_u64(st, tag | (ST_AGG << 32) | total);
            // The chain of this digit, LB predecessors per round: their status words are loaded
            // together (independent loads, one round trip) and consumed nearest first, up to the
            // first PREFIX or the first word that is not written yet. One word per round made a
            // pass a chain of up to n_tiles dependent L2 round trips (all tiles of a batch this
            // size run in one wave, so nobody finds a PREFIX nearby).
            constexpr int LB = CHRONOS_RADIX_LB;
            const uint64_t *q = st - RADIX_BINS;  // nearest predecessor not consumed yet
            uint32_t left = tile;                 // predecessors not consumed yet (tile 0 always ends the walk)
            bool done = false;
            while (!done) {
                uint64_t v[LB];
                const int nb = (int)min((uint32_t)LB, left);
#pragma unroll
                for (int k = 0; k < LB; k++)
                    if (k < nb) v[k] = ld_relaxed_u64(q - (size_t)k * RADIX_BINS);
                int used = 0;
#pragma unroll
                for (int k = 0; k < LB; k++) {
                    if (k < nb && used == k && !done) {
                        const uint32_t stt = (uint32_t)(v[k] >> 32) & 3u;
                        if ((v[k] >> 34) == epoch && stt != ST_NONE) {
                            excl += (uint32_t)v[k];
                            used = k + 1;
                            done = stt == ST_PREFIX;
                        }
                    }
                }
                q -= (size_t)used * RADIX_BINS;
                left -= (uint32_t)used;
            }
            st_relaxed_u64(st, tag | (ST_PREFIX << 32) | (uint64_t)(excl + total));
        }
        __syncwarp();  // the chains of a warp's 32 digits end at different depths
        {
            uint32_t run = gbase[tid] + excl;
#pragma unroll
            for (int w = 0; w < RADIX_WARPS; w++) {
                const uint32_t c = whist[w][tid];
                whist[w][tid] = run;
                run += c;
            }
        }
        __syncthreads();
        // 3) stable placement, 32 keys per step, in index order
#pragma unroll
        for (int it = 0; it < PER_LANE; it++) {
            const uint32_t i = w_lo + it * 32 + lane;
            if (w_lo + it * 32 >= w_hi) break;  // (uniform over the warp)
            const bool valid = i < w_hi;
            uint32_t key = 0, d = RADIX_BINS;  // invalid lanes form their own group
            if (valid) {
                key = kreg[it];
                d = (key >> shift) & mask;
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
                CH_ASSERT(dst < n);
                keys_out[dst] = key;
                vals_out[dst] = vreg[it];
            }
        }
    }
}

__global__ void __launch_bounds__(RADIX_THREADS) onesweep_pass(const SortParams p) {
    onesweep_pass_body(p, p.pass, p.ticket, p.epoch);
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

}  // namespace chronos
