// window.cuh -- kernels 3/4: per-PID context windows over the pid-grouped entries.
//
// Reference semantics (chronos_sensor.py): append to process_memory[pid] (:138); if the
// event string holds a trigger needle (:141) and the list now has >= 2 entries (:142)
// the whole list is emitted as one window and the list is cleared (:157).
// Per PID that is a 2-state automaton over its events:
//     state E (list empty) / N (non-empty);   non-trigger: -> N
//     trigger in E: -> N (no flush)           trigger in N: FLUSH, -> E
// The first event of a PID always ends in N. Events still pending from earlier batches
// re-enter the grouped array as non-trigger entries ahead of the batch's survivors, so
// no separate per-PID state lookup is needed for the window logic.
//
// Both directions are single-pass scans with decoupled look-back whose carried value is
// a FUNCTION over the 2-state automaton (what the tile does for either entry state),
// which makes the composition associative:
//   window_forward  : flush bits, window ordinals, group starts -> window descriptors
//   window_reverse  : "is my group ever closed?" -> pending-list compaction
//   window_order    : windows re-ordered by arrival index of their trigger event (:143-147)
//   gather_records  : 288-byte records copied once into the pid-grouped arena
#pragma once
#include "common.cuh"
#include "filter.cuh"

namespace chronos {

constexpr uint32_t NONE32 = 0xffffffffu;
constexpr int WIN_THREADS = 256;
constexpr int WIN_ITEMS = 16;
constexpr int WIN_TILE = WIN_THREADS * WIN_ITEMS;

// What a run of elements does to the automaton, for either entry state s in {0,1}:
//   out bit s  : exit state
//   a[s]       : counter increment (flushes / open elements)
//   b[s]       : last "marker" position set inside the run, NONE32 if it inherits
struct Fn2 {
    uint32_t out;
    uint32_t a[2];
    uint32_t b[2];
};
__device__ __forceinline__ Fn2 fn_identity() { return Fn2{0x2u, {0u, 0u}, {NONE32, NONE32}}; }
// first x, then y
__device__ __forceinline__ Fn2 fn_compose(const Fn2 &x, const Fn2 &y) {
    Fn2 r;
    uint32_t m0 = x.out & 1u, m1 = (x.out >> 1) & 1u;
    r.out = ((y.out >> m0) & 1u) | (((y.out >> m1) & 1u) << 1);
    r.a[0] = x.a[0] + y.a[m0];
    r.a[1] = x.a[1] + y.a[m1];
    r.b[0] = y.b[m0] != NONE32 ? y.b[m0] : x.b[0];
    r.b[1] = y.b[m1] != NONE32 ? y.b[m1] : x.b[1];
    return r;
}
struct St2 {  // concrete automaton state + carried counter + carried marker
    uint32_t s, a, b;
};
__device__ __forceinline__ St2 fn_apply(const Fn2 &f, const St2 &c) {
    St2 r;
    r.s = (f.out >> c.s) & 1u;
    r.a = c.a + f.a[c.s];
    r.b = f.b[c.s] != NONE32 ? f.b[c.s] : c.b;
    return r;
}
// status word for entry state s: [63:62] status [61] out_s [60:32] a_s (29 bits) [31:0] b_s
__device__ __forceinline__ uint64_t fn_pack(uint64_t status, uint32_t out_s, uint32_t a_s, uint32_t b_s) {
    return (status << 62) | ((uint64_t)(out_s & 1u) << 61) | ((uint64_t)(a_s & 0x1fffffffu) << 32) | b_s;
}
__device__ __forceinline__ Fn2 fn_shfl(const Fn2 &f, int src) {
    Fn2 r;
    r.out = __shfl_sync(0xffffffffu, f.out, src);
    r.a[0] = __shfl_sync(0xffffffffu, f.a[0], src);
    r.a[1] = __shfl_sync(0xffffffffu, f.a[1], src);
    r.b[0] = __shfl_sync(0xffffffffu, f.b[0], src);
    r.b[1] = __shfl_sync(0xffffffffu, f.b[1], src);
    return r;
}

// Block-wide exclusive scan of per-thread Fn2 (WIN_THREADS threads). Returns the
// exclusive prefix of the calling thread; *total (valid in every thread) is the tile's Fn2.
__device__ __forceinline__ Fn2 fn_block_exclusive(const Fn2 &mine, Fn2 *warp_tot /*[8] smem*/, Fn2 *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Fn2 inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Fn2 o = fn_shfl(inc, lane >= d ? lane - d : lane);
        if (lane >= d) inc = fn_compose(o, inc);
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    Fn2 prev = fn_shfl(inc, lane > 0 ? lane - 1 : 0);
    Fn2 excl = lane > 0 ? prev : fn_identity();
    Fn2 run = fn_identity();
    Fn2 wpre = fn_identity();
#pragma unroll
    for (int w = 0; w < WIN_THREADS / 32; w++) {
        if (w == warp) wpre = run;
        run = fn_compose(run, warp_tot[w]);
    }
    *total = run;
    return fn_compose(wpre, excl);
}

// Decoupled look-back over Fn2 tile summaries; one full warp. status has 2 words per
// tile. Publishes this tile's summary / inclusive state; returns the concrete state in
// front of the tile. `init` is the state in front of tile 0.
__device__ __forceinline__ St2 fn_lookback(uint64_t *status, uint32_t tile, const Fn2 &agg, St2 init, int lane) {
    St2 in = init;
    if (tile != 0) {
        if (lane == 0) {
            st_relaxed_u64(status + 2 * (size_t)tile, fn_pack(ST_AGG, agg.out, agg.a[0], agg.b[0]));
            st_relaxed_u64(status + 2 * (size_t)tile + 1, fn_pack(ST_AGG, agg.out >> 1, agg.a[1], agg.b[1]));
        }
        Fn2 H = fn_identity();  // everything between the tiles inspected so far and us
        int64_t look = (int64_t)tile - 1;
        while (true) {
            int64_t idx = look - lane;
            uint64_t w0 = fn_pack(ST_PREFIX, init.s, init.a, init.b), w1 = w0;  // virtual tiles below 0
            if (idx >= 0) {
                // the owner rewrites both words when it upgrades AGG -> PREFIX: only accept a
                // pair whose status fields agree (a torn pair mixes two encodings)
                do {
                    w0 = ld_relaxed_u64(status + 2 * (size_t)idx);
                    w1 = ld_relaxed_u64(status + 2 * (size_t)idx + 1);
                } while ((w0 >> 62) == ST_NONE || (w0 >> 62) != (w1 >> 62));
            }
            Fn2 mine;
            mine.out = (uint32_t)((w0 >> 61) & 1u) | ((uint32_t)((w1 >> 61) & 1u) << 1);
            mine.a[0] = (uint32_t)(w0 >> 32) & 0x1fffffffu;
            mine.a[1] = (uint32_t)(w1 >> 32) & 0x1fffffffu;
            mine.b[0] = (uint32_t)w0;
            mine.b[1] = (uint32_t)w1;
            const uint32_t is_p = __ballot_sync(0xffffffffu, (w0 >> 62) == ST_PREFIX);
            const int first = is_p ? __ffs((int)is_p) - 1 : 32;
            // G = tiles (look-first, look] composed in ascending tile order (= descending
            // lane order): ordered tree reduction over the lanes below `first`
            Fn2 G = (lane < first) ? mine : fn_identity();
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const Fn2 older = fn_shfl(G, (lane + d) & 31);
                if (lane + d < 32) G = fn_compose(older, G);
            }
            G = fn_shfl(G, 0);
            H = fn_compose(G, H);
            if (is_p) {
                Fn2 pf = fn_shfl(mine, first);
                St2 c{pf.out & 1u, pf.a[0], pf.b[0]};
                in = fn_apply(H, c);
                break;
            }
            look -= 32;
        }
    }
    St2 outc = fn_apply(agg, in);
    if (lane == 0) {
        uint64_t w = fn_pack(ST_PREFIX, outc.s, outc.a, outc.b);
        st_relaxed_u64(status + 2 * (size_t)tile, w);
        st_relaxed_u64(status + 2 * (size_t)tile + 1, w);
    }
    return in;
}

struct WinSorted {  // window descriptor in pid-grouped order
    uint32_t pid, trig_entry, start, len;
};

struct WinParams {
    const uint32_t *skey;   // M sorted pids
    const uint32_t *sval;   // M entry ids: [0,P) carried-in pending, [P,M) this batch's survivors
    uint32_t M, P;
    const uint8_t *surv_meta;       // S
    const uint32_t *surv_src;       // S batch-local record index
    const uint64_t *surv_arrival;   // S explicit arrival indices, or nullptr => base_arrival + surv_src
    uint64_t base_arrival;
    const uint8_t *pend_meta;       // P (previous batch)
    const uint64_t *pend_arrival;   // P
    const uint32_t *pend_slot;      // P record index in the previous arena
    // forward outputs
    uint8_t *flush;                 // M
    WinSorted *wsorted;             // <= S/1 descriptors in grouped order
    uint32_t *wi_of_entry;          // S: window ordinal flushed by survivor j, or NONE32
    uint32_t *n_windows;
    // reverse outputs (right-aligned in arrays of capacity pend_cap)
    uint32_t pend_cap;
    uint32_t *npend_key;
    uint32_t *npend_slot;
    uint64_t *npend_arrival;
    uint8_t *npend_meta;
    uint32_t *n_pending;
    // chain labeller outputs consumed here (chain.cuh); nullptr when not run
    const uint8_t *match;           // M: this batch's event at pos completed a curl->chmod->EXEC chain
    const uint32_t *mcnt;           // M: inclusive count of chain-complete events up to pos
    // ordering outputs
    chronos_window *windows;        // n_windows, ascending trigger arrival
    // scheduling
    uint32_t *ticket;               // one counter per kernel
    uint64_t *status;               // look-back words for this kernel
};

__device__ __forceinline__ uint8_t win_meta_of(const WinParams &p, uint32_t e) {
    return e < p.P ? (uint8_t)(p.pend_meta[e] & ~META_TRIG) : p.surv_meta[e - p.P];
}
__device__ __forceinline__ uint64_t win_arrival_of(const WinParams &p, uint32_t e) {
    if (e < p.P) return p.pend_arrival[e];
    uint32_t j = e - p.P;
    return p.surv_arrival ? p.surv_arrival[j] : p.base_arrival + p.surv_src[j];
}

// ---------------------------------------------------------------------------------------
// forward: automaton state E=0 / N=1. Marker b = start position of the current group.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(WIN_THREADS) window_forward(const WinParams p) {
    __shared__ Fn2 warp_tot[WIN_THREADS / 32];
    __shared__ uint32_t s_tile;
    __shared__ St2 s_in;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_tiles = (p.M + WIN_TILE - 1) / WIN_TILE;
    while (true) {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
        const uint32_t p0 = tile * WIN_TILE + tid * WIN_ITEMS;
        uint32_t key[WIN_ITEMS], ent[WIN_ITEMS];
        bool head[WIN_ITEMS], trig[WIN_ITEMS];
        uint32_t prevk = (p0 > 0 && p0 - 1 < p.M) ? p.skey[p0 - 1] : 0;
        Fn2 f = fn_identity();
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            const uint32_t pos = p0 + k;
            if (pos < p.M) {
                key[k] = p.skey[pos];
                ent[k] = p.sval[pos];
                head[k] = (pos == 0) || key[k] != prevk;
                trig[k] = (win_meta_of(p, ent[k]) & META_TRIG) != 0;
                prevk = key[k];
                Fn2 e;
                if (head[k]) e = Fn2{0x3u, {0u, 0u}, {pos, pos}};
                else if (!trig[k]) e = Fn2{0x3u, {0u, 0u}, {pos, NONE32}};
                else e = Fn2{0x1u, {0u, 1u}, {pos, NONE32}};
                f = fn_compose(f, e);
            } else {
                key[k] = 0; ent[k] = 0; head[k] = false; trig[k] = false;
            }
        }
        Fn2 total;
        Fn2 excl = fn_block_exclusive(f, warp_tot, &total);
        if (warp == 0) {
            St2 in = fn_lookback(p.status, tile, total, St2{0u, 0u, NONE32}, lane);
            if (lane == 0) {
                s_in = in;
                if (tile == n_tiles - 1) *p.n_windows = fn_apply(total, in).a;
            }
        }
        __syncthreads();
        St2 st = fn_apply(excl, s_in);
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            const uint32_t pos = p0 + k;
            if (pos >= p.M) break;
            bool fl = false;
            uint32_t wi = NONE32;
            if (head[k]) { st.b = pos; st.s = 1; }
            else if (!trig[k]) { if (st.s == 0) st.b = pos; st.s = 1; }
            else if (st.s == 0) { st.b = pos; st.s = 1; }
            else {
                fl = true;
                wi = st.a++;
                st.s = 0;
                p.wsorted[wi] = WinSorted{key[k], ent[k], st.b, pos - st.b + 1};
            }
            p.flush[pos] = fl ? 1 : 0;
            if (ent[k] >= p.P) p.wi_of_entry[ent[k] - p.P] = wi;
        }
    }
}

// ---------------------------------------------------------------------------------------
// reverse: walking positions downward, state C=0 (group closed by a later flush) /
// O=1 (open: reaches its PID's end unflushed => stays in process_memory).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(WIN_THREADS) window_reverse(const WinParams p) {
    __shared__ Fn2 warp_tot[WIN_THREADS / 32];
    __shared__ uint32_t s_tile;
    __shared__ St2 s_in;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_tiles = (p.M + WIN_TILE - 1) / WIN_TILE;
    while (true) {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
        const uint32_t q0 = tile * WIN_TILE + tid * WIN_ITEMS;  // reversed index q = M-1-pos
        uint32_t key[WIN_ITEMS];
        bool setter[WIN_ITEMS], fl[WIN_ITEMS];
        uint32_t nextk = 0;  // key at pos+1
        bool have_next = false;
        if (q0 > 0 && q0 <= p.M) { nextk = p.skey[p.M - q0]; have_next = true; }
        Fn2 f = fn_identity();
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            const uint32_t q = q0 + k;
            if (q < p.M) {
                const uint32_t pos = p.M - 1 - q;
                key[k] = p.skey[pos];
                fl[k] = p.flush[pos] != 0;
                const bool tail = !have_next || key[k] != nextk;
                setter[k] = tail || fl[k];
                nextk = key[k];
                have_next = true;
                Fn2 e;
                if (setter[k]) e = fl[k] ? Fn2{0x0u, {0u, 0u}, {NONE32, NONE32}} : Fn2{0x3u, {1u, 1u}, {NONE32, NONE32}};
                else e = Fn2{0x2u, {0u, 1u}, {NONE32, NONE32}};
                f = fn_compose(f, e);
            } else {
                key[k] = 0; setter[k] = false; fl[k] = false;
            }
        }
        Fn2 total;
        Fn2 excl = fn_block_exclusive(f, warp_tot, &total);
        if (warp == 0) {
            St2 in = fn_lookback(p.status, tile, total, St2{0u, 0u, NONE32}, lane);
            if (lane == 0) {
                s_in = in;
                if (tile == n_tiles - 1) *p.n_pending = fn_apply(total, in).a;
            }
        }
        __syncthreads();
        St2 st = fn_apply(excl, s_in);
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            const uint32_t q = q0 + k;
            if (q >= p.M) break;
            const uint32_t pos = p.M - 1 - q;
            if (setter[k]) st.s = fl[k] ? 0u : 1u;
            if (st.s == 1u) {
                const uint32_t slot = p.pend_cap - 1u - st.a;
                st.a++;
                const uint32_t e = p.sval[pos];
                p.npend_key[slot] = key[k];
                p.npend_slot[slot] = pos;
                p.npend_arrival[slot] = win_arrival_of(p, e);
                uint8_t pm = (uint8_t)(win_meta_of(p, e) & ~META_TRIG);
                if (p.match && p.match[pos]) pm |= (uint8_t)META_CHAIN;
                p.npend_meta[slot] = pm;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// order: windows sorted by the arrival index of their trigger event = compaction of
// wi_of_entry in entry order (entry order of this batch's survivors IS arrival order).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(WIN_THREADS) window_order(const WinParams p) {
    __shared__ uint32_t warp_sum[WIN_THREADS / 32];
    __shared__ uint32_t s_tile, s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t S = p.M - p.P;
    const uint32_t n_tiles = (S + WIN_TILE - 1) / WIN_TILE;
    while (true) {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
        const uint32_t j0 = tile * WIN_TILE + tid * WIN_ITEMS;
        uint32_t wi[WIN_ITEMS];
        uint32_t cnt = 0;
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            wi[k] = (j0 + k < S) ? p.wi_of_entry[j0 + k] : NONE32;
            cnt += wi[k] != NONE32;
        }
        uint32_t inc = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += t;
        }
        if (lane == 31) warp_sum[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t tot = 0;
#pragma unroll
            for (int w = 0; w < WIN_THREADS / 32; w++) tot += warp_sum[w];
            uint32_t excl = lookback_count(p.status, tile, tot, lane);
            if (lane == 0) s_base = excl;
        }
        __syncthreads();
        uint32_t r = s_base + inc - cnt;
#pragma unroll
        for (int w = 0; w < WIN_THREADS / 32; w++) r += (w < warp) ? warp_sum[w] : 0u;
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            if (wi[k] == NONE32) continue;
            const WinSorted d = p.wsorted[wi[k]];
            chronos_window o;
            o.trigger_idx = win_arrival_of(p, d.trig_entry);
            o.start = d.start;
            o.len = d.len;
            o.pid = d.pid;
            o.flags = 0;
            if (p.mcnt) {
                const uint32_t hi = p.mcnt[d.start + d.len - 1], lo = d.start ? p.mcnt[d.start - 1] : 0u;
                o.flags = hi != lo ? 1u : 0u;
            }
            o.reserved = 0;
            p.windows[r++] = o;
        }
    }
}

// ---------------------------------------------------------------------------------------
// gather: arena[pos] = record of entry sval[pos]; every surviving event is written
// exactly once (the 288*S term of the algorithmic bytes).
// ---------------------------------------------------------------------------------------
struct GatherParams {
    const uint32_t *sval;
    uint32_t M, P;
    const uint32_t *surv_src;
    const uint8_t *batch;        // this batch's records (device)
    const uint32_t *pend_slot;   // P
    const uint8_t *prev_arena;
    uint8_t *arena;
};

__global__ void __launch_bounds__(256) gather_records(const GatherParams g) {
    const int lane = threadIdx.x & 31;
    const uint32_t warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t n_warps = (gridDim.x * blockDim.x) >> 5;
    constexpr uint32_t GROUP = 16;  // records per warp step: 16*18 = 288 chunks = 9 per lane
    for (uint32_t base = warp_global * GROUP; base < g.M; base += n_warps * GROUP) {
        const uint32_t cnt = min(GROUP, g.M - base);
        // lane r (< cnt) resolves the source of record base+r
        const uint8_t *src = nullptr;
        if ((uint32_t)lane < cnt) {
            uint32_t e = g.sval[base + lane];
            src = e < g.P ? g.prev_arena + (size_t)g.pend_slot[e] * CH_REC_BYTES
                          : g.batch + (size_t)g.surv_src[e - g.P] * CH_REC_BYTES;
        }
        uint4 *dst = reinterpret_cast<uint4 *>(g.arena + (size_t)base * CH_REC_BYTES);
#pragma unroll
        for (int it = 0; it < 9; it++) {
            const uint32_t c = it * 32 + lane;        // chunk id within the group
            const uint32_t r = c / CH_REC_CHUNKS;     // record within the group
            const uint32_t ch = c - r * CH_REC_CHUNKS;
            const unsigned long long sp =
                __shfl_sync(0xffffffffu, (unsigned long long)(uintptr_t)src, (int)(r & 31u));
            if (r < cnt) dst[c] = reinterpret_cast<const uint4 *>((uintptr_t)sp)[ch];
        }
    }
}

}  // namespace chronos
