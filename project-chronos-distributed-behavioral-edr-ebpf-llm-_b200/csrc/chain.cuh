// chain.cuh -- the kill-chain labeller and the persistent per-PID table.
//
// EXTENSION, not reference behaviour: chronos_sensor.py only *describes* the dropper
// pattern to the LLM ("'curl' -> 'chmod' -> 'exec' is a Dropper", :112; README.md:32-37).
// Here it is a coded per-PID DFA over the surviving events in arrival order:
//     d=0 --event_str has "curl"--> 1 --has "chmod"--> 2 --type is EXEC--> match, back to 0
// one step per event, state kept per PID across window flushes and across batches. It never
// changes which events survive or which windows are emitted ("parity unpinned":
// its semantics are defined by oracle/chronos_oracle.c::co_feed).
//
// The same table gives the key set of process_memory (:105): every PID that ever had a
// surviving event, including those whose list is currently [] (:157).
//
// pid table: open addressing, u64 slots: bit 63 valid | bits 33:32 DFA state | low 32 pid.
// chain_scan: single-pass scan over the pid-grouped entries whose carried value is the
// tile's function over the three DFA states (exit state and match count for each entry
// state); a PID's first entry of the batch replaces the carried state by the table's.
#pragma once
#include "common.cuh"
#include "window.cuh"

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

// ---- function over the 3 DFA states ----------------------------------------------------
struct Fn3 {
    uint32_t out;   // 2 bits per entry state: exit state
    uint32_t a[3];  // chain-complete events counted for each entry state
};
__device__ __forceinline__ Fn3 fn3_identity() { return Fn3{0x24u, {0u, 0u, 0u}}; }  // 0->0, 1->1, 2->2
__device__ __forceinline__ uint32_t fn3_out(const Fn3 &f, uint32_t s) { return (f.out >> (2 * s)) & 3u; }
__device__ __forceinline__ uint32_t fn3_a(const Fn3 &f, uint32_t s) { return s == 0 ? f.a[0] : (s == 1 ? f.a[1] : f.a[2]); }
// first x, then y
__device__ __forceinline__ Fn3 fn3_compose(const Fn3 &x, const Fn3 &y) {
    Fn3 r;
    const uint32_t m0 = fn3_out(x, 0), m1 = fn3_out(x, 1), m2 = fn3_out(x, 2);
    r.out = fn3_out(y, m0) | (fn3_out(y, m1) << 2) | (fn3_out(y, m2) << 4);
    r.a[0] = x.a[0] + fn3_a(y, m0);
    r.a[1] = x.a[1] + fn3_a(y, m1);
    r.a[2] = x.a[2] + fn3_a(y, m2);
    return r;
}
__device__ __forceinline__ Fn3 fn3_shfl(const Fn3 &f, int src) {
    Fn3 r;
    r.out = __shfl_sync(0xffffffffu, f.out, src);
    r.a[0] = __shfl_sync(0xffffffffu, f.a[0], src);
    r.a[1] = __shfl_sync(0xffffffffu, f.a[1], src);
    r.a[2] = __shfl_sync(0xffffffffu, f.a[2], src);
    return r;
}
struct St3 {
    uint32_t s, a;
};
__device__ __forceinline__ St3 fn3_apply(const Fn3 &f, const St3 &c) { return St3{fn3_out(f, c.s), c.a + fn3_a(f, c.s)}; }

// what one surviving event does to the DFA (meta = CHRONOS_META_* of the event)
__device__ __forceinline__ Fn3 fn3_event(uint32_t meta) {
    Fn3 e;
    const uint32_t o0 = (meta & META_CURL) ? 1u : 0u;
    const uint32_t o1 = (meta & META_CHMOD) ? 2u : 1u;
    const uint32_t o2 = (meta & META_EXEC) ? 0u : 2u;
    e.out = o0 | (o1 << 2) | (o2 << 4);
    e.a[0] = 0;
    e.a[1] = 0;
    e.a[2] = (meta & META_EXEC) ? 1u : 0u;
    return e;
}
// constant function: whatever the entry state, exit in `s` having counted `a`
__device__ __forceinline__ Fn3 fn3_const(uint32_t s, uint32_t a) { return Fn3{s | (s << 2) | (s << 4), {a, a, a}}; }

__device__ __forceinline__ Fn3 fn3_block_exclusive(const Fn3 &mine, Fn3 *warp_tot, Fn3 *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Fn3 inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Fn3 o = fn3_shfl(inc, lane >= d ? lane - d : lane);
        if (lane >= d) inc = fn3_compose(o, inc);
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    Fn3 prev = fn3_shfl(inc, lane > 0 ? lane - 1 : 0);
    Fn3 excl = lane > 0 ? prev : fn3_identity();
    Fn3 run = fn3_identity(), wpre = fn3_identity();
#pragma unroll
    for (int w = 0; w < WIN_THREADS / 32; w++) {
        if (w == warp) wpre = run;
        run = fn3_compose(run, warp_tot[w]);
    }
    *total = run;
    return fn3_compose(wpre, excl);
}

// status word for entry state s: [63:62] status [61:60] out_s [31:0] a_s ; 3 words per tile
__device__ __forceinline__ uint64_t fn3_pack(uint64_t status, uint32_t out_s, uint32_t a_s) {
    return (status << 62) | ((uint64_t)(out_s & 3u) << 60) | a_s;
}
__device__ __forceinline__ St3 fn3_lookback(uint64_t *status, uint32_t tile, const Fn3 &agg, int lane) {
    St3 in{0u, 0u};
    if (tile != 0) {
        if (lane == 0)
            for (int s = 0; s < 3; s++) st_relaxed_u64(status + 3 * (size_t)tile + s, fn3_pack(ST_AGG, fn3_out(agg, s), fn3_a(agg, s)));
        Fn3 H = fn3_identity();
        int64_t look = (int64_t)tile - 1;
        while (true) {
            const int64_t idx = look - lane;
            uint64_t w0 = fn3_pack(ST_PREFIX, 0, 0), w1 = w0, w2 = w0;  // virtual tiles below 0
            if (idx >= 0) {
                do {  // all three words must carry the same status (see window.cuh on torn pairs)
                    w0 = ld_relaxed_u64(status + 3 * (size_t)idx);
                    w1 = ld_relaxed_u64(status + 3 * (size_t)idx + 1);
                    w2 = ld_relaxed_u64(status + 3 * (size_t)idx + 2);
                } while ((w0 >> 62) == ST_NONE || (w0 >> 62) != (w1 >> 62) || (w0 >> 62) != (w2 >> 62));
            }
            Fn3 mine;
            mine.out = (uint32_t)((w0 >> 60) & 3u) | ((uint32_t)((w1 >> 60) & 3u) << 2) | ((uint32_t)((w2 >> 60) & 3u) << 4);
            mine.a[0] = (uint32_t)w0;
            mine.a[1] = (uint32_t)w1;
            mine.a[2] = (uint32_t)w2;
            const uint32_t is_p = __ballot_sync(0xffffffffu, (w0 >> 62) == ST_PREFIX);
            const int first = is_p ? __ffs((int)is_p) - 1 : 32;
            Fn3 G = (lane < first) ? mine : fn3_identity();
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const Fn3 older = fn3_shfl(G, (lane + d) & 31);
                if (lane + d < 32) G = fn3_compose(older, G);
            }
            G = fn3_shfl(G, 0);
            H = fn3_compose(G, H);
            if (is_p) {
                const Fn3 pf = fn3_shfl(mine, first);
                in = fn3_apply(H, St3{pf.out & 3u, pf.a[0]});
                break;
            }
            look -= 32;
        }
    }
    const St3 outc = fn3_apply(agg, in);
    if (lane == 0)
        for (int s = 0; s < 3; s++) st_relaxed_u64(status + 3 * (size_t)tile + s, fn3_pack(ST_PREFIX, outc.s, outc.a));
    return in;
}

struct ChainParams {
    const uint32_t *skey;
    const uint32_t *sval;
    uint32_t M, P;
    const uint8_t *surv_meta;
    const uint8_t *pend_meta;
    uint64_t *pid_tab;
    uint32_t pid_mask;
    uint32_t *pid_count;
    uint8_t *match;     // M: 1 where this batch's event completed a chain (pid-grouped order)
    uint8_t *match_entry;  // S: the same flag in entry (= arrival) order
    uint32_t *mcnt;     // M: inclusive count of chain-complete events (new + carried) up to pos
    uint32_t *ticket;
    uint64_t *status;   // 3 words per tile
};

__global__ void __launch_bounds__(WIN_THREADS) chain_scan(const ChainParams p) {
    __shared__ Fn3 warp_tot[WIN_THREADS / 32];
    __shared__ uint32_t s_tile;
    __shared__ St3 s_in;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_tiles = (p.M + WIN_TILE - 1) / WIN_TILE;
    while (true) {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
        const uint32_t p0 = tile * WIN_TILE + tid * WIN_ITEMS;
        uint32_t key[WIN_ITEMS];
        uint8_t meta[WIN_ITEMS], kind[WIN_ITEMS];  // kind: bit0 head, bit1 carried-in entry, bits 3:2 table state at a head
        uint32_t prevk = (p0 > 0 && p0 - 1 < p.M) ? p.skey[p0 - 1] : 0;
        Fn3 f = fn3_identity();
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            const uint32_t pos = p0 + k;
            key[k] = 0; meta[k] = 0; kind[k] = 0;
            if (pos < p.M) {
                key[k] = p.skey[pos];
                const uint32_t e = p.sval[pos];
                const bool carried = e < p.P;
                meta[k] = carried ? p.pend_meta[e] : p.surv_meta[e - p.P];
                const bool head = (pos == 0) || key[k] != prevk;
                prevk = key[k];
                // a carried-in entry was stepped through the DFA when it arrived; it only
                // contributes its "completed a chain" mark to the window label
                Fn3 ev = carried ? Fn3{0x24u, {(meta[k] & META_CHAIN) ? 1u : 0u, (meta[k] & META_CHAIN) ? 1u : 0u,
                                               (meta[k] & META_CHAIN) ? 1u : 0u}}
                                 : fn3_event(meta[k]);
                uint32_t kd = (head ? 1u : 0u) | (carried ? 2u : 0u);
                if (head) {
                    const uint32_t slot = pt_find_or_insert(p.pid_tab, p.pid_mask, key[k], p.pid_count);
                    const uint32_t d0 = (uint32_t)(ld_relaxed_u64(p.pid_tab + slot) >> 32) & 3u;
                    kd |= d0 << 2;
                    ev = fn3_const(fn3_out(ev, d0), fn3_a(ev, d0));  // the PID's stored state replaces the carried one
                }
                kind[k] = (uint8_t)kd;
                f = fn3_compose(f, ev);
            }
        }
        Fn3 total;
        Fn3 excl = fn3_block_exclusive(f, warp_tot, &total);
        if (warp == 0) {
            St3 in = fn3_lookback(p.status, tile, total, lane);
            if (lane == 0) s_in = in;
        }
        __syncthreads();
        St3 st = fn3_apply(excl, s_in);
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            const uint32_t pos = p0 + k;
            if (pos >= p.M) break;
            if (kind[k] & 1u) st.s = (kind[k] >> 2) & 3u;
            bool m = false;
            if (kind[k] & 2u) {
                st.a += (meta[k] & META_CHAIN) ? 1u : 0u;
            } else {
                const uint32_t mt = meta[k];
                if (st.s == 2u) { if (mt & META_EXEC) { m = true; st.s = 0; st.a++; } }
                else if (st.s == 1u) { if (mt & META_CHMOD) st.s = 2; }
                else if (mt & META_CURL) st.s = 1;
            }
            p.match[pos] = m ? 1 : 0;
            p.mcnt[pos] = st.a;
            if (!(kind[k] & 2u)) p.match_entry[p.sval[pos] - p.P] = m ? 1 : 0;
            const bool tail = (pos + 1 == p.M) || p.skey[pos + 1] != key[k];
            if (tail) {  // the PID's DFA state after this batch
                const uint32_t slot = pt_find_or_insert(p.pid_tab, p.pid_mask, key[k], p.pid_count);
                st_relaxed_u64(p.pid_tab + slot, PT_VALID | ((uint64_t)st.s << 32) | key[k]);
            }
        }
    }
}

// Completed chains in arrival order: compaction of match_entry over this batch's survivors.
struct ChainOrderParams {
    const uint8_t *match_entry;  // S
    const uint8_t *batch;        // this batch's records: the pid is read back from the record
    const uint32_t *surv_src;
    const uint64_t *surv_arrival;  // or nullptr => base + src
    uint64_t base_arrival;
    uint32_t S;
    uint32_t *chain_pid;
    uint64_t *chain_idx;
    uint32_t *n_chains;
    uint32_t *ticket;
    uint64_t *status;
};

__global__ void __launch_bounds__(WIN_THREADS) chain_order(const ChainOrderParams p) {
    __shared__ uint32_t warp_sum[WIN_THREADS / 32];
    __shared__ uint32_t s_tile, s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_tiles = (p.S + WIN_TILE - 1) / WIN_TILE;
    while (true) {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
        const uint32_t j0 = tile * WIN_TILE + tid * WIN_ITEMS;
        uint32_t mask = 0, cnt = 0;
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            const bool m = (j0 + k < p.S) && p.match_entry[j0 + k];
            mask |= m ? (1u << k) : 0u;
            cnt += m;
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
            const uint32_t excl = lookback_count(p.status, tile, tot, lane);
            if (lane == 0) {
                s_base = excl;
                if (tile == n_tiles - 1) *p.n_chains = excl + tot;
            }
        }
        __syncthreads();
        uint32_t r = s_base + inc - cnt;
#pragma unroll
        for (int w = 0; w < WIN_THREADS / 32; w++) r += (w < warp) ? warp_sum[w] : 0u;
        while (mask) {
            const int k = __ffs((int)mask) - 1;
            mask &= mask - 1;
            const uint32_t j = j0 + k;
            p.chain_pid[r] = *reinterpret_cast<const uint32_t *>(p.batch + (size_t)p.surv_src[j] * CH_REC_BYTES);
            p.chain_idx[r] = p.surv_arrival ? p.surv_arrival[j] : p.base_arrival + p.surv_src[j];
            r++;
        }
    }
}

}  // namespace chronos
