// window.cuh -- kernels 3/4: per-PID context windows (and the kill-chain labeller) over the
// pid-grouped entries.
//
// Reference semantics (chronos_sensor.py): append to process_memory[pid] (:138); if the
// event string holds a trigger needle (:141) and the list now has >= 2 entries (:142)
// the whole list is emitted as one window and the list is cleared (:157).
// Per PID that is a 2-state automaton over its events:
//     state E (list empty) / N (non-empty);   non-trigger: -> N
//     trigger in E: -> N (no flush)           trigger in N: FLUSH, -> E
// The first event of a PID always ends in N. Events still pending from earlier batches
// re-enter the grouped array as non-trigger entries ahead of the batch's survivors, so
// the window logic needs no per-PID state lookup.
//
// EXTENSION (no reference semantics, SURVEY 8a "parity unpinned"; defined by
// oracle/chronos_oracle.c::co_feed): chronos_sensor.py only *describes* the dropper
// pattern to the LLM ("'curl' -> 'chmod' -> 'exec' is a Dropper", :112). Here it is a
// per-PID DFA over the surviving events in arrival order,
//     d=0 --"curl" in event_str--> 1 --"chmod"--> 2 --type is EXEC--> match, back to 0,
// one step per event, state kept per PID across flushes and batches (pid table, chain.cuh).
// It never changes which events survive or which windows are emitted.
//
//   group_forward   : both automata in ONE single-pass function scan (fnscan.cuh):
//                     flush bits, window descriptors, chain matches, per-PID table update
//   window_reverse  : "is my group ever closed?" -> pending-list compaction (process_memory)
//   emit_ordered    : windows and chain matches in arrival order of their trigger event
//                     (:143-147): two compactions over the batch's entries in one pass
//   gather_records  : 288-byte records copied once into the pid-grouped arena
//
// Sorted values carry the event's class bits so that no kernel after the sort has to
// gather them again: value = entry id (27 bits) | META bits << 27.
#pragma once
#include "chain.cuh"
#include "common.cuh"
#include "filter.cuh"
#include "fnscan.cuh"

namespace chronos {

constexpr int VAL_ID_BITS = 27;
constexpr uint32_t VAL_ID_MASK = (1u << VAL_ID_BITS) - 1u;
constexpr int WIN_THREADS = 256;
#ifndef CHRONOS_WIN_ITEMS
#define CHRONOS_WIN_ITEMS 8   // 16 -> 8: same throughput on 10 M-event batches, 64K-event batches 0.21 -> 0.15 ms
#endif
constexpr int WIN_ITEMS = CHRONOS_WIN_ITEMS;
constexpr int WIN_TILE = WIN_THREADS * WIN_ITEMS;
// per-position flags written by group_forward: bit0 flush, bit1 chain match, bits 31:2 =
// inclusive count of chain-complete events up to this position
constexpr uint32_t PF_FLUSH = 1u, PF_MATCH = 2u;
// per-entry result: bit 31 = this event completed a chain, low 31 bits = ordinal of the
// window it flushed (EO_NONE if it flushed nothing)
constexpr uint32_t EO_NONE = 0x7fffffffu, EO_MATCH = 0x80000000u;

struct WinSorted {  // window descriptor in pid-grouped order
    uint32_t pid, trig_entry, start, len;
};

struct WinParams {
    const uint32_t *skey;   // M sorted pids
    const uint32_t *sval;   // M values: entry id | meta << 27; ids [0,P) carried-in pending, [P,M) survivors
    uint32_t M, P;
    const Plan *plan;               // when set, skey / sval / M / P are taken from it on the device
    const uint32_t *surv_src;       // S batch-local record index
    const uint64_t *surv_arrival;   // S explicit arrival indices, or nullptr => base_arrival + surv_src
    uint64_t base_arrival;
    const uint64_t *pend_arrival;   // previous batch's pending arrival indices, right-aligned in
    uint32_t pend_cap_prev;         //   an array of pend_cap_prev entries (P of them, device-side count)
    Carry *carry;                   // window_reverse leaves the new pending count here for the next batch
    const uint32_t *surv_pid;       // S pids in arrival order (the sort consumed the key array)
    // per-PID table (chain.cuh)
    uint64_t *pid_tab;
    uint32_t pid_mask;
    uint32_t pid_epoch;
    uint32_t *pid_count;
    uint32_t *n_pids_out;           // emit_ordered leaves the table's count here for the host
    // forward outputs
    uint32_t *pflag;                // M
    uint32_t *entry_out;            // S
    WinSorted *wsorted;             // window descriptors in grouped order
    uint32_t *n_windows;
    // reverse outputs (right-aligned in arrays of capacity pend_cap)
    uint32_t pend_cap;
    uint32_t *npend_key;
    uint32_t *npend_slot;
    uint64_t *npend_arrival;
    uint8_t *npend_meta;
    uint32_t *n_pending;
    // ordered outputs
    chronos_window *windows;        // n_windows, ascending trigger arrival
    uint32_t *chain_pid;            // n_chains, ascending arrival
    uint64_t *chain_idx;
    uint32_t *n_chains;
    // scheduling
    uint32_t *ticket;               // one counter per kernel
    uint64_t *status;               // look-back words for this kernel
};

// M, P and the sorted arrays are resolved into plain locals at the top of every kernel.
// (Patching a by-value copy of the parameter struct from the plan did not work: nvcc 12.9
// kept reading the ORIGINAL skey pointer out of the parameter bank)
__device__ __forceinline__ uint64_t win_arrival_of(const WinParams &p, uint32_t P, uint32_t e) {
    if (e < P) return p.pend_arrival[p.pend_cap_prev - P + e];
    const uint32_t j = e - P;
    return p.surv_arrival ? p.surv_arrival[j] : p.base_arrival + p.surv_src[j];
}

// what one surviving event does to the chain DFA
__device__ __forceinline__ Fn3 chain_event(uint32_t meta) {
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

// ---------------------------------------------------------------------------------------
// forward. Flush automaton: E=0 / N=1, marker b = start position of the current group,
// counter a = windows flushed so far. Chain DFA: counter a = chain-complete events so far.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void group_forward_body(const WinParams &p) {
    const uint32_t M = p.plan ? p.plan->M : p.M, P = p.plan ? p.plan->P : p.P;
    const uint32_t *const skey = p.plan ? p.plan->skey : p.skey;
    const uint32_t *const sval = p.plan ? p.plan->sval : p.sval;
    __shared__ FnG warp_tot[WIN_THREADS / 32];
    __shared__ uint32_t s_tile;
    __shared__ StG s_in;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_tiles = (M + WIN_TILE - 1) / WIN_TILE;
    while (true) {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
        const uint32_t p0 = tile * WIN_TILE + tid * WIN_ITEMS;
        uint32_t key[WIN_ITEMS], val[WIN_ITEMS];
        uint8_t kind[WIN_ITEMS];  // bit0 head, bits 2:1 the PID's stored DFA state at a head
        uint32_t prevk = (p0 > 0 && p0 - 1 < M) ? skey[p0 - 1] : 0;
        FnG f = FnG::identity();
        uint32_t new_pids = 0;  // PIDs this thread inserted into the per-PID table
        // pass 1: keys, values, group heads
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            const uint32_t pos = p0 + k;
            key[k] = 0; val[k] = 0; kind[k] = 0;
            if (pos < M) {
                key[k] = skey[pos];
                val[k] = sval[pos];
                kind[k] = ((pos == 0) || key[k] != prevk) ? 1u : 0u;
                prevk = key[k];
            }
        }
        // pass 2: the stored chain state of every head. The home slots of all heads are read
        // first (independent loads, in flight together); only a miss -- a new PID or a collision
        // -- takes the probing path.
        uint64_t home[WIN_ITEMS];
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            home[k] = 0;
            if (kind[k] & 1u) home[k] = ld_relaxed_u64(p.pid_tab + pt_home(key[k], p.pid_mask));
        }
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            if (kind[k] & 1u) {
                uint32_t d0 = 0;
                if (pt_live(home[k], p.pid_epoch) && (uint32_t)home[k] == key[k]) d0 = (uint32_t)(home[k] >> 32) & 3u;
                else pt_find_or_insert(p.pid_tab, p.pid_mask, p.pid_epoch, key[k], new_pids, &d0);
                kind[k] |= (uint8_t)(d0 << 1);
            }
        }
        // pass 3: what the thread's items do to the two automata
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            const uint32_t pos = p0 + k;
            if (pos < M) {
                const uint32_t meta = val[k] >> VAL_ID_BITS;
                const bool carried = (val[k] & VAL_ID_MASK) < P;
                const bool head = kind[k] & 1u;
                const bool trig = (meta & META_TRIG) != 0;  // carried entries were stored without it
                FnG e;
                if (head) e.w = Fn2{0x3u, {0u, 0u}, {pos, pos}};
                else if (!trig) e.w = Fn2{0x3u, {0u, 0u}, {pos, NONE32}};
                else e.w = Fn2{0x1u, {0u, 1u}, {pos, NONE32}};
                // a carried-in entry was stepped through the DFA when it arrived; it only brings
                // its "completed a chain" mark along for the window label
                const uint32_t cm = (meta & META_CHAIN) ? 1u : 0u;
                e.c = carried ? Fn3{0x24u, {cm, cm, cm}} : chain_event(meta);
                if (head) {  // the PID's stored DFA state replaces whatever is carried over
                    const uint32_t d0 = (kind[k] >> 1) & 3u;
                    e.c = Fn3::constant(Fn3::o(e.c, d0), Fn3::av(e.c, d0));
                }
                f = FnG::compose(f, e);
            }
        }
        FnG total;
        FnG excl = fn_block_exclusive<FnG, WIN_THREADS>(f, warp_tot, &total);
        if (warp == 0) {
            StG in = fn_lookback<FnG>(p.status, tile, total, StG{St2{0u, 0u, NONE32}, St3{0u, 0u}}, lane);
            if (lane == 0) {
                s_in = in;
                if (tile == n_tiles - 1) *p.n_windows = FnG::apply(total, in).w.a;
            }
        }
        __syncthreads();
        StG st = FnG::apply(excl, s_in);
        uint32_t stored_dfa = NONE32;  // what the table holds for the current group, if its head is ours
        uint32_t pf[WIN_ITEMS];
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            const uint32_t pos = p0 + k;
            pf[k] = 0;
            if (pos >= M) continue;
            const uint32_t meta = val[k] >> VAL_ID_BITS, e = val[k] & VAL_ID_MASK;
            const bool carried = e < P, head = kind[k] & 1u, trig = (meta & META_TRIG) != 0;
            // ---- flush automaton
            bool fl = false;
            uint32_t wi = EO_NONE;
            if (head) { st.w.b = pos; st.w.s = 1; }
            else if (!trig) { if (st.w.s == 0) st.w.b = pos; st.w.s = 1; }
            else if (st.w.s == 0) { st.w.b = pos; st.w.s = 1; }
            else {
                fl = true;
                wi = st.w.a++;
                st.w.s = 0;
                CH_ASSERT(wi < p.plan->cap_w);
                p.wsorted[wi] = WinSorted{key[k], e, st.w.b, pos - st.w.b + 1};
            }
            // ---- chain DFA
            if (head) {
                st.c.s = (kind[k] >> 1) & 3u;
                stored_dfa = st.c.s;
            }
            bool m = false;
            if (carried) {
                st.c.a += (meta & META_CHAIN) ? 1u : 0u;
            } else {
                if (st.c.s == 2u) { if (meta & META_EXEC) { m = true; st.c.s = 0; st.c.a++; } }
                else if (st.c.s == 1u) { if (meta & META_CHMOD) st.c.s = 2; }
                else if (meta & META_CURL) st.c.s = 1;
                CH_ASSERT(e >= P && e - P < p.plan->cap_s);
                p.entry_out[e - P] = wi | (m ? EO_MATCH : 0u);
            }
            pf[k] = (fl ? PF_FLUSH : 0u) | (m ? PF_MATCH : 0u) | (st.c.a << 2);
            const bool tail = (pos + 1 == M) || skey[pos + 1] != key[k];
            if (tail) {
                // the PID's DFA state after this batch. Only chain events move it, so when this
                // thread also saw the group's head and the state is what the table already
                // holds, there is nothing to write and no second probe.
                if (stored_dfa != st.c.s) {
                    const uint32_t slot = pt_find_or_insert(p.pid_tab, p.pid_mask, p.pid_epoch, key[k], new_pids);
                    st_relaxed_u64(p.pid_tab + slot, pt_make(p.pid_epoch, st.c.s, key[k]));
                }
                stored_dfa = NONE32;
            }
        }
        new_pids = __reduce_add_sync(0xffffffffu, new_pids);
        if (lane == 0 && new_pids) atomicAdd(p.pid_count, new_pids);
        if (p0 + WIN_ITEMS <= M) {
            uint4 *dst = reinterpret_cast<uint4 *>(p.pflag + p0);
#pragma unroll
            for (int k = 0; k < WIN_ITEMS; k += 4) dst[k / 4] = make_uint4(pf[k], pf[k + 1], pf[k + 2], pf[k + 3]);
        } else {
#pragma unroll
            for (int k = 0; k < WIN_ITEMS; k++)
                if (p0 + k < M) p.pflag[p0 + k] = pf[k];
        }
    }
}

__global__ void __launch_bounds__(WIN_THREADS) group_forward(const WinParams p) { group_forward_body(p); }

// ---------------------------------------------------------------------------------------
// reverse: walking positions downward, state C=0 (group closed by a later flush) /
// O=1 (open: reaches its PID's end unflushed => stays in process_memory).
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void window_reverse_body(const WinParams &p) {
    const uint32_t M = p.plan ? p.plan->M : p.M, P = p.plan ? p.plan->P : p.P;
    const uint32_t *const skey = p.plan ? p.plan->skey : p.skey;
    const uint32_t *const sval = p.plan ? p.plan->sval : p.sval;
    __shared__ Fn2 warp_tot[WIN_THREADS / 32];
    __shared__ uint32_t s_tile;
    __shared__ St2 s_in;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_tiles = (M + WIN_TILE - 1) / WIN_TILE;
    while (true) {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
        const uint32_t q0 = tile * WIN_TILE + tid * WIN_ITEMS;  // reversed index q = M-1-pos
        uint32_t key[WIN_ITEMS];
        uint8_t kd[WIN_ITEMS];  // bit0 setter, bit1 flush, bit2 chain match
        uint32_t nextk = 0;     // key at pos+1
        bool have_next = false;
        if (q0 > 0 && q0 <= M) { nextk = skey[M - q0]; have_next = true; }
        Fn2 f = Fn2::identity();
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            const uint32_t q = q0 + k;
            key[k] = 0; kd[k] = 0;
            if (q < M) {
                const uint32_t pos = M - 1 - q;
                key[k] = skey[pos];
                const uint32_t pfl = p.pflag[pos];
                const bool fl = (pfl & PF_FLUSH) != 0;
                const bool tail = !have_next || key[k] != nextk;
                const bool setter = tail || fl;
                nextk = key[k];
                have_next = true;
                kd[k] = (uint8_t)((setter ? 1u : 0u) | (fl ? 2u : 0u) | ((pfl & PF_MATCH) ? 4u : 0u));
                Fn2 e;
                if (setter) e = fl ? Fn2{0x0u, {0u, 0u}, {NONE32, NONE32}} : Fn2{0x3u, {1u, 1u}, {NONE32, NONE32}};
                else e = Fn2{0x2u, {0u, 1u}, {NONE32, NONE32}};
                f = Fn2::compose(f, e);
            }
        }
        Fn2 total;
        Fn2 excl = fn_block_exclusive<Fn2, WIN_THREADS>(f, warp_tot, &total);
        if (warp == 0) {
            St2 in = fn_lookback<Fn2>(p.status, tile, total, St2{0u, 0u, NONE32}, lane);
            if (lane == 0) {
                s_in = in;
                if (tile == n_tiles - 1) {
                    const uint32_t np = Fn2::apply(total, in).a;
                    *p.n_pending = np;
                    p.carry->P = np;  // every kernel of this batch takes P from the plan, not from here
                }
            }
        }
        __syncthreads();
        St2 st = Fn2::apply(excl, s_in);
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            const uint32_t q = q0 + k;
            if (q >= M) break;
            const uint32_t pos = M - 1 - q;
            if (kd[k] & 1u) st.s = (kd[k] & 2u) ? 0u : 1u;
            if (st.s == 1u) {
                CH_ASSERT(st.a < p.pend_cap && pos < M);
                const uint32_t slot = p.pend_cap - 1u - st.a;
                st.a++;
                const uint32_t v = sval[pos];
                p.npend_key[slot] = key[k];
                p.npend_slot[slot] = pos;
                p.npend_arrival[slot] = win_arrival_of(p, P, v & VAL_ID_MASK);
                // stored without the trigger bit (a pending event never triggers again), with
                // the chain mark if it completed a chain, for the label of its future window
                p.npend_meta[slot] = (uint8_t)(((v >> VAL_ID_BITS) & ~META_TRIG) | ((kd[k] & 4u) ? META_CHAIN : 0u));
            }
        }
    }
}

__global__ void __launch_bounds__(WIN_THREADS) window_reverse(const WinParams p) { window_reverse_body(p); }

// ---------------------------------------------------------------------------------------
// ordered emission: compaction of entry_out over this batch's survivors, whose entry order
// IS arrival order: windows by trigger arrival (:143-147) and chain matches by arrival.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void emit_ordered_body(const WinParams &p) {
    const uint32_t M = p.plan ? p.plan->M : p.M, P = p.plan ? p.plan->P : p.P;
    __shared__ uint32_t warp_w[WIN_THREADS / 32], warp_c[WIN_THREADS / 32];
    __shared__ uint32_t s_tile, s_base_w, s_base_c;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t S = M - P;
    const uint32_t n_tiles = (S + WIN_TILE - 1) / WIN_TILE;
    if (blockIdx.x == 0 && threadIdx.x == 0 && p.n_pids_out) *p.n_pids_out = *p.pid_count;
    while (true) {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
        const uint32_t j0 = tile * WIN_TILE + tid * WIN_ITEMS;
        uint32_t eo[WIN_ITEMS];
        uint32_t cw = 0, cc = 0;
        if (j0 + WIN_ITEMS <= S) {
            const uint4 *src = reinterpret_cast<const uint4 *>(p.entry_out + j0);
#pragma unroll
            for (int k = 0; k < WIN_ITEMS; k += 4) {
                const uint4 v = src[k / 4];
                eo[k] = v.x; eo[k + 1] = v.y; eo[k + 2] = v.z; eo[k + 3] = v.w;
            }
        } else {
#pragma unroll
            for (int k = 0; k < WIN_ITEMS; k++) eo[k] = (j0 + k < S) ? p.entry_out[j0 + k] : EO_NONE;
        }
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            cw += (eo[k] & EO_NONE) != EO_NONE;
            cc += (eo[k] & EO_MATCH) != 0;
        }
        uint32_t iw = cw, ic = cc;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t tw = __shfl_up_sync(0xffffffffu, iw, d), tc = __shfl_up_sync(0xffffffffu, ic, d);
            if (lane >= d) { iw += tw; ic += tc; }
        }
        if (lane == 31) { warp_w[warp] = iw; warp_c[warp] = ic; }
        __syncthreads();
        if (warp == 0) {
            uint32_t tw = 0, tc = 0;
#pragma unroll
            for (int w = 0; w < WIN_THREADS / 32; w++) { tw += warp_w[w]; tc += warp_c[w]; }
            const uint2 ex = lookback_count2(p.status, tile, tw, tc, lane);
            if (lane == 0) {
                s_base_w = ex.x;
                s_base_c = ex.y;
                if (tile == n_tiles - 1) *p.n_chains = ex.y + tc;
            }
        }
        __syncthreads();
        uint32_t rw = s_base_w + iw - cw, rc = s_base_c + ic - cc;
#pragma unroll
        for (int w = 0; w < WIN_THREADS / 32; w++) {
            rw += (w < warp) ? warp_w[w] : 0u;
            rc += (w < warp) ? warp_c[w] : 0u;
        }
        // window descriptors, four at a time: the three levels of dependent random reads
        // (descriptor -> trigger arrival / chain counters) of four windows are in flight together
#pragma unroll
        for (int k0 = 0; k0 < WIN_ITEMS; k0 += 4) {
            WinSorted d[4];
            bool has[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const uint32_t wi = eo[k0 + u] & EO_NONE;
                has[u] = wi != EO_NONE;
                d[u] = WinSorted{0u, 0u, 0u, 1u};
                if (has[u]) d[u] = p.wsorted[wi];
            }
            uint64_t arr[4];
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                arr[u] = 0; hi[u] = 0; lo[u] = 0;
                if (has[u]) {
                    arr[u] = win_arrival_of(p, P, d[u].trig_entry);
                    hi[u] = p.pflag[d[u].start + d[u].len - 1] >> 2;
                    lo[u] = d[u].start ? (p.pflag[d[u].start - 1] >> 2) : 0u;
                }
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                if (has[u]) {
                    chronos_window o;
                    o.trigger_idx = arr[u];
                    o.start = d[u].start;
                    o.len = d[u].len;
                    o.pid = d[u].pid;
                    o.flags = hi[u] != lo[u] ? 1u : 0u;  // a chain-complete event lies inside the window
                    o.reserved = 0;
                    CH_ASSERT(rw < p.plan->cap_w && d[u].start + d[u].len <= M);
                    p.windows[rw++] = o;
                }
            }
        }
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; k++) {
            if (eo[k] & EO_MATCH) {
                const uint32_t j = j0 + k;
                CH_ASSERT(rc < p.plan->cap_s);
                p.chain_pid[rc] = p.surv_pid[j];
                p.chain_idx[rc] = p.surv_arrival ? p.surv_arrival[j] : p.base_arrival + p.surv_src[j];
                rc++;
            }
        }
    }
}

__global__ void __launch_bounds__(WIN_THREADS) emit_ordered(const WinParams p) { emit_ordered_body(p); }

// ---------------------------------------------------------------------------------------
// gather: arena[pos] = record of entry sval[pos]; every surviving event is written
// exactly once (the 288*S term of the algorithmic bytes).
// ---------------------------------------------------------------------------------------
struct GatherParams {
    const uint32_t *sval;        // sorted values (entry id in the low 27 bits)
    uint32_t M, P;
    const uint32_t *surv_src;
    const uint8_t *batch;        // this batch's records (device)
    const uint32_t *pend_slot;   // previous batch's pending arena slots, right-aligned in an array of
    uint32_t pend_cap_prev;      //   pend_cap_prev entries
    const uint8_t *prev_arena;
    uint8_t *arena;
    // transport form (chronos_submit_survivors): entry j's record is slot_ns[j] 16-byte slots
    // starting at slot_off[j] of `slots`; nullptr => `batch` holds plain 288-byte records
    const uint4 *slots;
    const uint32_t *slot_off;
    const uint8_t *slot_ns;
    const Plan *plan;            // when set, sval / M / P are taken from it on the device
};

// One warp moves 16 records (288 chunks of 16 bytes, 9 per lane) per step. The chunks travel
// global -> shared memory with cp.async (LDGSTS): the data in flight sits in shared memory, not
// in registers, so the kernel is small enough (about 32 registers a thread) to share every SM
// with the window scans that run beside it; two stages per warp keep the next group's loads
// in flight while the current one is stored.
constexpr int GATHER_THREADS = 256;
constexpr int GATHER_WARPS = GATHER_THREADS / 32;
constexpr uint32_t GATHER_GROUP = 16;
constexpr int GATHER_STAGES = 2;
constexpr size_t GATHER_SMEM = (size_t)GATHER_WARPS * GATHER_STAGES * GATHER_GROUP * CH_REC_BYTES;

__device__ __forceinline__ void gather_records_body(const GatherParams &g, uint4 *gather_smem) {
    const uint32_t gM = g.plan ? g.plan->M : g.M, gP = g.plan ? g.plan->P : g.P;
    const uint32_t *const gsval = g.plan ? g.plan->sval : g.sval;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t warp_global = blockIdx.x * GATHER_WARPS + warp;
    const uint32_t n_warps = gridDim.x * GATHER_WARPS;
    constexpr uint32_t GROUP_CHUNKS = GATHER_GROUP * CH_REC_CHUNKS;  // 288
    uint4 *const my = gather_smem + (size_t)warp * GATHER_STAGES * GROUP_CHUNKS;

    // all chunks of the group starting at `base` into stage `s` (asynchronous)
    auto issue = [&](uint32_t base, int s) {
        const uint32_t cnt = min(GATHER_GROUP, gM - base);
        const uint8_t *src = nullptr;  // lane r (< cnt) resolves the source of record base + r
        uint32_t ns = 0;               // 0: plain 288-byte record at src; else transport form of ns slots
        if ((uint32_t)lane < cnt) {
            const uint32_t e = gsval[base + lane] & VAL_ID_MASK;
            CH_ASSERT(e < gM && gM <= g.plan->cap_m);
            if (e < gP) src = g.prev_arena + (size_t)g.pend_slot[g.pend_cap_prev - gP + e] * CH_REC_BYTES;
            else if (g.slots) {
                src = reinterpret_cast<const uint8_t *>(g.slots + g.slot_off[e - gP]);
                ns = g.slot_ns[e - gP];
            } else src = g.batch + (size_t)g.surv_src[e - gP] * CH_REC_BYTES;
        }
        uint4 *stage = my + (size_t)s * GROUP_CHUNKS;
#pragma unroll
        for (int it = 0; it < 9; it++) {
            const uint32_t c = it * 32 + lane;        // chunk id within the group
            const uint32_t r = c / CH_REC_CHUNKS;     // record within the group
            const uint32_t ch = c - r * CH_REC_CHUNKS;
            const unsigned long long sp = __shfl_sync(0xffffffffu, (unsigned long long)(uintptr_t)src, (int)(r & 31u));
            const uint32_t rns = __shfl_sync(0xffffffffu, ns, (int)(r & 31u));
            // transport form: slot 0,1 = chunks 0,1; slot 2 = chunk 17; slots 3.. = chunks 2..;
            // chunks the entry does not carry are zeros (nothing but NUL padding was there)
            const uint32_t sl = rns == 0 ? ch : (ch < 2 ? ch : (ch == 17 ? 2u : ch + 1u));
            const bool have = rns == 0 || sl < rns;
            if (r < cnt) cp_async16(stage + c, reinterpret_cast<const uint4 *>((uintptr_t)sp) + (have ? sl : 0u), have ? 16u : 0u);
        }
        cp_async_commit();
    };

    // software pipeline: step k issues group k (if any) and then stores group k-1
    const uint32_t stride = n_warps * GATHER_GROUP;
    uint32_t prev = 0xffffffffu;
    int s = 0;
#pragma unroll 1
    for (uint32_t base = warp_global * GATHER_GROUP;; base += stride) {
        const bool have = base < gM;
        if (have) issue(base, s);
        else cp_async_commit();  // (an empty group keeps the wait count uniform)
        if (prev != 0xffffffffu) {
            cp_async_wait<1>();  // everything but the group just issued has landed
            __syncwarp();
            const uint32_t cnt = min(GATHER_GROUP, gM - prev);
            const uint4 *stage = my + (size_t)(s ^ 1) * GROUP_CHUNKS;
            uint4 *dst = reinterpret_cast<uint4 *>(g.arena + (size_t)prev * CH_REC_BYTES);
#pragma unroll 3
            for (int it = 0; it < 9; it++) {
                const uint32_t c = it * 32 + lane;
                if (c < cnt * CH_REC_CHUNKS) dst[c] = stage[c];
            }
            __syncwarp();  // the stage is refilled by the next issue
        }
        if (!have) break;
        prev = base;
        s ^= 1;
    }
}

__global__ void __launch_bounds__(GATHER_THREADS) gather_records(const GatherParams g) {
    extern __shared__ uint4 gather_smem_dyn[];
    gather_records_body(g, gather_smem_dyn);
}

// ---------------------------------------------------------------------------------------
// The whole tail in ONE launch, for batches small enough that launches, not bytes, are what
// they cost (a 64 K-event micro-batch groups about 12 K entries: every kernel above is a few
// microseconds of work behind a launch and a dependency). The kernel bodies are the ones
// above, run one after the other by a grid of at most one CTA per SM, with a grid-wide barrier
// where a phase needs everything of the previous one; window_reverse, emit_ordered and
// gather_records only depend on group_forward and follow without a barrier between them.
// All CTAs must be resident at the same time (they spin at the barriers): the host launches
// no more CTAs than there are SMs.
// ---------------------------------------------------------------------------------------
struct FusedTailParams {
    SortParams sort;
    uint32_t *pass_ticket[RADIX_MAX_PASSES];
    uint32_t pass_epoch[RADIX_MAX_PASSES];
    WinParams fwd, rev, emit;
    GatherParams gather;
    uint32_t *barrier;  // zero at launch
};

__device__ __forceinline__ void grid_barrier(uint32_t *counter, uint32_t target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1u);
        while (ld_acquire_u32(counter) < target) {
        }
        __threadfence();
    }
    __syncthreads();
}

__global__ void __launch_bounds__(WIN_THREADS) fused_tail(const FusedTailParams f) {
    extern __shared__ uint4 fused_smem_dyn[];
    uint32_t target = 0;
    sort_prepare_body(f.sort);
    grid_barrier(f.barrier, target += gridDim.x);
    const uint32_t n_passes = f.sort.plan->n_passes;
    for (uint32_t i = 0; i < n_passes; i++) {
        onesweep_pass_body(f.sort, (int)i, f.pass_ticket[i], f.pass_epoch[i]);
        grid_barrier(f.barrier, target += gridDim.x);
    }
    group_forward_body(f.fwd);
    grid_barrier(f.barrier, target += gridDim.x);
    window_reverse_body(f.rev);
    emit_ordered_body(f.emit);
    gather_records_body(f.gather, fused_smem_dyn);
}

}  // namespace chronos
