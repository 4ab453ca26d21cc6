// fnscan.cuh -- single-pass scans whose carried value is a FUNCTION over a small automaton.
//
// The per-PID logic of chronos_sensor.py is sequential per PID (append :138, test :141-142,
// clear :157). To run it as a data-parallel scan over the pid-grouped entries, a run of
// elements is summarised by what it does to the automaton for EVERY possible entry state
// (exit state + counters + markers per entry state). Such summaries compose associatively,
// so the standard machinery applies: thread-serial -> warp shuffle scan -> block scan ->
// decoupled look-back across tiles (status words in global memory, one 64-bit word per
// entry state so that each word is written atomically).
//
//   Fn2  two states  (flush automaton E/N; also the reverse "open/closed" pass)
//   Fn3  three states (kill-chain DFA)
//   FnG  both at once (the fused forward pass)
#pragma once
#include "common.cuh"

namespace chronos {

constexpr uint32_t NONE32 = 0xffffffffu;

// ---- two states: out bit s = exit state, a[s] = counter increment, b[s] = last marker ----
struct St2 {
    uint32_t s, a, b;
};
struct Fn2 {
    uint32_t out;
    uint32_t a[2];
    uint32_t b[2];
    typedef St2 State;
    static constexpr int NW = 2;
    __device__ __forceinline__ static Fn2 identity() { return Fn2{0x2u, {0u, 0u}, {NONE32, NONE32}}; }
    // first x, then y
    __device__ __forceinline__ static Fn2 compose(const Fn2 &x, const Fn2 &y) {
        Fn2 r;
        const uint32_t m0 = x.out & 1u, m1 = (x.out >> 1) & 1u;
        r.out = ((y.out >> m0) & 1u) | (((y.out >> m1) & 1u) << 1);
        r.a[0] = x.a[0] + y.a[m0];
        r.a[1] = x.a[1] + y.a[m1];
        r.b[0] = y.b[m0] != NONE32 ? y.b[m0] : x.b[0];
        r.b[1] = y.b[m1] != NONE32 ? y.b[m1] : x.b[1];
        return r;
    }
    __device__ __forceinline__ static Fn2 shfl(const Fn2 &f, int src) {
        Fn2 r;
        r.out = __shfl_sync(0xffffffffu, f.out, src);
        r.a[0] = __shfl_sync(0xffffffffu, f.a[0], src);
        r.a[1] = __shfl_sync(0xffffffffu, f.a[1], src);
        r.b[0] = __shfl_sync(0xffffffffu, f.b[0], src);
        r.b[1] = __shfl_sync(0xffffffffu, f.b[1], src);
        return r;
    }
    __device__ __forceinline__ static St2 apply(const Fn2 &f, const St2 &c) {
        St2 r;
        r.s = (f.out >> c.s) & 1u;
        r.a = c.a + f.a[c.s];
        r.b = f.b[c.s] != NONE32 ? f.b[c.s] : c.b;
        return r;
    }
    // word s: [61] out_s [60:32] a_s (29 bits) [31:0] b_s   (status goes to [63:62])
    __device__ __forceinline__ static void pack(const Fn2 &f, uint64_t *w) {
        w[0] = ((uint64_t)(f.out & 1u) << 61) | ((uint64_t)(f.a[0] & 0x1fffffffu) << 32) | f.b[0];
        w[1] = ((uint64_t)((f.out >> 1) & 1u) << 61) | ((uint64_t)(f.a[1] & 0x1fffffffu) << 32) | f.b[1];
    }
    __device__ __forceinline__ static Fn2 unpack(const uint64_t *w) {
        Fn2 f;
        f.out = (uint32_t)((w[0] >> 61) & 1u) | ((uint32_t)((w[1] >> 61) & 1u) << 1);
        f.a[0] = (uint32_t)(w[0] >> 32) & 0x1fffffffu;
        f.a[1] = (uint32_t)(w[1] >> 32) & 0x1fffffffu;
        f.b[0] = (uint32_t)w[0];
        f.b[1] = (uint32_t)w[1];
        return f;
    }
    __device__ __forceinline__ static void pack_state(const St2 &c, uint64_t *w) {
        w[0] = ((uint64_t)(c.s & 1u) << 61) | ((uint64_t)(c.a & 0x1fffffffu) << 32) | c.b;
        w[1] = w[0];
    }
    __device__ __forceinline__ static St2 unpack_state(const uint64_t *w) {
        return St2{(uint32_t)(w[0] >> 61) & 1u, (uint32_t)(w[0] >> 32) & 0x1fffffffu, (uint32_t)w[0]};
    }
};

// ---- three states: 2 bits of exit state per entry state, one counter ------------------
struct St3 {
    uint32_t s, a;
};
struct Fn3 {
    uint32_t out;
    uint32_t a[3];
    typedef St3 State;
    static constexpr int NW = 3;
    __device__ __forceinline__ static Fn3 identity() { return Fn3{0x24u, {0u, 0u, 0u}}; }  // 0->0 1->1 2->2
    __device__ __forceinline__ static uint32_t o(const Fn3 &f, uint32_t s) { return (f.out >> (2 * s)) & 3u; }
    __device__ __forceinline__ static uint32_t av(const Fn3 &f, uint32_t s) { return s == 0 ? f.a[0] : (s == 1 ? f.a[1] : f.a[2]); }
    __device__ __forceinline__ static Fn3 constant(uint32_t s, uint32_t a) { return Fn3{s | (s << 2) | (s << 4), {a, a, a}}; }
    __device__ __forceinline__ static Fn3 compose(const Fn3 &x, const Fn3 &y) {
        Fn3 r;
        const uint32_t m0 = o(x, 0), m1 = o(x, 1), m2 = o(x, 2);
        r.out = o(y, m0) | (o(y, m1) << 2) | (o(y, m2) << 4);
        r.a[0] = x.a[0] + av(y, m0);
        r.a[1] = x.a[1] + av(y, m1);
        r.a[2] = x.a[2] + av(y, m2);
        return r;
    }
    __device__ __forceinline__ static Fn3 shfl(const Fn3 &f, int src) {
        Fn3 r;
        r.out = __shfl_sync(0xffffffffu, f.out, src);
        r.a[0] = __shfl_sync(0xffffffffu, f.a[0], src);
        r.a[1] = __shfl_sync(0xffffffffu, f.a[1], src);
        r.a[2] = __shfl_sync(0xffffffffu, f.a[2], src);
        return r;
    }
    __device__ __forceinline__ static St3 apply(const Fn3 &f, const St3 &c) { return St3{o(f, c.s), c.a + av(f, c.s)}; }
    // word s: [61:60] out_s [31:0] a_s
    __device__ __forceinline__ static void pack(const Fn3 &f, uint64_t *w) {
        for (int s = 0; s < 3; s++) w[s] = ((uint64_t)o(f, s) << 60) | av(f, s);
    }
    __device__ __forceinline__ static Fn3 unpack(const uint64_t *w) {
        Fn3 f;
        f.out = (uint32_t)((w[0] >> 60) & 3u) | ((uint32_t)((w[1] >> 60) & 3u) << 2) | ((uint32_t)((w[2] >> 60) & 3u) << 4);
        f.a[0] = (uint32_t)w[0];
        f.a[1] = (uint32_t)w[1];
        f.a[2] = (uint32_t)w[2];
        return f;
    }
    __device__ __forceinline__ static void pack_state(const St3 &c, uint64_t *w) {
        w[0] = ((uint64_t)(c.s & 3u) << 60) | c.a;
        w[1] = w[0];
        w[2] = w[0];
    }
    __device__ __forceinline__ static St3 unpack_state(const uint64_t *w) {
        return St3{(uint32_t)(w[0] >> 60) & 3u, (uint32_t)w[0]};
    }
};

// ---- both automata side by side (independent, so the product composes component-wise) ----
struct StG {
    St2 w;
    St3 c;
};
struct FnG {
    Fn2 w;
    Fn3 c;
    typedef StG State;
    static constexpr int NW = 5;
    __device__ __forceinline__ static FnG identity() { return FnG{Fn2::identity(), Fn3::identity()}; }
    __device__ __forceinline__ static FnG compose(const FnG &x, const FnG &y) {
        return FnG{Fn2::compose(x.w, y.w), Fn3::compose(x.c, y.c)};
    }
    __device__ __forceinline__ static FnG shfl(const FnG &f, int src) { return FnG{Fn2::shfl(f.w, src), Fn3::shfl(f.c, src)}; }
    __device__ __forceinline__ static StG apply(const FnG &f, const StG &s) { return StG{Fn2::apply(f.w, s.w), Fn3::apply(f.c, s.c)}; }
    __device__ __forceinline__ static void pack(const FnG &f, uint64_t *w) { Fn2::pack(f.w, w); Fn3::pack(f.c, w + 2); }
    __device__ __forceinline__ static FnG unpack(const uint64_t *w) { return FnG{Fn2::unpack(w), Fn3::unpack(w + 2)}; }
    __device__ __forceinline__ static void pack_state(const StG &s, uint64_t *w) { Fn2::pack_state(s.w, w); Fn3::pack_state(s.c, w + 2); }
    __device__ __forceinline__ static StG unpack_state(const uint64_t *w) { return StG{Fn2::unpack_state(w), Fn3::unpack_state(w + 2)}; }
};

// Block-wide exclusive scan of per-thread functions (THREADS threads). Returns the calling
// thread's exclusive prefix; *total (same in every thread) is the tile's function.
template <class F, int THREADS>
__device__ __forceinline__ F fn_block_exclusive(const F &mine, F *warp_tot /* [THREADS/32] smem */, F *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    F inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        F o = F::shfl(inc, lane >= d ? lane - d : lane);
        if (lane >= d) inc = F::compose(o, inc);
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    F prev = F::shfl(inc, lane > 0 ? lane - 1 : 0);
    F excl = lane > 0 ? prev : F::identity();
    F run = F::identity(), wpre = F::identity();
#pragma unroll
    for (int w = 0; w < THREADS / 32; w++) {
        if (w == warp) wpre = run;
        run = F::compose(run, warp_tot[w]);
    }
    *total = run;
    return F::compose(wpre, excl);
}

// Decoupled look-back over tile summaries; one full warp; F::NW status words per tile, each
// [63:62] = ST_NONE / ST_AGG (function) / ST_PREFIX (concrete state after the tile).
// Returns the concrete state in front of the tile; `init` is the state in front of tile 0.
template <class F>
__device__ __forceinline__ typename F::State fn_lookback(uint64_t *status, uint32_t tile, const F &agg,
                                                         typename F::State init, int lane) {
    typename F::State in = init;
    uint64_t w[F::NW];
    if (tile != 0) {
        if (lane == 0) {
            F::pack(agg, w);
#pragma unroll
            for (int k = 0; k < F::NW; k++) st_relaxed_u64(status + (size_t)F::NW * tile + k, (ST_AGG << 62) | w[k]);
        }
        F H = F::identity();  // everything between the tiles inspected so far and us
        int64_t look = (int64_t)tile - 1;
        while (true) {
            const int64_t idx = look - lane;
            bool is_prefix = true;  // virtual tiles below 0 carry `init`
            if (idx >= 0) {
                // the owner rewrites all its words when it upgrades AGG -> PREFIX: only accept
                // a set whose status fields agree (a torn set would mix two encodings)
                bool ok;
                do {
                    ok = true;
#pragma unroll
                    for (int k = 0; k < F::NW; k++) w[k] = ld_relaxed_u64(status + (size_t)F::NW * idx + k);
                    const uint64_t s0 = w[0] >> 62;
                    ok = s0 != ST_NONE;
#pragma unroll
                    for (int k = 1; k < F::NW; k++) ok = ok && (w[k] >> 62) == s0;
                } while (!ok);
                is_prefix = (w[0] >> 62) == ST_PREFIX;
#pragma unroll
                for (int k = 0; k < F::NW; k++) w[k] &= ~(3ull << 62);
            } else {
                F::pack_state(init, w);
            }
            const uint32_t is_p = __ballot_sync(0xffffffffu, is_prefix);
            const int first = is_p ? __ffs((int)is_p) - 1 : 32;
            // G = tiles (look-first, look] in ascending tile order (= descending lane order):
            // ordered shuffle-tree reduction over the lanes below `first`
            F G = (lane < first) ? F::unpack(w) : F::identity();
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const F older = F::shfl(G, (lane + d) & 31);
                if (lane + d < 32) G = F::compose(older, G);
            }
            G = F::shfl(G, 0);
            H = F::compose(G, H);
            if (is_p) {
                uint64_t pw[F::NW];
#pragma unroll
                for (int k = 0; k < F::NW; k++) pw[k] = __shfl_sync(0xffffffffu, w[k], first);
                in = F::apply(H, F::unpack_state(pw));
                break;
            }
            look -= 32;
        }
    }
    const typename F::State outc = F::apply(agg, in);
    if (lane == 0) {
        F::pack_state(outc, w);
#pragma unroll
        for (int k = 0; k < F::NW; k++) st_relaxed_u64(status + (size_t)F::NW * tile + k, (ST_PREFIX << 62) | w[k]);
    }
    return in;
}

// Plain exclusive count across tiles (one status word per tile): [63:62] status, [31:0] count.
__device__ __forceinline__ uint32_t lookback_count(uint64_t *status, uint32_t tile, uint32_t agg, int lane) {
    if (tile == 0) {
        if (lane == 0) st_relaxed_u64(status, (ST_PREFIX << 62) | agg);
        __syncwarp();
        return 0;
    }
    if (lane == 0) st_relaxed_u64(status + tile, (ST_AGG << 62) | agg);
    uint32_t excl = 0;
    int64_t look = (int64_t)tile - 1;
    while (true) {
        int64_t idx = look - lane;
        uint64_t v = ST_PREFIX << 62;  // virtual tiles below 0: prefix 0
        if (idx >= 0) {
            do {
                v = ld_relaxed_u64(status + idx);
            } while ((v >> 62) == ST_NONE);
        }
        uint32_t is_p = __ballot_sync(0xffffffffu, (v >> 62) == ST_PREFIX);
        if (is_p) {
            int first = __ffs((int)is_p) - 1;
            excl += __reduce_add_sync(0xffffffffu, lane <= first ? (uint32_t)v : 0u);
            break;
        }
        excl += __reduce_add_sync(0xffffffffu, (uint32_t)v);
        look -= 32;
    }
    if (lane == 0) st_relaxed_u64(status + tile, (ST_PREFIX << 62) | (uint64_t)(excl + agg));
    __syncwarp();  // reconverge before the caller's __syncthreads() (see fn_lookback)
    return excl;
}

// Two counts at once: [63:62] status, [61:31] count a, [30:0] count b (31 bits each).
__device__ __forceinline__ uint2 lookback_count2(uint64_t *status, uint32_t tile, uint32_t a, uint32_t b, int lane) {
    const uint64_t agg = ((uint64_t)a << 31) | b;
    if (tile == 0) {
        if (lane == 0) st_relaxed_u64(status, (ST_PREFIX << 62) | agg);
        __syncwarp();
        return make_uint2(0u, 0u);
    }
    if (lane == 0) st_relaxed_u64(status + tile, (ST_AGG << 62) | agg);
    uint64_t excl = 0;
    int64_t look = (int64_t)tile - 1;
    while (true) {
        int64_t idx = look - lane;
        uint64_t v = ST_PREFIX << 62;
        if (idx >= 0) {
            do {
                v = ld_relaxed_u64(status + idx);
            } while ((v >> 62) == ST_NONE);
        }
        const uint32_t is_p = __ballot_sync(0xffffffffu, (v >> 62) == ST_PREFIX);
        const int first = is_p ? __ffs((int)is_p) - 1 : 31;
        const bool use = lane <= first;
        const uint32_t sa = __reduce_add_sync(0xffffffffu, use ? (uint32_t)((v >> 31) & 0x7fffffffu) : 0u);
        const uint32_t sb = __reduce_add_sync(0xffffffffu, use ? (uint32_t)(v & 0x7fffffffu) : 0u);
        excl += ((uint64_t)sa << 31) + sb;
        if (is_p) break;
        look -= 32;
    }
    if (lane == 0) st_relaxed_u64(status + tile, (ST_PREFIX << 62) | ((excl + agg) & ~(3ull << 62)));
    __syncwarp();
    return make_uint2((uint32_t)((excl >> 31) & 0x7fffffffu), (uint32_t)(excl & 0x7fffffffu));
}

}  // namespace chronos
