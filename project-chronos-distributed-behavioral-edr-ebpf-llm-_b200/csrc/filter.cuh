// filter.cuh -- kernel 1 of the hot path: noise filter + validity gate + ignore-list +
// trigger bits over raw `struct data_t` records, with order-preserving compaction of the
// survivors.
//
// Reference semantics (all in /root/reference/chronos_sensor.py):
//   openat hook filter  :74-92  (starts_with :41-47, ends_with :28-38)   -> prefix / suffix drop
//   ctypes field read   :125    (bytes before first NUL)                 -> field lengths
//   strict utf-8 decode :128-131                                          -> validity gate
//   comm ignore-list    :134-135
//   trigger needles     :137, :141 (a match always lies inside one field)
//
// Execution model: persistent CTAs pull tiles of TILE records from an atomic ticket
// counter; each tile is fetched with ONE 1-D TMA bulk copy (cp.async.bulk, UBLKCP) into
// a ring of NSTAGE shared-memory stages, completion tracked by mbarrier transaction
// bytes, so the HBM stream is fully coalesced and asynchronous.  Per tile:
//   A  thread-per-record: pid, "type is OPEN", the 8 anchored prefixes (first 12 argv
//      bytes) -> most noise dies here touching 3 of the 18 16-byte chunks in smem;
//      the still-alive records are compacted into a work list,
//   B  thread-per-alive-record: NUL scan (16-byte LDS), suffix table, SWAR needle
//      search, slow-path UTF-8 DFA only when a high bit was seen,
//   C  ordered compaction inside the tile (ballot ranks); compact_survivors then applies
//      the exclusive scan of the per-tile counts, so CTAs never wait on each other.
#pragma once
#include "common.cuh"

namespace chronos {

constexpr uint32_t META_TRIG = 1, META_CURL = 2, META_CHMOD = 4, META_EXEC = 8, META_CHAIN = 16;

struct FilterParams {
    const uint8_t *recs;     // batch base (device, 16-byte aligned)
    uint64_t n;              // records in the batch
    uint32_t tile_begin;     // this launch covers tiles [tile_begin, tile_end)
    uint32_t tile_end;
    uint32_t *ticket;        // per-launch tile ticket counter (zeroed)
    uint32_t *tile_count;    // survivors per tile (one word per tile of the batch)
    uint32_t *sp_key;        // tile-local compaction: slot tile*TILE + rank holds the pid,
    uint32_t *sp_src;        //   the batch-local record index
    uint8_t *sp_meta;        //   and the META_* bits of the tile's rank-th survivor
    uint32_t *out_count;     // total survivors (atomicAdd, one per CTA)
    uint32_t *key_or;        // OR / AND over survivor pids (radix pass planning)
    uint32_t *key_and;
};

// ---- needle tests on one 4-byte window `t` (bytes k..k+3) and the byte after it ------
__device__ __forceinline__ uint32_t trigger_bits(uint32_t t, uint32_t nb) {
    uint32_t m = 0;
    m |= (t == lit4('c', 'u', 'r', 'l')) ? (META_TRIG | META_CURL) : 0u;
    m |= (t == lit4('c', 'h', 'm', 'o') && (nb & 0xFFu) == (uint32_t)'d') ? (META_TRIG | META_CHMOD) : 0u;
    m |= (t == lit4('b', 'a', 's', 'h')) ? META_TRIG : 0u;
    m |= ((t & 0xFFFFu) == lit2('n', 'c')) ? META_TRIG : 0u;
    m |= ((t & 0xFFFFFFu) == lit3('c', 'a', 't')) ? META_TRIG : 0u;
    return m;
}
__device__ __forceinline__ bool ignore_hit(uint32_t t, uint32_t t4) {
    bool h = (t == lit4('n', 'o', 'd', 'e')) | (t == lit4('c', 'o', 'd', 'e')) |
             ((t & 0xFFFFFFu) == lit3('g', 'i', 't'));
    h |= (t == lit4('o', 'l', 'l', 'a')) & ((t4 & 0xFFFFu) == lit2('m', 'a'));
    h |= (t == lit4('p', 'y', 't', 'h')) & ((t4 & 0xFFFFu) == lit2('o', 'n'));
    h |= (t == lit4('c', 'h', 'r', 'o')) & ((t4 & 0xFFFFu) == lit2('m', 'e'));
    h |= (t == lit4('v', 'm', 't', 'o')) & ((t4 & 0xFFFFFFu) == lit3('o', 'l', 's'));
    return h;
}

// NUL-truncate a field held in up to 4 words (already limited to the field's bytes);
// returns its length and zeroes everything from the NUL on.
__device__ __forceinline__ int truncate_field(uint32_t &a0, uint32_t &a1, uint32_t &a2, uint32_t &a3, int cap) {
    uint32_t z;
    if ((z = zero_bytes(a0)) != 0) { int k = first_zero_index(z); a0 = keep_low_bytes(a0, k); a1 = a2 = a3 = 0; return min(k, cap); }
    if ((z = zero_bytes(a1)) != 0) { int k = first_zero_index(z); a1 = keep_low_bytes(a1, k); a2 = a3 = 0; return min(4 + k, cap); }
    if ((z = zero_bytes(a2)) != 0) { int k = first_zero_index(z); a2 = keep_low_bytes(a2, k); a3 = 0; return min(8 + k, cap); }
    if ((z = zero_bytes(a3)) != 0) { int k = first_zero_index(z); a3 = keep_low_bytes(a3, k); return min(12 + k, cap); }
    return cap;
}

// Slide a byte window over a truncated <=16-byte field: trigger bits, optional ignore-list.
__device__ __forceinline__ void scan_small_field(uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, int len,
                                                 bool with_ignore, bool &ign, uint32_t &meta) {
#pragma unroll 1
    for (int k = 0; k < len; k++) {
        meta |= trigger_bits(a0, a1);
        if (with_ignore) ign |= ignore_hit(a0, a1);
        a0 = __funnelshift_r(a0, a1, 8);
        a1 = __funnelshift_r(a1, a2, 8);
        a2 = __funnelshift_r(a2, a3, 8);
        a3 >>= 8;
    }
}

// CPython's strict UTF-8 acceptance (Unicode table 3-7) on n bytes in shared memory.
__device__ __noinline__ bool utf8_valid_smem(const uint8_t *s, int n) {
    int i = 0;
    while (i < n) {
        uint32_t b = s[i];
        if (b < 0x80u) { i++; continue; }
        int need;
        uint32_t lo = 0x80u, hi = 0xBFu;
        if (b >= 0xC2u && b <= 0xDFu) need = 1;
        else if (b == 0xE0u) { need = 2; lo = 0xA0u; }
        else if (b >= 0xE1u && b <= 0xECu) need = 2;
        else if (b == 0xEDu) { need = 2; hi = 0x9Fu; }
        else if (b == 0xEEu || b == 0xEFu) need = 2;
        else if (b == 0xF0u) { need = 3; lo = 0x90u; }
        else if (b >= 0xF1u && b <= 0xF3u) need = 3;
        else if (b == 0xF4u) { need = 3; hi = 0x8Fu; }
        else return false;
        if (i + need >= n) return false;
        uint32_t c = s[i + 1];
        if (c < lo || c > hi) return false;
        for (int k = 2; k <= need; k++)
            if ((s[i + k] & 0xC0u) != 0x80u) return false;
        i += need + 1;
    }
    return true;
}

__host__ __device__ constexpr uint64_t lit64(const char *s, int n) {
    uint64_t v = 0;
    for (int i = 0; i < n; i++) v |= (uint64_t)(uint8_t)s[i] << (8 * i);
    return v;
}

// Anchored prefixes of the openat hook (:76-78, :90) on the first 12 argv bytes.
__device__ __forceinline__ bool prefix_noise(uint32_t a0, uint32_t a1, uint32_t a2) {
    bool usr = a0 == lit4('/', 'u', 's', 'r');
    bool etc = a0 == lit4('/', 'e', 't', 'c');
    bool d = a0 == lit4('/', 'l', 'i', 'b');
    d |= usr & (a1 == lit4('/', 'l', 'i', 'b'));
    d |= usr & (a1 == lit4('/', 's', 'h', 'a')) & ((a2 & 0xFFFFu) == lit2('r', 'e'));
    d |= etc & (a1 == lit4('/', 's', 's', 'l'));
    d |= etc & (a1 == lit4('/', 'f', 'o', 'n')) & ((a2 & 0xFFFFu) == lit2('t', 's'));
    d |= etc & (a1 == lit4('/', 'h', 'o', 's')) & ((a2 & 0xFFu) == (uint32_t)'t');
    d |= (a0 == lit4('/', 'd', 'e', 'v')) & ((a1 & 0xFFu) == (uint32_t)'/');
    d |= (a0 == lit4('/', 'p', 'r', 'o')) & ((a1 & 0xFFFFu) == lit2('c', '/'));
    return d;
}

// Full analysis of one record that passed the prefix test. `w` = 72 words in shared
// memory (16-byte aligned). Returns 0 when the event is dropped, else 0x80 | META bits.
__device__ __forceinline__ uint32_t analyze_record(const uint32_t *w) {
    uint32_t meta = 0;
    // ---- comm: words 1..4 -----------------------------------------------------------
    uint32_t c0 = w[1], c1 = w[2], c2 = w[3], c3 = w[4];
    int clen = truncate_field(c0, c1, c2, c3, 16);
    bool ign = false;
    scan_small_field(c0, c1, c2, c3, clen, true, ign, meta);
    if (ign) return 0;  // :134-135
    bool hi_comm = ((c0 | c1 | c2 | c3) & 0x80808080u) != 0;

    // ---- type: words 69, 70 and half of 71 ------------------------------------------
    uint32_t t0 = w[69], t1 = w[70], t2 = w[71] & 0xFFFFu, t3 = 0;
    bool nul4 = (t1 & 0xFFu) == 0;
    bool is_open = (t0 == lit4('O', 'P', 'E', 'N')) & nul4;
    bool is_exec = (t0 == lit4('E', 'X', 'E', 'C')) & nul4;
    bool hi_type = false;
    int tlen = 4;
    if (!(is_open | is_exec)) {  // replayed records may carry any <=10 bytes here
        tlen = truncate_field(t0, t1, t2, t3, 10);
        bool dummy = false;
        scan_small_field(t0, t1, t2, t3, tlen, false, dummy, meta);
        hi_type = ((t0 | t1 | t2) & 0x80808080u) != 0;
    }
    if (is_exec) meta |= META_EXEC;

    // ---- argv pass 1: length (first NUL, cap 256) + "any byte >= 0x80 before it" -----
    const uint4 *v4 = reinterpret_cast<const uint4 *>(w);
    int alen = 256;
    uint32_t hi = 0;
#pragma unroll 1
    for (int q = 1; q <= 17; q++) {
        uint4 v = v4[q];
        if (q == 1) v.x = 0x01010101u;               // word 4 is comm's tail, not argv
        if (q == 17) { v.y = 0; v.z = 0; v.w = 0; }  // words 69.. are `type`: act as the cap
        uint32_t zx = zero_bytes(v.x), zy = zero_bytes(v.y), zz = zero_bytes(v.z), zw = zero_bytes(v.w);
        if ((zx | zy | zz | zw) != 0) {
            int base = (4 * q - 5) * 4;  // argv byte position of v.x
            if (zx) { int k = first_zero_index(zx); alen = base + k; hi |= keep_low_bytes(v.x, k); }
            else if (zy) { int k = first_zero_index(zy); alen = base + 4 + k; hi |= v.x | keep_low_bytes(v.y, k); }
            else if (zz) { int k = first_zero_index(zz); alen = base + 8 + k; hi |= v.x | v.y | keep_low_bytes(v.z, k); }
            else { int k = first_zero_index(zw); alen = base + 12 + k; hi |= v.x | v.y | v.z | keep_low_bytes(v.w, k); }
            break;
        }
        hi |= v.x | v.y | v.z | v.w;
    }
    bool hi_argv = (hi & 0x80808080u) != 0;

    // ---- suffix table of the openat hook (:83-85) on the last <=8 bytes --------------
    if (is_open) {
        uint64_t T = 0;
        if (alen > 0) {
            int o = 12 + alen;  // record byte offset of argv[alen-8]
            int i = o >> 2, s = (o & 3) * 8;
            uint32_t W0 = w[i], W1 = w[i + 1], W2 = w[i + 2];
            uint32_t lo = __funnelshift_r(W0, W1, s), hi32 = __funnelshift_r(W1, W2, s);
            T = ((uint64_t)hi32 << 32) | lo;
            if (alen < 8) T &= ~0ULL << (8 * (8 - alen));
        }
        bool d = (T >> 40) == lit64(".so", 3);
        d |= (T >> 16) == lit64(".cache", 6);
        d |= (T >> 40) == lit64(".mo", 3);
        d |= (T >> 24) == lit64(".conf", 5);
        d |= (T >> 32) == lit64(".crt", 4);
        d |= (T >> 8) == lit64(".curlrc", 7);
        if (d) return 0;
    }

    // ---- strict UTF-8 gate (:128-131); ASCII-only fields are valid by construction ----
    const uint8_t *b = reinterpret_cast<const uint8_t *>(w);
    if (hi_comm && !utf8_valid_smem(b + 4, clen)) return 0;
    if (hi_type && !utf8_valid_smem(b + 276, tlen)) return 0;
    if (hi_argv && !utf8_valid_smem(b + 20, alen)) return 0;

    // ---- argv pass 2 (survivors only): trigger needles (:141) ------------------------
    {
        const uint32_t *a = w + 5;
        auto ldm = [&](int j) -> uint32_t {
            int rem = alen - 4 * j;
            if (rem <= 0) return 0u;
            uint32_t x = a[j];
            return rem >= 4 ? x : keep_low_bytes(x, rem);
        };
        uint32_t cur = ldm(0);
#pragma unroll 1
        for (int j = 0; 4 * j < alen; j++) {
            uint32_t nxt = ldm(j + 1);
            meta |= trigger_bits(cur, nxt);
            meta |= trigger_bits(__funnelshift_r(cur, nxt, 8), nxt >> 8);
            meta |= trigger_bits(__funnelshift_r(cur, nxt, 16), nxt >> 16);
            meta |= trigger_bits(__funnelshift_r(cur, nxt, 24), nxt >> 24);
            cur = nxt;
        }
    }
    return 0x80u | meta;
}

// Single-pass exclusive prefix over per-tile counts (decoupled look-back). Called by one
// full warp; tile 0 of the batch seeds the chain. Returns the exclusive prefix.
__device__ __forceinline__ uint32_t lookback_count(uint64_t *status, uint32_t tile, uint32_t agg, int lane) {
    if (tile == 0) {
        if (lane == 0) st_relaxed_u64(status, (ST_PREFIX << 62) | agg);
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
    return excl;
}

template <int TILE, int NSTAGE>
struct FilterSmemTail {
    uint64_t bar[NSTAGE];
    uint32_t tile[NSTAGE];
    uint32_t n_alive;
    uint32_t base;
    uint32_t k_or, k_and;
    uint32_t wcnt[TILE / 32];
    uint16_t alive[TILE];
    uint8_t res[TILE];
};

template <int TILE, int NSTAGE>
constexpr size_t filter_smem_bytes() {
    return (size_t)NSTAGE * TILE * CH_REC_BYTES + sizeof(FilterSmemTail<TILE, NSTAGE>) + 128;
}

template <int TILE, int NSTAGE>
__global__ void __launch_bounds__(TILE) filter_kernel(const FilterParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    uint8_t *stages = smem;
    auto &S = *reinterpret_cast<FilterSmemTail<TILE, NSTAGE> *>(smem + (size_t)NSTAGE * TILE * CH_REC_BYTES);
    constexpr int NW = TILE / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    auto claim_and_issue = [&](int s) {  // thread 0 only
        uint32_t t = p.tile_begin + atomicAdd(p.ticket, 1u);
        S.tile[s] = t;
        if (t < p.tile_end) {
            uint64_t first = (uint64_t)t * TILE;
            uint32_t cnt = (uint32_t)min((uint64_t)TILE, p.n - first);
            uint32_t bytes = cnt * CH_REC_BYTES;
            mbar_arrive_expect_tx(&S.bar[s], bytes);
            bulk_g2s(stages + (size_t)s * TILE * CH_REC_BYTES, p.recs + first * CH_REC_BYTES, bytes, &S.bar[s]);
        }
    };

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; s++) mbar_init(&S.bar[s], 1);
        S.n_alive = 0;
        S.base = 0;
        S.k_or = 0;
        S.k_and = 0xffffffffu;
        fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0)
        for (int s = 0; s < NSTAGE; s++) claim_and_issue(s);
    __syncthreads();

    for (uint32_t it = 0;; it++) {
        const int s = it % NSTAGE;
        const uint32_t tile = S.tile[s];
        if (tile >= p.tile_end) break;
        const uint64_t first = (uint64_t)tile * TILE;
        const uint32_t n_in = (uint32_t)min((uint64_t)TILE, p.n - first);
        mbar_wait(&S.bar[s], (it / NSTAGE) & 1);
        const uint8_t *tile_base = stages + (size_t)s * TILE * CH_REC_BYTES;

        // ---- phase A: one record per thread ------------------------------------------
        uint32_t pid = 0;
        bool alive = false;
        if (tid < n_in) {
            const uint4 *r4 = reinterpret_cast<const uint4 *>(tile_base + (size_t)tid * CH_REC_BYTES);
            uint4 q0 = r4[0];    // pid, comm[0:12]
            uint4 q1 = r4[1];    // comm[12:16], argv[0:12]
            uint4 q17 = r4[17];  // argv[252:256], type[0:10], pad
            pid = q0.x;
            bool is_open = (q17.y == lit4('O', 'P', 'E', 'N')) & ((q17.z & 0xFFu) == 0);
            alive = !(is_open && prefix_noise(q1.y, q1.z, q1.w));
        }
        S.res[tid] = 0;
        {
            uint32_t bal = __ballot_sync(0xffffffffu, alive);
            uint32_t pos = 0;
            if (lane == 0 && bal) pos = atomicAdd(&S.n_alive, (uint32_t)__popc(bal));
            pos = __shfl_sync(0xffffffffu, pos, 0);
            if (alive) S.alive[pos + __popc(bal & lanemask_lt())] = (uint16_t)tid;
        }
        __syncthreads();

        // ---- phase B: one alive record per thread, lanes packed ------------------------
        const uint32_t n_alive = S.n_alive;
        if ((uint32_t)tid < n_alive) {
            uint32_t r = S.alive[tid];
            const uint32_t *w = reinterpret_cast<const uint32_t *>(tile_base + (size_t)r * CH_REC_BYTES);
            S.res[r] = (uint8_t)analyze_record(w);
        }
        __syncthreads();
        // the stage is dead now: refill it while we compact
        if (tid == 0) {
            S.n_alive = 0;
            claim_and_issue(s);
        }

        // ---- phase C: tile-local, arrival-ordered compaction -----------------------------
        // (the cross-tile prefix is taken by compact_survivors afterwards: a blocking
        //  look-back here would chain every CTA's stage ring to its neighbour's)
        const uint32_t code = S.res[tid];
        const bool keep = (code & 0x80u) != 0;
        const uint32_t kb = __ballot_sync(0xffffffffu, keep);
        const uint32_t rank = __popc(kb & lanemask_lt());
        {
            uint32_t o = __reduce_or_sync(0xffffffffu, keep ? pid : 0u);
            uint32_t a = __reduce_and_sync(0xffffffffu, keep ? pid : 0xffffffffu);
            if (lane == 0) {
                S.wcnt[warp] = __popc(kb);
                if (kb) {
                    atomicOr(&S.k_or, o);
                    atomicAnd(&S.k_and, a);
                }
            }
        }
        __syncthreads();
        {
            uint32_t before = rank, tot = 0;
#pragma unroll
            for (int k = 0; k < NW; k++) {
                const uint32_t c = S.wcnt[k];
                before += (k < warp) ? c : 0u;
                tot += c;
            }
            if (keep) {
                const size_t slot = (size_t)tile * TILE + before;
                p.sp_key[slot] = pid;
                p.sp_src[slot] = (uint32_t)first + tid;
                p.sp_meta[slot] = (uint8_t)(code & 0x7Fu);
            }
            if (tid == 0) {
                p.tile_count[tile] = tot;
                S.base += tot;
            }
        }
    }
    if (tid == 0) {
        if (S.base) atomicAdd(p.out_count, S.base);
        if (S.k_or != 0 || S.k_and != 0xffffffffu) {
            atomicOr(p.key_or, S.k_or);
            atomicAnd(p.key_and, S.k_and);
        }
    }
}

// Cross-tile step of the ordered compaction: tile_off[] = exclusive scan of tile_count[]
// (radix_scan); one warp per filter tile moves its run to the dense arrays.
struct CompactParams {
    const uint32_t *tile_off;
    const uint32_t *sp_key;
    const uint32_t *sp_src;
    const uint8_t *sp_meta;
    uint32_t *out_key;
    uint32_t *out_src;
    uint8_t *out_meta;
    uint32_t n_tiles;
    uint32_t tile;       // records per filter tile
    const uint32_t *total;  // device scalar: survivor count == tile_off[n_tiles]
};

__global__ void __launch_bounds__(256) compact_survivors(const CompactParams c) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t n_warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t t = warp; t < c.n_tiles; t += n_warps) {
        const uint32_t lo = c.tile_off[t];
        const uint32_t hi = (t + 1 < c.n_tiles) ? c.tile_off[t + 1] : *c.total;
        const size_t base = (size_t)t * c.tile;
        for (uint32_t j = lane; j < hi - lo; j += 32) {
            c.out_key[lo + j] = c.sp_key[base + j];
            c.out_src[lo + j] = c.sp_src[base + j];
            c.out_meta[lo + j] = c.sp_meta[base + j];
        }
    }
}

}  // namespace chronos
