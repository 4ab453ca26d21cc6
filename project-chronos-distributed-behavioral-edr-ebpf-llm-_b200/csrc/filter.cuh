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
//   B1 thread-per-alive-record (lanes packed): comm through a per-warp memo table
//      (ignore-list, trigger bits, UTF-8 of the 16 comm bytes are a pure function of
//      them and process names repeat), type, argv NUL search over the first 60 bytes,
//      suffix table, UTF-8 DFA only when a high bit was seen,
//   B1b argv longer than 60 bytes: 16 lanes per record, one 16-byte chunk per lane,
//   B2 survivors only: SWAR trigger-needle search as (record, 16-byte chunk) work items
//      spread over all threads, so long paths do not serialise a warp,
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

// ---- per-field classifiers --------------------------------------------------------------
constexpr uint32_t CR_IGN = 1u << 8, CR_INVALID = 1u << 9, CR_VALID = 1u << 31;
constexpr uint32_t RW_KEEP = 0x80u, RW_OPEN = 1u << 16, RW_HI = 1u << 17;

// comm (16 raw bytes at record words 1..4): trigger bits | CR_IGN | CR_INVALID.
// A pure function of the 16 bytes, which is what makes the per-warp memo table exact.
__device__ __noinline__ uint32_t comm_classify(const uint32_t *w) {
    uint32_t c0 = w[1], c1 = w[2], c2 = w[3], c3 = w[4];
    int clen = truncate_field(c0, c1, c2, c3, 16);
    uint32_t meta = 0;
    bool ign = false;
    scan_small_field(c0, c1, c2, c3, clen, true, ign, meta);
    uint32_t r = meta | (ign ? CR_IGN : 0u);  // :134-135
    if (((c0 | c1 | c2 | c3) & 0x80808080u) != 0 &&
        !utf8_valid_smem(reinterpret_cast<const uint8_t *>(w) + 4, clen))
        r |= CR_INVALID;  // :128, :131
    return r;
}

// type field that is neither "OPEN" nor "EXEC" (replays only): trigger bits | CR_INVALID
__device__ __noinline__ uint32_t type_classify(const uint32_t *w) {
    uint32_t t0 = w[69], t1 = w[70], t2 = w[71] & 0xFFFFu, t3 = 0;
    int tlen = truncate_field(t0, t1, t2, t3, 10);
    uint32_t meta = 0;
    bool dummy = false;
    scan_small_field(t0, t1, t2, t3, tlen, false, dummy, meta);
    if (((t0 | t1 | t2) & 0x80808080u) != 0 &&
        !utf8_valid_smem(reinterpret_cast<const uint8_t *>(w) + 276, tlen))
        meta |= CR_INVALID;  // :130-131
    return meta;
}

// NUL search in one 16-byte chunk q (1..17) of the record, argv-relative.
// Returns true when the chunk holds argv's end; pos = argv length, hi = OR of the bytes
// before the end (only bit 7 of each byte matters to the caller).
__device__ __forceinline__ bool argv_chunk_end(const uint4 *v4, int q, int &pos, uint32_t &hi) {
    uint4 v = v4[q];
    if (q == 1) v.x = 0x01010101u;               // word 4 is comm's tail, not argv
    if (q == 17) { v.y = 0; v.z = 0; v.w = 0; }  // words 69.. are `type`: they act as the 256 cap
    const uint32_t zx = zero_bytes(v.x), zy = zero_bytes(v.y), zz = zero_bytes(v.z), zw = zero_bytes(v.w);
    if ((zx | zy | zz | zw) == 0) {
        hi |= v.x | v.y | v.z | v.w;
        return false;
    }
    const int base = (4 * q - 5) * 4;  // argv byte position of v.x
    if (zx) { int k = first_zero_index(zx); pos = base + k; hi |= keep_low_bytes(v.x, k); }
    else if (zy) { int k = first_zero_index(zy); pos = base + 4 + k; hi |= v.x | keep_low_bytes(v.y, k); }
    else if (zz) { int k = first_zero_index(zz); pos = base + 8 + k; hi |= v.x | v.y | keep_low_bytes(v.z, k); }
    else { int k = first_zero_index(zw); pos = base + 12 + k; hi |= v.x | v.y | v.z | keep_low_bytes(v.w, k); }
    return true;
}

// Suffix table of the openat hook (:83-85) on the last <= 8 bytes of argv.
__device__ __forceinline__ bool suffix_noise(const uint32_t *w, int alen) {
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
    return d;
}

// Trigger needles (:141) starting in argv bytes [16j, 16j+16) of a survivor.
__device__ __forceinline__ uint32_t argv_trigger_chunk(const uint32_t *w, int alen, int j) {
    const uint32_t *a = w + 5;
    auto ldm = [&](int k) -> uint32_t {
        int rem = alen - 4 * k;
        if (rem <= 0) return 0u;
        uint32_t x = a[k];
        return rem >= 4 ? x : keep_low_bytes(x, rem);
    };
    uint32_t meta = 0;
    uint32_t cur = ldm(4 * j);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        uint32_t nxt = ldm(4 * j + k + 1);
        meta |= trigger_bits(cur, nxt);
        meta |= trigger_bits(__funnelshift_r(cur, nxt, 8), nxt >> 8);
        meta |= trigger_bits(__funnelshift_r(cur, nxt, 16), nxt >> 16);
        meta |= trigger_bits(__funnelshift_r(cur, nxt, 24), nxt >> 24);
        cur = nxt;
    }
    return meta;
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

constexpr int MEMO_SLOTS = 64;   // per-warp comm memo entries
constexpr int SHORT_CHUNKS = 4;  // argv NUL search done thread-per-record over chunks 1..4 (60 bytes)

template <int TILE, int NSTAGE>
struct FilterSmemTail {
    uint4 memo_key[TILE / 32][MEMO_SLOTS];   // raw 16 comm bytes
    uint32_t memo_val[TILE / 32][MEMO_SLOTS];
    uint64_t bar[NSTAGE];
    uint32_t tile[NSTAGE];
    uint32_t n_alive, n_long, n_items;
    uint32_t base;
    uint32_t k_or, k_and;
    uint32_t wcnt[TILE / 32];
    uint32_t resw[TILE];         // per record: RW_KEEP | META bits (+ RW_OPEN/RW_HI while pending)
    uint16_t alive[TILE];        // records that passed the prefix test
    uint16_t longl[TILE];        // alive records whose argv has no NUL in its first 60 bytes
    uint16_t alen[TILE];         // argv length of survivors
    uint16_t items[TILE * 16];   // (record << 4 | 16-byte argv chunk) trigger-scan work items
};

template <int TILE, int NSTAGE>
constexpr size_t filter_smem_bytes() {
    return (size_t)NSTAGE * TILE * CH_REC_BYTES + sizeof(FilterSmemTail<TILE, NSTAGE>) + 128;
}

// Last stage of the per-record decision once argv's length is known: suffix drop,
// argv UTF-8 gate, then the survivor's trigger-scan work items.
template <int TILE, int NSTAGE>
__device__ __forceinline__ void finalize_record(FilterSmemTail<TILE, NSTAGE> &S, const uint32_t *w, uint32_t r,
                                                int alen, bool hi_argv, uint32_t meta, bool is_open) {
    if (is_open && suffix_noise(w, alen)) { S.resw[r] = 0; return; }
    if (hi_argv && !utf8_valid_smem(reinterpret_cast<const uint8_t *>(w) + 20, alen)) { S.resw[r] = 0; return; }
    S.resw[r] = RW_KEEP | (meta & 0x7Fu);
    S.alen[r] = (uint16_t)alen;
    const uint32_t nch = (uint32_t)(alen + 15) >> 4;
    if (nch) {
        uint32_t at = atomicAdd(&S.n_items, nch);
        for (uint32_t j = 0; j < nch; j++) S.items[at + j] = (uint16_t)((r << 4) | j);
    }
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
        S.n_long = 0;
        S.n_items = 0;
        S.base = 0;
        S.k_or = 0;
        S.k_and = 0xffffffffu;
        fence_mbar_init();
    }
    for (int i = lane; i < MEMO_SLOTS; i += 32) S.memo_val[warp][i] = 0;
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

        // ---- phase A: one record per thread: pid + "OPEN with a noise prefix" -----------
        uint32_t pid = 0;
        bool alive = false;
        if (tid < n_in) {
            const uint4 *r4 = reinterpret_cast<const uint4 *>(tile_base + (size_t)tid * CH_REC_BYTES);
            uint4 q1 = r4[1];    // comm[12:16], argv[0:12]
            uint4 q17 = r4[17];  // argv[252:256], type[0:10], pad
            pid = reinterpret_cast<const uint32_t *>(r4)[0];
            bool is_open = (q17.y == lit4('O', 'P', 'E', 'N')) & ((q17.z & 0xFFu) == 0);
            alive = !(is_open && prefix_noise(q1.y, q1.z, q1.w));
        }
        S.resw[tid] = 0;
        {
            uint32_t bal = __ballot_sync(0xffffffffu, alive);
            uint32_t pos = 0;
            if (lane == 0 && bal) pos = atomicAdd(&S.n_alive, (uint32_t)__popc(bal));
            pos = __shfl_sync(0xffffffffu, pos, 0);
            if (alive) S.alive[pos + __popc(bal & lanemask_lt())] = (uint16_t)tid;
        }
        __syncthreads();

        // ---- phase B1: one alive record per thread, lanes packed --------------------------
        const uint32_t n_alive = S.n_alive;
        if ((uint32_t)(warp * 32) < n_alive) {  // warp-uniform
            const bool act = (uint32_t)tid < n_alive;
            const uint32_t r = act ? S.alive[tid] : 0u;
            const uint32_t *w = reinterpret_cast<const uint32_t *>(tile_base + (size_t)r * CH_REC_BYTES);
            // comm through the per-warp memo table (process names repeat massively)
            const uint32_t c0 = w[1], c1 = w[2], c2 = w[3], c3 = w[4];
            const uint32_t h =
                ((c0 * 0x9E3779B1u) ^ (c1 * 0x85EBCA77u) ^ (c2 * 0xC2B2AE3Du) ^ (c3 * 0x27D4EB2Fu)) >> 26;
            const uint4 mk = S.memo_key[warp][h];
            uint32_t cres = S.memo_val[warp][h];
            const bool hit = (cres & CR_VALID) && mk.x == c0 && mk.y == c1 && mk.z == c2 && mk.w == c3;
            const bool miss = act && !hit;
            if (miss) cres = comm_classify(w) | CR_VALID;
            {
                const uint32_t peers = __match_any_sync(0xffffffffu, miss ? h : 0x100u + lane);
                if (miss && (peers & lanemask_lt()) == 0) {
                    S.memo_key[warp][h] = make_uint4(c0, c1, c2, c3);
                    S.memo_val[warp][h] = cres;
                }
                __syncwarp();
            }
            bool live = act && (cres & (CR_IGN | CR_INVALID)) == 0;
            uint32_t meta = cres & 7u;
            bool is_open = false;
            if (live) {
                const uint32_t t0 = w[69], t1 = w[70];
                const bool nul4 = (t1 & 0xFFu) == 0;
                is_open = (t0 == lit4('O', 'P', 'E', 'N')) & nul4;
                const bool is_exec = (t0 == lit4('E', 'X', 'E', 'C')) & nul4;
                if (is_exec) meta |= META_EXEC;
                if (!(is_open | is_exec)) {
                    const uint32_t tr = type_classify(w);
                    if (tr & CR_INVALID) live = false;
                    meta |= tr & 7u;
                }
            }
            if (live) {
                const uint4 *v4 = reinterpret_cast<const uint4 *>(w);
                int alen = -1;
                uint32_t hi = 0;
#pragma unroll
                for (int q = 1; q <= SHORT_CHUNKS; q++) {
                    if (alen < 0) {
                        int pos = 0;
                        if (argv_chunk_end(v4, q, pos, hi)) alen = pos;
                    }
                }
                const bool hi_argv = (hi & 0x80808080u) != 0;
                if (alen >= 0) {
                    finalize_record<TILE, NSTAGE>(S, w, r, alen, hi_argv, meta, is_open);
                } else {  // long path: finish cooperatively in phase B1b
                    S.resw[r] = meta | (is_open ? RW_OPEN : 0u) | (hi_argv ? RW_HI : 0u);
                    S.longl[atomicAdd(&S.n_long, 1u)] = (uint16_t)r;
                }
            }
        }
        __syncthreads();

        // ---- phase B1b: long argv, 16 lanes per record, one 16-byte chunk per lane -------
        const uint32_t n_long = S.n_long;
        if (n_long) {
            for (uint32_t l0 = (uint32_t)warp * 2; l0 < n_long; l0 += NW * 2) {
                const uint32_t li = l0 + (lane >> 4);
                const bool valid = li < n_long;
                const uint32_t r = valid ? S.longl[li] : 0u;
                const uint32_t *w = reinterpret_cast<const uint32_t *>(tile_base + (size_t)r * CH_REC_BYTES);
                const int sub = lane & 15, q = SHORT_CHUNKS + 1 + sub;
                bool has = false;
                int pos = 0;
                uint32_t hi = 0;
                if (valid && q <= 17) has = argv_chunk_end(reinterpret_cast<const uint4 *>(w), q, pos, hi);
                const uint32_t hb = (__ballot_sync(0xffffffffu, has) >> (lane & 16)) & 0xFFFFu;
                const int fz = hb ? __ffs((int)hb) - 1 : 0;  // chunk 17 always ends argv, so hb != 0 when valid
                const int alen = __shfl_sync(0xffffffffu, pos, (lane & 16) + fz);
                const uint32_t contrib = (valid && sub <= fz) ? hi : 0u;
                const uint32_t o_lo = __reduce_or_sync(0xffffffffu, lane < 16 ? contrib : 0u);
                const uint32_t o_hi = __reduce_or_sync(0xffffffffu, lane >= 16 ? contrib : 0u);
                if (valid && sub == 0) {
                    const uint32_t st = S.resw[r];
                    const bool hi_argv = (((lane < 16 ? o_lo : o_hi) & 0x80808080u) != 0) || (st & RW_HI);
                    finalize_record<TILE, NSTAGE>(S, w, r, alen, hi_argv, st & 0x7Fu, (st & RW_OPEN) != 0);
                }
            }
            __syncthreads();
        }

        // ---- phase B2: trigger needles over survivors' argv, one 16-byte chunk per thread --
        const uint32_t n_items = S.n_items;
        for (uint32_t k = tid; k < n_items; k += TILE) {
            const uint32_t itv = S.items[k];
            const uint32_t r = itv >> 4;
            const uint32_t *w = reinterpret_cast<const uint32_t *>(tile_base + (size_t)r * CH_REC_BYTES);
            const uint32_t m = argv_trigger_chunk(w, (int)S.alen[r], (int)(itv & 15u));
            if (m) atomicOr(&S.resw[r], m);
        }
        __syncthreads();
        // the stage is dead now: refill it while we compact
        if (tid == 0) {
            S.n_alive = 0;
            S.n_long = 0;
            S.n_items = 0;
            claim_and_issue(s);
        }

        // ---- phase C: tile-local, arrival-ordered compaction -----------------------------
        // (the cross-tile prefix is taken by compact_survivors afterwards: a blocking
        //  look-back here would chain every CTA's stage ring to its neighbour's)
        const uint32_t code = S.resw[tid];
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
