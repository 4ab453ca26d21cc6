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
// Execution model: persistent single-warp CTAs pull 32-record tiles from an atomic ticket
// counter; each tile is fetched with ONE 1-D TMA bulk copy (cp.async.bulk, UBLKCP) into
// shared memory, completion tracked by mbarrier transaction bytes, so the HBM stream is
// fully coalesced and asynchronous. ~19 such CTAs are resident per SM; their independent
// progress is what overlaps copy latency with the byte work. Per tile, lane i <-> record i:
//   A   pid, "type is OPEN", the 8 anchored prefixes (first 12 argv bytes): most noise
//       dies here after touching 3 of the record's 18 16-byte chunks,
//   B1  comm through a per-CTA memo table (ignore-list, trigger bits and UTF-8 validity are
//       a pure function of the 16 comm bytes, and process names repeat), type, argv NUL
//       search over the first 60 bytes,
//   B1b argv longer than that: 16 lanes per record, one 16-byte chunk per lane,
//       then suffix table and (only when a byte >= 0x80 was seen) the UTF-8 DFA,
//   B2  survivors only: SWAR trigger-needle search as (record, 8-byte chunk) work items
//       spread over all 32 lanes with a shuffle binary search, so long paths do not
//       serialise the warp,
//   C   ballot-ranked, arrival-ordered compaction inside the tile; compact_survivors then
//       applies the exclusive scan of the per-tile counts, so CTAs never wait on each other.
#pragma once
#include "common.cuh"
#include "sort.cuh"  // lookback_count_epoch

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
    uint8_t *sp_meta;        //   the META_* bits of the tile's rank-th survivor
    uint8_t *sp_ns;          //   and (optional, may be nullptr) its size in 16-byte transport slots
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
        // every field starts 4-byte aligned in the record: skip pure-ASCII words whole
        if ((i & 3) == 0 && i + 4 <= n && (*reinterpret_cast<const uint32_t *>(s + i) & 0x80808080u) == 0) {
            i += 4;
            continue;
        }
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

// Transport form of a surviving record (multi-GPU exchange, chronos_filter_partition): the
// 16-byte chunks that can hold anything before argv's NUL -- chunk 0 (pid, comm[0:12]),
// chunk 1 (comm[12:16], argv[0:12]), chunk 17 (argv[252:256], type) and as many of the
// chunks 2..16 as argv reaches into. Bytes after the NUL are not transported.
__host__ __device__ __forceinline__ uint32_t transport_slots(int alen) {
    const int extra = alen <= 12 ? 0 : (alen - 12 + 15) >> 4;
    return 3u + (uint32_t)(extra > 15 ? 15 : extra);
}

constexpr int FILTER_TILE = 32;  // records per tile: one per lane of the (single) warp of a CTA
#ifndef CHRONOS_MEMO_SETS_LOG2
#define CHRONOS_MEMO_SETS_LOG2 4
#endif
constexpr int MEMO_WAYS = 4;
constexpr int MEMO_SETS_LOG2 = CHRONOS_MEMO_SETS_LOG2;
constexpr int MEMO_SLOTS = (1 << MEMO_SETS_LOG2) * MEMO_WAYS;   // comm memo entries per CTA
constexpr int SHORT_CHUNKS = 4;  // argv NUL search done lane-per-record over chunks 1..4 (60 bytes)

template <int NSTAGE>
struct FilterSmemTail {
    uint4 memo_key[MEMO_SLOTS];  // raw 16 comm bytes
    uint32_t memo_val[MEMO_SLOTS];
    uint64_t bar[NSTAGE];
    uint32_t tile[NSTAGE];
    uint32_t resw[FILTER_TILE];  // trigger bits found by other lanes for lane i's record
};

template <int NSTAGE>
constexpr size_t filter_smem_bytes() {
    return (size_t)NSTAGE * FILTER_TILE * CH_REC_BYTES + sizeof(FilterSmemTail<NSTAGE>) + 128;
}

// Trigger needles (:141) starting in argv bytes [8j, 8j+8) of a survivor.
__device__ __forceinline__ uint32_t argv_trigger_item(const uint32_t *w, int alen, int j) {
    const uint32_t *a = w + 5;
    auto ldm = [&](int k) -> uint32_t {
        int rem = alen - 4 * k;
        if (rem <= 0) return 0u;
        uint32_t x = a[k];
        return rem >= 4 ? x : keep_low_bytes(x, rem);
    };
    const uint32_t x0 = ldm(2 * j), x1 = ldm(2 * j + 1), x2 = ldm(2 * j + 2);
    uint32_t meta = trigger_bits(x0, x1);
    meta |= trigger_bits(__funnelshift_r(x0, x1, 8), x1 >> 8);
    meta |= trigger_bits(__funnelshift_r(x0, x1, 16), x1 >> 16);
    meta |= trigger_bits(__funnelshift_r(x0, x1, 24), x1 >> 24);
    meta |= trigger_bits(x1, x2);
    meta |= trigger_bits(__funnelshift_r(x1, x2, 8), x2 >> 8);
    meta |= trigger_bits(__funnelshift_r(x1, x2, 16), x2 >> 16);
    meta |= trigger_bits(__funnelshift_r(x1, x2, 24), x2 >> 24);
    return meta;
}

// One warp per CTA, one 32-record tile per step, everything warp-synchronous: no block
// barriers, no shared work lists. Many such CTAs are resident per SM (about 9 KB of
// staged records + a small tail each), which is what hides the TMA and LDS latencies.
template <int NSTAGE>
__global__ void __launch_bounds__(FILTER_TILE) filter_kernel(const FilterParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    uint8_t *stages = smem;
    auto &S = *reinterpret_cast<FilterSmemTail<NSTAGE> *>(smem + (size_t)NSTAGE * FILTER_TILE * CH_REC_BYTES);
    constexpr uint32_t FULL = 0xffffffffu;
    constexpr uint32_t STAGE_BYTES = FILTER_TILE * CH_REC_BYTES;
    const int lane = threadIdx.x;

    // No L2 eviction hint on the tile loads: an evict-first policy (the batch is read once) costs
    // nothing when the kernel is in its fast regime, but makes the slow one (profiles/README.md,
    // "The filter is bistable") more likely -- with hinted loads in the gather before it, certain.
#ifdef CHRONOS_L2_HINT
    const uint64_t l2_evict_first = l2_policy_evict_first();
#endif
    auto issue = [&](int s, uint32_t t) {  // lane 0 only
        S.tile[s] = t;
        if (t < p.tile_end) {
            const uint64_t first = (uint64_t)t * FILTER_TILE;
            const uint32_t bytes = (uint32_t)min((uint64_t)FILTER_TILE, p.n - first) * CH_REC_BYTES;
            mbar_arrive_expect_tx(&S.bar[s], bytes);
#ifdef CHRONOS_L2_HINT
            bulk_g2s_hint(stages + (size_t)s * STAGE_BYTES, p.recs + first * CH_REC_BYTES, bytes, &S.bar[s], l2_evict_first);
#else
            bulk_g2s(stages + (size_t)s * STAGE_BYTES, p.recs + first * CH_REC_BYTES, bytes, &S.bar[s]);
#endif
        }
    };

    for (int i = lane; i < MEMO_SLOTS; i += 32) S.memo_val[i] = 0;
    if (lane == 0) {
        for (int s = 0; s < NSTAGE; s++) mbar_init(&S.bar[s], 1);
        fence_mbar_init();
        for (int s = 0; s < NSTAGE; s++) issue(s, p.tile_begin + atomicAdd(p.ticket, 1u));
    }
    __syncwarp();

    uint32_t cta_total = 0, k_or = 0, k_and = 0xffffffffu;
    for (uint32_t it = 0;; it++) {
        const int s = (NSTAGE == 1) ? 0 : (int)(it % NSTAGE);
        const uint32_t tile = S.tile[s];
        if (tile >= p.tile_end) break;
        // ticket for this stage's NEXT tile: claimed now, needed only after phase B2
        uint32_t next_ticket = 0;
        if (lane == 0) next_ticket = atomicAdd(p.ticket, 1u);
        const uint64_t first = (uint64_t)tile * FILTER_TILE;
        const uint32_t n_in = (uint32_t)min((uint64_t)FILTER_TILE, p.n - first);
        mbar_wait(&S.bar[s], (it / NSTAGE) & 1);
        const uint8_t *tile_base = stages + (size_t)s * STAGE_BYTES;
        const uint32_t *w = reinterpret_cast<const uint32_t *>(tile_base + (size_t)lane * CH_REC_BYTES);
        const uint4 *v4 = reinterpret_cast<const uint4 *>(w);

        // ---- phase A: lane i <-> record i: pid + "OPEN with a noise prefix" ----------------
        uint32_t pid = 0;
        bool alive = false, is_open = false, is_exec = false;
        if ((uint32_t)lane < n_in) {
            const uint4 q1 = v4[1];    // comm[12:16], argv[0:12]
            const uint4 q17 = v4[17];  // argv[252:256], type[0:10], pad
            pid = w[0];
            const bool nul4 = (q17.z & 0xFFu) == 0;
            is_open = (q17.y == lit4('O', 'P', 'E', 'N')) & nul4;
            is_exec = (q17.y == lit4('E', 'X', 'E', 'C')) & nul4;
            alive = !(is_open && prefix_noise(q1.y, q1.z, q1.w));
        }
        S.resw[lane] = 0;
        __syncwarp();

        // ---- phase B1 (in place): comm memo, type, argv NUL search in the first 60 bytes ----
        bool live = false;
        int alen = -1;
        uint32_t meta = 0, hi = 0;
        if (__any_sync(FULL, alive)) {
            const uint32_t c0 = w[1], c1 = w[2], c2 = w[3], c3 = w[4];
            // 4-way set-associative memo in shared memory: a CTA sees a few dozen process
            // names, and in a direct-mapped table two names that share a slot evict each other on
            // every occurrence.
            const uint32_t set =
                (((c0 * 0x9E3779B1u) ^ (c1 * 0x85EBCA77u) ^ (c2 * 0xC2B2AE3Du) ^ (c3 * 0x27D4EB2Fu)) >> (32 - MEMO_SETS_LOG2)) * MEMO_WAYS;
            uint32_t cres = 0;
            bool hit = false;
            int free_way = -1;
#pragma unroll
            for (int wy = 0; wy < MEMO_WAYS; wy++) {
                const uint32_t mv = S.memo_val[set + wy];
                const uint4 mk = S.memo_key[set + wy];
                if ((mv & CR_VALID) && mk.x == c0 && mk.y == c1 && mk.z == c2 && mk.w == c3) { hit = true; cres = mv; }
                if (!(mv & CR_VALID) && free_way < 0) free_way = wy;
            }
            const bool miss = alive && !hit;
            if (miss) cres = comm_classify(w) | CR_VALID;
            {
                // one writer per set and tile (the lowest missing lane of the set); a full set loses
                // the way the tile number points at
                const uint32_t peers = __match_any_sync(FULL, miss ? set : 0x1000u + lane);
                if (miss && (peers & lanemask_lt()) == 0) {
                    const uint32_t victim = set + (free_way >= 0 ? (uint32_t)free_way : (tile & (MEMO_WAYS - 1)));
                    S.memo_key[victim] = make_uint4(c0, c1, c2, c3);
                    S.memo_val[victim] = cres;
                }
                __syncwarp();
            }
            live = alive && (cres & (CR_IGN | CR_INVALID)) == 0;
            meta = cres & 7u;
            if (live) {
                if (is_exec) meta |= META_EXEC;
                if (!(is_open | is_exec)) {
                    const uint32_t tr = type_classify(w);
                    if (tr & CR_INVALID) live = false;
                    meta |= tr & 7u;
                }
            }
            if (live) {
                // which of the chunks 1..4 holds argv's end? (cheap: zero-byte masks only) ...
                int qe = 0;
#pragma unroll
                for (int q = 1; q <= SHORT_CHUNKS; q++) {
                    if (qe == 0) {
                        uint4 v = v4[q];
                        if (q == 1) v.x = 0x01010101u;
                        if ((zero_bytes(v.x) | zero_bytes(v.y) | zero_bytes(v.z) | zero_bytes(v.w)) != 0) qe = q;
                        else hi |= v.x | v.y | v.z | v.w;
                    }
                }
                // ... then the exact position once, whatever chunk each lane ended in
                if (qe) {
                    int pos = 0;
                    argv_chunk_end(v4, qe, pos, hi);
                    alen = pos;
                }
            }
        }

        // ---- phase B1b: argv longer than 60 bytes: 16 lanes per record, a chunk per lane ----
        uint32_t lm = __ballot_sync(FULL, live && alen < 0);
        while (lm) {
            const int ra = __ffs((int)lm) - 1;
            lm &= lm - 1;
            int rb = -1;
            if (lm) {
                rb = __ffs((int)lm) - 1;
                lm &= lm - 1;
            }
            const int r = (lane < 16) ? ra : rb;
            const bool valid = r >= 0;
            const uint4 *rv4 = reinterpret_cast<const uint4 *>(tile_base + (size_t)(valid ? r : 0) * CH_REC_BYTES);
            const int sub = lane & 15, q = SHORT_CHUNKS + 1 + sub;
            bool has = false;
            int pos = 0;
            uint32_t h2 = 0;
            if (valid && q <= 17) has = argv_chunk_end(rv4, q, pos, h2);
            const uint32_t bal = __ballot_sync(FULL, has);
            const uint32_t hbA = bal & 0xFFFFu, hbB = bal >> 16;
            const int fzA = hbA ? __ffs((int)hbA) - 1 : 0;  // chunk 17 always ends argv: hb != 0 when valid
            const int fzB = hbB ? __ffs((int)hbB) - 1 : 0;
            const int alenA = __shfl_sync(FULL, pos, fzA), alenB = __shfl_sync(FULL, pos, 16 + fzB);
            const uint32_t oA = __reduce_or_sync(FULL, (lane < 16 && sub <= fzA) ? h2 : 0u);
            const uint32_t oB = __reduce_or_sync(FULL, (lane >= 16 && valid && sub <= fzB) ? h2 : 0u);
            if (lane == ra) { alen = alenA; hi |= oA; }
            if (lane == rb) { alen = alenB; hi |= oB; }
        }

        // ---- finalize (in place): suffix table, argv UTF-8 gate ----------------------------
        bool keep = false;
        if (live) {
            const bool hi_argv = (hi & 0x80808080u) != 0;
            if (is_open && suffix_noise(w, alen)) keep = false;
            else if (hi_argv && !utf8_valid_smem(reinterpret_cast<const uint8_t *>(w) + 20, alen)) keep = false;
            else keep = true;
        }

        // ---- phase B2: trigger needles over survivors' argv as 8-byte work items ------------
        {
            const uint32_t nitems = keep ? (uint32_t)(alen + 7) >> 3 : 0u;
            uint32_t incl = nitems;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, incl, d);
                if (lane >= d) incl += t;
            }
            const uint32_t excl = incl - nitems;
            const uint32_t total = __shfl_sync(FULL, incl, 31);
            for (uint32_t b0 = 0; b0 < total; b0 += 32) {
                const uint32_t k = b0 + lane;
                int sidx = 0;  // owner of item k = number of lanes whose inclusive count is <= k
#pragma unroll
                for (int step = 16; step >= 1; step >>= 1) {
                    const uint32_t vv = __shfl_sync(FULL, incl, sidx + step - 1);
                    if (vv <= k) sidx += step;
                }
                const uint32_t ex = __shfl_sync(FULL, excl, sidx);
                const int al = __shfl_sync(FULL, alen, sidx);
                if (k < total) {
                    const uint32_t *rw = reinterpret_cast<const uint32_t *>(tile_base + (size_t)sidx * CH_REC_BYTES);
                    const uint32_t m = argv_trigger_item(rw, al, (int)(k - ex));
                    if (m) atomicOr(&S.resw[sidx], m);
                }
            }
        }
        __syncwarp();
        const uint32_t found = S.resw[lane];

        // ---- the stage is dead: refill it, then compact while the copy is in flight -------
        next_ticket = __shfl_sync(FULL, next_ticket, 0);
        if (lane == 0) issue(s, p.tile_begin + next_ticket);
        __syncwarp();

        // ---- phase C: tile-local, arrival-ordered compaction ---------------------------------
        // (compact_survivors applies the cross-tile prefix afterwards, so CTAs never wait
        //  on each other)
        const uint32_t kb = __ballot_sync(FULL, keep);
        if (kb) {
            const uint32_t rank = __popc(kb & lanemask_lt());
            if (keep) {
                const size_t slot = (size_t)tile * FILTER_TILE + rank;
                p.sp_key[slot] = pid;
                p.sp_src[slot] = (uint32_t)first + lane;
                p.sp_meta[slot] = (uint8_t)((meta | found) & 0x7Fu);
                if (p.sp_ns) p.sp_ns[slot] = (uint8_t)transport_slots(alen);
            }
            k_or |= __reduce_or_sync(FULL, keep ? pid : 0u);
            k_and &= __reduce_and_sync(FULL, keep ? pid : 0xffffffffu);
        }
        if (lane == 0) p.tile_count[tile] = __popc(kb);
        cta_total += __popc(kb);
    }
    if (lane == 0 && cta_total) {
        atomicAdd(p.out_count, cta_total);
        atomicOr(p.key_or, k_or);
        atomicAnd(p.key_and, k_and);
    }
}

// Cross-tile step of the ordered compaction: tile_off[] = exclusive scan of tile_count[]
// (scan_exclusive_u32); one thread per filter tile moves its short run to the dense arrays.
struct CompactParams {
    const uint32_t *tile_off;
    const uint32_t *sp_key;
    const uint32_t *sp_src;
    const uint8_t *sp_meta;
    const uint8_t *sp_ns;    // optional
    uint32_t *out_key;
    uint32_t *out_src;
    uint8_t *out_meta;
    uint8_t *out_ns;         // optional
    uint32_t *out_pid;       // optional second copy of the pids (the sort consumes out_key)
    uint32_t n_tiles;
    uint32_t tile;       // records per filter tile
    const uint32_t *total;  // device scalar: survivor count == tile_off[n_tiles]
};

__global__ void __launch_bounds__(256) compact_survivors(const CompactParams c) {
    // a tile holds 32 records and on average only a few survivors: one thread per tile
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < c.n_tiles; t += gridDim.x * blockDim.x) {
        const uint32_t lo = c.tile_off[t];
        const uint32_t hi = (t + 1 < c.n_tiles) ? c.tile_off[t + 1] : *c.total;
        const size_t base = (size_t)t * c.tile;
        for (uint32_t j = 0; j < hi - lo; j++) {
            c.out_key[lo + j] = c.sp_key[base + j];
            c.out_src[lo + j] = c.sp_src[base + j];
            c.out_meta[lo + j] = c.sp_meta[base + j];
            if (c.out_ns) c.out_ns[lo + j] = c.sp_ns[base + j];
            if (c.out_pid) c.out_pid[lo + j] = c.sp_key[base + j];
        }
    }
}

// The same step in ONE kernel for the batch pipeline: a chunk of CS_CHUNK filter tiles per
// CTA iteration -- scan of the tile counts inside the chunk, chunk base by decoupled look-back
// (sort.cuh lookback_count_epoch), then every thread places survivors by output position
// (binary search of the position in the chunk's offsets), so the writes are coalesced.
constexpr int CS_THREADS = 256;
constexpr int CS_ITEMS = 4;
constexpr int CS_CHUNK = CS_THREADS * CS_ITEMS;
static_assert(CS_CHUNK == 1024, "the position search below is unrolled for 2^10 tiles");

struct CompactScanParams {
    const uint32_t *tile_count;  // survivors per filter tile (filter_kernel), left untouched
    uint32_t n_tiles;
    uint32_t tile;               // records (= slots) per filter tile
    const uint32_t *sp_key, *sp_src;
    const uint8_t *sp_meta, *sp_ns;
    uint32_t *out_key, *out_src, *out_pid;  // out_key: optional second copy of the pids (may be nullptr)
    uint8_t *out_meta, *out_ns;
    uint32_t *ticket;            // zeroed
    uint64_t *status;            // ceil(n_tiles / CS_CHUNK) epoch-tagged words
    uint32_t epoch;
};

__global__ void __launch_bounds__(CS_THREADS) compact_scan(const CompactScanParams p) {
    __shared__ uint32_t off[CS_CHUNK + 1];
    __shared__ uint32_t warp_sums[CS_THREADS / 32];
    __shared__ uint32_t s_chunk, s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n_chunks = (p.n_tiles + CS_CHUNK - 1) / CS_CHUNK;
    while (true) {
        __syncthreads();
        if (tid == 0) s_chunk = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const uint32_t chunk = s_chunk;
        if (chunk >= n_chunks) break;
        const uint32_t t0 = chunk * CS_CHUNK + tid * CS_ITEMS;
        const uint32_t *tc = p.tile_count;
        uint32_t v[CS_ITEMS];
        if (t0 + CS_ITEMS <= p.n_tiles) {
            const uint4 a = *reinterpret_cast<const uint4 *>(tc + t0);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
        } else {
#pragma unroll
            for (int k = 0; k < CS_ITEMS; k++) v[k] = (t0 + k < p.n_tiles) ? tc[t0 + k] : 0u;
        }
        const uint32_t sum = v[0] + v[1] + v[2] + v[3];
        uint32_t inc = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += t;
        }
        if (lane == 31) warp_sums[warp] = inc;
        __syncthreads();
        uint32_t before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < CS_THREADS / 32; w++) {
            const uint32_t ws = warp_sums[w];
            if (w < warp) before += ws;
            total += ws;
        }
        uint32_t run = before + inc - sum;
#pragma unroll
        for (int k = 0; k < CS_ITEMS; k++) {
            off[tid * CS_ITEMS + k] = run;
            run += v[k];
        }
        if (tid == CS_THREADS - 1) off[CS_CHUNK] = total;
        if (warp == 0) {
            const uint32_t b = lookback_count_epoch(p.status, chunk, total, lane, p.epoch);
            if (lane == 0) s_base = b;
        }
        __syncthreads();
        const uint32_t base = s_base;
        const size_t slot0 = (size_t)chunk * CS_CHUNK * p.tile;
        // four independent positions per thread and step: the search and the four loads of one
        // position are a dependent chain, so the chains are interleaved
        constexpr int CS_ILP = 4;
        for (uint32_t j0 = tid; j0 < total; j0 += CS_ILP * CS_THREADS) {
            size_t src[CS_ILP];
            uint32_t k[CS_ILP], sr[CS_ILP];
            uint8_t me[CS_ILP], nn[CS_ILP];
#pragma unroll
            for (int u = 0; u < CS_ILP; u++) {
                const uint32_t j = min(j0 + u * CS_THREADS, total - 1);
                uint32_t lo = 0, hi = CS_CHUNK;  // off[lo] <= j < off[hi]
#pragma unroll
                for (int it = 0; it < 10; it++) {  // CS_CHUNK = 1024 = 2^10
                    const uint32_t mid = (lo + hi) >> 1;
                    if (off[mid] <= j) lo = mid; else hi = mid;
                }
                src[u] = slot0 + (size_t)lo * p.tile + (j - off[lo]);
            }
#pragma unroll
            for (int u = 0; u < CS_ILP; u++) {
                k[u] = ld_sparse_u32(p.sp_key + src[u]);  // (one 128-byte line per filter tile, its head in use)
                sr[u] = ld_sparse_u32(p.sp_src + src[u]);
                me[u] = p.sp_meta[src[u]];
                nn[u] = p.sp_ns[src[u]];
            }
#pragma unroll
            for (int u = 0; u < CS_ILP; u++) {
                const uint32_t j = j0 + u * CS_THREADS;
                if (j < total) {
                    const uint32_t dst = base + j;
                    CH_ASSERT(dst < p.n_tiles * p.tile);
                    if (p.out_key) p.out_key[dst] = k[u];
                    p.out_pid[dst] = k[u];
                    p.out_src[dst] = sr[u];
                    p.out_meta[dst] = me[u];
                    p.out_ns[dst] = nn[u];
                }
            }
        }
    }
}

}  // namespace chronos
