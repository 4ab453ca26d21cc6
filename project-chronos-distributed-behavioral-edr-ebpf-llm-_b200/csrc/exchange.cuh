// exchange.cuh -- repartition of survivors by owner rank, sender and receiver.
//
// process_memory is keyed by pid and PIDs never interact (chronos_sensor.py:105, :138, :142,
// :157), so the window logic shards by owner = hash32(pid) % world. Every rank filters a
// contiguous slice of the arrival-ordered batch; its survivors then travel to their owners
// in the transport form (filter.cuh transport_slots): one header word
// (arrival 48 bits | META << 48 | slot count << 56) and 3..18 16-byte slots per survivor.
//
// The partition is one stable counting pass over the survivors (arrival order is kept inside
// every owner's stream, which is all the reference's semantics need, :18-23 has no timestamp):
//   xchg_hist   per tile of XCHG_TILE survivors: entries and slots per owner
//   scan        exclusive scan of that table (sort.cuh)
//   xchg_pack   stable ranks inside the tile, then the headers and slots are written
//               STRAIGHT INTO THE OWNER'S RECEIVE WINDOW over NVLink (peer pointers), and the
//               last CTA to finish publishes the counts and raises this rank's flag there.
// There is no staging buffer, no count exchange and no host round trip on that path; the same
// kernel can instead write one local stream ordered by (owner, arrival) for a host-driven
// all-to-all (chronos_filter_partition).
//
// Receive window of a rank (one allocation, exported with cudaIpcGetMemHandle), two halves
// used alternately (epoch & 1); per half:
//   ctrl   u32 entries[16] | u32 slots[16] | u32 flag[16]      (one word per SOURCE rank)
//   desc   world x cap_e uint4    region s = what source rank s sent, arrival order:
//                                 {header word lo, hi, first slot of the entry relative to region s, 0}
//   slots  world x cap_s uint4
// Why two halves are enough: a rank packs step k+1 only after its own stage 2 of step k, which
// waited for every peer's flag of step k; a peer raises that flag after ITS stage 2 of step
// k-1 (stream order), the last reader of the half that step k+1 overwrites.
#pragma once
#include "common.cuh"

namespace chronos {

constexpr int XCHG_MAX_WORLD = 16;
constexpr uint32_t XCHG_TILE = 1024;
constexpr uint32_t XCHG_THREADS = 256;
constexpr uint32_t XCHG_WARPS = XCHG_THREADS / 32;
constexpr uint32_t XCHG_OVERFLOW = 0xFFFFFFFFu;
constexpr uint32_t XCHG_CTRL_BYTES = 256;
constexpr uint32_t XERR_OVERFLOW = 1u, XERR_TIMEOUT = 2u, XERR_PEER_OVERFLOW = 4u;

struct XchgLayout {
    uint32_t world, rank;
    uint32_t cap_e, cap_s;  // per (source, owner) pair: entries, slots
    uint64_t off_desc, off_slots, half_bytes;
};
__host__ __device__ inline uint64_t xchg_align(uint64_t v) { return (v + 255) & ~255ull; }
__host__ __device__ inline XchgLayout xchg_layout(uint32_t world, uint32_t rank, uint32_t cap_e, uint32_t cap_s) {
    XchgLayout L;
    L.world = world; L.rank = rank; L.cap_e = cap_e; L.cap_s = cap_s;
    L.off_desc = XCHG_CTRL_BYTES;
    L.off_slots = L.off_desc + xchg_align((uint64_t)world * cap_e * 16);
    L.half_bytes = L.off_slots + xchg_align((uint64_t)world * cap_s * 16);
    return L;
}

__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint64_t global_timer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ---------------------------------------------------------------------------------------
// per tile: how many entries / transport slots go to each owner.
// table is owner-major with row stride T (an upper bound of the tile count, known to the host):
// rows [0, W) entries, rows [W, 2W) slots, and one closing element table[2W*T] so that the
// exclusive scan leaves the grand total there. Persistent CTAs; tiles past the survivors get
// zeros (the scan runs over the whole table).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(XCHG_THREADS) xchg_hist(const uint32_t *pid, const uint8_t *ns, const uint32_t *n_ptr,
                                                         uint32_t world, uint32_t T, uint32_t *table) {
    __shared__ uint32_t he[XCHG_MAX_WORLD], hs[XCHG_MAX_WORLD];
    const uint32_t S = *n_ptr;
    const uint32_t used = (S + XCHG_TILE - 1) / XCHG_TILE;
    if (blockIdx.x == 0 && threadIdx.x == 0) table[(size_t)2 * world * T] = 0;
    // columns of tiles that hold no survivor: zeros
    for (uint32_t i = blockIdx.x * XCHG_THREADS + threadIdx.x; i < T * 2 * world; i += gridDim.x * XCHG_THREADS)
        if (i % T >= used) table[i] = 0;
    for (uint32_t tile = blockIdx.x; tile < used; tile += gridDim.x) {
        __syncthreads();
        if (threadIdx.x < XCHG_MAX_WORLD) { he[threadIdx.x] = 0; hs[threadIdx.x] = 0; }
        __syncthreads();
        const uint32_t lo = tile * XCHG_TILE, hi = min(S, lo + XCHG_TILE);
        for (uint32_t i = lo + threadIdx.x; i < hi; i += XCHG_THREADS) {
            const uint32_t o = hash32(pid[i]) % world;
            atomicAdd(&he[o], 1u);
            atomicAdd(&hs[o], (uint32_t)ns[i]);
        }
        __syncthreads();
        if (threadIdx.x < world) {
            table[(size_t)threadIdx.x * T + tile] = he[threadIdx.x];
            table[(size_t)(world + threadIdx.x) * T + tile] = hs[threadIdx.x];
        }
    }
}

struct XchgPackParams {
    // survivors of this rank, arrival order
    const uint32_t *pid, *surv_src;
    const uint8_t *surv_meta, *surv_ns;
    const uint32_t *n_ptr;        // survivor count (device scalar)
    const uint8_t *batch;         // the slice's records (device)
    uint64_t base_arrival;
    const uint32_t *table;        // scanned
    uint32_t T;                   // row stride of the table
    uint32_t world;
    // direct mode: peer[d] = base of owner d's receive window (nullptr => local mode)
    uint8_t *peer[XCHG_MAX_WORLD];
    XchgLayout L;
    uint32_t epoch;
    // local mode: one stream ordered by (owner, arrival)
    uint4 *send_slots;
    uint64_t *send_hdr;
    uint64_t cap_entries, cap_slots;
    // both: per-owner totals for the host (entries[16] | slots[16]), error bits, CTA counter
    uint32_t *totals;
    uint32_t *err;
    uint32_t *done;
    bool direct;
};

constexpr uint32_t XCHG_SUB = XCHG_TILE / XCHG_WARPS;   // contiguous sub-range of a warp
// Staging of one warp step: a WINDOW of XCHG_STAGE_SLOTS transport slots (a step of 32 survivors
// holds 96..576, about 180 on the bench workloads) + its 32 descriptors. A full-size buffer
// (576 slots) made the kernel shared-memory bound at 2 CTAs per SM, 24 % of the warp slots:
// the kernel is a chain of dependent loads, it needs the warps. Steps with more slots than the
// window take several rounds through it.
constexpr uint32_t XCHG_STAGE_SLOTS = 224;
constexpr uint32_t XCHG_STAGE_PER_WARP = XCHG_STAGE_SLOTS + 32;
constexpr size_t XCHG_PACK_SMEM = (size_t)XCHG_WARPS * XCHG_STAGE_PER_WARP * 16;
constexpr uint32_t NONE_AT = 0xffffffffu;
constexpr int XCHG_STEPS = XCHG_SUB / 32;               // 32 survivors per step

__global__ void __launch_bounds__(XCHG_THREADS) xchg_pack(const XchgPackParams g) {
    __shared__ uint32_t rs_e[XCHG_MAX_WORLD + 1], rs_s[XCHG_MAX_WORLD + 1];  // row starts = per-owner stream starts
    __shared__ uint32_t wb_e[XCHG_WARPS][XCHG_MAX_WORLD], wb_s[XCHG_WARPS][XCHG_MAX_WORLD];
    __shared__ uint32_t s_skip, s_last;
    // owner-major order of the tile's survivors (stable inside an owner): perm[j] = index in the tile
    __shared__ uint16_t perm[XCHG_TILE];
    __shared__ uint32_t o_start[XCHG_MAX_WORLD + 1], o_run[XCHG_MAX_WORLD];
    extern __shared__ uint4 stage[];  // XCHG_PACK_SMEM bytes
    const uint32_t T = g.T, W = g.world, S = *g.n_ptr;
    const uint32_t used = (S + XCHG_TILE - 1) / XCHG_TILE;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid <= (int)W) {
        rs_e[tid] = g.table[(size_t)tid * T];
        rs_s[tid] = g.table[(size_t)(W + tid) * T];
    }
    if (tid == 0) s_skip = 0;
    __syncthreads();
    // owners whose stream does not fit are skipped as a whole (reported, never truncated)
    if (tid < (int)W) {
        const uint32_t te = rs_e[tid + 1] - rs_e[tid], ts = rs_s[tid + 1] - rs_s[tid];
        const bool over = g.direct ? (te > g.L.cap_e || ts > g.L.cap_s)
                                   : ((uint64_t)rs_e[W] > g.cap_entries || (uint64_t)(rs_s[W] - rs_s[0]) > g.cap_slots);
        if (over) atomicOr(&s_skip, 1u << tid);
    }
    __syncthreads();
    const uint32_t skip = s_skip;
    const uint32_t half = g.epoch & 1u;
    for (uint32_t tile = blockIdx.x; tile < used; tile += gridDim.x) {
        __syncwarp();  // (the copy loop below ends in predicated stores: reconverge before the barrier)
        __syncthreads();
        for (int i = tid; i < (int)(XCHG_WARPS * XCHG_MAX_WORLD); i += XCHG_THREADS) {
            (&wb_e[0][0])[i] = 0;
            (&wb_s[0][0])[i] = 0;
        }
        __syncthreads();
        const uint32_t t_lo = tile * XCHG_TILE, t_hi = min(S, t_lo + XCHG_TILE);
        // 0) The tile's survivors in OWNER-MAJOR order (a stable counting sort of 1024 indices in
        //    shared memory). The warps then walk that order, so the 32 survivors of a warp step
        //    nearly always share one owner and leave as ONE run of a few KB: in arrival order a
        //    step's run per owner was about 360 bytes of slots (eight owners), too short for NVLink.
        if (tid == 0) {
            uint32_t run = 0;
            for (uint32_t d = 0; d < W; d++) {  // entries of this tile per owner, from the scanned table
                o_start[d] = run;
                run += g.table[(size_t)d * T + tile + 1] - g.table[(size_t)d * T + tile];
            }
            o_start[W] = run;
        }
        if (tid < XCHG_MAX_WORLD) o_run[tid] = 0;
        __syncthreads();
        for (uint32_t r0 = 0; r0 < XCHG_TILE; r0 += XCHG_THREADS) {
            const uint32_t i = t_lo + r0 + tid;
            const uint32_t o = i < t_hi ? hash32(g.pid[i]) % W : 0xFFu;
            const uint32_t peers = __match_any_sync(0xffffffffu, o);
            const uint32_t below = __popc(peers & lanemask_lt());
            if (o != 0xFFu && below == 0) wb_e[warp][o] = __popc(peers);  // (wb_e is zero here: see the reset above)
            __syncthreads();
            if (tid < (int)W) {  // where each warp's group of owner `tid` starts
                uint32_t run = o_run[tid];
                for (int w = 0; w < (int)XCHG_WARPS; w++) {
                    const uint32_t cnt = wb_e[w][tid];
                    wb_e[w][tid] = run;
                    run += cnt;
                }
                o_run[tid] = run;
            }
            __syncthreads();
            if (o != 0xFFu) perm[o_start[o] + wb_e[warp][o] + below] = (uint16_t)(r0 + tid);
            __syncthreads();
            for (int k = tid; k < (int)(XCHG_WARPS * XCHG_MAX_WORLD); k += XCHG_THREADS) (&wb_e[0][0])[k] = 0;
            __syncthreads();
        }
        const uint32_t n_tile = t_hi - t_lo;
        const uint32_t w_lo = min(n_tile, warp * XCHG_SUB), w_hi = min(n_tile, w_lo + XCHG_SUB);  // in owner-major positions
        // 1) this warp's survivors, all steps at once (the loads of the later steps are in
        //    flight while the first step is placed), and the per-warp counts per owner
        uint32_t ow[XCHG_STEPS], nsv[XCHG_STEPS], recv[XCHG_STEPS], metav[XCHG_STEPS];
#pragma unroll
        for (int q = 0; q < XCHG_STEPS; q++) {
            const uint32_t j = w_lo + q * 32 + lane;
            const uint32_t i = j < w_hi ? t_lo + perm[j] : 0xffffffffu;
            ow[q] = 0xFFu; nsv[q] = 0; recv[q] = 0; metav[q] = 0;
            if (j < w_hi) {
                ow[q] = hash32(g.pid[i]) % W;
                nsv[q] = g.surv_ns[i];
                recv[q] = g.surv_src[i];
                metav[q] = g.surv_meta[i];
            }
        }
#pragma unroll
        for (int q = 0; q < XCHG_STEPS; q++) {
            if (ow[q] != 0xFFu) {
                atomicAdd(&wb_e[warp][ow[q]], 1u);
                atomicAdd(&wb_s[warp][ow[q]], nsv[q]);
            }
        }
        __syncwarp();
        __syncthreads();
        // 2) owner d: running base over the warps, seeded with this tile's offset inside d's stream
        if (tid < (int)W) {
            uint32_t re = g.table[(size_t)tid * T + tile], rsl = g.table[(size_t)(W + tid) * T + tile];
            if (g.direct) { re -= rs_e[tid]; rsl -= rs_s[tid]; }   // relative to the (source, owner) region
            else rsl -= rs_s[0];                                   // slot rows come after the entry rows in the scan
#pragma unroll
            for (int w = 0; w < (int)XCHG_WARPS; w++) {
                const uint32_t ce = wb_e[w][tid], cs = wb_s[w][tid];
                wb_e[w][tid] = re; wb_s[w][tid] = rsl;
                re += ce; rsl += cs;
            }
        }
        __syncwarp();
        __syncthreads();
        // 3) stable placement, 32 survivors per step (every step is executed by the whole warp;
        //    lanes past the end carry owner 0xFF and nothing to copy). The warp walks the tile in
        //    owner-major order, so the lanes of one owner are NEIGHBOURS: a group is a lane range
        //    [leader, last], and what a group needs -- ranks, slots in front of an entry, where the
        //    group's run starts in the staging buffer and at the owner -- falls out of two warp
        //    scans. (Per-owner ballots and a search per copied slot cost 41 warp instructions per
        //    survivor, 60 % of the kernel: ncu, profiles/README.md.)
#pragma unroll
        for (int q = 0; q < XCHG_STEPS; q++) {
            constexpr uint32_t FULL = 0xffffffffu;
            const uint32_t o = ow[q], ns = nsv[q], rec = recv[q];
            const bool valid = o != 0xFFu;
            const uint32_t peers = __match_any_sync(FULL, o);
            const int leader = __ffs((int)peers) - 1, last = 31 - __clz((int)peers);
            CH_ASSERT(__popc(peers) == last - leader + 1);  // the group is a lane range
            const uint32_t rank_e = (uint32_t)(lane - leader);
            const bool live = valid && !((skip >> (o & 31u)) & 1u);  // (an owner is skipped as a whole)
            // slots in front of this lane: over all valid lanes (the books), over live lanes (the staging)
            const uint32_t nv = valid ? ns : 0u, cnt = live ? ns : 0u;
            uint32_t inc_v = nv, inc_l = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t1 = __shfl_up_sync(FULL, inc_v, d), t2 = __shfl_up_sync(FULL, inc_l, d);
                if (lane >= d) { inc_v += t1; inc_l += t2; }
            }
            const uint32_t exc_v = inc_v - nv, exc_l = inc_l - cnt;
            const uint32_t pre_s = exc_v - __shfl_sync(FULL, exc_v, leader);  // slots of my owner's lower lanes
            uint32_t be = 0, bs = 0;
            if (valid) { be = wb_e[warp][o]; bs = wb_s[warp][o]; }
            __syncwarp();
            if (valid && lane == last) {  // highest lane of its group: it knows the group totals
                wb_e[warp][o] = be + rank_e + 1;
                wb_s[warp][o] = bs + pre_s + ns;
            }
            __syncwarp();
            // ---- owner-major staging in shared memory ----------------------------------------
            // Written entry by entry, a survivor is an 80-byte fragment at some place in its
            // owner's window; NVLink moves such scattered fragments at about 375 GB/s, runs of
            // 128 bytes and more at 660 GB/s (tools/nvlink_probe.cu). All entries a warp step
            // sends to ONE owner are neighbours at the destination, so the step is staged in lane
            // (= owner-major) order and every group's run then leaves with consecutive lanes on
            // consecutive slots.
            const uint32_t live_mask = __ballot_sync(FULL, live);
            const uint32_t e_idx = __popc(live_mask & lanemask_lt());  // staged descriptor of this lane
            const uint32_t total = __shfl_sync(FULL, inc_l, 31);
            const uint32_t g_s0 = __shfl_sync(FULL, exc_l, leader), g_s1 = __shfl_sync(FULL, inc_l, last);  // the group's staged slots
            const uint32_t g_e0 = __shfl_sync(FULL, e_idx, leader);
            const uint32_t g_ne = live ? (uint32_t)(last - leader + 1) : 0u;
            uint4 *run_dst = nullptr, *desc_dst = nullptr;  // where the group's run / descriptors go (same on all its lanes)
            if (live) {
                if (g.direct) {
                    uint8_t *hb = g.peer[o] + (size_t)half * g.L.half_bytes;
                    run_dst = reinterpret_cast<uint4 *>(hb + g.L.off_slots) + (size_t)g.L.rank * g.L.cap_s + bs;
                    desc_dst = reinterpret_cast<uint4 *>(hb + g.L.off_desc) + (size_t)g.L.rank * g.L.cap_e + be;
                } else {
                    run_dst = g.send_slots + bs;
                }
            }
            uint4 *stg = stage + (size_t)warp * XCHG_STAGE_PER_WARP;  // [0, STAGE_SLOTS) slots, then 32 descriptors
            const uint4 *src = nullptr;
            if (live) {
                const uint64_t hdr = ((g.base_arrival + rec) & 0xFFFFFFFFFFFFull) | ((uint64_t)metav[q] << 48) | ((uint64_t)ns << 56);
                src = reinterpret_cast<const uint4 *>(g.batch + (size_t)rec * CH_REC_BYTES);
                if (g.direct) stg[XCHG_STAGE_SLOTS + e_idx] = make_uint4((uint32_t)hdr, (uint32_t)(hdr >> 32), bs + pre_s, 0u);
                else g.send_hdr[be + rank_e] = hdr;
            }
            const uint32_t group_leaders = __ballot_sync(FULL, live && lane == leader);
            for (uint32_t w0 = 0; w0 < total; w0 += XCHG_STAGE_SLOTS) {  // staged slots [w0, w1) this round
                const uint32_t w1 = min(total, w0 + XCHG_STAGE_SLOTS);
                // every lane stages the slots [sa, sb) of its own entry that fall into the window (chunks
                // 0, 1, 17, then 2..), global -> shared memory without passing through registers: all
                // slots of the step are in flight together, one round trip to the batch per step (with
                // register loads, four per trip, the wait for them was 40 % of the kernel's stall samples)
                const uint32_t sa = exc_l < w0 ? min(cnt, w0 - exc_l) : 0u;
                const uint32_t sb = exc_l >= w1 ? 0u : min(cnt, w1 - exc_l);
                for (uint32_t sl = sa; sl < sb; sl++) {
                    CH_ASSERT(exc_l + sl >= w0 && exc_l + sl - w0 < XCHG_STAGE_SLOTS);
                    cp_async16(stg + (exc_l + sl - w0), src + (sl < 2 ? sl : (sl == 2 ? 17u : sl - 1u)), 16u);
                }
                cp_async_commit();
                cp_async_wait<0>();
                __syncwarp();
                // every group's run leaves with consecutive lanes on consecutive 16-byte slots
                for (uint32_t lm = group_leaders; lm; lm &= lm - 1) {
                    const int L = __ffs((int)lm) - 1;
                    const uint32_t s0 = __shfl_sync(FULL, g_s0, L), s1 = __shfl_sync(FULL, g_s1, L);
                    const unsigned long long dp = __shfl_sync(FULL, (unsigned long long)(uintptr_t)run_dst, L);
                    const uint32_t hi = min(s1, w1);
                    for (uint32_t j = max(s0, w0) + (uint32_t)lane; j < hi; j += 32) {
                        CH_ASSERT(j - w0 < XCHG_STAGE_SLOTS);
                        reinterpret_cast<uint4 *>((uintptr_t)dp)[j - s0] = stg[j - w0];
                    }
                }
                __syncwarp();  // the window is refilled by the next round
            }
            if (g.direct) {  // descriptors: at most 32 per step, group by group
                for (uint32_t lm = group_leaders; lm; lm &= lm - 1) {
                    const int L = __ffs((int)lm) - 1;
                    const uint32_t e0 = __shfl_sync(FULL, g_e0, L), ne = __shfl_sync(FULL, g_ne, L);
                    const unsigned long long dp = __shfl_sync(FULL, (unsigned long long)(uintptr_t)desc_dst, L);
                    if ((uint32_t)lane < ne) reinterpret_cast<uint4 *>((uintptr_t)dp)[lane] = stg[XCHG_STAGE_SLOTS + e0 + lane];
                }
            }
            __syncwarp();  // the staging buffer is reused by the next step
        }
    }
    // 4) the last CTA to finish publishes counts and flag. One system-scope fence per CTA, by
    //    the thread that then counts the CTA in: the block barrier orders every thread's
    //    stores before it (the pattern of a grid barrier); a fence per thread made this
    //    kernel membar-bound.
    __syncwarp();
    __syncthreads();
    if (tid == 0) {
        __threadfence_system();
        s_last = (atomicAdd(g.done, 1u) == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence_system();
    if (tid < (int)W) {
        const uint32_t te = rs_e[tid + 1] - rs_e[tid], ts = rs_s[tid + 1] - rs_s[tid];
        const bool over = (skip >> tid) & 1u;
        g.totals[tid] = te;
        g.totals[XCHG_MAX_WORLD + tid] = ts;
        if (over) atomicOr(g.err, XERR_OVERFLOW);
        if (g.direct) {
            uint32_t *ctrl = reinterpret_cast<uint32_t *>(g.peer[tid] + (size_t)half * g.L.half_bytes);
            ctrl[g.L.rank] = over ? XCHG_OVERFLOW : te;
            ctrl[XCHG_MAX_WORLD + g.L.rank] = over ? 0u : ts;
            __threadfence_system();
            st_release_sys_u32(ctrl + 2 * XCHG_MAX_WORLD + g.L.rank, g.epoch);
        }
    }
}

// ---------------------------------------------------------------------------------------
// receiver: wait for every source's flag, then turn the regions of this half into the dense,
// source-ordered (= arrival-ordered) entry arrays the window stage reads.
// ---------------------------------------------------------------------------------------
struct XchgUnpackParams {
    const uint8_t *window;   // this rank's receive window
    XchgLayout L;
    uint32_t epoch;
    uint64_t timeout_ns;
    // outputs, one per received entry j
    uint64_t *arrival;
    uint8_t *meta, *ns;
    uint32_t *slot_off;      // index of the entry's first slot in the half's slot area
    uint32_t *key, *pid_copy, *src;
    uint32_t *n_out, *k_or, *k_and, *err;
    uint32_t n_cap;          // capacity of the output arrays
};

__global__ void __launch_bounds__(256) xchg_unpack(const XchgUnpackParams g) {
    __shared__ uint32_t es[XCHG_MAX_WORLD + 1];
    __shared__ uint32_t s_err;
    const uint32_t W = g.L.world, half = g.epoch & 1u;
    const uint8_t *hb = g.window + (size_t)half * g.L.half_bytes;
    const uint32_t *ctrl = reinterpret_cast<const uint32_t *>(hb);
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        uint32_t e = 0, bad = 0;
        if (lane < (int)W) {
            const uint64_t t0 = global_timer_ns();
            while (ld_acquire_sys_u32(ctrl + 2 * XCHG_MAX_WORLD + lane) != g.epoch) {
                if (global_timer_ns() - t0 > g.timeout_ns) { bad = XERR_TIMEOUT; break; }
                __nanosleep(200);
            }
            if (!bad) {
                e = ctrl[lane];
                if (e == XCHG_OVERFLOW) { e = 0; bad = XERR_PEER_OVERFLOW; }
                else if (e > g.L.cap_e) { e = 0; bad = XERR_PEER_OVERFLOW; }
            }
        }
        bad = __reduce_or_sync(0xffffffffu, bad);
        if (bad & XERR_TIMEOUT) e = 0;  // nothing of this step is trusted
        uint32_t incl = e;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        if (lane <= (int)W) es[lane] = incl - e;  // es[W] = total
        if (lane == 0) s_err = bad;
    }
    __syncthreads();
    uint32_t total = es[W];
    if (total > g.n_cap) total = g.n_cap;  // cannot happen when the context was sized for world * cap_e
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *g.n_out = total;
        if (s_err) atomicOr(g.err, s_err);
    }
    const uint4 *desc = reinterpret_cast<const uint4 *>(hb + g.L.off_desc);
    const uint4 *slots = reinterpret_cast<const uint4 *>(hb + g.L.off_slots);
    uint32_t o = 0, a = 0xffffffffu;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < total; j += gridDim.x * blockDim.x) {
        uint32_t s = 0;
#pragma unroll
        for (int k = 1; k < XCHG_MAX_WORLD; k++)
            if (k < (int)W && j >= es[k]) s = k;
        CH_ASSERT(j - es[s] < g.L.cap_e);
        const size_t r = (size_t)s * g.L.cap_e + (j - es[s]);
        const uint4 dsc = desc[r];
        const uint64_t h = (uint64_t)dsc.x | ((uint64_t)dsc.y << 32);
        const uint32_t so = s * g.L.cap_s + dsc.z;
        const uint32_t k = reinterpret_cast<const uint32_t *>(slots + so)[0];  // pid = first word of slot 0
        g.arrival[j] = h & 0xFFFFFFFFFFFFull;
        g.meta[j] = (uint8_t)(h >> 48);
        g.ns[j] = (uint8_t)(h >> 56);
        g.slot_off[j] = so;
        if (g.key) g.key[j] = k;
        g.pid_copy[j] = k;
        g.src[j] = j;
        o |= k;
        a &= k;
    }
    o = __reduce_or_sync(0xffffffffu, o);
    a = __reduce_and_sync(0xffffffffu, a);
    if ((threadIdx.x & 31) == 0 && (o | ~a)) {
        atomicOr(g.k_or, o);
        atomicAnd(g.k_and, a);
    }
}

}  // namespace chronos
