// chronos.cu -- context, batch pipeline and the C ABI (include/chronos.h) of the
// B200-native CHRONOS hot path. Kernels live in filter.cuh / sort.cuh / window.cuh.
//
// Per batch (one perf_buffer_poll() worth of events, chronos_sensor.py:160-162):
//   filter_kernel      raw records -> tile-local survivors (pid, src index, meta)
//   compact_scan       dense, arrival-ordered survivor arrays
//   sort_prepare + onesweep_pass x4
//                      stable sort of [carried-in pending ++ survivors] by pid
//   gather_records     every grouped event copied once into the arena (second stream)
//   group_forward      flush automaton + chain labeller -> window descriptors
//   window_reverse     unflushed tails -> next batch's pending list (process_memory)
//   emit_ordered       windows in trigger-arrival order (the order analyze_sequence
//                      is called in, :143-147)
// A submit only enqueues: extents of everything after the filter live in a device-side plan,
// grids are upper bounds; the first host wait is chronos_fetch.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <new>
#include <vector>

#include "chronos.h"
#include "common.cuh"
#include "filter.cuh"
#include "sort.cuh"
#include "tracegen.cuh"
#include "chain.cuh"
#include "fnscan.cuh"
#include "window.cuh"
#include "exchange.cuh"

using namespace chronos;

namespace {

struct DevBuf {
    void *p = nullptr;
    size_t bytes = 0;
};

struct DeviceScalars {  // one 64-byte block on the device, mirrored in pinned memory
    uint32_t surv_count;
    uint32_t key_or;
    uint32_t key_and;
    uint32_t n_windows;
    uint32_t n_pending;
    uint32_t n_chains;
    uint32_t xchg_err;  // XERR_* bits of the exchange (exchange.cuh); cleared by stage 1, not per batch
    uint32_t n_pids_seen;  // copy of the per-PID table's count at the end of the batch
    uint32_t pad[8];
};
static_assert(sizeof(DeviceScalars) == 64, "DeviceScalars layout");
// the Plan of the batch in flight lives right behind the scalars: ONE copy brings both back
constexpr size_t RESULT_BLOCK_BYTES = 64 + ((sizeof(Plan) + 63) / 64) * 64;

constexpr int N_TICKETS = 64;
// ticket counters followed by the global digit histograms of the sort (sort.cuh), cleared together
constexpr size_t TICKET_BYTES = (N_TICKETS + RADIX_MAX_PASSES * RADIX_BINS) * 4;
constexpr uint32_t COMM_MEMO_SLOTS = 8192;  // device-wide comm memo (insert-only), 192 KB

}  // namespace

struct chronos_ctx;
namespace {
void mark(chronos_ctx *c, const char *name);
}
struct chronos_ctx {
    int device = 0;
    int sm_count = 0;
    uint64_t max_batch = 0;
    cudaStream_t stream = nullptr, stream2 = nullptr;  // stream2: record gather, overlapped with the window scans
    cudaEvent_t ev_sorted = nullptr, ev_gathered = nullptr;
    cudaEvent_t ev_start = nullptr, ev_filter = nullptr, ev_end = nullptr, ev_f0 = nullptr, ev_f1 = nullptr;
    char err[256] = {0};

    // batch staging for host-pointer submits
    DevBuf d_batch;
    // entry arrays (capacity in entries = ecap)
    DevBuf keyA, keyB, valA, valB, surv_src, surv_meta, pflag, entry_out, wsorted, windows, owner_table;
    DevBuf sp_key, sp_src, sp_meta, sp_ns, tile_count;  // filter's tile-local compaction
    DevBuf surv_pid, surv_ns, slot_off;                  // arrival-ordered pid copy; transport-form bookkeeping
    const uint4 *b_slots = nullptr;                      // transport-form batch (chronos_submit_survivors)
    DevBuf comm_memo;                                    // second-level comm memo: lock[N] | val[N] | key[N]
    // kill-chain labeller + persistent per-PID table (chain.cuh)
    DevBuf pid_tab, pt_count, chain_pid, chain_idx;
    uint32_t pt_cap = 0;
    uint32_t pt_epoch = 1;
    Plan *h_plan = nullptr;      // pinned copy of the batch's device Plan, read at fetch (behind h_scalars)
    uint64_t b_Mmax = 0;         // P + n: what the buffers of the batch in flight are sized for
    uint64_t n_pids_seen = 0, b_chains = 0;
    DevBuf surv_arrival;  // only for chronos_submit_survivors
    // double-buffered persistent state
    DevBuf arena[2];
    DevBuf pend_key[2], pend_slot[2], pend_arrival[2], pend_meta[2];
    uint64_t pend_cap[2] = {0, 0};  // capacity (entries) the right-aligned arrays were laid out for
    int cur = 0;                    // index written by the batch in flight / last fetched
    uint64_t n_pending = 0;         // entries pending after the last fetched batch (in buffer `cur`)
    // scheduling scratch
    DevBuf status, tickets, scalars, scan_status, sweep_status;
    uint32_t scan_epoch = 0, sweep_epoch = 0;
    int scan_slot = 0;  // next free ticket counter for scans in this batch
    DeviceScalars *h_scalars = nullptr;  // pinned
    DevBuf pack;                         // scratch for chronos_copy_window_records / chronos_state
    DevBuf pack_off;

    // direct exchange over peer memory (exchange.cuh)
    DevBuf x_window, x_totals;
    XchgLayout x_L = {};
    bool x_created = false, x_connected = false, x_sent = false, b_exchanged = false;
    uint8_t *x_peer[XCHG_MAX_WORLD] = {nullptr};
    bool x_opened[XCHG_MAX_WORLD] = {false};
    uint32_t x_epoch = 0;
    uint64_t x_launches = 0;
    bool x_filter_timed = false;  // stage 1 of this step ran the filter (non-empty slice)
    uint64_t x_timeout_ns = 20ull * 1000000000ull;
    cudaEvent_t ev_x0 = nullptr;
    // asynchronous result download (chronos_download_begin)
    cudaStream_t dl_stream = nullptr;
    cudaEvent_t dl_ready = nullptr, dl_done[2] = {nullptr, nullptr};
    DevBuf dl_stage[2];
    bool dl_inflight[2] = {false, false};
    int dl_k = 0;

    // last batch
    enum { IDLE, SUBMITTED, FETCHED } phase = IDLE;
    uint64_t b_n = 0, b_base = 0, b_S = 0, b_P = 0, b_M = 0, b_W = 0;
    const uint8_t *b_records = nullptr;  // device pointer of the batch records
    bool b_explicit_arrival = false;
    uint64_t launches = 0;
    float t_total = 0.f, t_filter = 0.f, t_filter_kernel = 0.f;
    bool filter_timed = false;
    int filter_variant = 0;
    // CHRONOS_TRACE=1: CUDA events between the stages of a submit, printed at fetch
    bool trace = false;
    cudaEvent_t tev[12] = {nullptr};
    const char *tname[12] = {nullptr};
    int n_tev = 0;
};

namespace {

const char *g_static_err = "chronos: invalid context";

// CHRONOS_DEBUG_SYNC=1: wait after every stage so that a device fault names the stage that raised it
int dbg_sync(chronos_ctx *c, const char *what);

void mark(chronos_ctx *c, const char *name) {
    if (!c->trace || c->n_tev >= 12) return;
    if (!c->tev[c->n_tev]) cudaEventCreate(&c->tev[c->n_tev]);
    cudaEventRecord(c->tev[c->n_tev], c->stream);
    c->tname[c->n_tev++] = name;
}

int fail(chronos_ctx *c, int code, const char *fmt, ...) {
    if (c) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(c->err, sizeof(c->err), fmt, ap);
        va_end(ap);
    }
    return code;
}

#define CK(call)                                                                                      \
    do {                                                                                              \
        cudaError_t e__ = (call);                                                                     \
        if (e__ != cudaSuccess)                                                                       \
            return fail(c, e__ == cudaErrorMemoryAllocation ? CHRONOS_E_NOMEM : CHRONOS_E_CUDA,       \
                        "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

int dbg_sync(chronos_ctx *c, const char *what) {
    static const bool on = getenv("CHRONOS_DEBUG_SYNC") != nullptr;
    if (!on) return CHRONOS_OK;
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream2);
    if (e != cudaSuccess) return fail(c, CHRONOS_E_CUDA, "after %s: %s", what, cudaGetErrorString(e));
    return CHRONOS_OK;
}

int ensure(chronos_ctx *c, DevBuf &b, size_t bytes) {
    if (bytes <= b.bytes && b.p) return CHRONOS_OK;
    if (b.p) {
        CK(cudaFree(b.p));
        b.p = nullptr;
        b.bytes = 0;
    }
    size_t want = std::max<size_t>(bytes + bytes / 8, 256);
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess) {
        cudaGetLastError();
        want = std::max<size_t>(bytes, 256);
        e = cudaMalloc(&b.p, want);
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        b.p = nullptr;
        return fail(c, CHRONOS_E_NOMEM, "cudaMalloc(%zu bytes) failed: %s", want, cudaGetErrorString(e));
    }
    b.bytes = want;
    return CHRONOS_OK;
}

void release(DevBuf &b) {
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.bytes = 0;
}

template <typename T>
T *ptr(DevBuf &b) {
    return reinterpret_cast<T *>(b.p);
}

bool is_device_pointer(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// ---- small helper kernels ---------------------------------------------------------------
__global__ void fold_keys_kernel(const uint32_t *keys, uint32_t n, uint32_t *k_or, uint32_t *k_and) {
    uint32_t o = 0, a = 0xffffffffu;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        uint32_t k = keys[i];
        o |= k;
        a &= k;
    }
    o = __reduce_or_sync(0xffffffffu, o);
    a = __reduce_and_sync(0xffffffffu, a);
    if ((threadIdx.x & 31) == 0) {
        atomicOr(k_or, o);
        atomicAnd(k_and, a);
    }
}

__global__ void set_u32_kernel(uint32_t *p, uint32_t v) { *p = v; }

// start of a batch: scalars to their neutral values, ticket counters and digit histograms to 0
__global__ void init_scalars_kernel(DeviceScalars *s, uint32_t *tickets) {
    for (uint32_t i = threadIdx.x; i < TICKET_BYTES / 4; i += blockDim.x) tickets[i] = 0;
    if (threadIdx.x) return;
    s->surv_count = 0;
    s->key_or = 0;
    s->key_and = 0xffffffffu;
    s->n_windows = 0;
    s->n_pending = 0;
    s->n_chains = 0;
}

// Stage 2 of the sharded path, step 1: split the header words of the received entries
// (arrival 48 bits | meta << 48 | slots << 56)
__global__ void unpack_hdr_kernel(const uint64_t *hdr, uint32_t n, uint64_t *arrival, uint8_t *meta, uint8_t *ns,
                                  uint32_t *ns32) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint64_t h = hdr[i];
        arrival[i] = h & 0xFFFFFFFFFFFFull;
        meta[i] = (uint8_t)(h >> 48);
        const uint32_t k = (uint32_t)(h >> 56);
        ns[i] = (uint8_t)k;
        ns32[i] = k;
    }
}

// step 2 (after the scan of the slot counts): the pid is the first word of an entry's slot 0
__global__ void extract_pid_kernel(const uint4 *slots, const uint32_t *slot_off, uint32_t n, uint32_t *key,
                                   uint32_t *pid_copy, uint32_t *src, uint32_t *k_or, uint32_t *k_and) {
    uint32_t o = 0, a = 0xffffffffu;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t k = reinterpret_cast<const uint32_t *>(slots + slot_off[i])[0];
        key[i] = k;
        pid_copy[i] = k;
        src[i] = i;
        o |= k;
        a &= k;
    }
    o = __reduce_or_sync(0xffffffffu, o);
    a = __reduce_and_sync(0xffffffffu, a);
    if ((threadIdx.x & 31) == 0) {
        atomicOr(k_or, o);
        atomicAnd(k_and, a);
    }
}

// one warp per packed record: dst[i] = arena[slot(i)]
__global__ void pack_by_slot_kernel(const uint8_t *arena, const uint32_t *slot, uint32_t n, uint8_t *dst) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const uint32_t n_warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = warp; i < n; i += n_warps) {
        const uint4 *s = reinterpret_cast<const uint4 *>(arena + (size_t)slot[i] * CH_REC_BYTES);
        uint4 *d = reinterpret_cast<uint4 *>(dst + (size_t)i * CH_REC_BYTES);
        if (lane < CH_REC_CHUNKS) d[lane] = s[lane];
    }
}

// one warp per window: packed[off[w] ..] = arena[start .. start+len)
__global__ void window_len_kernel(const chronos_window *win, uint32_t nw, uint32_t *len) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nw; i += gridDim.x * blockDim.x) len[i] = win[i].len;
}

// records of every window, window after window (off = exclusive scan of the window lengths)
__global__ void pack_windows_kernel(const uint8_t *arena, const chronos_window *win, const uint32_t *off, uint32_t nw,
                                    uint8_t *dst) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const uint32_t n_warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t w = warp; w < nw; w += n_warps) {
        const uint4 *s = reinterpret_cast<const uint4 *>(arena + (size_t)win[w].start * CH_REC_BYTES);
        uint4 *d = reinterpret_cast<uint4 *>(dst + (size_t)off[w] * CH_REC_BYTES);
        const uint32_t chunks = win[w].len * CH_REC_CHUNKS;
        for (uint32_t ch = lane; ch < chunks; ch += 32) d[ch] = s[ch];
    }
}

__global__ void add_base_kernel(const uint32_t *src, uint32_t n, uint64_t base, uint64_t *out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = base + src[i];
}

__global__ void l2_flush_kernel(uint4 *p, size_t n16, uint32_t v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
        p[i] = make_uint4(v, v, v, v);
}

__global__ void tracegen_kernel(chronos_tracegen_cfg cfg, const uint64_t *cdf, uint64_t first, uint64_t n,
                                uint8_t *dst) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint8_t rec[CH_REC_BYTES];
        chronos_gen::gen_record(cfg, first + i, cdf, rec);
        uint32_t *d = reinterpret_cast<uint32_t *>(dst + i * CH_REC_BYTES);
        const uint32_t *s = reinterpret_cast<const uint32_t *>(rec);
        for (int k = 0; k < CH_REC_WORDS; k++) d[k] = s[k];
    }
}

// ---- filter launch ------------------------------------------------------------------------
template <int NSTAGE>
int launch_filter_t(chronos_ctx *c, FilterParams fp) {
    static bool configured[16] = {false};
    auto kern = filter_kernel<NSTAGE>;
    const size_t smem = filter_smem_bytes<NSTAGE>();
    if (!configured[c->device & 15]) {
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[c->device & 15] = true;
    }
    int per_sm = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, FILTER_TILE, smem));
    if (per_sm < 1) return fail(c, CHRONOS_E_CUDA, "filter kernel does not fit on an SM (smem %zu)", smem);
    const uint32_t tiles = fp.tile_end - fp.tile_begin;
    const uint32_t grid = std::min<uint32_t>(tiles, (uint32_t)(per_sm * c->sm_count));
    kern<<<grid, FILTER_TILE, smem, c->stream>>>(fp);
    CK(cudaGetLastError());
    c->launches++;
    return CHRONOS_OK;
}

int filter_tile_size(const chronos_ctx *) { return FILTER_TILE; }

// CHRONOS_FILTER_VARIANT=2 selects the double-buffered build of the filter (tuning knob)
int launch_filter(chronos_ctx *c, const FilterParams &fp) {
    if (c->filter_variant == 2) return launch_filter_t<2>(c, fp);
    return launch_filter_t<1>(c, fp);
}

// exclusive scan in place (device array of n u32) on the context's stream
// (n_dev != nullptr: the real count lives on the device, n is only its upper bound)
int launch_scan(chronos_ctx *c, uint32_t *data, uint64_t n, const uint32_t *n_dev = nullptr, int own_ticket = -1) {
    if (n == 0) return CHRONOS_OK;
    cudaStream_t stream = c->stream;
    const uint32_t n_tiles = (uint32_t)((n + SCAN_TILE - 1) / SCAN_TILE);
    int rc;
    const void *before = c->scan_status.p;
    if ((rc = ensure(c, c->scan_status, std::max<size_t>((size_t)n_tiles * 8, 4096)))) return rc;
    if (c->scan_status.p != before || c->scan_epoch == 0)  // fresh memory or wrapped epoch: tag 0 = "never"
        CK(cudaMemsetAsync(c->scan_status.p, 0, c->scan_status.bytes, stream));
    if (own_ticket < 0 && c->scan_slot >= 32) return fail(c, CHRONOS_E_STATE, "too many scans in one batch");
    if (own_ticket >= 0) CK(cudaMemsetAsync(ptr<uint32_t>(c->tickets) + own_ticket, 0, 4, stream));
    ScanParams sp;
    sp.data = data;
    sp.n = (uint32_t)n;
    sp.n_ptr = n_dev;
    sp.ticket = own_ticket >= 0 ? ptr<uint32_t>(c->tickets) + own_ticket : ptr<uint32_t>(c->tickets) + 8 + c->scan_slot++;
    sp.status = ptr<uint64_t>(c->scan_status);
    sp.epoch = ++c->scan_epoch;
    if (c->scan_epoch >= (1u << 29)) c->scan_epoch = 0;  // re-zero on the next call
    scan_exclusive_u32<<<std::min<uint32_t>(n_tiles, (uint32_t)c->sm_count * 2), SCAN_THREADS, 0, stream>>>(sp);
    CK(cudaGetLastError());
    c->launches++;
    return CHRONOS_OK;
}

// filter + cross-tile compaction: dense, arrival-ordered survivors (pid -> out_key,
// batch-local index -> surv_src, META bits -> surv_meta, count -> scalars.surv_count)
// Buffers and parameters of the filter for a batch of n records (no launch).
int filter_setup(chronos_ctx *c, const uint8_t *d_recs, uint64_t n, FilterParams &fp) {
    int rc;
    DeviceScalars *ds = ptr<DeviceScalars>(c->scalars);
    const int TILE = filter_tile_size(c);
    fp.recs = d_recs;
    fp.n = n;
    fp.tile_begin = 0;
    fp.tile_end = (uint32_t)((n + TILE - 1) / TILE);
    fp.ticket = ptr<uint32_t>(c->tickets);
    const uint64_t slots = (uint64_t)fp.tile_end * TILE;
    if ((rc = ensure(c, c->sp_key, slots * 4))) return rc;
    if ((rc = ensure(c, c->sp_src, slots * 4))) return rc;
    if ((rc = ensure(c, c->sp_meta, slots))) return rc;
    if ((rc = ensure(c, c->sp_ns, slots))) return rc;
    if ((rc = ensure(c, c->surv_pid, (n + 1) * 4))) return rc;
    if ((rc = ensure(c, c->surv_ns, n + 1))) return rc;
    if ((rc = ensure(c, c->tile_count, ((uint64_t)fp.tile_end + 4) * 4))) return rc;
    fp.tile_count = ptr<uint32_t>(c->tile_count);
    fp.sp_key = ptr<uint32_t>(c->sp_key);
    fp.sp_src = ptr<uint32_t>(c->sp_src);
    fp.sp_meta = ptr<uint8_t>(c->sp_meta);
    fp.sp_ns = ptr<uint8_t>(c->sp_ns);
    fp.out_count = &ds->surv_count;
    fp.key_or = &ds->key_or;
    fp.key_and = &ds->key_and;
    fp.gm_lock = ptr<uint32_t>(c->comm_memo);
    fp.gm_val = fp.gm_lock + COMM_MEMO_SLOTS;
    fp.gm_key = reinterpret_cast<uint4 *>(fp.gm_val + COMM_MEMO_SLOTS);
    fp.gm_mask = COMM_MEMO_SLOTS - 1;
    return CHRONOS_OK;
}

// Cross-tile prefix + dense, arrival-ordered survivor arrays, one kernel.
int launch_compact(chronos_ctx *c, const FilterParams &fp, uint32_t *out_key) {
    int rc;
    cudaStream_t stream = c->stream;
    const uint32_t n_tiles = fp.tile_end;
    const uint32_t n_chunks = (n_tiles + CS_CHUNK - 1) / CS_CHUNK;
    const void *before = c->scan_status.p;
    if ((rc = ensure(c, c->scan_status, std::max<size_t>((size_t)n_chunks * 8, 4096)))) return rc;
    if (c->scan_status.p != before || c->scan_epoch == 0)
        CK(cudaMemsetAsync(c->scan_status.p, 0, c->scan_status.bytes, stream));
    if (c->scan_slot >= 32) return fail(c, CHRONOS_E_STATE, "too many scans in one batch");
    CompactScanParams cp;
    cp.tile_count = fp.tile_count;
    cp.n_tiles = n_tiles;
    cp.tile = (uint32_t)filter_tile_size(c);
    cp.sp_key = fp.sp_key;
    cp.sp_src = fp.sp_src;
    cp.sp_meta = fp.sp_meta;
    cp.sp_ns = fp.sp_ns;
    cp.out_key = out_key;
    cp.out_src = ptr<uint32_t>(c->surv_src);
    cp.out_meta = ptr<uint8_t>(c->surv_meta);
    cp.out_ns = ptr<uint8_t>(c->surv_ns);
    cp.out_pid = ptr<uint32_t>(c->surv_pid);
    cp.ticket = ptr<uint32_t>(c->tickets) + 8 + c->scan_slot++;
    cp.status = ptr<uint64_t>(c->scan_status);
    cp.epoch = ++c->scan_epoch;
    if (c->scan_epoch >= (1u << 29)) c->scan_epoch = 0;
    compact_scan<<<std::max(1u, std::min<uint32_t>(n_chunks, (uint32_t)c->sm_count * 4)), CS_THREADS, 0, stream>>>(cp);
    CK(cudaGetLastError());
    c->launches += 1;
    return CHRONOS_OK;
}

// filter + cross-tile compaction: dense, arrival-ordered survivors (pid -> out_key,
// batch-local index -> surv_src, META bits -> surv_meta, count -> scalars.surv_count)
int run_filter_stage(chronos_ctx *c, const uint8_t *d_recs, uint64_t n, uint32_t *out_key) {
    int rc;
    FilterParams fp;
    if ((rc = filter_setup(c, d_recs, n, fp))) return rc;
    CK(cudaEventRecord(c->ev_f0, c->stream));
    if ((rc = dbg_sync(c, "begin_batch"))) return rc;
    if ((rc = launch_filter(c, fp))) return rc;
    CK(cudaEventRecord(c->ev_f1, c->stream));
    if ((rc = dbg_sync(c, "filter_kernel"))) return rc;
    c->filter_timed = true;
    if ((rc = launch_compact(c, fp, out_key))) return rc;
    if ((rc = dbg_sync(c, "filter + scan + compact"))) return rc;
    return CHRONOS_OK;
}

int finish_batch(chronos_ctx *c);

// The per-PID table must stay at most half full: room for every PID seen so far plus one new
// PID per grouped entry of this batch.
int ensure_pid_table(chronos_ctx *c, uint64_t M) {
    const uint64_t need = 2 * (c->n_pids_seen + M) + 64;
    if (c->pt_cap >= need) return CHRONOS_OK;
    uint64_t cap = 1024;
    while (cap < need) cap <<= 1;
    if (cap > (1ull << 31)) return fail(c, CHRONOS_E_CAPACITY, "per-PID table would need %llu slots", (unsigned long long)cap);
    DevBuf fresh;
    int rc;
    if ((rc = ensure(c, fresh, cap * 8))) return rc;
    auto grow = [&]() -> int {
        CK(cudaMemsetAsync(fresh.p, 0, cap * 8, c->stream));
        if (c->pt_cap && c->n_pids_seen) {
            pt_rehash_kernel<<<(unsigned)std::min<uint64_t>((c->pt_cap + 255) / 256, 4096), 256, 0, c->stream>>>(
                ptr<uint64_t>(c->pid_tab), c->pt_cap, static_cast<uint64_t *>(fresh.p), (uint32_t)cap - 1, c->pt_epoch);
            CK(cudaGetLastError());
            c->launches++;
        }
        CK(cudaStreamSynchronize(c->stream));
        return CHRONOS_OK;
    };
    if ((rc = grow())) {
        release(fresh);  // the old table stays in place
        return rc;
    }
    release(c->pid_tab);
    c->pid_tab = fresh;
    c->pt_cap = (uint32_t)cap;
    return CHRONOS_OK;
}

// Everything after the survivors are known ON THE DEVICE (keyA[P..P+S), surv_src, surv_meta,
// scalars.surv_count / key_or / key_and). Nothing here waits for the GPU: buffers are sized
// for the worst case M = P + n, grids are upper bounds, extents come from the device plan.
int run_window_stage(chronos_ctx *c) {
    const uint64_t P = c->b_P;
    const uint64_t Mmax = c->b_Mmax;  // P + n
    DeviceScalars *ds = ptr<DeviceScalars>(c->scalars);
    Plan *plan = reinterpret_cast<Plan *>(ptr<uint8_t>(c->scalars) + 64);
    mark(c, "filter+compact done");
    CK(cudaEventRecord(c->ev_filter, c->stream));
    const int cur = c->cur, prev = cur ^ 1;
    const uint64_t Smax = Mmax - P;
    if (Mmax >= (1ull << VAL_ID_BITS))
        return fail(c, CHRONOS_E_CAPACITY, "more than 2^27 grouped entries possible (%llu)", (unsigned long long)Mmax);
    int rc;
    if ((rc = ensure(c, c->arena[cur], (Mmax + 1) * CH_REC_BYTES))) return rc;
    if ((rc = ensure(c, c->pend_key[cur], (Mmax + 1) * 4))) return rc;
    if ((rc = ensure(c, c->pend_slot[cur], (Mmax + 1) * 4))) return rc;
    if ((rc = ensure(c, c->pend_arrival[cur], (Mmax + 1) * 8))) return rc;
    if ((rc = ensure(c, c->pend_meta[cur], Mmax + 1))) return rc;
    c->pend_cap[cur] = Mmax;
    if ((rc = ensure(c, c->pflag, (Mmax + 4) * 4))) return rc;
    if ((rc = ensure(c, c->entry_out, (Smax + 4) * 4))) return rc;
    if ((rc = ensure(c, c->wsorted, (Smax / 2 + 1) * sizeof(WinSorted)))) return rc;
    if ((rc = ensure(c, c->windows, (Smax / 2 + 1) * sizeof(chronos_window)))) return rc;
    if ((rc = ensure(c, c->chain_pid, (Smax + 1) * 4))) return rc;
    if ((rc = ensure(c, c->chain_idx, (Smax + 1) * 8))) return rc;
    if ((rc = ensure_pid_table(c, Mmax))) return rc;
    const uint32_t max_rtiles = (uint32_t)((Mmax + RADIX_TILE_KEYS - 1) / RADIX_TILE_KEYS);

    // ---- stable sort by pid: plan + global digit histograms, then one kernel per pass --------
    SortParams sp;
    memset(&sp, 0, sizeof(sp));
    sp.plan = plan;
    sp.in.surv_count = &ds->surv_count;
    sp.in.key_or = &ds->key_or;
    sp.in.key_and = &ds->key_and;
    sp.in.P = (uint32_t)P;
    sp.in.m_max = (uint32_t)Mmax;
    sp.in.win_tile = WIN_TILE;
    sp.keyA = ptr<uint32_t>(c->keyA);
    sp.keyB = ptr<uint32_t>(c->keyB);
    sp.valA = ptr<uint32_t>(c->valA);
    sp.valB = ptr<uint32_t>(c->valB);
    sp.meta_lo = P ? ptr<uint8_t>(c->pend_meta[prev]) + (c->pend_cap[prev] - P) : nullptr;
    sp.meta_hi = ptr<uint8_t>(c->surv_meta);
    sp.ghist = ptr<uint32_t>(c->tickets) + N_TICKETS;  // zeroed with the tickets
    {
        const void *before = c->sweep_status.p;
        const size_t need = std::max<size_t>((size_t)max_rtiles, 1) * RADIX_BINS * 8;
        if ((rc = ensure(c, c->sweep_status, need))) return rc;
        if (c->sweep_status.p != before || c->sweep_epoch == 0)  // fresh memory or wrapped epoch: tag 0 = "never"
            CK(cudaMemsetAsync(c->sweep_status.p, 0, c->sweep_status.bytes, c->stream));
    }
    sp.status = ptr<uint64_t>(c->sweep_status);
    const unsigned sort_blocks = std::max(1u, std::min(max_rtiles, (uint32_t)c->sm_count * 4));
    sort_prepare<<<sort_blocks, RADIX_THREADS, 0, c->stream>>>(sp);
    c->launches++;
    if ((rc = dbg_sync(c, "sort_prepare"))) return rc;
    if (Mmax == 0) {  // nothing pending, nothing submitted: only the (empty) counts go back
        CK(cudaMemcpyAsync(c->h_scalars, ds, RESULT_BLOCK_BYTES, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaEventRecord(c->ev_end, c->stream));
        return CHRONOS_OK;
    }
    {
        for (int i = 0; i < RADIX_MAX_PASSES; i++) {  // passes beyond plan->n_passes return at once
            sp.pass = i;
            sp.ticket = ptr<uint32_t>(c->tickets) + 40 + i;
            sp.epoch = ++c->sweep_epoch;
            if (c->sweep_epoch >= (1u << 29)) c->sweep_epoch = 0;  // re-zero on the next batch
            onesweep_pass<<<sort_blocks, RADIX_THREADS, 0, c->stream>>>(sp);
            c->launches++;
            if ((rc = dbg_sync(c, "onesweep_pass"))) return rc;
        }
        CK(cudaGetLastError());
    }
    mark(c, "sort done");

    // ---- records into the grouped arena ----------------------------------------------
    {
        GatherParams g;
        memset(&g, 0, sizeof(g));
        g.plan = plan;
        g.surv_src = ptr<uint32_t>(c->surv_src);
        g.batch = c->b_records;
        const uint64_t pc = c->pend_cap[prev];
        g.pend_slot = P ? ptr<uint32_t>(c->pend_slot[prev]) + (pc - P) : nullptr;
        g.prev_arena = ptr<uint8_t>(c->arena[prev]);
        g.arena = ptr<uint8_t>(c->arena[cur]);
        g.slots = c->b_slots;
        g.slot_off = c->b_slots ? ptr<uint32_t>(c->slot_off) : nullptr;
        g.slot_ns = c->b_slots ? ptr<uint8_t>(c->surv_ns) : nullptr;
        // bandwidth-bound copy on a second stream; it runs while the host enqueues the window scans
        uint32_t blocks = (uint32_t)std::min<uint64_t>((Mmax + 127) / 128, (uint64_t)c->sm_count * 8);
        CK(cudaEventRecord(c->ev_sorted, c->stream));
        CK(cudaStreamWaitEvent(c->stream2, c->ev_sorted, 0));
        gather_records<<<blocks, 256, 0, c->stream2>>>(g);
        CK(cudaGetLastError());
        CK(cudaEventRecord(c->ev_gathered, c->stream2));
        c->launches++;
        if ((rc = dbg_sync(c, "gather_records"))) return rc;
    }

    // ---- window automaton + chain labeller -------------------------------------------------
    {
        WinParams w;
        memset(&w, 0, sizeof(w));
        w.plan = plan;
        w.surv_src = ptr<uint32_t>(c->surv_src);
        w.surv_arrival = c->b_explicit_arrival ? ptr<uint64_t>(c->surv_arrival) : nullptr;
        w.base_arrival = c->b_base;
        const uint64_t pc = c->pend_cap[prev];
        w.pend_arrival = P ? ptr<uint64_t>(c->pend_arrival[prev]) + (pc - P) : nullptr;
        w.surv_pid = ptr<uint32_t>(c->surv_pid);
        w.pid_tab = ptr<uint64_t>(c->pid_tab);
        w.pid_mask = c->pt_cap - 1;
        w.pid_epoch = c->pt_epoch;
        w.pid_count = ptr<uint32_t>(c->pt_count);
        w.n_pids_out = &ds->n_pids_seen;
        w.pflag = ptr<uint32_t>(c->pflag);
        w.entry_out = ptr<uint32_t>(c->entry_out);
        w.wsorted = ptr<WinSorted>(c->wsorted);
        w.n_windows = &ds->n_windows;
        w.pend_cap = (uint32_t)Mmax;
        w.npend_key = ptr<uint32_t>(c->pend_key[cur]);
        w.npend_slot = ptr<uint32_t>(c->pend_slot[cur]);
        w.npend_arrival = ptr<uint64_t>(c->pend_arrival[cur]);
        w.npend_meta = ptr<uint8_t>(c->pend_meta[cur]);
        w.n_pending = &ds->n_pending;
        w.windows = ptr<chronos_window>(c->windows);
        w.chain_pid = ptr<uint32_t>(c->chain_pid);
        w.chain_idx = ptr<uint64_t>(c->chain_idx);
        w.n_chains = &ds->n_chains;
        const uint32_t tiles_m = (uint32_t)((Mmax + WIN_TILE - 1) / WIN_TILE);  // upper bounds
        const uint32_t tiles_s = (uint32_t)((Smax + WIN_TILE - 1) / WIN_TILE);
        uint64_t *st = ptr<uint64_t>(c->status);
        uint32_t *tk = ptr<uint32_t>(c->tickets);
        const uint32_t cap_blocks = (uint32_t)c->sm_count * (WIN_ITEMS >= 16 ? 2 : 4);
        w.ticket = tk + 60;
        w.status = st;                              // 5 words per tile
        group_forward<<<std::max(1u, std::min(tiles_m, cap_blocks)), WIN_THREADS, 0, c->stream>>>(w);
        if ((rc = dbg_sync(c, "group_forward"))) return rc;
        w.ticket = tk + 61;
        w.status = st + 5 * (size_t)tiles_m;        // 2 words per tile
        window_reverse<<<std::max(1u, std::min(tiles_m, cap_blocks)), WIN_THREADS, 0, c->stream>>>(w);
        if ((rc = dbg_sync(c, "window_reverse"))) return rc;
        w.ticket = tk + 62;
        w.status = st + 7 * (size_t)tiles_m;        // 1 word per tile
        emit_ordered<<<std::max(1u, std::min(tiles_s, cap_blocks)), WIN_THREADS, 0, c->stream>>>(w);
        c->launches += 3;
        CK(cudaGetLastError());
        if ((rc = dbg_sync(c, "emit_ordered"))) return rc;
    }
    mark(c, "window kernels done");
    CK(cudaStreamWaitEvent(c->stream, c->ev_gathered, 0));
    mark(c, "gather joined");
    CK(cudaMemcpyAsync(c->h_scalars, ds, RESULT_BLOCK_BYTES, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaEventRecord(c->ev_end, c->stream));
    return CHRONOS_OK;
}

// status words needed for a batch of n raw records producing at most n survivors on top of P pending
size_t status_words(uint64_t n, uint64_t P) {
    const uint64_t M = n + P;
    const uint64_t tiles_m = (M + WIN_TILE - 1) / WIN_TILE;
    return (size_t)(8 * tiles_m + 64);
}

int begin_batch(chronos_ctx *c, uint64_t n, uint64_t base) {
    CK(cudaSetDevice(c->device));
    if (c->phase == chronos_ctx::SUBMITTED) {
        int rc = finish_batch(c);  // results dropped, state kept
        if (rc) return rc;
    }
    if (n > c->max_batch) return fail(c, CHRONOS_E_CAPACITY, "batch of %llu events > max_batch_events %llu",
                                      (unsigned long long)n, (unsigned long long)c->max_batch);
    if (n >= (1ull << 32)) return fail(c, CHRONOS_E_CAPACITY, "batch too large");
    // the batch in flight writes buffer set cur^1; the previous state stays in cur until fetch
    const uint64_t P = c->n_pending;
    c->cur ^= 1;
    c->b_n = n;
    c->b_base = base;
    c->b_P = P;
    c->b_S = c->b_M = c->b_W = 0;
    c->launches = 0;
    c->scan_slot = 0;
    c->filter_timed = false;
    const uint64_t ecap = P + n;
    c->b_Mmax = ecap;
    int rc;
    if ((rc = ensure(c, c->keyA, (ecap + 1) * 4))) return rc;
    if ((rc = ensure(c, c->keyB, (ecap + 1) * 4))) return rc;
    if ((rc = ensure(c, c->valA, (ecap + 1) * 4))) return rc;
    if ((rc = ensure(c, c->valB, (ecap + 1) * 4))) return rc;
    if ((rc = ensure(c, c->surv_src, (n + 1) * 4))) return rc;
    if ((rc = ensure(c, c->surv_meta, n + 1))) return rc;
    const size_t sw = status_words(n, P);
    if ((rc = ensure(c, c->status, sw * 8))) return rc;
    CK(cudaEventRecord(c->ev_start, c->stream));
    if (!c->x_sent) c->n_tev = 0;  // (stage 2 of an exchange step keeps the marks of its stage 1)
    mark(c, "start");
    CK(cudaMemsetAsync(c->status.p, 0, sw * 8, c->stream));
    init_scalars_kernel<<<1, 256, 0, c->stream>>>(ptr<DeviceScalars>(c->scalars), ptr<uint32_t>(c->tickets));
    c->launches++;
    if (P) {
        const int prev = c->cur ^ 1;
        const uint64_t pc = c->pend_cap[prev];
        const uint32_t *pk = ptr<uint32_t>(c->pend_key[prev]) + (pc - P);
        CK(cudaMemcpyAsync(c->keyA.p, pk, P * 4, cudaMemcpyDeviceToDevice, c->stream));
        DeviceScalars *ds = ptr<DeviceScalars>(c->scalars);
        fold_keys_kernel<<<(unsigned)std::min<uint64_t>((P + 255) / 256, 1024), 256, 0, c->stream>>>(
            pk, (uint32_t)P, &ds->key_or, &ds->key_and);
        c->launches++;
    }
    CK(cudaGetLastError());
    return CHRONOS_OK;
}

int finish_batch(chronos_ctx *c) {
    CK(cudaStreamSynchronize(c->stream));
    c->b_S = c->h_plan->S;
    c->b_M = c->h_plan->M;
    c->b_W = c->b_M ? c->h_scalars->n_windows : 0;
    c->n_pending = c->b_M ? c->h_scalars->n_pending : 0;
    c->b_chains = c->b_M ? c->h_scalars->n_chains : 0;
    if (c->h_plan->overflow) return fail(c, CHRONOS_E_CAPACITY, "more survivors than the batch was sized for");
    c->n_pids_seen = c->b_M ? c->h_scalars->n_pids_seen : c->n_pids_seen;
    if (c->b_exchanged) {
        c->b_n = c->b_S;  // the upper bound the buffers were sized for is of no interest to the caller
        const uint32_t xe = c->h_scalars->xchg_err;
        c->phase = chronos_ctx::FETCHED;
        if (xe & XERR_TIMEOUT)
            return fail(c, CHRONOS_E_STATE, "exchange: a peer rank did not deliver its survivors within %.0f s",
                        (double)c->x_timeout_ns * 1e-9);
        if (xe & (XERR_OVERFLOW | XERR_PEER_OVERFLOW))
            return fail(c, CHRONOS_E_CAPACITY, "exchange: survivors for one owner rank exceed the window (%u entries / %u slots per peer)",
                        c->x_L.cap_e, c->x_L.cap_s);
    }
    cudaEventElapsedTime(&c->t_total, c->b_exchanged ? c->ev_x0 : c->ev_start, c->ev_end);
    if (c->trace) {
        for (int i = 1; i < c->n_tev; i++) {
            float ms = 0;
            cudaEventElapsedTime(&ms, c->tev[i - 1], c->tev[i]);
            fprintf(stderr, "[chronos trace] %-28s %8.3f ms\n", c->tname[i], ms);
        }
    }
    cudaEventElapsedTime(&c->t_filter, c->ev_start, c->ev_filter);
    // (a stage-2-only batch keeps the filter time measured by chronos_filter_partition)
    if (c->filter_timed || (c->b_exchanged && c->x_filter_timed)) cudaEventElapsedTime(&c->t_filter_kernel, c->ev_f0, c->ev_f1);
    cudaGetLastError();
    c->phase = chronos_ctx::FETCHED;
    return CHRONOS_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Fixed scratch for chronos_filter_noalloc (the one entry point that never allocates).
// It lives in the library image: every GPU that loads the library holds these ~2.3 MiB
// until the process exits. In HOST code the symbol below is only a stand-in, not a device
// address; the real address is fetched once per device with cudaGetSymbolAddress.
// ---------------------------------------------------------------------------------------------
namespace noalloc {
constexpr uint64_t MAX_EVENTS = CHRONOS_NOALLOC_MAX_EVENTS;
constexpr uint64_t MAX_TILES = MAX_EVENTS / FILTER_TILE;
constexpr size_t OFF_HEADER = 0;        // u32[64]: [0] filter ticket [1] scan ticket [2] survivors [3] pid OR [4] pid AND
constexpr size_t OFF_STATUS = 256;      // look-back words of the tile-count scan
constexpr size_t STATUS_BYTES = ((MAX_TILES + SCAN_TILE - 1) / SCAN_TILE * 8 + 255) / 256 * 256;
constexpr size_t OFF_TILE_COUNT = OFF_STATUS + STATUS_BYTES;
constexpr size_t TILE_COUNT_BYTES = ((MAX_TILES + 1) * 4 + 255) / 256 * 256;
constexpr size_t OFF_SP_KEY = OFF_TILE_COUNT + TILE_COUNT_BYTES;
constexpr size_t OFF_SP_SRC = OFF_SP_KEY + MAX_EVENTS * 4;
constexpr size_t OFF_SP_META = OFF_SP_SRC + MAX_EVENTS * 4;
constexpr size_t TOTAL_BYTES = OFF_SP_META + MAX_EVENTS;
constexpr size_t CLEAR_BYTES = OFF_TILE_COUNT;  // tickets, counters and scan status must start at zero
}  // namespace noalloc

static __device__ __align__(256) unsigned char g_noalloc_scratch[noalloc::TOTAL_BYTES];

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

int chronos_device_ok(int device) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) {
        cudaGetLastError();
        return CHRONOS_E_CUDA;
    }
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, device) != cudaSuccess) return CHRONOS_E_CUDA;
    return p.major == 10 ? CHRONOS_OK : CHRONOS_E_CUDA;
}

const char *chronos_last_error(const chronos_ctx *ctx) { return ctx ? ctx->err : g_static_err; }

int chronos_create(int device, uint64_t max_batch_events, chronos_ctx **out) {
    if (!out || max_batch_events == 0 || max_batch_events >= (1ull << 32)) return CHRONOS_E_ARG;
    *out = nullptr;
    if (chronos_device_ok(device) != CHRONOS_OK) return CHRONOS_E_CUDA;
    chronos_ctx *c = new (std::nothrow) chronos_ctx();
    if (!c) return CHRONOS_E_NOMEM;
    c->device = device;
    c->max_batch = max_batch_events;
    const char *fv = getenv("CHRONOS_FILTER_VARIANT");
    c->filter_variant = fv ? atoi(fv) : -1;  // -1: the tuned default (TILE 32, 1 stage)
    c->trace = getenv("CHRONOS_TRACE") != nullptr;
    auto boot = [&]() -> int {
        CK(cudaSetDevice(device));
        cudaDeviceProp p;
        CK(cudaGetDeviceProperties(&p, device));
        c->sm_count = p.multiProcessorCount;
        int prio_lo = 0, prio_hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        // the latency-bound scans get priority over the bandwidth-bound record gather
        CK(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio_hi));
        CK(cudaStreamCreateWithPriority(&c->stream2, cudaStreamNonBlocking, prio_lo));
        CK(cudaEventCreateWithFlags(&c->ev_sorted, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&c->ev_gathered, cudaEventDisableTiming));
        CK(cudaEventCreate(&c->ev_start));
        CK(cudaEventCreate(&c->ev_filter));
        CK(cudaEventCreate(&c->ev_end));
        CK(cudaEventCreate(&c->ev_f0));
        CK(cudaEventCreate(&c->ev_f1));
        CK(cudaMallocHost(reinterpret_cast<void **>(&c->h_scalars), RESULT_BLOCK_BYTES));
        memset(c->h_scalars, 0, RESULT_BLOCK_BYTES);
        c->h_plan = reinterpret_cast<Plan *>(reinterpret_cast<uint8_t *>(c->h_scalars) + 64);
        int rc;
        if ((rc = ensure(c, c->tickets, TICKET_BYTES))) return rc;
        if ((rc = ensure(c, c->scalars, RESULT_BLOCK_BYTES))) return rc;
        if ((rc = ensure(c, c->pt_count, 256))) return rc;
        if ((rc = ensure(c, c->comm_memo, (size_t)COMM_MEMO_SLOTS * 24))) return rc;
        CK(cudaMemsetAsync(c->comm_memo.p, 0, (size_t)COMM_MEMO_SLOTS * 24, c->stream));
        CK(cudaMemsetAsync(c->pt_count.p, 0, 256, c->stream));
        return CHRONOS_OK;
    };
    int rc = boot();
    if (rc != CHRONOS_OK) {
        chronos_destroy(c);
        return rc;
    }
    *out = c;
    return CHRONOS_OK;
}

int chronos_destroy(chronos_ctx *c) {
    if (!c) return CHRONOS_E_ARG;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    DevBuf *all[] = {&c->d_batch, &c->keyA, &c->keyB, &c->valA, &c->valB, &c->surv_src, &c->surv_meta, &c->pflag,
                     &c->entry_out, &c->wsorted, &c->comm_memo, &c->sp_ns, &c->surv_pid, &c->surv_ns, &c->slot_off, &c->pid_tab, &c->pt_count, &c->chain_pid, &c->chain_idx, &c->sp_key, &c->sp_src, &c->sp_meta, &c->tile_count, &c->windows, &c->owner_table, &c->surv_arrival, &c->arena[0],
                     &c->arena[1], &c->pend_key[0], &c->pend_key[1], &c->pend_slot[0], &c->pend_slot[1],
                     &c->pend_arrival[0], &c->pend_arrival[1], &c->pend_meta[0], &c->pend_meta[1], &c->status,
                     &c->tickets, &c->scalars, &c->scan_status, &c->sweep_status, &c->pack, &c->pack_off};
    for (DevBuf *b : all) release(*b);
    for (int d = 0; d < XCHG_MAX_WORLD; d++)
        if (c->x_opened[d] && c->x_peer[d]) cudaIpcCloseMemHandle(c->x_peer[d]);
    release(c->x_window);
    release(c->x_totals);
    if (c->dl_stream) {
        cudaStreamSynchronize(c->dl_stream);
        cudaStreamDestroy(c->dl_stream);
        cudaEventDestroy(c->dl_ready);
        cudaEventDestroy(c->dl_done[0]);
        cudaEventDestroy(c->dl_done[1]);
    }
    release(c->dl_stage[0]);
    release(c->dl_stage[1]);
    if (c->ev_x0) cudaEventDestroy(c->ev_x0);
    if (c->h_scalars) cudaFreeHost(c->h_scalars);
    if (c->ev_start) cudaEventDestroy(c->ev_start);
    if (c->ev_filter) cudaEventDestroy(c->ev_filter);
    if (c->ev_end) cudaEventDestroy(c->ev_end);
    if (c->ev_f0) cudaEventDestroy(c->ev_f0);
    if (c->ev_f1) cudaEventDestroy(c->ev_f1);
    if (c->ev_sorted) cudaEventDestroy(c->ev_sorted);
    if (c->ev_gathered) cudaEventDestroy(c->ev_gathered);
    if (c->stream2) cudaStreamDestroy(c->stream2);
    if (c->stream) cudaStreamDestroy(c->stream);
    cudaGetLastError();
    delete c;
    return CHRONOS_OK;
}

int chronos_submit(chronos_ctx *c, const void *records, uint64_t n, uint64_t base_arrival_index) {
    if (!c) return CHRONOS_E_ARG;
    if (n && !records) return fail(c, CHRONOS_E_ARG, "records is NULL");
    int rc = begin_batch(c, n, base_arrival_index);
    if (rc) return rc;
    const uint8_t *d_recs = nullptr;
    if (n) {
        if (is_device_pointer(records)) {
            if ((uintptr_t)records & 15) return fail(c, CHRONOS_E_ARG, "device records pointer must be 16-byte aligned");
            d_recs = static_cast<const uint8_t *>(records);
        } else {
            if ((rc = ensure(c, c->d_batch, n * CH_REC_BYTES))) return rc;
            CK(cudaMemcpyAsync(c->d_batch.p, records, n * CH_REC_BYTES, cudaMemcpyHostToDevice, c->stream));
            d_recs = ptr<uint8_t>(c->d_batch);
        }
    }
    c->b_records = d_recs;
    c->b_slots = nullptr;
    c->b_explicit_arrival = false;
    c->b_exchanged = false;
    if (n && (rc = run_filter_stage(c, d_recs, n, ptr<uint32_t>(c->keyA) + c->b_P))) return rc;
    if ((rc = run_window_stage(c))) return rc;
    c->phase = chronos_ctx::SUBMITTED;
    return CHRONOS_OK;
}

int chronos_submit_survivors(chronos_ctx *c, const void *slots, const uint64_t *hdr, uint64_t n, uint64_t n_slots) {
    if (!c) return CHRONOS_E_ARG;
    if (n && (!slots || !hdr)) return fail(c, CHRONOS_E_ARG, "NULL argument");
    if (n && (!is_device_pointer(slots) || ((uintptr_t)slots & 15) || !is_device_pointer(hdr)))
        return fail(c, CHRONOS_E_ARG, "slots and hdr must be device pointers (slots 16-byte aligned)");
    if (n_slots < 3 * n || n_slots > 18 * n) return fail(c, CHRONOS_E_ARG, "slot count inconsistent with entry count");
    int rc = begin_batch(c, n, 0);
    if (rc) return rc;
    c->b_records = nullptr;
    c->b_slots = static_cast<const uint4 *>(slots);
    c->b_explicit_arrival = true;
    c->b_exchanged = false;
    DeviceScalars *ds = ptr<DeviceScalars>(c->scalars);
    if (n) {
        if ((rc = ensure(c, c->surv_arrival, n * 8))) return rc;
        if ((rc = ensure(c, c->surv_pid, (n + 1) * 4))) return rc;
        if ((rc = ensure(c, c->surv_ns, n + 1))) return rc;
        if ((rc = ensure(c, c->slot_off, (n + 8) * 4))) return rc;
        const unsigned blocks = (unsigned)std::min<uint64_t>((n + 255) / 256, 2048);
        unpack_hdr_kernel<<<blocks, 256, 0, c->stream>>>(hdr, (uint32_t)n, ptr<uint64_t>(c->surv_arrival),
                                                         ptr<uint8_t>(c->surv_meta), ptr<uint8_t>(c->surv_ns),
                                                         ptr<uint32_t>(c->slot_off));
        if ((rc = launch_scan(c, ptr<uint32_t>(c->slot_off), n))) return rc;
        extract_pid_kernel<<<blocks, 256, 0, c->stream>>>(c->b_slots, ptr<uint32_t>(c->slot_off), (uint32_t)n,
                                                          ptr<uint32_t>(c->keyA) + c->b_P, ptr<uint32_t>(c->surv_pid),
                                                          ptr<uint32_t>(c->surv_src), &ds->key_or, &ds->key_and);
        CK(cudaGetLastError());
        c->launches += 2;
        set_u32_kernel<<<1, 1, 0, c->stream>>>(&ds->surv_count, (uint32_t)n);
        c->launches++;
    }
    if ((rc = run_window_stage(c))) return rc;
    c->phase = chronos_ctx::SUBMITTED;
    return CHRONOS_OK;
}

int chronos_fetch(chronos_ctx *c, chronos_counts *out) {
    if (!c || !out) return CHRONOS_E_ARG;
    if (c->phase == chronos_ctx::IDLE) return fail(c, CHRONOS_E_STATE, "nothing submitted");
    CK(cudaSetDevice(c->device));
    if (c->phase == chronos_ctx::SUBMITTED) {
        int rc = finish_batch(c);
        if (rc) return rc;
    }
    memset(out, 0, sizeof(*out));
    out->n_events = c->b_n;
    out->n_survivors = c->b_S;
    out->n_windows = c->b_W;
    out->n_arena = c->b_M;
    out->n_pending = c->n_pending;
    out->n_chains = c->b_chains;
    out->n_pids_seen = c->n_pids_seen;
    return CHRONOS_OK;
}

static int need_fetched(chronos_ctx *c) {
    if (!c) return CHRONOS_E_ARG;
    if (c->phase != chronos_ctx::FETCHED) return fail(c, CHRONOS_E_STATE, "call chronos_fetch first");
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return fail(c, CHRONOS_E_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    return CHRONOS_OK;
}

int chronos_copy_survivors(chronos_ctx *c, uint64_t *global_idx, uint8_t *meta, uint64_t cap) {
    int rc = need_fetched(c);
    if (rc) return rc;
    const uint64_t S = c->b_S;
    if (cap < S || (S && !global_idx)) return fail(c, CHRONOS_E_ARG, "survivor buffer too small (%llu < %llu)",
                                                   (unsigned long long)cap, (unsigned long long)S);
    if (!S) return CHRONOS_OK;
    if (c->b_explicit_arrival) {
        CK(cudaMemcpyAsync(global_idx, c->surv_arrival.p, S * 8, cudaMemcpyDeviceToHost, c->stream));
    } else {
        if ((rc = ensure(c, c->pack, S * 8))) return rc;
        add_base_kernel<<<(unsigned)std::min<uint64_t>((S + 255) / 256, 2048), 256, 0, c->stream>>>(
            ptr<uint32_t>(c->surv_src), (uint32_t)S, c->b_base, ptr<uint64_t>(c->pack));
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(global_idx, c->pack.p, S * 8, cudaMemcpyDeviceToHost, c->stream));
    }
    if (meta) CK(cudaMemcpyAsync(meta, c->surv_meta.p, S, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return CHRONOS_OK;
}

int chronos_copy_windows(chronos_ctx *c, chronos_window *dst, uint64_t cap) {
    int rc = need_fetched(c);
    if (rc) return rc;
    if (cap < c->b_W || (c->b_W && !dst)) return fail(c, CHRONOS_E_ARG, "window buffer too small");
    if (!c->b_W) return CHRONOS_OK;
    CK(cudaMemcpyAsync(dst, c->windows.p, c->b_W * sizeof(chronos_window), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return CHRONOS_OK;
}

int chronos_copy_arena(chronos_ctx *c, void *dst, uint64_t cap_records) {
    int rc = need_fetched(c);
    if (rc) return rc;
    if (cap_records < c->b_M || (c->b_M && !dst)) return fail(c, CHRONOS_E_ARG, "arena buffer too small");
    if (!c->b_M) return CHRONOS_OK;
    CK(cudaMemcpyAsync(dst, c->arena[c->cur].p, c->b_M * CH_REC_BYTES, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return CHRONOS_OK;
}

// records of all windows of the fetched batch, window after window, into a DEVICE buffer
// (every grouped entry is in a window or pending, so the total is known without a read-back)
static int pack_window_records(chronos_ctx *c, uint8_t *dst_dev) {
    const uint64_t W = c->b_W;
    if (!W) return CHRONOS_OK;
    int rc;
    if ((rc = ensure(c, c->pack_off, (W + 8) * 4))) return rc;
    const unsigned blocks = (unsigned)std::min<uint64_t>((W + 255) / 256, 2048);
    window_len_kernel<<<blocks, 256, 0, c->stream>>>(ptr<chronos_window>(c->windows), (uint32_t)W, ptr<uint32_t>(c->pack_off));
    if ((rc = launch_scan(c, ptr<uint32_t>(c->pack_off), W, nullptr, 56))) return rc;
    pack_windows_kernel<<<(unsigned)std::min<uint64_t>((W + 7) / 8, (uint64_t)c->sm_count * 8), 256, 0, c->stream>>>(
        ptr<uint8_t>(c->arena[c->cur]), ptr<chronos_window>(c->windows), ptr<uint32_t>(c->pack_off), (uint32_t)W, dst_dev);
    CK(cudaGetLastError());
    return CHRONOS_OK;
}

int chronos_copy_window_records(chronos_ctx *c, void *dst, uint64_t cap_records) {
    int rc = need_fetched(c);
    if (rc) return rc;
    const uint64_t tot = c->b_M - c->n_pending;  // grouped entries that left in a window
    if (!c->b_W) return CHRONOS_OK;
    if (cap_records < tot || !dst) return fail(c, CHRONOS_E_ARG, "window record buffer too small (%llu < %llu)",
                                               (unsigned long long)cap_records, (unsigned long long)tot);
    if ((rc = ensure(c, c->pack, tot * CH_REC_BYTES))) return rc;
    if ((rc = pack_window_records(c, ptr<uint8_t>(c->pack)))) return rc;
    CK(cudaMemcpyAsync(dst, c->pack.p, tot * CH_REC_BYTES, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return CHRONOS_OK;
}

// ---- asynchronous download of a fetched batch's results ---------------------------------------
int chronos_download_begin(chronos_ctx *c, void *host_dst, uint64_t cap_bytes, chronos_download *out) {
    int rc = need_fetched(c);
    if (rc) return rc;
    if (!out) return fail(c, CHRONOS_E_ARG, "out is NULL");
    auto al = [](uint64_t v) { return (v + 255) & ~255ull; };
    const uint64_t S = c->b_S, W = c->b_W, R = c->b_W ? c->b_M - c->n_pending : 0;
    memset(out, 0, sizeof(*out));
    out->n_survivors = S;
    out->n_windows = W;
    out->n_window_records = R;
    out->off_survivor_idx = 0;
    out->off_survivor_meta = al(S * 8);
    out->off_windows = out->off_survivor_meta + al(S);
    out->off_window_records = out->off_windows + al(W * sizeof(chronos_window));
    out->total_bytes = out->off_window_records + R * CH_REC_BYTES;
    if (out->total_bytes == 0) return CHRONOS_OK;
    if (!host_dst || cap_bytes < out->total_bytes)
        return fail(c, CHRONOS_E_ARG, "download buffer too small (%llu < %llu bytes)", (unsigned long long)cap_bytes,
                    (unsigned long long)out->total_bytes);
    if (!c->dl_stream) {
        CK(cudaStreamCreateWithFlags(&c->dl_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&c->dl_ready, cudaEventDisableTiming));
        for (int k = 0; k < 2; k++) CK(cudaEventCreateWithFlags(&c->dl_done[k], cudaEventDisableTiming));
    }
    const int k = c->dl_k;
    if (c->dl_inflight[k]) {  // the copy started two calls ago still reads this staging buffer
        CK(cudaEventSynchronize(c->dl_done[k]));
        c->dl_inflight[k] = false;
    }
    if ((rc = ensure(c, c->dl_stage[k], out->total_bytes))) return rc;
    uint8_t *st = ptr<uint8_t>(c->dl_stage[k]);
    if (S) {
        if (c->b_explicit_arrival) {
            CK(cudaMemcpyAsync(st + out->off_survivor_idx, c->surv_arrival.p, S * 8, cudaMemcpyDeviceToDevice, c->stream));
        } else {
            add_base_kernel<<<(unsigned)std::min<uint64_t>((S + 255) / 256, 2048), 256, 0, c->stream>>>(
                ptr<uint32_t>(c->surv_src), (uint32_t)S, c->b_base, reinterpret_cast<uint64_t *>(st + out->off_survivor_idx));
            CK(cudaGetLastError());
        }
        CK(cudaMemcpyAsync(st + out->off_survivor_meta, c->surv_meta.p, S, cudaMemcpyDeviceToDevice, c->stream));
    }
    if (W) {
        CK(cudaMemcpyAsync(st + out->off_windows, c->windows.p, W * sizeof(chronos_window), cudaMemcpyDeviceToDevice, c->stream));
        if ((rc = pack_window_records(c, st + out->off_window_records))) return rc;
    }
    CK(cudaEventRecord(c->dl_ready, c->stream));
    CK(cudaStreamWaitEvent(c->dl_stream, c->dl_ready, 0));
    CK(cudaMemcpyAsync(host_dst, st, out->total_bytes, cudaMemcpyDeviceToHost, c->dl_stream));
    CK(cudaEventRecord(c->dl_done[k], c->dl_stream));
    c->dl_inflight[k] = true;
    c->dl_k ^= 1;
    return CHRONOS_OK;
}

int chronos_download_wait(chronos_ctx *c) {
    if (!c) return CHRONOS_E_ARG;
    if (!c->dl_stream) return CHRONOS_OK;
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->dl_stream));
    c->dl_inflight[0] = c->dl_inflight[1] = false;
    return CHRONOS_OK;
}

int chronos_copy_chains(chronos_ctx *c, uint32_t *pid, uint64_t *global_idx, uint64_t cap) {
    int rc = need_fetched(c);
    if (rc) return rc;
    const uint64_t K = c->b_chains;
    if (cap < K) return fail(c, CHRONOS_E_ARG, "chain buffer too small (%llu < %llu)", (unsigned long long)cap,
                             (unsigned long long)K);
    if (!K) return CHRONOS_OK;
    if (pid) CK(cudaMemcpyAsync(pid, c->chain_pid.p, K * 4, cudaMemcpyDeviceToHost, c->stream));
    if (global_idx) CK(cudaMemcpyAsync(global_idx, c->chain_idx.p, K * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return CHRONOS_OK;
}

int chronos_state(chronos_ctx *c, uint32_t *pid, uint64_t *global_idx, void *records, uint64_t cap) {
    if (!c) return CHRONOS_E_ARG;
    if (c->phase == chronos_ctx::SUBMITTED) return fail(c, CHRONOS_E_STATE, "call chronos_fetch first");
    CK(cudaSetDevice(c->device));
    const uint64_t P = c->n_pending;
    if (cap < P) return fail(c, CHRONOS_E_ARG, "state buffer too small (%llu < %llu)", (unsigned long long)cap,
                             (unsigned long long)P);
    if (!P) return CHRONOS_OK;
    const int cur = c->cur;
    const uint64_t pc = c->pend_cap[cur];
    if (pid) CK(cudaMemcpyAsync(pid, ptr<uint32_t>(c->pend_key[cur]) + (pc - P), P * 4, cudaMemcpyDeviceToHost, c->stream));
    if (global_idx)
        CK(cudaMemcpyAsync(global_idx, ptr<uint64_t>(c->pend_arrival[cur]) + (pc - P), P * 8, cudaMemcpyDeviceToHost,
                           c->stream));
    if (records) {
        int rc;
        if ((rc = ensure(c, c->pack, P * CH_REC_BYTES))) return rc;
        pack_by_slot_kernel<<<(unsigned)std::min<uint64_t>((P + 7) / 8, (uint64_t)c->sm_count * 8), 256, 0, c->stream>>>(
            ptr<uint8_t>(c->arena[cur]), ptr<uint32_t>(c->pend_slot[cur]) + (pc - P), (uint32_t)P, ptr<uint8_t>(c->pack));
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(records, c->pack.p, P * CH_REC_BYTES, cudaMemcpyDeviceToHost, c->stream));
    }
    CK(cudaStreamSynchronize(c->stream));
    return CHRONOS_OK;
}

int chronos_state_pids(chronos_ctx *c, uint32_t *pid, uint64_t cap) {
    if (!c) return CHRONOS_E_ARG;
    if (c->phase == chronos_ctx::SUBMITTED) return fail(c, CHRONOS_E_STATE, "call chronos_fetch first");
    CK(cudaSetDevice(c->device));
    const uint64_t K = c->n_pids_seen;
    if (cap < K || (K && !pid)) return fail(c, CHRONOS_E_ARG, "pid buffer too small (%llu < %llu)",
                                            (unsigned long long)cap, (unsigned long long)K);
    if (!K) return CHRONOS_OK;
    int rc;
    if ((rc = ensure(c, c->pack, K * 4 + 256))) return rc;
    if ((rc = ensure(c, c->pack_off, 256))) return rc;
    uint32_t *cnt = ptr<uint32_t>(c->pack_off);
    CK(cudaMemsetAsync(cnt, 0, 4, c->stream));
    pt_dump_kernel<<<(unsigned)std::min<uint32_t>((c->pt_cap + 255) / 256, 4096), 256, 0, c->stream>>>(
        ptr<uint64_t>(c->pid_tab), c->pt_cap, c->pt_epoch, ptr<uint32_t>(c->pack), cnt);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(pid, c->pack.p, K * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return CHRONOS_OK;
}

// ---- checkpoint of the per-PID state ----------------------------------------------------------
namespace {
struct CkptHeader {
    char magic[8];
    uint32_t version, record_bytes;
    uint64_t n_pending, n_pids;
    uint64_t reserved[4];
};
static_assert(sizeof(CkptHeader) == 64, "checkpoint header");
const char CKPT_MAGIC[8] = {'C', 'H', 'R', 'N', 'C', 'K', 'P', '1'};
inline uint64_t al8(uint64_t v) { return (v + 7) & ~7ull; }
struct CkptLayout { uint64_t off_tab, off_pid, off_arr, off_meta, off_rec, total; };
CkptLayout ckpt_layout(uint64_t P, uint64_t K) {
    CkptLayout L;
    L.off_tab = sizeof(CkptHeader);
    L.off_pid = L.off_tab + K * 8;
    L.off_arr = L.off_pid + al8(P * 4);
    L.off_meta = L.off_arr + P * 8;
    L.off_rec = L.off_meta + al8(P);
    L.total = L.off_rec + P * CH_REC_BYTES;
    return L;
}
__global__ void iota_u32_kernel(uint32_t *p, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = i;
}
}  // namespace

int chronos_checkpoint_bytes(chronos_ctx *c, uint64_t *bytes) {
    if (!c || !bytes) return CHRONOS_E_ARG;
    if (c->phase == chronos_ctx::SUBMITTED) return fail(c, CHRONOS_E_STATE, "call chronos_fetch first");
    *bytes = ckpt_layout(c->n_pending, c->n_pids_seen).total;
    return CHRONOS_OK;
}

int chronos_checkpoint_save(chronos_ctx *c, void *dst, uint64_t cap) {
    if (!c || !dst) return CHRONOS_E_ARG;
    if (c->phase == chronos_ctx::SUBMITTED) return fail(c, CHRONOS_E_STATE, "call chronos_fetch first");
    CK(cudaSetDevice(c->device));
    const uint64_t P = c->n_pending, K = c->n_pids_seen;
    const CkptLayout L = ckpt_layout(P, K);
    if (cap < L.total) return fail(c, CHRONOS_E_ARG, "checkpoint buffer too small (%llu < %llu)", (unsigned long long)cap,
                                   (unsigned long long)L.total);
    uint8_t *out = static_cast<uint8_t *>(dst);
    memset(out, 0, L.off_pid);
    CkptHeader h;
    memset(&h, 0, sizeof(h));
    memcpy(h.magic, CKPT_MAGIC, 8);
    h.version = 1;
    h.record_bytes = CH_REC_BYTES;
    h.n_pending = P;
    h.n_pids = K;
    memcpy(out, &h, sizeof(h));
    int rc;
    if (K) {
        if ((rc = ensure(c, c->pack, K * 8 + 256))) return rc;
        if ((rc = ensure(c, c->pack_off, 256))) return rc;
        uint32_t *cnt = ptr<uint32_t>(c->pack_off);
        CK(cudaMemsetAsync(cnt, 0, 4, c->stream));
        pt_dump64_kernel<<<(unsigned)std::min<uint32_t>((c->pt_cap + 255) / 256, 4096), 256, 0, c->stream>>>(
            ptr<uint64_t>(c->pid_tab), c->pt_cap, c->pt_epoch, ptr<uint64_t>(c->pack), cnt);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(out + L.off_tab, c->pack.p, K * 8, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaStreamSynchronize(c->stream));
        uint64_t *e = reinterpret_cast<uint64_t *>(out + L.off_tab);
        std::sort(e, e + K, [](uint64_t a, uint64_t b) { return (uint32_t)a < (uint32_t)b; });  // by pid: files compare equal
    }
    if (P) {
        memset(out + L.off_pid, 0, L.off_arr - L.off_pid);
        memset(out + L.off_meta, 0, L.off_rec - L.off_meta);
        const int cur = c->cur;
        const uint64_t pc = c->pend_cap[cur];
        CK(cudaMemcpyAsync(out + L.off_pid, ptr<uint32_t>(c->pend_key[cur]) + (pc - P), P * 4, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaMemcpyAsync(out + L.off_arr, ptr<uint64_t>(c->pend_arrival[cur]) + (pc - P), P * 8, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaMemcpyAsync(out + L.off_meta, ptr<uint8_t>(c->pend_meta[cur]) + (pc - P), P, cudaMemcpyDeviceToHost, c->stream));
        if ((rc = ensure(c, c->pack, P * CH_REC_BYTES))) return rc;
        pack_by_slot_kernel<<<(unsigned)std::min<uint64_t>((P + 7) / 8, (uint64_t)c->sm_count * 8), 256, 0, c->stream>>>(
            ptr<uint8_t>(c->arena[cur]), ptr<uint32_t>(c->pend_slot[cur]) + (pc - P), (uint32_t)P, ptr<uint8_t>(c->pack));
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(out + L.off_rec, c->pack.p, P * CH_REC_BYTES, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaStreamSynchronize(c->stream));
    }
    return CHRONOS_OK;
}

int chronos_checkpoint_load(chronos_ctx *c, const void *src, uint64_t bytes) {
    if (!c || !src) return CHRONOS_E_ARG;
    if (bytes < sizeof(CkptHeader)) return fail(c, CHRONOS_E_ARG, "not a checkpoint (too short)");
    const uint8_t *in = static_cast<const uint8_t *>(src);
    CkptHeader h;
    memcpy(&h, in, sizeof(h));
    if (memcmp(h.magic, CKPT_MAGIC, 8) != 0 || h.version != 1 || h.record_bytes != CH_REC_BYTES)
        return fail(c, CHRONOS_E_ARG, "not a version-1 checkpoint of 288-byte records");
    const uint64_t P = h.n_pending, K = h.n_pids;
    if (P >= (1ull << VAL_ID_BITS) || K >= (1ull << 31)) return fail(c, CHRONOS_E_CAPACITY, "checkpoint too large");
    const CkptLayout L = ckpt_layout(P, K);
    if (bytes < L.total) return fail(c, CHRONOS_E_ARG, "checkpoint truncated (%llu < %llu bytes)", (unsigned long long)bytes,
                                     (unsigned long long)L.total);
    // the pending list must be what window_reverse would have written: ascending pid, ascending
    // arrival inside a pid; every pending pid must be in the key set
    const uint32_t *ppid = reinterpret_cast<const uint32_t *>(in + L.off_pid);
    const uint64_t *parr = reinterpret_cast<const uint64_t *>(in + L.off_arr);
    for (uint64_t i = 1; i < P; i++)
        if (ppid[i] < ppid[i - 1] || (ppid[i] == ppid[i - 1] && parr[i] <= parr[i - 1]))
            return fail(c, CHRONOS_E_ARG, "checkpoint: pending entry %llu out of order", (unsigned long long)i);
    int rc = chronos_reset(c);
    if (rc) return rc;
    if (K) {
        if ((rc = ensure_pid_table(c, K))) return rc;
        if ((rc = ensure(c, c->pack, K * 8))) return rc;
        CK(cudaMemcpyAsync(c->pack.p, in + L.off_tab, K * 8, cudaMemcpyHostToDevice, c->stream));
        pt_load_kernel<<<(unsigned)std::min<uint64_t>((K + 255) / 256, 2048), 256, 0, c->stream>>>(
            ptr<uint64_t>(c->pid_tab), c->pt_cap - 1, c->pt_epoch, ptr<uint64_t>(c->pack), (uint32_t)K);
        CK(cudaGetLastError());
        const uint32_t k32 = (uint32_t)K;
        CK(cudaMemcpyAsync(c->pt_count.p, &k32, 4, cudaMemcpyHostToDevice, c->stream));
        CK(cudaStreamSynchronize(c->stream));
    }
    const int cur = c->cur;
    if (P) {
        if ((rc = ensure(c, c->arena[cur], (P + 1) * CH_REC_BYTES))) return rc;
        if ((rc = ensure(c, c->pend_key[cur], (P + 1) * 4))) return rc;
        if ((rc = ensure(c, c->pend_slot[cur], (P + 1) * 4))) return rc;
        if ((rc = ensure(c, c->pend_arrival[cur], (P + 1) * 8))) return rc;
        if ((rc = ensure(c, c->pend_meta[cur], P + 1))) return rc;
        c->pend_cap[cur] = P;  // right-aligned in a layout of exactly P entries
        CK(cudaMemcpyAsync(c->pend_key[cur].p, ppid, P * 4, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->pend_arrival[cur].p, parr, P * 8, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->pend_meta[cur].p, in + L.off_meta, P, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->arena[cur].p, in + L.off_rec, P * CH_REC_BYTES, cudaMemcpyHostToDevice, c->stream));
        iota_u32_kernel<<<(unsigned)std::min<uint64_t>((P + 255) / 256, 2048), 256, 0, c->stream>>>(
            ptr<uint32_t>(c->pend_slot[cur]), (uint32_t)P);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(c->stream));
    }
    c->n_pending = P;
    c->n_pids_seen = K;
    c->phase = chronos_ctx::IDLE;
    return CHRONOS_OK;
}

int chronos_get_device_view(chronos_ctx *c, chronos_device_view *out) {
    int rc = need_fetched(c);
    if (rc) return rc;
    if (!out) return CHRONOS_E_ARG;
    const int cur = c->cur;
    const uint64_t P = c->n_pending, pc = c->pend_cap[cur];
    out->arena = c->arena[cur].p;
    out->windows = ptr<chronos_window>(c->windows);
    out->survivor_idx = ptr<uint32_t>(c->surv_src);
    out->survivor_meta = ptr<uint8_t>(c->surv_meta);
    out->pending_pid = P ? ptr<uint32_t>(c->pend_key[cur]) + (pc - P) : nullptr;
    out->pending_slot = P ? ptr<uint32_t>(c->pend_slot[cur]) + (pc - P) : nullptr;
    out->pending_arrival = P ? ptr<uint64_t>(c->pend_arrival[cur]) + (pc - P) : nullptr;
    return CHRONOS_OK;
}

int chronos_last_timing(chronos_ctx *c, float *total_ms, float *filter_ms) {
    int rc = need_fetched(c);
    if (rc) return rc;
    if (total_ms) *total_ms = c->t_total;
    if (filter_ms) *filter_ms = c->t_filter_kernel;
    return CHRONOS_OK;
}

int chronos_reset(chronos_ctx *c) {
    if (!c) return CHRONOS_E_ARG;
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    c->n_pending = 0;
    c->n_pids_seen = 0;
    // slots written under an older epoch count as empty: a reset costs nothing on the device
    c->pt_epoch = (c->pt_epoch + 1) & 0xFFu;
    if (c->pt_epoch == 0) {  // tags wrapped: really clear the table once every 255 resets
        c->pt_epoch = 1;
        if (c->pt_cap) CK(cudaMemsetAsync(c->pid_tab.p, 0, (size_t)c->pt_cap * 8, c->stream));
    }
    CK(cudaMemsetAsync(c->pt_count.p, 0, 256, c->stream));
    c->phase = chronos_ctx::IDLE;
    return CHRONOS_OK;
}

int chronos_last_launches(chronos_ctx *c, uint64_t *launches) {
    if (!c || !launches) return CHRONOS_E_ARG;
    *launches = c->launches;
    return CHRONOS_OK;
}

void *chronos_stream(chronos_ctx *c) { return c ? (void *)c->stream : nullptr; }

uint32_t chronos_owner_of(uint32_t pid, uint32_t world) { return world ? hash32(pid) % world : 0; }

// Stage 1 of the sharded path: filter the slice, partition the survivors by owner rank and
// deliver them (pk.direct: into the owners' windows; else into the caller's send buffers):
// filter -> compact_scan -> xchg_hist -> scan -> xchg_pack, all on the main stream. The
// per-PID state of the context is not touched.
// (Cutting the slice into pieces so that the delivery of piece k overlaps the filter of piece
//  k+1 was built and measured on 2xB200, 12.5 M events per rank: stage 1 got SLOWER, 0.80 ->
//  1.15 ms with four pieces -- the filter loses its ~19 single-warp CTAs per SM to the pack
//  kernel -- so it was removed again.)
static int stage1(chronos_ctx *c, const void *records, uint64_t n, uint64_t base, XchgPackParams &pk) {
    int rc;
    const uint8_t *d_recs = nullptr;
    if (n == 0) {
        // an empty slice still takes part in the exchange (zero counts, flag raised)
    } else if (is_device_pointer(records)) {
        if ((uintptr_t)records & 15) return fail(c, CHRONOS_E_ARG, "device records pointer must be 16-byte aligned");
        d_recs = static_cast<const uint8_t *>(records);
    } else {
        if ((rc = ensure(c, c->d_batch, n * CH_REC_BYTES))) return rc;
        CK(cudaMemcpyAsync(c->d_batch.p, records, n * CH_REC_BYTES, cudaMemcpyHostToDevice, c->stream));
        d_recs = ptr<uint8_t>(c->d_batch);
    }
    c->launches = 0;
    c->scan_slot = 0;
    c->n_tev = 0;
    if ((rc = ensure(c, c->keyA, (n + 1) * 4))) return rc;
    if ((rc = ensure(c, c->surv_src, (n + 1) * 4))) return rc;
    if ((rc = ensure(c, c->surv_meta, n + 1))) return rc;
    if ((rc = ensure(c, c->surv_pid, (n + 1) * 4))) return rc;
    if ((rc = ensure(c, c->surv_ns, n + 1))) return rc;
    if ((rc = ensure(c, c->x_totals, 512))) return rc;
    DeviceScalars *ds = ptr<DeviceScalars>(c->scalars);
    init_scalars_kernel<<<1, 256, 0, c->stream>>>(ds, ptr<uint32_t>(c->tickets));
    set_u32_kernel<<<1, 1, 0, c->stream>>>(&ds->xchg_err, 0u);
    c->launches += 2;
    mark(c, "stage 1 start");
    if (n && (rc = run_filter_stage(c, d_recs, n, ptr<uint32_t>(c->keyA)))) return rc;
    c->x_filter_timed = n != 0;
    mark(c, "filter + compact");

    const uint32_t W = pk.world;
    const uint32_t T = (uint32_t)std::max<uint64_t>((n + XCHG_TILE - 1) / XCHG_TILE, 1);  // upper bound of the tile count
    const uint64_t table_len = (uint64_t)2 * W * T + 1;
    if ((rc = ensure(c, c->owner_table, table_len * 4))) return rc;
    uint32_t *table = ptr<uint32_t>(c->owner_table);
    const unsigned xblocks = std::min<uint32_t>(T, (uint32_t)c->sm_count * 4);  // persistent CTAs
    xchg_hist<<<xblocks, XCHG_THREADS, 0, c->stream>>>(ptr<uint32_t>(c->surv_pid), ptr<uint8_t>(c->surv_ns), &ds->surv_count, W, T,
                                                     table);
    if ((rc = dbg_sync(c, "xchg_hist"))) return rc;
    if ((rc = launch_scan(c, table, table_len))) return rc;
    if ((rc = dbg_sync(c, "owner table scan"))) return rc;
    mark(c, "owner histogram + scan");
    pk.pid = ptr<uint32_t>(c->surv_pid);
    pk.surv_src = ptr<uint32_t>(c->surv_src);
    pk.surv_meta = ptr<uint8_t>(c->surv_meta);
    pk.surv_ns = ptr<uint8_t>(c->surv_ns);
    pk.n_ptr = &ds->surv_count;
    pk.batch = d_recs;
    pk.base_arrival = base;
    pk.table = table;
    pk.T = T;
    pk.totals = ptr<uint32_t>(c->x_totals);
    pk.err = &ds->xchg_err;
    pk.done = ptr<uint32_t>(c->tickets) + 58;
    static bool pack_configured[16] = {false};
    if (!pack_configured[c->device & 15]) {
        CK(cudaFuncSetAttribute(xchg_pack, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)XCHG_PACK_SMEM));
        pack_configured[c->device & 15] = true;
    }
    xchg_pack<<<xblocks, XCHG_THREADS, XCHG_PACK_SMEM, c->stream>>>(pk);
    CK(cudaGetLastError());
    c->launches += 2;
    if ((rc = dbg_sync(c, "xchg_pack"))) return rc;
    mark(c, "pack into owners' windows");
    return CHRONOS_OK;
}

int chronos_filter_partition(chronos_ctx *c, const void *records, uint64_t n, uint64_t base_arrival_index,
                             uint32_t world, void *send_slots, uint64_t cap_slots, uint64_t *send_hdr,
                             uint64_t cap_entries, uint64_t *entry_counts_host, uint64_t *slot_counts_host) {
    if (!c || !entry_counts_host || !slot_counts_host || world == 0 || world > XCHG_MAX_WORLD)
        return c ? fail(c, CHRONOS_E_ARG, "bad argument (world must be 1..%d)", XCHG_MAX_WORLD) : CHRONOS_E_ARG;
    if (n && (!records || !send_slots || !send_hdr)) return fail(c, CHRONOS_E_ARG, "NULL argument");
    CK(cudaSetDevice(c->device));
    if (c->phase == chronos_ctx::SUBMITTED) {
        int rc0 = finish_batch(c);
        if (rc0) return rc0;
    }
    if (n > c->max_batch) return fail(c, CHRONOS_E_CAPACITY, "batch of %llu events > max_batch_events %llu",
                                      (unsigned long long)n, (unsigned long long)c->max_batch);
    for (uint32_t d = 0; d < world; d++) entry_counts_host[d] = slot_counts_host[d] = 0;
    if (!n) return CHRONOS_OK;
    int rc;
    CK(cudaEventRecord(c->ev_start, c->stream));
    XchgPackParams pk;
    memset(&pk, 0, sizeof(pk));
    pk.world = world;
    pk.direct = false;
    pk.send_slots = static_cast<uint4 *>(send_slots);
    pk.send_hdr = send_hdr;
    pk.cap_entries = cap_entries;
    pk.cap_slots = cap_slots;
    if ((rc = stage1(c, records, n, base_arrival_index, pk))) return rc;
    uint32_t hc[2 * XCHG_MAX_WORLD];
    DeviceScalars *ds = ptr<DeviceScalars>(c->scalars);
    CK(cudaMemcpyAsync(hc, c->x_totals.p, sizeof(hc), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(c->h_scalars, ds, sizeof(DeviceScalars), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaEventRecord(c->ev_end, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    uint64_t tot_e = 0, tot_s = 0;
    for (uint32_t d = 0; d < world; d++) {
        entry_counts_host[d] = hc[d];
        slot_counts_host[d] = hc[XCHG_MAX_WORLD + d];
        tot_e += hc[d];
        tot_s += hc[XCHG_MAX_WORLD + d];
    }
    cudaEventElapsedTime(&c->t_filter_kernel, c->ev_f0, c->ev_f1);
    cudaGetLastError();
    if (c->h_scalars->xchg_err & XERR_OVERFLOW)
        return fail(c, CHRONOS_E_CAPACITY, "%llu survivors / %llu transport slots do not fit the send buffers (%llu / %llu)",
                    (unsigned long long)tot_e, (unsigned long long)tot_s, (unsigned long long)cap_entries,
                    (unsigned long long)cap_slots);
    return CHRONOS_OK;
}

// ---- direct exchange over peer memory ---------------------------------------------------------
int chronos_exchange_create(chronos_ctx *c, uint32_t world, uint32_t rank, uint64_t cap_entries_per_peer,
                            uint64_t cap_slots_per_peer, void **window_out, uint8_t *ipc_handle_out) {
    if (!c) return CHRONOS_E_ARG;
    if (world == 0 || world > XCHG_MAX_WORLD || rank >= world || cap_entries_per_peer == 0)
        return fail(c, CHRONOS_E_ARG, "bad argument (world must be 1..%d, rank < world)", XCHG_MAX_WORLD);
    if (c->x_created) return fail(c, CHRONOS_E_STATE, "exchange window already created");
    if (cap_slots_per_peer == 0) cap_slots_per_peer = cap_entries_per_peer * 8;
    if (cap_slots_per_peer < 3 * cap_entries_per_peer) cap_slots_per_peer = 3 * cap_entries_per_peer;
    if ((uint64_t)world * cap_slots_per_peer >= (1ull << 32) || (uint64_t)world * cap_entries_per_peer > c->max_batch)
        return fail(c, CHRONOS_E_CAPACITY, "world x capacity per peer too large (entries must fit max_batch_events %llu)",
                    (unsigned long long)c->max_batch);
    CK(cudaSetDevice(c->device));
    c->x_L = xchg_layout(world, rank, (uint32_t)cap_entries_per_peer, (uint32_t)cap_slots_per_peer);
    int rc;
    if ((rc = ensure(c, c->x_window, 2 * c->x_L.half_bytes))) return rc;
    if ((rc = ensure(c, c->x_totals, 512))) return rc;
    CK(cudaMemsetAsync(c->x_window.p, 0, 2 * c->x_L.half_bytes, c->stream));
    CK(cudaStreamSynchronize(c->stream));  // peers may write as soon as they hold the address
    if (!c->ev_x0) CK(cudaEventCreate(&c->ev_x0));
    if (ipc_handle_out) {
        cudaIpcMemHandle_t h;
        static_assert(sizeof(h) == CHRONOS_IPC_HANDLE_BYTES, "IPC handle size");
        CK(cudaIpcGetMemHandle(&h, c->x_window.p));
        memcpy(ipc_handle_out, &h, sizeof(h));
    }
    if (window_out) *window_out = c->x_window.p;
    for (int d = 0; d < XCHG_MAX_WORLD; d++) { c->x_peer[d] = nullptr; c->x_opened[d] = false; }
    c->x_peer[rank] = ptr<uint8_t>(c->x_window);
    c->x_created = true;
    c->x_connected = (world == 1);
    c->x_epoch = 0;
    c->x_sent = false;
    return CHRONOS_OK;
}

int chronos_exchange_connect(chronos_ctx *c, void *const *peer_windows, const uint8_t *peer_ipc_handles) {
    if (!c) return CHRONOS_E_ARG;
    if (!c->x_created) return fail(c, CHRONOS_E_STATE, "call chronos_exchange_create first");
    if (!peer_windows && !peer_ipc_handles && c->x_L.world > 1) return fail(c, CHRONOS_E_ARG, "no peer addresses given");
    CK(cudaSetDevice(c->device));
    for (uint32_t d = 0; d < c->x_L.world; d++) {
        if (d == c->x_L.rank || c->x_peer[d]) continue;
        if (peer_windows && peer_windows[d]) {
            c->x_peer[d] = static_cast<uint8_t *>(peer_windows[d]);
            continue;
        }
        if (!peer_ipc_handles) return fail(c, CHRONOS_E_ARG, "no address for rank %u", d);
        cudaIpcMemHandle_t h;
        memcpy(&h, peer_ipc_handles + (size_t)d * CHRONOS_IPC_HANDLE_BYTES, sizeof(h));
        void *p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return fail(c, CHRONOS_E_CUDA, "cudaIpcOpenMemHandle for rank %u failed: %s", d, cudaGetErrorString(e));
        }
        c->x_peer[d] = static_cast<uint8_t *>(p);
        c->x_opened[d] = true;
    }
    c->x_connected = true;
    return CHRONOS_OK;
}

int chronos_filter_exchange(chronos_ctx *c, const void *records, uint64_t n, uint64_t base_arrival_index) {
    if (!c) return CHRONOS_E_ARG;
    if (!c->x_connected) return fail(c, CHRONOS_E_STATE, "exchange not connected");
    if (c->x_sent) return fail(c, CHRONOS_E_STATE, "previous exchange not consumed: call chronos_submit_exchanged");
    if (n && !records) return fail(c, CHRONOS_E_ARG, "records is NULL");
    CK(cudaSetDevice(c->device));
    if (c->phase == chronos_ctx::SUBMITTED) {
        int rc0 = finish_batch(c);
        if (rc0) return rc0;
    }
    if (n > c->max_batch) return fail(c, CHRONOS_E_CAPACITY, "batch of %llu events > max_batch_events %llu",
                                      (unsigned long long)n, (unsigned long long)c->max_batch);
    int rc;
    CK(cudaEventRecord(c->ev_x0, c->stream));
    XchgPackParams pk;
    memset(&pk, 0, sizeof(pk));
    pk.world = c->x_L.world;
    pk.direct = true;
    pk.L = c->x_L;
    pk.epoch = c->x_epoch + 1;
    for (uint32_t d = 0; d < c->x_L.world; d++) pk.peer[d] = c->x_peer[d];
    if ((rc = stage1(c, records, n, base_arrival_index, pk))) return rc;
    c->x_epoch++;
    c->x_sent = true;
    c->x_launches = c->launches;
    return CHRONOS_OK;
}

int chronos_submit_exchanged(chronos_ctx *c) {
    if (!c) return CHRONOS_E_ARG;
    if (!c->x_sent) return fail(c, CHRONOS_E_STATE, "nothing sent: call chronos_filter_exchange first");
    const XchgLayout &L = c->x_L;
    const uint64_t n_max = (uint64_t)L.world * L.cap_e;
    const uint64_t stage1_launches = c->x_launches;
    int rc = begin_batch(c, n_max, 0);
    if (rc) return rc;
    c->launches += stage1_launches;
    const uint8_t *hb = ptr<uint8_t>(c->x_window) + (size_t)(c->x_epoch & 1u) * L.half_bytes;
    c->b_records = nullptr;
    c->b_slots = reinterpret_cast<const uint4 *>(hb + L.off_slots);
    c->b_explicit_arrival = true;
    c->b_exchanged = true;
    c->x_sent = false;
    DeviceScalars *ds = ptr<DeviceScalars>(c->scalars);
    if ((rc = ensure(c, c->surv_arrival, n_max * 8))) return rc;
    if ((rc = ensure(c, c->surv_pid, (n_max + 1) * 4))) return rc;
    if ((rc = ensure(c, c->surv_ns, n_max + 1))) return rc;
    if ((rc = ensure(c, c->slot_off, (n_max + 8) * 4))) return rc;
    XchgUnpackParams u;
    memset(&u, 0, sizeof(u));
    u.window = ptr<uint8_t>(c->x_window);
    u.L = L;
    u.epoch = c->x_epoch;
    u.timeout_ns = c->x_timeout_ns;
    u.arrival = ptr<uint64_t>(c->surv_arrival);
    u.meta = ptr<uint8_t>(c->surv_meta);
    u.ns = ptr<uint8_t>(c->surv_ns);
    u.slot_off = ptr<uint32_t>(c->slot_off);
    u.key = ptr<uint32_t>(c->keyA) + c->b_P;
    u.pid_copy = ptr<uint32_t>(c->surv_pid);
    u.src = ptr<uint32_t>(c->surv_src);
    u.n_out = &ds->surv_count;
    u.k_or = &ds->key_or;
    u.k_and = &ds->key_and;
    u.err = &ds->xchg_err;
    u.n_cap = (uint32_t)n_max;
    xchg_unpack<<<(unsigned)std::min<uint64_t>((n_max + 255) / 256, (uint64_t)c->sm_count * 4), 256, 0, c->stream>>>(u);
    CK(cudaGetLastError());
    c->launches++;
    if ((rc = run_window_stage(c))) return rc;
    c->phase = chronos_ctx::SUBMITTED;
    return CHRONOS_OK;
}

int chronos_exchange_set_timeout(chronos_ctx *c, double seconds) {
    if (!c || !(seconds > 0)) return CHRONOS_E_ARG;
    c->x_timeout_ns = (uint64_t)(seconds * 1e9);
    return CHRONOS_OK;
}

// ---- trace generator ------------------------------------------------------------------------
int chronos_tracegen_host(const chronos_tracegen_cfg *cfg, const uint64_t *cdf, uint64_t first, uint64_t n, void *dst) {
    if (!cfg || (n && !dst) || cfg->n_pids == 0 || (cfg->use_cdf && !cdf)) return CHRONOS_E_ARG;
    uint8_t *d = static_cast<uint8_t *>(dst);
    for (uint64_t i = 0; i < n; i++) chronos_gen::gen_record(*cfg, first + i, cdf, d + i * CH_REC_BYTES);
    return CHRONOS_OK;
}

int chronos_tracegen_device(int device, const chronos_tracegen_cfg *cfg, const uint64_t *cdf_dev, uint64_t first,
                            uint64_t n, void *dst_dev) {
    if (!cfg || (n && !dst_dev) || cfg->n_pids == 0 || (cfg->use_cdf && !cdf_dev)) return CHRONOS_E_ARG;
    if (cudaSetDevice(device) != cudaSuccess) return CHRONOS_E_CUDA;
    if (!n) return CHRONOS_OK;
    tracegen_kernel<<<(unsigned)std::min<uint64_t>((n + 127) / 128, 148 * 16), 128>>>(*cfg, cdf_dev, first, n,
                                                                                      static_cast<uint8_t *>(dst_dev));
    if (cudaGetLastError() != cudaSuccess) return CHRONOS_E_CUDA;
    return cudaDeviceSynchronize() == cudaSuccess ? CHRONOS_OK : CHRONOS_E_CUDA;
}

// ---- plain memory helpers -------------------------------------------------------------------
int chronos_dev_alloc(int device, uint64_t bytes, void **out) {
    if (!out) return CHRONOS_E_ARG;
    if (cudaSetDevice(device) != cudaSuccess) return CHRONOS_E_CUDA;
    cudaError_t e = cudaMalloc(out, bytes ? bytes : 16);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return e == cudaErrorMemoryAllocation ? CHRONOS_E_NOMEM : CHRONOS_E_CUDA;
    }
    return CHRONOS_OK;
}
int chronos_dev_free(int device, void *p) {
    if (cudaSetDevice(device) != cudaSuccess) return CHRONOS_E_CUDA;
    return cudaFree(p) == cudaSuccess ? CHRONOS_OK : CHRONOS_E_CUDA;
}
int chronos_host_alloc_pinned(uint64_t bytes, void **out) {
    if (!out) return CHRONOS_E_ARG;
    cudaError_t e = cudaMallocHost(out, bytes ? bytes : 16);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return e == cudaErrorMemoryAllocation ? CHRONOS_E_NOMEM : CHRONOS_E_CUDA;
    }
    return CHRONOS_OK;
}
int chronos_host_free_pinned(void *p) { return cudaFreeHost(p) == cudaSuccess ? CHRONOS_OK : CHRONOS_E_CUDA; }
int chronos_memcpy_h2d(int device, void *dst_dev, const void *src_host, uint64_t bytes) {
    if (cudaSetDevice(device) != cudaSuccess) return CHRONOS_E_CUDA;
    return cudaMemcpy(dst_dev, src_host, bytes, cudaMemcpyHostToDevice) == cudaSuccess ? CHRONOS_OK : CHRONOS_E_CUDA;
}
int chronos_memcpy_d2h(int device, void *dst_host, const void *src_dev, uint64_t bytes) {
    if (cudaSetDevice(device) != cudaSuccess) return CHRONOS_E_CUDA;
    return cudaMemcpy(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost) == cudaSuccess ? CHRONOS_OK : CHRONOS_E_CUDA;
}
int chronos_flush_l2(int device) {
    static void *buf[16] = {nullptr};
    const size_t bytes = 512ull << 20;
    if (device < 0 || device >= 16) return CHRONOS_E_ARG;
    if (cudaSetDevice(device) != cudaSuccess) return CHRONOS_E_CUDA;
    if (!buf[device] && cudaMalloc(&buf[device], bytes) != cudaSuccess) {
        cudaGetLastError();
        return CHRONOS_E_NOMEM;
    }
    static uint32_t v = 1;
    l2_flush_kernel<<<148 * 8, 256>>>(static_cast<uint4 *>(buf[device]), bytes / 16, v++);
    return cudaDeviceSynchronize() == cudaSuccess ? CHRONOS_OK : CHRONOS_E_CUDA;
}

// ---- the allocation-free entry point ------------------------------------------------------------
int chronos_filter_noalloc(int device, void *stream, const void *records, uint64_t n, uint32_t *survivor_idx,
                           uint32_t *survivor_pid, uint8_t *survivor_meta, uint32_t *count, char *errbuf,
                           uint64_t errbuf_len) {
    auto say = [&](int code, const char *fmt, ...) -> int {
        if (errbuf && errbuf_len) {
            va_list ap;
            va_start(ap, fmt);
            vsnprintf(errbuf, (size_t)errbuf_len, fmt, ap);
            va_end(ap);
        }
        return code;
    };
    if (n > noalloc::MAX_EVENTS)
        return say(CHRONOS_E_CAPACITY,
                   "chronos_filter_noalloc: %llu events exceed the fixed scratch, which is sized for at most %llu "
                   "(CHRONOS_NOALLOC_MAX_EVENTS); split the batch or use chronos_submit",
                   (unsigned long long)n, (unsigned long long)noalloc::MAX_EVENTS);
    if (!count || (n && (!records || !survivor_idx || !survivor_pid || !survivor_meta)))
        return say(CHRONOS_E_ARG, "chronos_filter_noalloc: NULL argument");
    if ((uintptr_t)records & 15) return say(CHRONOS_E_ARG, "chronos_filter_noalloc: records must be 16-byte aligned");
    if (device < 0 || device >= 16) return say(CHRONOS_E_ARG, "chronos_filter_noalloc: device index out of range");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return say(CHRONOS_E_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));

    // one-time set-up per device, with that device current: turn the host-side stand-in
    // symbol into the array's real device address, and learn the launch geometry
    static unsigned char *scratch_of[16] = {nullptr};
    static int grid_of[16] = {0};
    static std::mutex setup_lock;
    unsigned char *base;
    int grid_cap;
    {
        std::lock_guard<std::mutex> g(setup_lock);
        if (!scratch_of[device]) {
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p)
                return say(CHRONOS_E_CUDA, "cudaGetSymbolAddress(scratch) on device %d: %s", device,
                           cudaGetErrorString(e));
            int sms = 0, per_sm = 0;
            const size_t smem = filter_smem_bytes<1>();
            if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess ||
                (e = cudaFuncSetAttribute(filter_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) !=
                    cudaSuccess ||
                (e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, filter_kernel<1>, FILTER_TILE, smem)) !=
                    cudaSuccess ||
                per_sm < 1)
                return say(CHRONOS_E_CUDA, "filter kernel set-up on device %d: %s", device, cudaGetErrorString(e));
            grid_of[device] = sms * per_sm;
            scratch_of[device] = static_cast<unsigned char *>(p);
        }
        base = scratch_of[device];
        grid_cap = grid_of[device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    uint32_t *hdr = reinterpret_cast<uint32_t *>(base + noalloc::OFF_HEADER);
    // the array keeps whatever the previous call left: tickets, counters, scan status start at zero
    if ((e = cudaMemsetAsync(base, 0, noalloc::CLEAR_BYTES, st)) != cudaSuccess)
        return say(CHRONOS_E_CUDA, "cudaMemsetAsync(scratch): %s", cudaGetErrorString(e));
    if (n) {
        FilterParams fp;
        fp.recs = static_cast<const uint8_t *>(records);
        fp.n = n;
        fp.tile_begin = 0;
        fp.tile_end = (uint32_t)((n + FILTER_TILE - 1) / FILTER_TILE);
        fp.ticket = hdr + 0;
        fp.tile_count = reinterpret_cast<uint32_t *>(base + noalloc::OFF_TILE_COUNT);
        fp.sp_key = reinterpret_cast<uint32_t *>(base + noalloc::OFF_SP_KEY);
        fp.sp_src = reinterpret_cast<uint32_t *>(base + noalloc::OFF_SP_SRC);
        fp.sp_meta = base + noalloc::OFF_SP_META;
        fp.sp_ns = nullptr;
        fp.out_count = hdr + 2;
        fp.key_or = hdr + 3;
        fp.key_and = hdr + 4;
        fp.gm_lock = nullptr;
        fp.gm_val = nullptr;
        fp.gm_key = nullptr;
        fp.gm_mask = 0;
        filter_kernel<1><<<std::min<uint32_t>(fp.tile_end, (uint32_t)grid_cap), FILTER_TILE, filter_smem_bytes<1>(), st>>>(fp);
        ScanParams sp;
        sp.data = fp.tile_count;
        sp.n = fp.tile_end;
        sp.n_ptr = nullptr;
        sp.ticket = hdr + 1;
        sp.status = reinterpret_cast<uint64_t *>(base + noalloc::OFF_STATUS);
        sp.epoch = 1;  // the status words were just zeroed, tag 0 means "not written yet"
        scan_exclusive_u32<<<std::min<uint32_t>((fp.tile_end + SCAN_TILE - 1) / SCAN_TILE, 64u), SCAN_THREADS, 0, st>>>(sp);
        CompactParams cp;
        cp.tile_off = fp.tile_count;
        cp.sp_key = fp.sp_key;
        cp.sp_src = fp.sp_src;
        cp.sp_meta = fp.sp_meta;
        cp.sp_ns = nullptr;
        cp.out_key = survivor_pid;
        cp.out_src = survivor_idx;
        cp.out_meta = survivor_meta;
        cp.out_ns = nullptr;
        cp.out_pid = nullptr;
        cp.n_tiles = fp.tile_end;
        cp.tile = FILTER_TILE;
        cp.total = hdr + 2;
        compact_survivors<<<std::min<uint32_t>((fp.tile_end + 255) / 256, 1024u), 256, 0, st>>>(cp);
        if ((e = cudaGetLastError()) != cudaSuccess)
            return say(CHRONOS_E_CUDA, "kernel launch: %s", cudaGetErrorString(e));
    }
    if ((e = cudaMemcpyAsync(count, hdr + 2, 4, cudaMemcpyDeviceToDevice, st)) != cudaSuccess)
        return say(CHRONOS_E_CUDA, "cudaMemcpyAsync(count): %s", cudaGetErrorString(e));
    return CHRONOS_OK;
}

}  // extern "C"
