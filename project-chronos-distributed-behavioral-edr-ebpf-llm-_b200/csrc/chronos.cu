// chronos.cu -- context, batch pipeline and the C ABI (include/chronos.h) of the
// B200-native CHRONOS hot path. Kernels live in filter.cuh / sort.cuh / window.cuh.
//
// Per batch (one perf_buffer_poll() worth of events, chronos_sensor.py:160-162):
//   stage 1 (needs nothing from earlier batches)
//     filter_kernel      raw records -> tile-local survivors (pid, src index, meta)
//     compact_scan       dense, arrival-ordered survivor arrays
//   tail (reads and writes the per-PID state, so tails run one after the other)
//     sort_prepare + onesweep_pass
//                        stable sort of [carried-in pending ++ survivors] by pid
//     gather_records     every grouped event copied once into the arena (own stream)
//     group_forward      flush automaton + chain labeller -> window descriptors
//     window_reverse     unflushed tails -> next batch's pending list (process_memory)
//     emit_ordered       windows in trigger-arrival order (the order analyze_sequence
//                        is called in, :143-147)
// A submit only enqueues: extents of everything after the filter -- including the number of
// pending entries the previous batch left -- live on the device (Plan, Carry), grids are upper
// bounds; the first host wait is chronos_fetch. Two batches may be in flight: the context
// keeps two buffer sets, and stage 1 of batch k+1 runs beside the tail of batch k.
// All device memory for batches of up to max_batch_events is allocated by chronos_create.
#include <cuda_runtime.h>
#include <cuda_profiler_api.h>
#include <emmintrin.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <new>
#include <atomic>
#include <thread>
#include <vector>

#include "chronos.h"
#include "common.cuh"
#include "filter.cuh"
#include "sort.cuh"
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
// the Plan of the batch lives right behind its scalars: ONE copy brings both back
constexpr size_t RESULT_BLOCK_BYTES = 64 + ((sizeof(Plan) + 63) / 64) * 64;

constexpr int N_TICKETS = 64;
// ticket counters followed by the global digit histograms of the sort (sort.cuh), cleared together
constexpr size_t TICKET_WORDS = N_TICKETS + RADIX_MAX_PASSES * RADIX_BINS;
constexpr size_t TICKET_BYTES = TICKET_WORDS * 4;
constexpr uint64_t MAX_GROUPED = (1ull << VAL_ID_BITS) - 1;  // entry ids are 27 bits of the sort value
constexpr uint64_t MAX_ARRIVAL = 1ull << 48;                  // the transport header keeps 48 bits

int device_sm_count(int device) {
    static int cached[64] = {0};
    if (device >= 0 && device < 64 && cached[device]) return cached[device];
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || sms <= 0) {
        cudaGetLastError();
        sms = 148;
    }
    if (device >= 0 && device < 64) cached[device] = sms;
    return sms;
}

// Everything one batch owns: the state it leaves behind (arena + pending lists, read by the
// next batch), its survivors and results (read by the host after fetch) and its bookkeeping.
// The context has two of these; batch k uses set k & 1.
struct BatchSet {
    DevBuf arena, pend_key, pend_slot, pend_arrival, pend_meta;
    uint64_t pend_cap = 0;  // entries the right-aligned pending arrays are laid out for
    uint64_t cap_M = 0;     // grouped entries (pending + survivors) this set has room for
    uint64_t cap_n = 0;     // raw events of one batch this set has room for
    DevBuf surv_src, surv_meta, surv_pid, surv_ns, surv_arrival, slot_off;
    DevBuf windows, chain_pid, chain_idx;
    DevBuf d_batch;         // upload staging for host-pointer submits (allocated on first use)
    DevBuf result;          // DeviceScalars + Plan
    DeviceScalars *h_scalars = nullptr;  // pinned mirror of `result`
    Plan *h_plan = nullptr;
    cudaEvent_t ev_start = nullptr, ev_stage1 = nullptr, ev_end = nullptr, ev_f0 = nullptr, ev_f1 = nullptr;
    cudaEvent_t ev_packed = nullptr;  // last read of this set's results by a download
    cudaEvent_t ev_prepack = nullptr, ev_s1end = nullptr, ev_sort = nullptr;  // stage boundaries (timing only)
    cudaEvent_t ev_x0 = nullptr;   // start of the exchange step this set serves
    bool x_filter_timed = false;   // stage 1 of that step ran the filter (non-empty slice)
    float stage_ms[8] = {0};
    enum State { FREE, SUBMITTED, DONE, FETCHED } state = FREE;
    // the batch
    uint64_t n = 0, base = 0, P_ub = 0, pids_ub = 0, Mmax = 0;
    uint64_t S = 0, M = 0, W = 0, chains = 0, pending_after = 0, pids_after = 0, launches = 0, seq = 0;
    const uint8_t *records = nullptr;  // device pointer of the batch records
    const uint4 *slots = nullptr;      // transport-form batch (chronos_submit_survivors / exchange)
    bool explicit_arrival = false, exchanged = false, filter_timed = false, pipelined = false;
    uint32_t x_epoch = 0;  // exchange step this batch consumed
    uint64_t gen = 0;      // state generation (chronos_reset starts a new one) the batch belongs to
    float t_total = 0.f, t_filter = 0.f, t_filter_kernel = 0.f;
    int rc = CHRONOS_OK;  // outcome found when the batch completed
};

}  // namespace

struct chronos_ctx {
    int device = 0;
    int sm_count = 0;
    uint64_t max_batch = 0, max_pending = 0;
    // stream: tails (and stage 1 when nothing else is in flight); stream_f: stage 1 of a batch
    // submitted while the previous one is still running; stream2: record gather; stream3: ordered
    // emission beside the reverse scan; dl_stream: everything the host asks for after a fetch
    cudaStream_t stream = nullptr, stream_f = nullptr, stream2 = nullptr, stream3 = nullptr, dl_stream = nullptr;
    cudaEvent_t ev_sorted = nullptr, ev_gathered = nullptr, ev_fwd = nullptr, ev_emitted = nullptr, ev_tailinit = nullptr;
    char err[256] = {0};

    BatchSet set[2];
    int last = 0;       // set of the most recently submitted batch (or of the loaded / reset state)
    int fetched = -1;   // set whose results the copy calls return
    uint64_t seq = 0;
    uint64_t gen = 0;  // bumped by chronos_reset: batches of an older generation no longer describe the state
    bool overlap = true;  // stage 1 of a batch may run beside the previous batch's tail
    uint64_t n_pending = 0, n_pids_seen = 0;  // exact once every submitted batch has completed

    // shared by all batches
    DevBuf carry;                                  // Carry: what a batch hands to the next on the device
    DevBuf keyA, keyB, valA, valB, pflag, entry_out, wsorted;   // tail temporaries
    DevBuf status, sweep_status;
    uint64_t cap_M = 0, cap_n = 0;
    DevBuf sp_key, sp_src, sp_meta, sp_ns, tile_count, scan_status;   // stage-1 temporaries
    DevBuf tickets_f, tickets_t;
    DevBuf owner_table;
    uint64_t owner_table_T = 0;
    DevBuf pid_tab, pt_count;                      // persistent per-PID table (chain.cuh)
    uint32_t pt_cap = 0;
    uint32_t pt_epoch = 1;
    uint32_t scan_epoch = 0, sweep_epoch = 0;
    int scan_slot = 0;  // next free ticket counter for scans of the stage 1 being enqueued
    // host-side copies after a fetch
    DevBuf pack, pack_off, dl_scan_status, dl_ticket;
    uint32_t dl_scan_epoch = 0;
    cudaEvent_t dl_ready = nullptr, dl_done[2] = {nullptr, nullptr};
    DevBuf dl_stage[2];
    bool dl_inflight[2] = {false, false};
    int dl_k = 0;

    // direct exchange over peer memory (exchange.cuh)
    DevBuf x_window, x_totals;
    XchgLayout x_L = {};
    bool x_created = false, x_connected = false, x_sent = false;
    uint8_t *x_peer[XCHG_MAX_WORLD] = {nullptr};
    bool x_opened[XCHG_MAX_WORLD] = {false};
    uint32_t x_epoch = 0;
    uint64_t x_launches = 0;
    uint64_t x_timeout_ns = 20ull * 1000000000ull;

    // host-side trimming of uploads (chronos_set_host_trim): per worker thread a stream, two pinned and
    // two device staging buffers and the events that guard their reuse
    int trim_threads = 0;
    uint64_t upload_bytes = 0;  // bytes the last host-memory submit sent over PCIe
    struct TrimWorker {
        cudaStream_t stream = nullptr;
        void *pinned[2] = {nullptr, nullptr};
        void *dev[2] = {nullptr, nullptr};
        cudaEvent_t done[2] = {nullptr, nullptr};
    };
    std::vector<TrimWorker> trim;
    cudaStream_t trim_dma = nullptr;    // the copy engine's share of a trimmed upload
    cudaEvent_t ev_trim_dma = nullptr;
    bool trim_no_dma = false;           // CHRONOS_TRIM_NO_DMA: the threads take everything

    uint64_t launches = 0;  // of the submit being enqueued
    int filter_variant = 0;
    int gather_ctas_per_sm = 0;
    int filter_ctas_per_sm = -1;  // > 0: that many; 0: as many as fit; -1 (default): one fewer than fit
    uint64_t fused_max = 0;  // batches whose grouped entries cannot exceed this run the tail as ONE kernel
    int prio_lo = 0;
    bool no_pipeline = false;
    int gather_serial = 0;  // experiments (CHRONOS_GATHER_SERIAL): 1 = gather on the main stream after window_reverse, 2 = before group_forward, 3 = own stream, after group_forward
    // CHRONOS_TRACE=1: CUDA events between the stages of a submit, printed at fetch
    bool trace = false;
    cudaEvent_t tev[24] = {nullptr};
    const char *tname[24] = {nullptr};
    int n_tev = 0;
};

namespace {

const char *g_static_err = "chronos: invalid context";

void mark(chronos_ctx *c, const char *name, cudaStream_t st = nullptr) {
    if (!c->trace || c->n_tev >= 24) return;
    if (!c->tev[c->n_tev]) cudaEventCreate(&c->tev[c->n_tev]);
    cudaEventRecord(c->tev[c->n_tev], st ? st : c->stream);
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

// CHRONOS_DEBUG_SYNC=1: wait after every stage so that a device fault names the stage that raised it
int dbg_sync(chronos_ctx *c, const char *what) {
    static const bool on = getenv("CHRONOS_DEBUG_SYNC") != nullptr;
    if (!on) return CHRONOS_OK;
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return fail(c, CHRONOS_E_CUDA, "after %s: %s", what, cudaGetErrorString(e));
    return CHRONOS_OK;
}

// grow-only: the buffer is re-allocated when it is too small (contents are NOT kept)
int ensure(chronos_ctx *c, DevBuf &b, size_t bytes) {
    if (bytes <= b.bytes && b.p) return CHRONOS_OK;
    if (b.p) {
        CK(cudaFree(b.p));
        b.p = nullptr;
        b.bytes = 0;
    }
    const size_t want = std::max<size_t>(bytes, 256);
    cudaError_t e = cudaMalloc(&b.p, want);
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

// Work the host asks for after a fetch (copies, packing) goes to the main stream while that is
// idle; when a younger batch is already queued there, to a stream of its own.
int host_stream(chronos_ctx *c, cudaStream_t *out) {
    bool busy = false;
    for (BatchSet &b : c->set) busy |= b.state == BatchSet::SUBMITTED;
    if (!busy) {
        *out = c->stream;
        return CHRONOS_OK;
    }
    if (!c->dl_stream) CK(cudaStreamCreateWithPriority(&c->dl_stream, cudaStreamNonBlocking, c->prio_lo));
    *out = c->dl_stream;
    return CHRONOS_OK;
}

DeviceScalars *dscal(BatchSet &b) { return ptr<DeviceScalars>(b.result); }
Plan *dplan(BatchSet &b) { return reinterpret_cast<Plan *>(ptr<uint8_t>(b.result) + 64); }

// ---- small helper kernels ---------------------------------------------------------------
__global__ void set_u32_kernel(uint32_t *p, uint32_t v) { *p = v; }
__global__ void copy_u32_kernel(uint32_t *dst, const uint32_t *src) { *dst = *src; }

// start of a batch: scalars to their neutral values, ticket counters and digit histograms to 0
// (what & 1: the stage-1 half, what & 2: the tail half)
__global__ void init_batch_kernel(DeviceScalars *s, uint32_t *tickets_f, uint32_t *tickets_t, int what) {
    if (what & 1) {
        for (uint32_t i = threadIdx.x; i < N_TICKETS; i += blockDim.x) tickets_f[i] = 0;
        if (threadIdx.x == 0) {
            s->surv_count = 0;
            s->key_or = 0;
            s->key_and = 0xffffffffu;
        }
    }
    if (what & 2) {
        for (uint32_t i = threadIdx.x; i < TICKET_WORDS; i += blockDim.x) tickets_t[i] = 0;
        if (threadIdx.x == 0) {
            s->n_windows = 0;
            s->n_pending = 0;
            s->n_chains = 0;
        }
    }
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
__global__ void extract_pid_kernel(const uint4 *slots, const uint32_t *slot_off, uint32_t n, uint32_t *pid_out,
                                   uint32_t *src, uint32_t *k_or, uint32_t *k_and) {
    uint32_t o = 0, a = 0xffffffffu;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t k = reinterpret_cast<const uint32_t *>(slots + slot_off[i])[0];
        pid_out[i] = k;
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

// One warp per window: a 64-bit digest of the window's records in CANONICAL form -- bytes behind
// a field's first NUL count as zero (that is all the reference ever reads of a field, ctypes
// c_char arrays stop at the NUL, chronos_sensor.py:125-130) and so do the two bytes of tail
// padding. Two runs produce equal digests exactly when analyze_sequence (:147) would be handed
// the same history, record by record, in the same order.
__device__ __forceinline__ uint64_t digest_mix(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
__global__ void window_digest_kernel(const uint8_t *arena, const chronos_window *win, uint32_t nw, uint64_t *out) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const uint32_t n_warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t w = warp; w < nw; w += n_warps) {
        uint64_t acc = 0;
        for (uint32_t r = 0; r < win[w].len; r++) {
            const uint32_t *rec = reinterpret_cast<const uint32_t *>(arena + (size_t)(win[w].start + r) * CH_REC_BYTES);
            uint32_t wd[3];
            uint32_t end_comm = 20, end_argv = 276, end_type = 286;  // byte offset of each field's first NUL
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const uint32_t i = lane + 32 * k;  // word index, 72 words per record
                wd[k] = i < CH_REC_WORDS ? rec[i] : 0u;
                if (i >= 1 && i < CH_REC_WORDS) {
                    uint32_t x = wd[k];
                    if (i == 71) x |= 0xFFFF0000u;  // the padding bytes are not part of `type`
                    const uint32_t z = zero_bytes(x);
                    if (z) {
                        const uint32_t pos = 4 * i + (uint32_t)first_zero_index(z);
                        if (i < 5) end_comm = min(end_comm, pos);
                        else if (i < 69) end_argv = min(end_argv, pos);
                        else end_type = min(end_type, pos);
                    }
                }
            }
            end_comm = __reduce_min_sync(0xffffffffu, end_comm);
            end_argv = __reduce_min_sync(0xffffffffu, end_argv);
            end_type = __reduce_min_sync(0xffffffffu, end_type);
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const uint32_t i = lane + 32 * k;
                if (i >= CH_REC_WORDS) continue;
                uint32_t x = wd[k];
                if (i >= 1) {
                    const uint32_t end = i < 5 ? end_comm : (i < 69 ? end_argv : end_type);
                    const uint32_t b0 = 4 * i;  // keep the bytes before `end`
                    if (end <= b0) x = 0;
                    else if (end < b0 + 4) x = keep_low_bytes(x, (int)(end - b0));
                }
                acc += digest_mix(((uint64_t)(r * CH_REC_WORDS + i) << 32) ^ x ^ 0x9E3779B97F4A7C15ULL);
            }
        }
        for (int d = 16; d >= 1; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
        if (lane == 0) out[w] = digest_mix(acc ^ ((uint64_t)win[w].len << 40) ^ win[w].pid);
    }
}

__global__ void add_base_kernel(const uint32_t *src, uint32_t n, uint64_t base, uint64_t *out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = base + src[i];
}

__global__ void l2_flush_kernel(uint4 *p, size_t n16, uint32_t v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
        p[i] = make_uint4(v, v, v, v);
}

// ---- one-time kernel attributes per device ------------------------------------------------
std::mutex g_attr_lock;
bool g_attr_done[64] = {false};

int configure_kernels(chronos_ctx *c) {
    std::lock_guard<std::mutex> g(g_attr_lock);
    if (c->device < 64 && g_attr_done[c->device]) return CHRONOS_OK;
    CK(cudaFuncSetAttribute(filter_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)filter_smem_bytes<1>()));
#ifndef CHRONOS_SMALL
    CK(cudaFuncSetAttribute(filter_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)filter_smem_bytes<2>()));
#endif
    CK(cudaFuncSetAttribute(xchg_pack, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)XCHG_PACK_SMEM));
    CK(cudaFuncSetAttribute(gather_records, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GATHER_SMEM));
    CK(cudaFuncSetAttribute(fused_tail, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GATHER_SMEM));
    if (!getenv("CHRONOS_NO_CARVEOUT")) {
        // The filter's residency (19 single-warp CTAs per SM) is bounded by shared memory: ask for
        // the largest shared-memory carve-out instead of leaving the split to the driver's guess.
        CK(cudaFuncSetAttribute(filter_kernel<1>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
#ifndef CHRONOS_SMALL
        CK(cudaFuncSetAttribute(filter_kernel<2>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
#endif
    }
    if (c->device < 64) g_attr_done[c->device] = true;
    return CHRONOS_OK;
}

// ---- filter launch ------------------------------------------------------------------------
template <int NSTAGE>
int launch_filter_t(chronos_ctx *c, FilterParams fp, cudaStream_t st) {
    auto kern = filter_kernel<NSTAGE>;
    const size_t smem = filter_smem_bytes<NSTAGE>();
    static int per_sm_of[64] = {0};
    int per_sm = c->device < 64 ? per_sm_of[c->device] : 0;
    if (per_sm == 0) {
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, FILTER_TILE, smem));
        if (per_sm < 1) return fail(c, CHRONOS_E_CUDA, "filter kernel does not fit on an SM (smem %zu)", smem);
        if (c->device < 64) per_sm_of[c->device] = per_sm;
    }
    const uint32_t tiles = fp.tile_end - fp.tile_begin;
    if (c->filter_ctas_per_sm > 0) per_sm = std::min(per_sm, c->filter_ctas_per_sm);
    else if (c->filter_ctas_per_sm < 0 && per_sm > 2) per_sm -= 1;
    const uint32_t grid = std::min<uint32_t>(tiles, (uint32_t)(per_sm * c->sm_count));
    kern<<<grid, FILTER_TILE, smem, st>>>(fp);
    CK(cudaGetLastError());
    c->launches++;
    return CHRONOS_OK;
}

// CHRONOS_FILTER_VARIANT=2 selects the double-buffered build of the filter (tuning knob)
int launch_filter(chronos_ctx *c, const FilterParams &fp, cudaStream_t st) {
#ifndef CHRONOS_SMALL
    if (c->filter_variant == 2) return launch_filter_t<2>(c, fp, st);
#endif
    return launch_filter_t<1>(c, fp, st);
}

// exclusive scan in place (device array of n u32) as part of the stage 1 being enqueued
int launch_scan(chronos_ctx *c, cudaStream_t stream, uint32_t *data, uint64_t n) {
    if (n == 0) return CHRONOS_OK;
    const uint32_t n_tiles = (uint32_t)((n + SCAN_TILE - 1) / SCAN_TILE);
    if ((size_t)n_tiles * 8 > c->scan_status.bytes) return fail(c, CHRONOS_E_STATE, "scan of %llu counters exceeds the status buffer", (unsigned long long)n);
    if (c->scan_epoch == 0) CK(cudaMemsetAsync(c->scan_status.p, 0, c->scan_status.bytes, stream));  // tag 0 = "never"
    if (c->scan_slot >= 32) return fail(c, CHRONOS_E_STATE, "too many scans in one batch");
    ScanParams sp;
    sp.data = data;
    sp.n = (uint32_t)n;
    sp.n_ptr = nullptr;
    sp.ticket = ptr<uint32_t>(c->tickets_f) + 8 + c->scan_slot++;
    sp.status = ptr<uint64_t>(c->scan_status);
    sp.epoch = ++c->scan_epoch;
    if (c->scan_epoch >= (1u << 29)) c->scan_epoch = 0;  // re-zero on the next call
    scan_exclusive_u32<<<std::min<uint32_t>(n_tiles, (uint32_t)c->sm_count * 2), SCAN_THREADS, 0, stream>>>(sp);
    CK(cudaGetLastError());
    c->launches++;
    return CHRONOS_OK;
}

// Parameters of the filter for a batch of n records (no launch); buffers exist since create.
void filter_setup(chronos_ctx *c, BatchSet &b, const uint8_t *d_recs, uint64_t n, FilterParams &fp) {
    DeviceScalars *ds = dscal(b);
    fp.recs = d_recs;
    fp.n = n;
    fp.tile_begin = 0;
    fp.tile_end = (uint32_t)((n + FILTER_TILE - 1) / FILTER_TILE);
    fp.ticket = ptr<uint32_t>(c->tickets_f);
    fp.tile_count = ptr<uint32_t>(c->tile_count);
    fp.sp_key = ptr<uint32_t>(c->sp_key);
    fp.sp_src = ptr<uint32_t>(c->sp_src);
    fp.sp_meta = ptr<uint8_t>(c->sp_meta);
    fp.sp_ns = ptr<uint8_t>(c->sp_ns);
    fp.out_count = &ds->surv_count;
    fp.key_or = &ds->key_or;
    fp.key_and = &ds->key_and;
}

// Cross-tile prefix + dense, arrival-ordered survivor arrays of the set, one kernel.
int launch_compact(chronos_ctx *c, BatchSet &b, const FilterParams &fp, cudaStream_t stream) {
    const uint32_t n_tiles = fp.tile_end;
    const uint32_t n_chunks = (n_tiles + CS_CHUNK - 1) / CS_CHUNK;
    if (c->scan_epoch == 0) CK(cudaMemsetAsync(c->scan_status.p, 0, c->scan_status.bytes, stream));
    if (c->scan_slot >= 32) return fail(c, CHRONOS_E_STATE, "too many scans in one batch");
    CompactScanParams cp;
    cp.tile_count = fp.tile_count;
    cp.n_tiles = n_tiles;
    cp.tile = (uint32_t)FILTER_TILE;
    cp.sp_key = fp.sp_key;
    cp.sp_src = fp.sp_src;
    cp.sp_meta = fp.sp_meta;
    cp.sp_ns = fp.sp_ns;
    cp.out_key = nullptr;
    cp.out_src = ptr<uint32_t>(b.surv_src);
    cp.out_meta = ptr<uint8_t>(b.surv_meta);
    cp.out_ns = ptr<uint8_t>(b.surv_ns);
    cp.out_pid = ptr<uint32_t>(b.surv_pid);
    cp.ticket = ptr<uint32_t>(c->tickets_f) + 8 + c->scan_slot++;
    cp.status = ptr<uint64_t>(c->scan_status);
    cp.epoch = ++c->scan_epoch;
    if (c->scan_epoch >= (1u << 29)) c->scan_epoch = 0;
    compact_scan<<<std::max(1u, std::min<uint32_t>(n_chunks, (uint32_t)c->sm_count * 4)), CS_THREADS, 0, stream>>>(cp);
    CK(cudaGetLastError());
    c->launches += 1;
    return CHRONOS_OK;
}

// filter + cross-tile compaction: dense, arrival-ordered survivors of the set (pid -> surv_pid,
// batch-local index -> surv_src, META bits -> surv_meta, count -> scalars.surv_count)
int run_filter_stage(chronos_ctx *c, BatchSet &b, const uint8_t *d_recs, uint64_t n, cudaStream_t st) {
    int rc;
    FilterParams fp;
    filter_setup(c, b, d_recs, n, fp);
    CK(cudaEventRecord(b.ev_f0, st));
    if ((rc = dbg_sync(c, "begin_batch"))) return rc;
    if ((rc = launch_filter(c, fp, st))) return rc;
    CK(cudaEventRecord(b.ev_f1, st));
    if ((rc = dbg_sync(c, "filter_kernel"))) return rc;
    b.filter_timed = true;
    if ((rc = launch_compact(c, b, fp, st))) return rc;
    if ((rc = dbg_sync(c, "filter + scan + compact"))) return rc;
    return CHRONOS_OK;
}

// ---- memory ----------------------------------------------------------------------------------
// status words needed by the window kernels for M grouped entries
size_t status_words(uint64_t M) {
    const uint64_t tiles_m = (M + WIN_TILE - 1) / WIN_TILE;
    return (size_t)(8 * tiles_m + 64);
}

// Buffers shared by all batches, for batches of up to n raw events and M grouped entries.
int reserve_shared(chronos_ctx *c, uint64_t n, uint64_t M) {
    int rc;
    n = std::max(n, c->cap_n);
    M = std::max(M, c->cap_M);
    if ((rc = ensure(c, c->keyA, (M + 1) * 4))) return rc;
    if ((rc = ensure(c, c->keyB, (M + 1) * 4))) return rc;
    if ((rc = ensure(c, c->valA, (M + 1) * 4))) return rc;
    if ((rc = ensure(c, c->valB, (M + 1) * 4))) return rc;
    if ((rc = ensure(c, c->pflag, (M + 4) * 4))) return rc;
    if ((rc = ensure(c, c->entry_out, (n + 4) * 4))) return rc;
    // a window holds at least two grouped entries and exactly one trigger survivor of its batch
    if ((rc = ensure(c, c->wsorted, (std::min(n, M / 2) + 1) * sizeof(WinSorted)))) return rc;
    const size_t sw_before = c->status.bytes;
    if ((rc = ensure(c, c->status, status_words(M) * 8))) return rc;
    (void)sw_before;
    {
        const uint64_t rtiles = std::max<uint64_t>((M + RADIX_TILE_KEYS - 1) / RADIX_TILE_KEYS, 1);
        const void *before = c->sweep_status.p;
        if ((rc = ensure(c, c->sweep_status, rtiles * RADIX_BINS * 8))) return rc;
        if (c->sweep_status.p != before) {  // fresh memory: tag 0 = "never written"
            CK(cudaMemset(c->sweep_status.p, 0, c->sweep_status.bytes));
            c->sweep_epoch = 0;
        }
    }
    const uint64_t tiles = (n + FILTER_TILE - 1) / FILTER_TILE;
    const uint64_t slots = tiles * FILTER_TILE;
    if ((rc = ensure(c, c->sp_key, slots * 4 + 16))) return rc;
    if ((rc = ensure(c, c->sp_src, slots * 4 + 16))) return rc;
    if ((rc = ensure(c, c->sp_meta, slots + 16))) return rc;
    if ((rc = ensure(c, c->sp_ns, slots + 16))) return rc;
    if ((rc = ensure(c, c->tile_count, (tiles + 4) * 4))) return rc;
    {
        // scans of stage 1: tile counts (chunks of CS_CHUNK tiles), slot counts of n entries,
        // the owner table of an exchange (2 * world * ceil(n / XCHG_TILE) + 1 counters)
        const uint64_t owner = 2ull * XCHG_MAX_WORLD * ((n + XCHG_TILE - 1) / XCHG_TILE + 1) + 1;
        const uint64_t words = std::max<uint64_t>({(tiles + CS_CHUNK - 1) / CS_CHUNK, (n + SCAN_TILE - 1) / SCAN_TILE,
                                                   (owner + SCAN_TILE - 1) / SCAN_TILE, 512});
        const void *before = c->scan_status.p;
        if ((rc = ensure(c, c->scan_status, words * 8))) return rc;
        if (c->scan_status.p != before) {
            CK(cudaMemset(c->scan_status.p, 0, c->scan_status.bytes));
            c->scan_epoch = 0;
        }
    }
    c->cap_n = n;
    c->cap_M = M;
    return CHRONOS_OK;
}

// One buffer set, for a batch of up to n raw events on top of carried-in state, M entries in all.
int reserve_set(chronos_ctx *c, BatchSet &b, uint64_t n, uint64_t M) {
    int rc;
    n = std::max(n, b.cap_n);
    M = std::max(M, b.cap_M);
    if ((rc = ensure(c, b.arena, (M + 1) * CH_REC_BYTES))) return rc;
    if ((rc = ensure(c, b.pend_key, (M + 1) * 4))) return rc;
    if ((rc = ensure(c, b.pend_slot, (M + 1) * 4))) return rc;
    if ((rc = ensure(c, b.pend_arrival, (M + 1) * 8))) return rc;
    if ((rc = ensure(c, b.pend_meta, M + 1))) return rc;
    if ((rc = ensure(c, b.surv_src, (n + 1) * 4))) return rc;
    if ((rc = ensure(c, b.surv_meta, n + 1))) return rc;
    if ((rc = ensure(c, b.surv_pid, (n + 1) * 4))) return rc;
    if ((rc = ensure(c, b.surv_ns, n + 1))) return rc;
    if ((rc = ensure(c, b.windows, (std::min(n, M / 2) + 1) * sizeof(chronos_window)))) return rc;
    if ((rc = ensure(c, b.chain_pid, (n + 1) * 4))) return rc;
    if ((rc = ensure(c, b.chain_idx, (n + 1) * 8))) return rc;
    b.cap_n = n;
    b.cap_M = M;
    b.pend_cap = M;
    return CHRONOS_OK;
}

// arrays only the transport-form submits need (chronos_submit_survivors, the exchange)
int reserve_transport(chronos_ctx *c, BatchSet &b) {
    int rc;
    if ((rc = ensure(c, b.surv_arrival, (b.cap_n + 1) * 8))) return rc;
    if ((rc = ensure(c, b.slot_off, (b.cap_n + 8) * 4))) return rc;
    return CHRONOS_OK;
}

// The per-PID table must stay at most half full: room for every PID seen so far plus one new
// PID per grouped entry of the batch about to run. Growing it waits for the GPU.
int ensure_pid_table(chronos_ctx *c, uint64_t pids_ub, uint64_t M) {
    const uint64_t need = 2 * (pids_ub + M) + 64;
    if (c->pt_cap >= need) return CHRONOS_OK;
    uint64_t cap = 1024;
    while (cap < need) cap <<= 1;
    if (cap > (1ull << 31)) return fail(c, CHRONOS_E_CAPACITY, "per-PID table would need %llu slots", (unsigned long long)cap);
    CK(cudaDeviceSynchronize());  // batches in flight use the old table
    DevBuf fresh;
    int rc;
    if ((rc = ensure(c, fresh, cap * 8))) return rc;
    auto grow = [&]() -> int {
        CK(cudaMemsetAsync(fresh.p, 0, cap * 8, c->stream));
        if (c->pt_cap) {
            pt_rehash_kernel<<<(unsigned)std::min<uint64_t>((c->pt_cap + 255) / 256, 4096), 256, 0, c->stream>>>(
                ptr<uint64_t>(c->pid_tab), c->pt_cap, static_cast<uint64_t *>(fresh.p), (uint32_t)cap - 1, c->pt_epoch);
            CK(cudaGetLastError());
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

__global__ void set_carry_kernel(Carry *c, uint32_t P, uint32_t k_or, uint32_t k_and) {
    c->P = P;
    c->key_or = k_or;
    c->key_and = k_and;
}

// (queued on the main stream, ahead of the next batch's tail)
int set_carry(chronos_ctx *c, uint32_t P, uint32_t k_or, uint32_t k_and) {
    set_carry_kernel<<<1, 1, 0, c->stream>>>(ptr<Carry>(c->carry), P, k_or, k_and);
    CK(cudaGetLastError());
    return CHRONOS_OK;
}

// ---- the tail -----------------------------------------------------------------------------------
// Everything after the survivors of set `b` are known ON THE DEVICE (surv_pid / surv_src /
// surv_meta, scalars.surv_count / key_or / key_and). Nothing here waits for the GPU: buffers are
// sized for the worst case, grids are upper bounds, extents come from the device plan.
//
//   main stream   sort_prepare, passes, group_forward, window_reverse
//   stream2       gather_records   (after the sort; bandwidth-bound)
//   stream3       emit_ordered     (after group_forward; independent of window_reverse)
// The main stream joins both before the result block goes to the host, so the next batch's
// tail (same stream) starts after everything of this one.
int run_window_stage(chronos_ctx *c, BatchSet &b, BatchSet &prev) {
    const uint64_t Mmax = b.Mmax;  // upper bound of P + n
    DeviceScalars *ds = dscal(b);
    Plan *plan = dplan(b);
    cudaStream_t st = c->stream;
    mark(c, "filter+compact done");
    static const bool prof_tail = getenv("CHRONOS_PROFILE_TAIL") != nullptr;  // ncu range replay over the concurrent tail
    if (prof_tail) { cudaDeviceSynchronize(); cudaProfilerStart(); }
    const uint64_t Smax = b.n;
    int rc;
    const uint32_t max_rtiles = (uint32_t)((Mmax + RADIX_TILE_KEYS - 1) / RADIX_TILE_KEYS);
    uint32_t *tk = ptr<uint32_t>(c->tickets_t);
    // The tail's counters and look-back words start at zero. Nothing of stage 1 is needed for
    // that, so it is done on stream3 while the filter runs (behind the previous batch's tail, which
    // uses the same words), and the main stream picks it up with an event.
    if (prev.state == BatchSet::SUBMITTED) CK(cudaStreamWaitEvent(c->stream3, prev.ev_end, 0));
    CK(cudaMemsetAsync(c->status.p, 0, status_words(Mmax) * 8, c->stream3));
    init_batch_kernel<<<1, 256, 0, c->stream3>>>(ds, ptr<uint32_t>(c->tickets_f), tk, 2);
    c->launches++;
    CK(cudaEventRecord(c->ev_tailinit, c->stream3));
    CK(cudaStreamWaitEvent(st, c->ev_tailinit, 0));

    // ---- parameters of every tail kernel --------------------------------------------------------
    SortParams sp;
    memset(&sp, 0, sizeof(sp));
    sp.plan = plan;
    sp.in.surv_count = &ds->surv_count;
    sp.in.key_or = &ds->key_or;
    sp.in.key_and = &ds->key_and;
    sp.in.carry = ptr<Carry>(c->carry);
    sp.in.m_max = (uint32_t)std::min<uint64_t>(Mmax, std::min(c->cap_M, b.cap_M));
    sp.in.win_tile = WIN_TILE;
    sp.in.s_max = (uint32_t)std::min<uint64_t>(b.n, b.cap_n);
    sp.in.w_max = (uint32_t)(std::min(std::min(c->cap_n, b.cap_n), std::min(c->cap_M, b.cap_M) / 2) + 1);
    sp.keyA = ptr<uint32_t>(c->keyA);
    sp.keyB = ptr<uint32_t>(c->keyB);
    sp.valA = ptr<uint32_t>(c->valA);
    sp.valB = ptr<uint32_t>(c->valB);
    sp.pend_key = ptr<uint32_t>(prev.pend_key);
    sp.pend_meta = ptr<uint8_t>(prev.pend_meta);
    sp.pend_cap = (uint32_t)prev.pend_cap;
    sp.surv_pid = ptr<uint32_t>(b.surv_pid);
    sp.surv_meta = ptr<uint8_t>(b.surv_meta);
    sp.ghist = tk + N_TICKETS;  // zeroed with the tickets
    if (c->sweep_epoch == 0) CK(cudaMemsetAsync(c->sweep_status.p, 0, c->sweep_status.bytes, st));  // wrapped epoch
    sp.status = ptr<uint64_t>(c->sweep_status);
    uint32_t pass_epoch[RADIX_MAX_PASSES];
    for (int i = 0; i < RADIX_MAX_PASSES; i++) {
        pass_epoch[i] = ++c->sweep_epoch;
        if (c->sweep_epoch >= (1u << 29)) c->sweep_epoch = 0;  // re-zero on the next batch
    }

    WinParams w;
    memset(&w, 0, sizeof(w));
    w.plan = plan;
    w.surv_src = ptr<uint32_t>(b.surv_src);
    w.surv_arrival = b.explicit_arrival ? ptr<uint64_t>(b.surv_arrival) : nullptr;
    w.base_arrival = b.base;
    w.pend_arrival = ptr<uint64_t>(prev.pend_arrival);
    w.pend_cap_prev = (uint32_t)prev.pend_cap;
    w.carry = ptr<Carry>(c->carry);
    w.surv_pid = ptr<uint32_t>(b.surv_pid);
    w.pid_tab = ptr<uint64_t>(c->pid_tab);
    w.pid_mask = c->pt_cap - 1;
    w.pid_epoch = c->pt_epoch;
    w.pid_count = ptr<uint32_t>(c->pt_count);
    w.n_pids_out = &ds->n_pids_seen;
    w.pflag = ptr<uint32_t>(c->pflag);
    w.entry_out = ptr<uint32_t>(c->entry_out);
    w.wsorted = ptr<WinSorted>(c->wsorted);
    w.n_windows = &ds->n_windows;
    w.pend_cap = (uint32_t)b.pend_cap;
    w.npend_key = ptr<uint32_t>(b.pend_key);
    w.npend_slot = ptr<uint32_t>(b.pend_slot);
    w.npend_arrival = ptr<uint64_t>(b.pend_arrival);
    w.npend_meta = ptr<uint8_t>(b.pend_meta);
    w.n_pending = &ds->n_pending;
    w.windows = ptr<chronos_window>(b.windows);
    w.chain_pid = ptr<uint32_t>(b.chain_pid);
    w.chain_idx = ptr<uint64_t>(b.chain_idx);
    w.n_chains = &ds->n_chains;
    const uint32_t tiles_m = (uint32_t)((Mmax + WIN_TILE - 1) / WIN_TILE);  // upper bounds
    const uint32_t tiles_s = (uint32_t)((Smax + WIN_TILE - 1) / WIN_TILE);
    uint64_t *stw = ptr<uint64_t>(c->status);
    WinParams w_fwd = w, w_rev = w, w_emit = w;
    w_fwd.ticket = tk + 60;
    w_fwd.status = stw;  // 5 words per tile
    w_rev.ticket = tk + 61;
    w_rev.status = stw + 5 * (size_t)tiles_m;  // 2 words per tile
    w_emit.ticket = tk + 62;
    w_emit.status = stw + 7 * (size_t)tiles_m;  // 1 word per tile

    GatherParams g;
    memset(&g, 0, sizeof(g));
    g.plan = plan;
    g.surv_src = ptr<uint32_t>(b.surv_src);
    g.batch = b.records;
    g.pend_slot = ptr<uint32_t>(prev.pend_slot);
    g.pend_cap_prev = (uint32_t)prev.pend_cap;
    g.prev_arena = ptr<uint8_t>(prev.arena);
    g.arena = ptr<uint8_t>(b.arena);
    g.slots = b.slots;
    g.slot_off = b.slots ? ptr<uint32_t>(b.slot_off) : nullptr;
    g.slot_ns = b.slots ? ptr<uint8_t>(b.surv_ns) : nullptr;

    auto finish = [&]() -> int {
        if (prof_tail) { cudaDeviceSynchronize(); cudaProfilerStop(); }
        CK(cudaMemcpyAsync(b.h_scalars, ds, RESULT_BLOCK_BYTES, cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(b.ev_end, st));
        return CHRONOS_OK;
    };

    // ---- small batches: the whole tail in one launch (window.cuh fused_tail) ----------------------
    if (Mmax != 0 && Mmax <= c->fused_max) {
        FusedTailParams f;
        memset(&f, 0, sizeof(f));
        f.sort = sp;
        for (int i = 0; i < RADIX_MAX_PASSES; i++) {
            f.pass_ticket[i] = tk + 40 + i;
            f.pass_epoch[i] = pass_epoch[i];
        }
        f.fwd = w_fwd;
        f.rev = w_rev;
        f.emit = w_emit;
        f.gather = g;
        f.barrier = tk + 63;
        // every CTA must be resident (they meet at grid barriers): never more CTAs than SMs
        const uint32_t grid = std::max(8u, std::min<uint32_t>((uint32_t)c->sm_count, (uint32_t)((Mmax + 1023) / 1024)));
        CK(cudaEventRecord(b.ev_sort, st));
        // a cooperative launch: the grid starts only when all of it fits on the GPU at once, so two
        // contexts on one GPU cannot end up holding half of the SMs each while both wait at a barrier
        void *args[] = {&f};
        CK(cudaLaunchCooperativeKernel(reinterpret_cast<const void *>(fused_tail), dim3(grid), dim3(WIN_THREADS), args,
                                       GATHER_SMEM, st));
        c->launches++;
        if ((rc = dbg_sync(c, "fused_tail"))) return rc;
        mark(c, "fused tail done");
        return finish();
    }

    // ---- stable sort by pid: plan + global digit histograms, then one kernel per pass --------
    const unsigned sort_blocks = std::max(1u, std::min(max_rtiles, (uint32_t)c->sm_count * 4));
    sort_prepare<<<sort_blocks, RADIX_THREADS, 0, st>>>(sp);
    c->launches++;
    if ((rc = dbg_sync(c, "sort_prepare"))) return rc;
    if (Mmax == 0) {  // nothing pending, nothing submitted: only the (empty) counts go back
        copy_u32_kernel<<<1, 1, 0, st>>>(&ds->n_pids_seen, ptr<uint32_t>(c->pt_count));
        return finish();
    }
    for (int i = 0; i < RADIX_MAX_PASSES; i++) {  // passes beyond plan->n_passes return at once
        sp.pass = i;
        sp.ticket = tk + 40 + i;
        sp.epoch = pass_epoch[i];
        onesweep_pass<<<sort_blocks, RADIX_THREADS, 0, st>>>(sp);
        c->launches++;
        if ((rc = dbg_sync(c, "onesweep_pass"))) return rc;
    }
    CK(cudaGetLastError());
    CK(cudaEventRecord(c->ev_sorted, st));
    CK(cudaEventRecord(b.ev_sort, st));
    mark(c, "sort done");
    const uint32_t cap_blocks = (uint32_t)c->sm_count * (WIN_ITEMS >= 16 ? 2 : 4);

    // ---- records into the grouped arena: a bandwidth-bound copy on its own stream. It keeps
    // its data in flight in shared memory and takes a bounded number of CTAs per SM, so that the
    // latency-bound scans find registers and warp slots on every SM beside it (a gather that
    // fills the SMs, or scans that fill the register file, serialise the two: profiles/README.md).
    auto launch_gather = [&](cudaStream_t gs) -> int {
        const uint32_t blocks = (uint32_t)std::min<uint64_t>((Mmax + GATHER_WARPS * GATHER_GROUP - 1) / (GATHER_WARPS * GATHER_GROUP),
                                                             (uint64_t)c->sm_count * (uint64_t)c->gather_ctas_per_sm);
        if (gs != st) CK(cudaStreamWaitEvent(gs, c->gather_serial == 3 ? c->ev_fwd : c->ev_sorted, 0));
        mark(c, "gather starts", gs);
        gather_records<<<std::max(1u, blocks), GATHER_THREADS, GATHER_SMEM, gs>>>(g);
        CK(cudaGetLastError());
        if (gs != st) CK(cudaEventRecord(c->ev_gathered, gs));
        mark(c, "gather done", gs);
        c->launches++;
        return dbg_sync(c, "gather_records");
    };
    if (c->gather_serial == 0 && (rc = launch_gather(c->stream2))) return rc;
    if (c->gather_serial == 2 && (rc = launch_gather(st))) return rc;

    // ---- flush automaton + chain labeller ----------------------------------------------------
    group_forward<<<std::max(1u, std::min(tiles_m, cap_blocks)), WIN_THREADS, 0, st>>>(w_fwd);
    c->launches++;
    CK(cudaGetLastError());
    if ((rc = dbg_sync(c, "group_forward"))) return rc;
    CK(cudaEventRecord(c->ev_fwd, st));
    mark(c, "group_forward done");
    if (c->gather_serial == 3 && (rc = launch_gather(c->stream2))) return rc;  // beside window_reverse / emit_ordered only

    // ---- windows in arrival order (stream3) beside the reverse scan (main stream) ----------
    CK(cudaStreamWaitEvent(c->stream3, c->ev_fwd, 0));
    emit_ordered<<<std::max(1u, std::min(tiles_s, cap_blocks)), WIN_THREADS, 0, c->stream3>>>(w_emit);
    CK(cudaGetLastError());
    CK(cudaEventRecord(c->ev_emitted, c->stream3));
    mark(c, "emit_ordered done", c->stream3);
    c->launches++;
    if ((rc = dbg_sync(c, "emit_ordered"))) return rc;
    window_reverse<<<std::max(1u, std::min(tiles_m, cap_blocks)), WIN_THREADS, 0, st>>>(w_rev);
    c->launches++;
    CK(cudaGetLastError());
    if ((rc = dbg_sync(c, "window_reverse"))) return rc;
    mark(c, "window_reverse done");
    if (c->gather_serial == 1 && (rc = launch_gather(st))) return rc;
    CK(cudaStreamWaitEvent(st, c->ev_emitted, 0));
    if (c->gather_serial == 0 || c->gather_serial == 3) CK(cudaStreamWaitEvent(st, c->ev_gathered, 0));
    mark(c, "all joined");
    return finish();
}

// ---- batch bookkeeping --------------------------------------------------------------------------
// Wait for a submitted batch and read its result block: sizes, what it left pending.
int complete_batch(chronos_ctx *c, BatchSet &b) {
    if (b.state != BatchSet::SUBMITTED) return CHRONOS_OK;
    CK(cudaEventSynchronize(b.ev_end));
    b.state = BatchSet::DONE;
    b.rc = CHRONOS_OK;
    const bool any = b.Mmax != 0;  // (else only sort_prepare ran: an empty plan)
    b.S = b.h_plan->S;
    b.M = b.h_plan->M;
    b.W = any && b.M ? b.h_scalars->n_windows : 0;
    b.chains = any && b.M ? b.h_scalars->n_chains : 0;
    b.pending_after = any && b.M ? b.h_scalars->n_pending : 0;
    b.pids_after = b.h_scalars->n_pids_seen;
    const bool newest = b.seq == c->seq;
    BatchSet &y = c->set[(&b == &c->set[0]) ? 1 : 0];
    if (newest) {
        if (b.gen == c->gen) {  // (else the state was reset after this batch was submitted)
            c->n_pending = b.pending_after;
            c->n_pids_seen = b.pids_after;
        }
    } else if (y.gen == b.gen) {  // a younger batch of the same state is in flight: what it carries in is now exact
        y.P_ub = std::min(y.P_ub, b.pending_after);
        y.pids_ub = std::min(y.pids_ub, b.pids_after);
    }
    if (any && b.h_plan->overflow)
        b.rc = fail(c, CHRONOS_E_CAPACITY, "more pending + surviving events than the context has room for (%llu)",
                    (unsigned long long)std::min(c->cap_M, b.cap_M));
    if (b.exchanged) {
        b.n = b.S;  // the upper bound the buffers were sized for is of no interest to the caller
        const uint32_t xe = b.h_scalars->xchg_err;
        if (xe & XERR_TIMEOUT) {
            // which sources are missing: the flag words of this step's half, read back for the message
            uint32_t flags[XCHG_MAX_WORLD] = {0};
            const uint8_t *hb = ptr<uint8_t>(c->x_window) + (size_t)(b.x_epoch & 1u) * c->x_L.half_bytes;
            cudaMemcpy(flags, hb + 2 * XCHG_MAX_WORLD * 4, sizeof(flags), cudaMemcpyDeviceToHost);
            char miss[96] = {0};
            size_t o = 0;
            for (uint32_t d = 0; d < c->x_L.world && o + 8 < sizeof(miss); d++)
                if (flags[d] != b.x_epoch) o += (size_t)snprintf(miss + o, sizeof(miss) - o, " %u(flag %u)", d, flags[d]);
            b.rc = fail(c, CHRONOS_E_STATE, "exchange step %u: no delivery within %.0f s from source rank(s)%s",
                        b.x_epoch, (double)c->x_timeout_ns * 1e-9, miss);
        }
        else if (xe & (XERR_OVERFLOW | XERR_PEER_OVERFLOW))
            b.rc = fail(c, CHRONOS_E_CAPACITY,
                        "exchange: survivors for one owner rank exceed the window (%u entries / %u slots per peer)",
                        c->x_L.cap_e, c->x_L.cap_s);
    }
    cudaEventElapsedTime(&b.t_total, b.exchanged ? b.ev_x0 : b.ev_start, b.ev_end);
    if (c->trace && !b.pipelined) {
        for (int i = 1; i < c->n_tev; i++) {  // time since the first mark (marks sit on several streams)
            float ms = 0;
            cudaEventSynchronize(c->tev[i]);
            cudaEventElapsedTime(&ms, c->tev[0], c->tev[i]);
            fprintf(stderr, "[chronos trace] %-28s at %8.3f ms\n", c->tname[i], ms);
        }
    }
    {   // stage boundaries of the batch (chronos_last_stage_times)
        float *t = b.stage_ms;
        for (int i = 0; i < 8; i++) t[i] = 0.f;
        const bool filtered = b.filter_timed || (b.exchanged && b.x_filter_timed);
        if (filtered) cudaEventElapsedTime(&t[0], b.ev_f0, b.ev_f1);
        if (b.exchanged) {
            if (filtered) cudaEventElapsedTime(&t[1], b.ev_f1, b.ev_prepack);
            cudaEventElapsedTime(&t[2], b.ev_prepack, b.ev_s1end);
            cudaEventElapsedTime(&t[3], b.ev_start, b.ev_stage1);
        } else if (filtered) {
            cudaEventElapsedTime(&t[1], b.ev_f1, b.ev_stage1);
        }
        if (any) {
            cudaEventElapsedTime(&t[4], b.ev_stage1, b.ev_sort);
            cudaEventElapsedTime(&t[5], b.ev_sort, b.ev_end);
        }
        cudaGetLastError();
    }
    cudaEventElapsedTime(&b.t_filter, b.ev_start, b.ev_stage1);
    // (a stage-2-only batch keeps the filter time measured by its stage 1)
    if (b.filter_timed || (b.exchanged && b.x_filter_timed)) cudaEventElapsedTime(&b.t_filter_kernel, b.ev_f0, b.ev_f1);
    cudaGetLastError();
    return CHRONOS_OK;
}

// Upper bounds of what the next batch carries in, from the host's point of view.
void carried_bounds(chronos_ctx *c, uint64_t &P_ub, uint64_t &pids_ub) {
    BatchSet &l = c->set[c->last];
    if (l.state == BatchSet::SUBMITTED && l.gen == c->gen) {
        P_ub = l.P_ub + l.n;
        pids_ub = l.pids_ub + l.n;
    } else {
        P_ub = c->n_pending;
        pids_ub = c->n_pids_seen;
    }
}

// Start of a submit of n events. Everything that can refuse the batch happens BEFORE any state
// of the context changes: a refused submit leaves process_memory exactly as it was.
// On success *out is the set the batch runs in and *sf the stream of its stage 1.
int begin_batch(chronos_ctx *c, uint64_t n, uint64_t base, bool transport, BatchSet **out, cudaStream_t *sf,
                bool main_stream_only = false) {
    CK(cudaSetDevice(c->device));
    if (n > c->max_batch) return fail(c, CHRONOS_E_CAPACITY, "batch of %llu events > max_batch_events %llu",
                                      (unsigned long long)n, (unsigned long long)c->max_batch);
    if (base >= MAX_ARRIVAL || base + n > MAX_ARRIVAL)
        return fail(c, CHRONOS_E_ARG, "arrival indices must stay below 2^48");
    const int t = c->last ^ 1;
    BatchSet &b = c->set[t], &prev = c->set[c->last];
    int rc;
    if (b.state == BatchSet::SUBMITTED) {  // two batches ago, never fetched: its results are dropped
        if ((rc = complete_batch(c, b))) return rc;
    }
    uint64_t P_ub, pids_ub;
    carried_bounds(c, P_ub, pids_ub);
    auto short_of = [&](uint64_t M) { return M > MAX_GROUPED || M > c->cap_M || M > b.cap_M || n > c->cap_n || n > b.cap_n; };
    if (short_of(P_ub + n) && prev.state == BatchSet::SUBMITTED) {
        // the bound is too coarse: wait for the batch in flight and use the exact count
        if ((rc = complete_batch(c, prev))) return rc;
        carried_bounds(c, P_ub, pids_ub);
    }
    const uint64_t Mmax = P_ub + n;
    if (Mmax > MAX_GROUPED)
        return fail(c, CHRONOS_E_CAPACITY, "%llu pending + %llu new events: more than 2^27 - 1 grouped entries possible",
                    (unsigned long long)P_ub, (unsigned long long)n);
    if (short_of(Mmax)) {
        // The pending lists outgrew what chronos_create reserved: the ONE place where a submit
        // allocates. Everything in flight finishes first (its results stay fetchable); the shared
        // temporaries and the set about to be written hold nothing that must survive.
        if (prev.state == BatchSet::SUBMITTED && (rc = complete_batch(c, prev))) return rc;
        CK(cudaDeviceSynchronize());
        const uint64_t want = std::min<uint64_t>(MAX_GROUPED, Mmax + Mmax / 4);
        if ((rc = reserve_shared(c, c->max_batch, want))) return rc;
        if ((rc = reserve_set(c, b, c->max_batch, want))) return rc;
    }
    if (transport && (rc = reserve_transport(c, b))) return rc;
    if ((rc = ensure_pid_table(c, pids_ub, Mmax))) return rc;

    // ---- from here on the batch is accepted ---------------------------------------------------
    // (main_stream_only: the caller's stage 1 already sits on the main stream, behind the previous batch)
    const bool pipelined = prev.state == BatchSet::SUBMITTED && !c->no_pipeline && !main_stream_only;
    if (prev.state == BatchSet::SUBMITTED && c->no_pipeline && !main_stream_only && (rc = complete_batch(c, prev))) return rc;
    if (c->fetched == t) c->fetched = -1;
    b.state = BatchSet::FREE;
    b.n = n;
    b.base = base;
    b.P_ub = P_ub;
    b.pids_ub = pids_ub;
    b.Mmax = Mmax;
    b.S = b.M = b.W = b.chains = 0;
    b.records = nullptr;
    b.slots = nullptr;
    b.explicit_arrival = b.exchanged = b.filter_timed = false;
    b.pipelined = pipelined;
    b.gen = c->gen;
    b.rc = CHRONOS_OK;
    c->launches = 0;
    c->scan_slot = 0;
    if (pipelined && !c->stream_f) CK(cudaStreamCreateWithPriority(&c->stream_f, cudaStreamNonBlocking, c->prio_lo));
    cudaStream_t s1 = pipelined ? c->stream_f : c->stream;
    // the stage-1 temporaries are shared; without overlap the batch starts when the previous one has ended
    if (pipelined) CK(cudaStreamWaitEvent(s1, c->overlap ? prev.ev_stage1 : prev.ev_end, 0));
    CK(cudaStreamWaitEvent(s1, b.ev_packed, 0));                    // a download may still read this set
    CK(cudaEventRecord(b.ev_start, s1));
    if (!c->x_sent && !pipelined) {  // (stage 2 of an exchange step keeps the marks of its stage 1)
        c->n_tev = 0;
        mark(c, "start");
    }
    *out = &b;
    *sf = s1;
    return CHRONOS_OK;
}

// The batch is fully enqueued: it becomes the newest one.
void commit_batch(chronos_ctx *c, BatchSet &b) {
    b.seq = ++c->seq;
    b.launches = c->launches;
    b.state = BatchSet::SUBMITTED;
    c->last = (&b == &c->set[0]) ? 0 : 1;
}

// stage 1 -> tail hand-over: the tail stream waits for the survivors when they were produced on
// the other stream
int join_stage1(chronos_ctx *c, BatchSet &b, cudaStream_t s1) {
    CK(cudaEventRecord(b.ev_stage1, s1));
    if (s1 != c->stream) CK(cudaStreamWaitEvent(c->stream, b.ev_stage1, 0));
    return CHRONOS_OK;
}

int no_batch_in_flight(chronos_ctx *c) {
    for (BatchSet &b : c->set)
        if (b.state == BatchSet::SUBMITTED || b.state == BatchSet::DONE) return fail(c, CHRONOS_E_STATE, "call chronos_fetch first");
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

int chronos_create_sized(int device, uint64_t max_batch_events, uint64_t max_pending_events, chronos_ctx **out) {
    if (!out || max_batch_events == 0 || max_batch_events > MAX_GROUPED) return CHRONOS_E_ARG;
    *out = nullptr;
    if (chronos_device_ok(device) != CHRONOS_OK) return CHRONOS_E_CUDA;
    chronos_ctx *c = new (std::nothrow) chronos_ctx();
    if (!c) return CHRONOS_E_NOMEM;
    c->device = device;
    c->max_batch = max_batch_events;
    c->max_pending = std::min<uint64_t>(max_pending_events, MAX_GROUPED - max_batch_events);
    const char *fv = getenv("CHRONOS_FILTER_VARIANT");
    c->filter_variant = fv ? atoi(fv) : -1;  // -1: the tuned default (TILE 32, 1 stage)
    const char *gc = getenv("CHRONOS_GATHER_CTAS");
    c->gather_ctas_per_sm = gc ? std::max(1, atoi(gc)) : 2;
    c->no_pipeline = getenv("CHRONOS_NO_PIPELINE") != nullptr;
    c->gather_serial = getenv("CHRONOS_GATHER_SERIAL") ? atoi(getenv("CHRONOS_GATHER_SERIAL")) : 0;
    // One CTA per SM fewer than shared memory would allow (18 instead of 19 on B200). At full
    // residency the filter is bistable: whole runs of batches take 0.72-0.87 ms instead of 0.49 ms
    // per 10 M events, depending on what the previous kernels left in L2; at 18 it is steady at
    // 0.50 ms, at 17 at 0.515 ms (sweep in profiles/README.md, "filter residency").
    c->filter_ctas_per_sm = getenv("CHRONOS_FILTER_CTAS") ? atoi(getenv("CHRONOS_FILTER_CTAS")) : -1;
    // (CHRONOS_FUSED_MAX=0 switches the one-launch tail off)
    c->fused_max = getenv("CHRONOS_FUSED_MAX") ? strtoull(getenv("CHRONOS_FUSED_MAX"), nullptr, 10) : 262144ull;
    const bool gather_low = getenv("CHRONOS_GATHER_LOW") != nullptr;
    c->trace = getenv("CHRONOS_TRACE") != nullptr;
    auto boot = [&]() -> int {
        CK(cudaSetDevice(device));
        c->sm_count = device_sm_count(device);
        int rc;
        if ((rc = configure_kernels(c))) return rc;
        int prio_lo = 0, prio_hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        // the latency-bound scans get priority over the bandwidth-bound work beside them
        CK(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio_hi));
        CK(cudaStreamCreateWithPriority(&c->stream3, cudaStreamNonBlocking, prio_hi));
        CK(cudaStreamCreateWithPriority(&c->stream2, cudaStreamNonBlocking, gather_low ? prio_lo : prio_hi));
        c->prio_lo = prio_lo;  // stream_f and dl_stream are created when first needed
        for (cudaEvent_t *e : {&c->ev_sorted, &c->ev_gathered, &c->ev_fwd, &c->ev_emitted, &c->ev_tailinit, &c->dl_ready,
                               &c->dl_done[0], &c->dl_done[1]})
            CK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
        for (BatchSet &b : c->set) {
            for (cudaEvent_t *e : {&b.ev_start, &b.ev_stage1, &b.ev_end, &b.ev_f0, &b.ev_f1, &b.ev_prepack, &b.ev_s1end, &b.ev_sort, &b.ev_x0})
                CK(cudaEventCreate(e));
            CK(cudaEventCreateWithFlags(&b.ev_packed, cudaEventDisableTiming));
            CK(cudaMallocHost(reinterpret_cast<void **>(&b.h_scalars), RESULT_BLOCK_BYTES));
            memset(b.h_scalars, 0, RESULT_BLOCK_BYTES);
            b.h_plan = reinterpret_cast<Plan *>(reinterpret_cast<uint8_t *>(b.h_scalars) + 64);
            if ((rc = ensure(c, b.result, RESULT_BLOCK_BYTES))) return rc;
            CK(cudaMemsetAsync(b.result.p, 0, RESULT_BLOCK_BYTES, c->stream));
        }
        if ((rc = ensure(c, c->tickets_f, TICKET_BYTES))) return rc;
        if ((rc = ensure(c, c->tickets_t, TICKET_BYTES))) return rc;
        if ((rc = ensure(c, c->carry, 256))) return rc;
        if ((rc = ensure(c, c->pt_count, 256))) return rc;
        if ((rc = ensure(c, c->x_totals, 512))) return rc;
        if ((rc = ensure(c, c->dl_ticket, 256))) return rc;
        if ((rc = ensure(c, c->dl_scan_status, 1 << 16))) return rc;
        if ((rc = ensure(c, c->pack_off, 1 << 16))) return rc;
        CK(cudaMemsetAsync(c->dl_scan_status.p, 0, c->dl_scan_status.bytes, c->stream));
        CK(cudaMemsetAsync(c->pt_count.p, 0, 256, c->stream));
        // every buffer a submit touches, for the largest batch on top of the pending budget
        const uint64_t M = c->max_batch + c->max_pending;
        if ((rc = reserve_shared(c, c->max_batch, M))) return rc;
        for (BatchSet &b : c->set)
            if ((rc = reserve_set(c, b, c->max_batch, M))) return rc;
        if ((rc = ensure_pid_table(c, 0, M))) return rc;
        return set_carry(c, 0u, 0u, 0xffffffffu);
    };
    int rc = boot();
    if (rc != CHRONOS_OK) {
        chronos_destroy(c);
        return rc;
    }
    *out = c;
    return CHRONOS_OK;
}

int chronos_create(int device, uint64_t max_batch_events, chronos_ctx **out) {
    // default pending budget: as many pending events as one full batch holds (at least 64 K)
    return chronos_create_sized(device, max_batch_events, std::max<uint64_t>(max_batch_events, 65536), out);
}

int chronos_destroy(chronos_ctx *c) {
    if (!c) return CHRONOS_E_ARG;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    for (BatchSet &b : c->set) {
        DevBuf *all[] = {&b.arena, &b.pend_key, &b.pend_slot, &b.pend_arrival, &b.pend_meta, &b.surv_src, &b.surv_meta,
                         &b.surv_pid, &b.surv_ns, &b.surv_arrival, &b.slot_off, &b.windows, &b.chain_pid, &b.chain_idx,
                         &b.d_batch, &b.result};
        for (DevBuf *d : all) release(*d);
        if (b.h_scalars) cudaFreeHost(b.h_scalars);
        for (cudaEvent_t e : {b.ev_start, b.ev_stage1, b.ev_end, b.ev_f0, b.ev_f1, b.ev_packed, b.ev_prepack, b.ev_s1end, b.ev_sort, b.ev_x0})
            if (e) cudaEventDestroy(e);
    }
    DevBuf *all[] = {&c->carry, &c->keyA, &c->keyB, &c->valA, &c->valB, &c->pflag, &c->entry_out, &c->wsorted, &c->status,
                     &c->sweep_status, &c->sp_key, &c->sp_src, &c->sp_meta, &c->sp_ns, &c->tile_count, &c->scan_status,
                     &c->tickets_f, &c->tickets_t, &c->owner_table, &c->pid_tab, &c->pt_count, &c->pack,
                     &c->pack_off, &c->dl_scan_status, &c->dl_ticket, &c->dl_stage[0], &c->dl_stage[1], &c->x_totals};
    for (DevBuf *b : all) release(*b);
    for (int d = 0; d < XCHG_MAX_WORLD; d++)
        if (c->x_opened[d] && c->x_peer[d]) cudaIpcCloseMemHandle(c->x_peer[d]);
    release(c->x_window);
    for (cudaEvent_t e : {c->ev_sorted, c->ev_gathered, c->ev_fwd, c->ev_emitted, c->ev_tailinit, c->dl_ready, c->dl_done[0], c->dl_done[1]})
        if (e) cudaEventDestroy(e);
    for (int i = 0; i < 24; i++)
        if (c->tev[i]) cudaEventDestroy(c->tev[i]);
    for (auto &w : c->trim) {
        for (int k = 0; k < 2; k++) {
            if (w.pinned[k]) cudaFreeHost(w.pinned[k]);
            if (w.dev[k]) cudaFree(w.dev[k]);
            if (w.done[k]) cudaEventDestroy(w.done[k]);
        }
        if (w.stream) cudaStreamDestroy(w.stream);
    }
    if (c->trim_dma) cudaStreamDestroy(c->trim_dma);
    if (c->ev_trim_dma) cudaEventDestroy(c->ev_trim_dma);
    for (cudaStream_t s : {c->stream, c->stream_f, c->stream2, c->stream3, c->dl_stream})
        if (s) cudaStreamDestroy(s);
    cudaGetLastError();
    delete c;
    return CHRONOS_OK;
}

// ---- trimmed upload -------------------------------------------------------------------------
// A record in host memory is mostly NUL padding (a 40-byte path in a 256-byte field). With
// chronos_set_host_trim the host threads cut every record, losslessly, to the 16-byte chunks that
// hold anything -- chunks 0, 1 and 17 always, the argv chunks up to the last one that is not all
// zero -- and only those cross PCIe (about a third of the bytes); expand_kernel rebuilds the
// 288-byte records in the staging buffer, byte for byte, and the pipeline runs on them unchanged.
// All DECISIONS about a record stay on the GPU; the host only drops zeros.
constexpr uint64_t TRIM_CHUNK = 32768;  // records per unit of work of a host thread
constexpr size_t TRIM_OFF_BYTES = (TRIM_CHUNK + 64) * 4;                 // u32 slot offset per record (+1)
constexpr size_t TRIM_BUF_BYTES = TRIM_OFF_BYTES + TRIM_CHUNK * CH_REC_BYTES;  // worst case: nothing to drop
static_assert(TRIM_OFF_BYTES % 16 == 0, "trim_chunk writes the slots with aligned 16-byte stores");

// dst[r] = record r of the chunk: slot order 0, 1, 17, 2.. (the transport form of filter.cuh)
__global__ void __launch_bounds__(256) expand_kernel(const uint8_t *stage, uint32_t cnt, uint8_t *dst) {
    const uint32_t *off = reinterpret_cast<const uint32_t *>(stage);
    const uint4 *slots = reinterpret_cast<const uint4 *>(stage + TRIM_OFF_BYTES);
    const int lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t base = warp * 16; base < cnt; base += n_warps * 16) {
        const uint32_t m = min(16u, cnt - base);
        uint32_t o = 0, ns = 0;
        if ((uint32_t)lane < m) {
            o = off[base + lane];
            ns = off[base + lane + 1] - o;
        }
        uint4 *d = reinterpret_cast<uint4 *>(dst + (size_t)base * CH_REC_BYTES);
        uint4 v[9];
#pragma unroll
        for (int it = 0; it < 9; it++) {
            const uint32_t c = it * 32 + lane, r = c / CH_REC_CHUNKS, ch = c - r * CH_REC_CHUNKS;
            const uint32_t ro = __shfl_sync(0xffffffffu, o, (int)(r & 31u)), rn = __shfl_sync(0xffffffffu, ns, (int)(r & 31u));
            const uint32_t sl = ch < 2 ? ch : (ch == 17 ? 2u : ch + 1u);
            v[it] = make_uint4(0u, 0u, 0u, 0u);
            if (r < m && sl < rn) v[it] = __ldg(slots + ro + sl);
        }
#pragma unroll
        for (int it = 0; it < 9; it++) {
            const uint32_t c = it * 32 + lane;
            if (c < m * CH_REC_CHUNKS) d[c] = v[it];
        }
    }
}

// one chunk on the host: offsets and slots into `out` (TRIM_BUF_BYTES); returns the bytes to copy
static size_t trim_chunk(const uint8_t *recs, uint32_t cnt, uint8_t *out) {
    // The slots are written with non-temporal stores: the staging buffer is written once and then
    // read by the copy engine, so the lines need not be fetched for ownership first (16 host
    // threads get through 98 instead of 72 GB/s of records, tools/trim_probe.cpp).
    uint32_t *off = reinterpret_cast<uint32_t *>(out);
    __m128i *slots = reinterpret_cast<__m128i *>(out + TRIM_OFF_BYTES);  // 16-byte aligned (pinned buffer, offset % 16 == 0)
    const __m128i zero = _mm_setzero_si128();
    uint32_t s = 0;
    for (uint32_t i = 0; i < cnt; i++) {
        const __m128i *r = reinterpret_cast<const __m128i *>(recs + (size_t)i * CH_REC_BYTES);  // (the caller's buffer: any alignment)
        int last = 1;  // last argv chunk (2..16) that is not all zero
        for (int ch = 16; ch >= 2; ch--) {
            if (_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128(r + ch), zero)) != 0xFFFF) { last = ch; break; }
        }
        off[i] = s;
        __m128i *o = slots + s;
        _mm_stream_si128(o, _mm_loadu_si128(r));
        _mm_stream_si128(o + 1, _mm_loadu_si128(r + 1));
        _mm_stream_si128(o + 2, _mm_loadu_si128(r + 17));
        for (int k = 2; k <= last; k++) _mm_stream_si128(o + 1 + k, _mm_loadu_si128(r + k));
        s += 2u + (uint32_t)last;
    }
    off[cnt] = s;
    _mm_sfence();  // the copy engine reads the buffer next
    return TRIM_OFF_BYTES + (size_t)s * 16;
}

static int trim_pools(chronos_ctx *c) {
    if (!c->trim_dma) {
        CK(cudaStreamCreateWithFlags(&c->trim_dma, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&c->ev_trim_dma, cudaEventDisableTiming));
        c->trim_no_dma = getenv("CHRONOS_TRIM_NO_DMA") != nullptr;
    }
    if ((int)c->trim.size() == c->trim_threads) return CHRONOS_OK;
    c->trim.resize((size_t)c->trim_threads);
    for (auto &w : c->trim) {
        if (w.stream) continue;
        CK(cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking));
        for (int k = 0; k < 2; k++) {
            CK(cudaMallocHost(&w.pinned[k], TRIM_BUF_BYTES));
            CK(cudaMalloc(&w.dev[k], TRIM_BUF_BYTES));
            CK(cudaEventCreateWithFlags(&w.done[k], cudaEventDisableTiming));
        }
    }
    return CHRONOS_OK;
}

// records (host memory) -> d_batch, host threads and the copy engine working side by side: the
// threads take chunks from the FRONT of the batch, trim them and send the trimmed form; the calling
// thread takes pieces from the BACK and copies them as they are (cudaMemcpyAsync), one piece at a
// time, claiming the next when the previous one has arrived. Both ends move until they meet, so
// the split follows what cores and bus deliver (pageable memory, which the driver copies slowly
// and synchronously, simply leaves more to the threads). On return everything is queued and
// stream `st` waits for it.
static int trimmed_upload(chronos_ctx *c, BatchSet &b, const uint8_t *records, uint64_t n, cudaStream_t st) {
    int rc;
    if ((rc = trim_pools(c))) return rc;
    const uint64_t n_chunks = (n + TRIM_CHUNK - 1) / TRIM_CHUNK;
    const int T = c->trim_threads;
    uint8_t *dst = ptr<uint8_t>(b.d_batch);
    // the worker streams must not run ahead of what `st` is ordered behind (the last readers of this set)
    CK(cudaEventRecord(c->ev_tailinit, st));
    for (auto &w : c->trim) CK(cudaStreamWaitEvent(w.stream, c->ev_tailinit, 0));
    CK(cudaStreamWaitEvent(c->trim_dma, c->ev_tailinit, 0));
    std::atomic<uint64_t> range{n_chunks};  // chunks [lo, hi) are unclaimed: lo in the high, hi in the low 32 bits
    std::atomic<uint64_t> sent{0};
    std::atomic<int> failed{0};
    const int device = c->device;
    auto take_front = [&]() -> int64_t {
        uint64_t r = range.load();
        while (true) {
            const uint64_t lo = r >> 32, hi = r & 0xffffffffu;
            if (lo >= hi) return -1;
            if (range.compare_exchange_weak(r, ((lo + 1) << 32) | hi)) return (int64_t)lo;
        }
    };
    auto take_back = [&](uint64_t want, uint64_t &first, uint64_t &cnt) -> bool {
        uint64_t r = range.load();
        while (true) {
            const uint64_t lo = r >> 32, hi = r & 0xffffffffu;
            if (lo >= hi) return false;
            cnt = std::min(want, hi - lo);
            first = hi - cnt;
            if (range.compare_exchange_weak(r, (lo << 32) | first)) return true;
        }
    };
    auto work = [&](int t) {
        if (cudaSetDevice(device) != cudaSuccess) { failed = 1; return; }
        chronos_ctx::TrimWorker &w = c->trim[(size_t)t];
        int k = 0;
        for (int64_t ch = take_front(); ch >= 0; ch = take_front(), k ^= 1) {
            const uint64_t first = (uint64_t)ch * TRIM_CHUNK;
            const uint32_t cnt = (uint32_t)std::min<uint64_t>(TRIM_CHUNK, n - first);
            if (cudaEventSynchronize(w.done[k]) != cudaSuccess) { failed = 1; return; }  // this buffer pair is free again
            const size_t bytes = trim_chunk(records + first * CH_REC_BYTES, cnt, static_cast<uint8_t *>(w.pinned[k]));
            sent += bytes;
            if (cudaMemcpyAsync(w.dev[k], w.pinned[k], bytes, cudaMemcpyHostToDevice, w.stream) != cudaSuccess) { failed = 1; return; }
            expand_kernel<<<64, 256, 0, w.stream>>>(static_cast<const uint8_t *>(w.dev[k]), cnt, dst + first * CH_REC_BYTES);
            if (cudaEventRecord(w.done[k], w.stream) != cudaSuccess) { failed = 1; return; }
        }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < T; t++) th.emplace_back(work, t);
    {   // the copy engine's share, paced by completion
        uint64_t first, cnt, pieces = 0;
        while (!c->trim_no_dma && take_back(4, first, cnt)) {
            const uint64_t r0 = first * TRIM_CHUNK, r1 = std::min<uint64_t>(n, (first + cnt) * TRIM_CHUNK);
            if (cudaMemcpyAsync(dst + r0 * CH_REC_BYTES, records + r0 * CH_REC_BYTES, (r1 - r0) * CH_REC_BYTES, cudaMemcpyHostToDevice,
                                c->trim_dma) != cudaSuccess ||
                cudaEventRecord(c->ev_trim_dma, c->trim_dma) != cudaSuccess || cudaEventSynchronize(c->ev_trim_dma) != cudaSuccess) {
                failed = 1;
                break;
            }
            sent += (r1 - r0) * CH_REC_BYTES;
            pieces++;
        }
        c->launches += pieces;
    }
    for (auto &x : th) x.join();
    if (failed) return fail(c, CHRONOS_E_CUDA, "trimmed upload: %s", cudaGetErrorString(cudaGetLastError()));
    for (auto &w : c->trim)
        for (int k = 0; k < 2; k++) CK(cudaStreamWaitEvent(st, w.done[k], 0));
    CK(cudaEventRecord(c->ev_trim_dma, c->trim_dma));
    CK(cudaStreamWaitEvent(st, c->ev_trim_dma, 0));
    c->launches += n_chunks;
    c->upload_bytes = sent.load();
    return CHRONOS_OK;
}

// upload of a host batch into the set's staging buffer (allocated on the first host-memory submit)
static int stage_records(chronos_ctx *c, BatchSet &b, const void *records, uint64_t n, cudaStream_t st,
                         const uint8_t **d_recs) {
    *d_recs = nullptr;
    if (!n) return CHRONOS_OK;
    if (is_device_pointer(records)) {
        *d_recs = static_cast<const uint8_t *>(records);
        return CHRONOS_OK;
    }
    int rc;
    if ((rc = ensure(c, b.d_batch, c->max_batch * CH_REC_BYTES))) return rc;
    *d_recs = ptr<uint8_t>(b.d_batch);
    if (c->trim_threads > 0 && n >= 4 * TRIM_CHUNK) {
        return trimmed_upload(c, b, static_cast<const uint8_t *>(records), n, st);
    }
    CK(cudaMemcpyAsync(b.d_batch.p, records, n * CH_REC_BYTES, cudaMemcpyHostToDevice, st));
    c->upload_bytes = n * CH_REC_BYTES;
    return CHRONOS_OK;
}

static int check_records_arg(chronos_ctx *c, const void *records, uint64_t n) {
    if (n && !records) return fail(c, CHRONOS_E_ARG, "records is NULL");
    if (n && is_device_pointer(records) && ((uintptr_t)records & 15))
        return fail(c, CHRONOS_E_ARG, "device records pointer must be 16-byte aligned");
    return CHRONOS_OK;
}

int chronos_submit(chronos_ctx *c, const void *records, uint64_t n, uint64_t base_arrival_index) {
    if (!c) return CHRONOS_E_ARG;
    int rc;
    if ((rc = check_records_arg(c, records, n))) return rc;
    if (c->x_sent) return fail(c, CHRONOS_E_STATE, "an exchange step is half done: call chronos_submit_exchanged");
    BatchSet *bp;
    cudaStream_t s1;
    if ((rc = begin_batch(c, n, base_arrival_index, false, &bp, &s1))) return rc;
    BatchSet &b = *bp, &prev = c->set[c->last];
    const uint8_t *d_recs = nullptr;
    if ((rc = stage_records(c, b, records, n, s1, &d_recs))) return rc;
    b.records = d_recs;
    init_batch_kernel<<<1, 256, 0, s1>>>(dscal(b), ptr<uint32_t>(c->tickets_f), ptr<uint32_t>(c->tickets_t), 1);
    c->launches++;
    if (n && (rc = run_filter_stage(c, b, d_recs, n, s1))) return rc;
    if ((rc = join_stage1(c, b, s1))) return rc;
    if ((rc = run_window_stage(c, b, prev))) return rc;
    commit_batch(c, b);
    return CHRONOS_OK;
}

int chronos_submit_survivors(chronos_ctx *c, const void *slots, const uint64_t *hdr, uint64_t n, uint64_t n_slots) {
    if (!c) return CHRONOS_E_ARG;
    if (n && (!slots || !hdr)) return fail(c, CHRONOS_E_ARG, "NULL argument");
    if (n && (!is_device_pointer(slots) || ((uintptr_t)slots & 15) || !is_device_pointer(hdr)))
        return fail(c, CHRONOS_E_ARG, "slots and hdr must be device pointers (slots 16-byte aligned)");
    if (n_slots < 3 * n || n_slots > 18 * n) return fail(c, CHRONOS_E_ARG, "slot count inconsistent with entry count");
    if (c->x_sent) return fail(c, CHRONOS_E_STATE, "an exchange step is half done: call chronos_submit_exchanged");
    int rc;
    BatchSet *bp;
    cudaStream_t s1;
    if ((rc = begin_batch(c, n, 0, true, &bp, &s1))) return rc;
    BatchSet &b = *bp, &prev = c->set[c->last];
    b.slots = static_cast<const uint4 *>(slots);
    b.explicit_arrival = true;
    DeviceScalars *ds = dscal(b);
    init_batch_kernel<<<1, 256, 0, s1>>>(ds, ptr<uint32_t>(c->tickets_f), ptr<uint32_t>(c->tickets_t), 1);
    c->launches++;
    if (n) {
        const unsigned blocks = (unsigned)std::min<uint64_t>((n + 255) / 256, 2048);
        unpack_hdr_kernel<<<blocks, 256, 0, s1>>>(hdr, (uint32_t)n, ptr<uint64_t>(b.surv_arrival), ptr<uint8_t>(b.surv_meta),
                                                  ptr<uint8_t>(b.surv_ns), ptr<uint32_t>(b.slot_off));
        if ((rc = launch_scan(c, s1, ptr<uint32_t>(b.slot_off), n))) return rc;
        extract_pid_kernel<<<blocks, 256, 0, s1>>>(b.slots, ptr<uint32_t>(b.slot_off), (uint32_t)n, ptr<uint32_t>(b.surv_pid),
                                                   ptr<uint32_t>(b.surv_src), &ds->key_or, &ds->key_and);
        CK(cudaGetLastError());
        set_u32_kernel<<<1, 1, 0, s1>>>(&ds->surv_count, (uint32_t)n);
        c->launches += 3;
    }
    if ((rc = join_stage1(c, b, s1))) return rc;
    if ((rc = run_window_stage(c, b, prev))) return rc;
    commit_batch(c, b);
    return CHRONOS_OK;
}

int chronos_fetch(chronos_ctx *c, chronos_counts *out) {
    if (!c || !out) return CHRONOS_E_ARG;
    CK(cudaSetDevice(c->device));
    // the oldest batch not fetched yet; else the one fetched last (its counts again)
    int k = -1;
    const int older = c->last ^ 1;
    if (c->set[older].state == BatchSet::SUBMITTED || c->set[older].state == BatchSet::DONE) k = older;
    else if (c->set[c->last].state == BatchSet::SUBMITTED || c->set[c->last].state == BatchSet::DONE) k = c->last;
    else if (c->fetched >= 0 && c->set[c->fetched].state == BatchSet::FETCHED) k = c->fetched;
    if (k < 0) return fail(c, CHRONOS_E_STATE, "nothing submitted");
    BatchSet &b = c->set[k];
    if (b.state != BatchSet::FETCHED) {
        int rc = complete_batch(c, b);
        if (rc) return rc;
        b.state = BatchSet::FETCHED;
        c->fetched = k;
        if (b.rc) return b.rc;  // (c->err was written when the batch completed)
    }
    memset(out, 0, sizeof(*out));
    out->n_events = b.n;
    out->n_survivors = b.S;
    out->n_windows = b.W;
    out->n_arena = b.M;
    out->n_pending = b.pending_after;
    out->n_chains = b.chains;
    out->n_pids_seen = b.pids_after;
    return CHRONOS_OK;
}

static int need_fetched(chronos_ctx *c, BatchSet **b) {
    if (!c) return CHRONOS_E_ARG;
    if (c->fetched < 0 || c->set[c->fetched].state != BatchSet::FETCHED) return fail(c, CHRONOS_E_STATE, "call chronos_fetch first");
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return fail(c, CHRONOS_E_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    *b = &c->set[c->fetched];
    return CHRONOS_OK;
}

// copies for the host run on dl_stream: a younger batch may be busy on the other streams
int chronos_copy_survivors(chronos_ctx *c, uint64_t *global_idx, uint8_t *meta, uint64_t cap) {
    BatchSet *bp;
    int rc = need_fetched(c, &bp);
    if (rc) return rc;
    BatchSet &b = *bp;
    const uint64_t S = b.S;
    if (cap < S || (S && !global_idx)) return fail(c, CHRONOS_E_ARG, "survivor buffer too small (%llu < %llu)",
                                                   (unsigned long long)cap, (unsigned long long)S);
    if (!S) return CHRONOS_OK;
    cudaStream_t st;
    if ((rc = host_stream(c, &st))) return rc;
    if (b.explicit_arrival) {
        CK(cudaMemcpyAsync(global_idx, b.surv_arrival.p, S * 8, cudaMemcpyDeviceToHost, st));
    } else {
        if ((rc = ensure(c, c->pack, S * 8))) return rc;
        add_base_kernel<<<(unsigned)std::min<uint64_t>((S + 255) / 256, 2048), 256, 0, st>>>(
            ptr<uint32_t>(b.surv_src), (uint32_t)S, b.base, ptr<uint64_t>(c->pack));
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(global_idx, c->pack.p, S * 8, cudaMemcpyDeviceToHost, st));
    }
    if (meta) CK(cudaMemcpyAsync(meta, b.surv_meta.p, S, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return CHRONOS_OK;
}

int chronos_copy_windows(chronos_ctx *c, chronos_window *dst, uint64_t cap) {
    BatchSet *bp;
    int rc = need_fetched(c, &bp);
    if (rc) return rc;
    if (cap < bp->W || (bp->W && !dst)) return fail(c, CHRONOS_E_ARG, "window buffer too small");
    if (!bp->W) return CHRONOS_OK;
    cudaStream_t st;
    if ((rc = host_stream(c, &st))) return rc;
    CK(cudaMemcpyAsync(dst, bp->windows.p, bp->W * sizeof(chronos_window), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return CHRONOS_OK;
}

int chronos_copy_arena(chronos_ctx *c, void *dst, uint64_t cap_records) {
    BatchSet *bp;
    int rc = need_fetched(c, &bp);
    if (rc) return rc;
    if (cap_records < bp->M || (bp->M && !dst)) return fail(c, CHRONOS_E_ARG, "arena buffer too small");
    if (!bp->M) return CHRONOS_OK;
    cudaStream_t st;
    if ((rc = host_stream(c, &st))) return rc;
    CK(cudaMemcpyAsync(dst, bp->arena.p, bp->M * CH_REC_BYTES, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return CHRONOS_OK;
}

// exclusive scan of n counters on dl_stream (own status words and ticket: a stage 1 may be
// scanning on another stream at the same time)
static int dl_scan(chronos_ctx *c, cudaStream_t st, uint32_t *data, uint64_t n) {
    if (n == 0) return CHRONOS_OK;
    const uint32_t n_tiles = (uint32_t)((n + SCAN_TILE - 1) / SCAN_TILE);
    int rc;
    const void *before = c->dl_scan_status.p;
    if ((rc = ensure(c, c->dl_scan_status, std::max<size_t>((size_t)n_tiles * 8, 1 << 16)))) return rc;
    if (c->dl_scan_status.p != before || c->dl_scan_epoch == 0)
        CK(cudaMemsetAsync(c->dl_scan_status.p, 0, c->dl_scan_status.bytes, st));
    CK(cudaMemsetAsync(c->dl_ticket.p, 0, 4, st));
    ScanParams sp;
    sp.data = data;
    sp.n = (uint32_t)n;
    sp.n_ptr = nullptr;
    sp.ticket = ptr<uint32_t>(c->dl_ticket);
    sp.status = ptr<uint64_t>(c->dl_scan_status);
    sp.epoch = ++c->dl_scan_epoch;
    if (c->dl_scan_epoch >= (1u << 29)) c->dl_scan_epoch = 0;
    scan_exclusive_u32<<<std::min<uint32_t>(n_tiles, (uint32_t)c->sm_count * 2), SCAN_THREADS, 0, st>>>(sp);
    CK(cudaGetLastError());
    return CHRONOS_OK;
}

// records of all windows of the fetched batch, window after window, into a DEVICE buffer
// (every grouped entry is in a window or pending, so the total is known without a read-back)
static int pack_window_records(chronos_ctx *c, cudaStream_t st, BatchSet &b, uint8_t *dst_dev) {
    const uint64_t W = b.W;
    if (!W) return CHRONOS_OK;
    int rc;
    if ((rc = ensure(c, c->pack_off, (W + 8) * 4))) return rc;
    const unsigned blocks = (unsigned)std::min<uint64_t>((W + 255) / 256, 2048);
    window_len_kernel<<<blocks, 256, 0, st>>>(ptr<chronos_window>(b.windows), (uint32_t)W, ptr<uint32_t>(c->pack_off));
    if ((rc = dl_scan(c, st, ptr<uint32_t>(c->pack_off), W))) return rc;
    pack_windows_kernel<<<(unsigned)std::min<uint64_t>((W + 7) / 8, (uint64_t)c->sm_count * 8), 256, 0, st>>>(
        ptr<uint8_t>(b.arena), ptr<chronos_window>(b.windows), ptr<uint32_t>(c->pack_off), (uint32_t)W, dst_dev);
    CK(cudaGetLastError());
    return CHRONOS_OK;
}

int chronos_copy_window_records(chronos_ctx *c, void *dst, uint64_t cap_records) {
    BatchSet *bp;
    int rc = need_fetched(c, &bp);
    if (rc) return rc;
    BatchSet &b = *bp;
    const uint64_t tot = b.M - b.pending_after;  // grouped entries that left in a window
    if (!b.W) return CHRONOS_OK;
    if (cap_records < tot || !dst) return fail(c, CHRONOS_E_ARG, "window record buffer too small (%llu < %llu)",
                                               (unsigned long long)cap_records, (unsigned long long)tot);
    if ((rc = ensure(c, c->pack, tot * CH_REC_BYTES))) return rc;
    cudaStream_t st;
    if ((rc = host_stream(c, &st))) return rc;
    if ((rc = pack_window_records(c, st, b, ptr<uint8_t>(c->pack)))) return rc;
    CK(cudaMemcpyAsync(dst, c->pack.p, tot * CH_REC_BYTES, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return CHRONOS_OK;
}

// ---- asynchronous download of a fetched batch's results ---------------------------------------
int chronos_download_begin(chronos_ctx *c, void *host_dst, uint64_t cap_bytes, chronos_download *out) {
    BatchSet *bp;
    int rc = need_fetched(c, &bp);
    if (rc) return rc;
    BatchSet &b = *bp;
    if (!out) return fail(c, CHRONOS_E_ARG, "out is NULL");
    auto al = [](uint64_t v) { return (v + 255) & ~255ull; };
    const uint64_t S = b.S, W = b.W, R = b.W ? b.M - b.pending_after : 0;
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
    const int k = c->dl_k;
    if (c->dl_inflight[k]) {  // the copy started two calls ago still reads this staging buffer
        CK(cudaEventSynchronize(c->dl_done[k]));
        c->dl_inflight[k] = false;
    }
    if ((rc = ensure(c, c->dl_stage[k], out->total_bytes))) return rc;
    uint8_t *st = ptr<uint8_t>(c->dl_stage[k]);
    if (!c->dl_stream) CK(cudaStreamCreateWithPriority(&c->dl_stream, cudaStreamNonBlocking, c->prio_lo));
    cudaStream_t ds;  // packing: main stream while idle; the copy to the host always on dl_stream
    if ((rc = host_stream(c, &ds))) return rc;
    if (S) {
        if (b.explicit_arrival) {
            CK(cudaMemcpyAsync(st + out->off_survivor_idx, b.surv_arrival.p, S * 8, cudaMemcpyDeviceToDevice, ds));
        } else {
            add_base_kernel<<<(unsigned)std::min<uint64_t>((S + 255) / 256, 2048), 256, 0, ds>>>(
                ptr<uint32_t>(b.surv_src), (uint32_t)S, b.base, reinterpret_cast<uint64_t *>(st + out->off_survivor_idx));
            CK(cudaGetLastError());
        }
        CK(cudaMemcpyAsync(st + out->off_survivor_meta, b.surv_meta.p, S, cudaMemcpyDeviceToDevice, ds));
    }
    if (W) {
        CK(cudaMemcpyAsync(st + out->off_windows, b.windows.p, W * sizeof(chronos_window), cudaMemcpyDeviceToDevice, ds));
        if ((rc = pack_window_records(c, ds, b, st + out->off_window_records))) return rc;
    }
    CK(cudaEventRecord(b.ev_packed, ds));  // the set itself may be written again from here on
    CK(cudaStreamWaitEvent(c->dl_stream, b.ev_packed, 0));
    CK(cudaMemcpyAsync(host_dst, st, out->total_bytes, cudaMemcpyDeviceToHost, c->dl_stream));
    CK(cudaEventRecord(c->dl_done[k], c->dl_stream));
    c->dl_inflight[k] = true;
    c->dl_k ^= 1;
    return CHRONOS_OK;
}

int chronos_download_wait(chronos_ctx *c) {
    if (!c) return CHRONOS_E_ARG;
    CK(cudaSetDevice(c->device));
    if (c->dl_stream) CK(cudaStreamSynchronize(c->dl_stream));
    c->dl_inflight[0] = c->dl_inflight[1] = false;
    return CHRONOS_OK;
}

int chronos_window_digests(chronos_ctx *c, uint64_t *dst, uint64_t cap) {
    BatchSet *bp;
    int rc = need_fetched(c, &bp);
    if (rc) return rc;
    BatchSet &b = *bp;
    if (cap < b.W || (b.W && !dst)) return fail(c, CHRONOS_E_ARG, "digest buffer too small");
    if (!b.W) return CHRONOS_OK;
    if ((rc = ensure(c, c->pack, b.W * 8))) return rc;
    cudaStream_t st;
    if ((rc = host_stream(c, &st))) return rc;
    window_digest_kernel<<<(unsigned)std::min<uint64_t>((b.W + 7) / 8, (uint64_t)c->sm_count * 8), 256, 0, st>>>(
        ptr<uint8_t>(b.arena), ptr<chronos_window>(b.windows), (uint32_t)b.W, ptr<uint64_t>(c->pack));
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(dst, c->pack.p, b.W * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return CHRONOS_OK;
}

int chronos_copy_chains(chronos_ctx *c, uint32_t *pid, uint64_t *global_idx, uint64_t cap) {
    BatchSet *bp;
    int rc = need_fetched(c, &bp);
    if (rc) return rc;
    const uint64_t K = bp->chains;
    if (cap < K) return fail(c, CHRONOS_E_ARG, "chain buffer too small (%llu < %llu)", (unsigned long long)cap,
                             (unsigned long long)K);
    if (!K) return CHRONOS_OK;
    cudaStream_t st;
    if ((rc = host_stream(c, &st))) return rc;
    if (pid) CK(cudaMemcpyAsync(pid, bp->chain_pid.p, K * 4, cudaMemcpyDeviceToHost, st));
    if (global_idx) CK(cudaMemcpyAsync(global_idx, bp->chain_idx.p, K * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return CHRONOS_OK;
}

// process_memory lives in the set of the newest batch; it can be read once that batch (and
// therefore every batch) has been fetched
int chronos_state(chronos_ctx *c, uint32_t *pid, uint64_t *global_idx, void *records, uint64_t cap) {
    if (!c) return CHRONOS_E_ARG;
    int rc;
    if ((rc = no_batch_in_flight(c))) return rc;
    CK(cudaSetDevice(c->device));
    const uint64_t P = c->n_pending;
    if (cap < P) return fail(c, CHRONOS_E_ARG, "state buffer too small (%llu < %llu)", (unsigned long long)cap,
                             (unsigned long long)P);
    if (!P) return CHRONOS_OK;
    BatchSet &b = c->set[c->last];
    const uint64_t pc = b.pend_cap;
    cudaStream_t st = c->stream;  // (nothing is in flight)
    if (pid) CK(cudaMemcpyAsync(pid, ptr<uint32_t>(b.pend_key) + (pc - P), P * 4, cudaMemcpyDeviceToHost, st));
    if (global_idx) CK(cudaMemcpyAsync(global_idx, ptr<uint64_t>(b.pend_arrival) + (pc - P), P * 8, cudaMemcpyDeviceToHost, st));
    if (records) {
        if ((rc = ensure(c, c->pack, P * CH_REC_BYTES))) return rc;
        pack_by_slot_kernel<<<(unsigned)std::min<uint64_t>((P + 7) / 8, (uint64_t)c->sm_count * 8), 256, 0, st>>>(
            ptr<uint8_t>(b.arena), ptr<uint32_t>(b.pend_slot) + (pc - P), (uint32_t)P, ptr<uint8_t>(c->pack));
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(records, c->pack.p, P * CH_REC_BYTES, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st));
    return CHRONOS_OK;
}

int chronos_state_pids(chronos_ctx *c, uint32_t *pid, uint64_t cap) {
    if (!c) return CHRONOS_E_ARG;
    int rc;
    if ((rc = no_batch_in_flight(c))) return rc;
    CK(cudaSetDevice(c->device));
    const uint64_t K = c->n_pids_seen;
    if (cap < K || (K && !pid)) return fail(c, CHRONOS_E_ARG, "pid buffer too small (%llu < %llu)",
                                            (unsigned long long)cap, (unsigned long long)K);
    if (!K) return CHRONOS_OK;
    if ((rc = ensure(c, c->pack, K * 4 + 256))) return rc;
    cudaStream_t st = c->stream;
    uint32_t *cnt = ptr<uint32_t>(c->dl_ticket) + 8;
    CK(cudaMemsetAsync(cnt, 0, 4, st));
    pt_dump_kernel<<<(unsigned)std::min<uint32_t>((c->pt_cap + 255) / 256, 4096), 256, 0, st>>>(
        ptr<uint64_t>(c->pid_tab), c->pt_cap, c->pt_epoch, ptr<uint32_t>(c->pack), cnt);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(pid, c->pack.p, K * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
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
    int rc;
    if ((rc = no_batch_in_flight(c))) return rc;
    *bytes = ckpt_layout(c->n_pending, c->n_pids_seen).total;
    return CHRONOS_OK;
}

int chronos_checkpoint_save(chronos_ctx *c, void *dst, uint64_t cap) {
    if (!c || !dst) return CHRONOS_E_ARG;
    int rc;
    if ((rc = no_batch_in_flight(c))) return rc;
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
    cudaStream_t st = c->stream;
    if (K) {
        if ((rc = ensure(c, c->pack, K * 8 + 256))) return rc;
        uint32_t *cnt = ptr<uint32_t>(c->dl_ticket) + 8;
        CK(cudaMemsetAsync(cnt, 0, 4, st));
        pt_dump64_kernel<<<(unsigned)std::min<uint32_t>((c->pt_cap + 255) / 256, 4096), 256, 0, st>>>(
            ptr<uint64_t>(c->pid_tab), c->pt_cap, c->pt_epoch, ptr<uint64_t>(c->pack), cnt);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(out + L.off_tab, c->pack.p, K * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        uint64_t *e = reinterpret_cast<uint64_t *>(out + L.off_tab);
        std::sort(e, e + K, [](uint64_t a, uint64_t b) { return (uint32_t)a < (uint32_t)b; });  // by pid: files compare equal
    }
    if (P) {
        memset(out + L.off_pid, 0, L.off_arr - L.off_pid);
        memset(out + L.off_meta, 0, L.off_rec - L.off_meta);
        BatchSet &b = c->set[c->last];
        const uint64_t pc = b.pend_cap;
        CK(cudaMemcpyAsync(out + L.off_pid, ptr<uint32_t>(b.pend_key) + (pc - P), P * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(out + L.off_arr, ptr<uint64_t>(b.pend_arrival) + (pc - P), P * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(out + L.off_meta, ptr<uint8_t>(b.pend_meta) + (pc - P), P, cudaMemcpyDeviceToHost, st));
        if ((rc = ensure(c, c->pack, P * CH_REC_BYTES))) return rc;
        pack_by_slot_kernel<<<(unsigned)std::min<uint64_t>((P + 7) / 8, (uint64_t)c->sm_count * 8), 256, 0, st>>>(
            ptr<uint8_t>(b.arena), ptr<uint32_t>(b.pend_slot) + (pc - P), (uint32_t)P, ptr<uint8_t>(c->pack));
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(out + L.off_rec, c->pack.p, P * CH_REC_BYTES, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
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
    if (P + c->max_batch > MAX_GROUPED || K >= (1ull << 30)) return fail(c, CHRONOS_E_CAPACITY, "checkpoint too large for this context");
    const CkptLayout L = ckpt_layout(P, K);
    if (bytes < L.total) return fail(c, CHRONOS_E_ARG, "checkpoint truncated (%llu < %llu bytes)", (unsigned long long)bytes,
                                     (unsigned long long)L.total);
    // Everything is validated before the context is touched (memcpy: the blob need not be aligned).
    // PID entries: (dfa << 32 | pid), strictly ascending pid, DFA state 0..2, nothing above bit 33.
    std::vector<uint64_t> tab(K);
    if (K) memcpy(tab.data(), in + L.off_tab, K * 8);
    for (uint64_t i = 0; i < K; i++) {
        if ((tab[i] >> 34) != 0 || ((tab[i] >> 32) & 3u) == 3u)
            return fail(c, CHRONOS_E_ARG, "checkpoint: PID entry %llu malformed", (unsigned long long)i);
        if (i && (uint32_t)tab[i] <= (uint32_t)tab[i - 1])
            return fail(c, CHRONOS_E_ARG, "checkpoint: PID entries not strictly ascending at %llu", (unsigned long long)i);
    }
    // Pending list: what window_reverse would have written -- ascending pid, ascending arrival
    // inside a pid -- and every pending pid is a key of process_memory.
    std::vector<uint32_t> ppid(P);
    std::vector<uint64_t> parr(P);
    if (P) {
        memcpy(ppid.data(), in + L.off_pid, P * 4);
        memcpy(parr.data(), in + L.off_arr, P * 8);
    }
    uint32_t k_or = 0, k_and = 0xffffffffu;
    for (uint64_t i = 0; i < K; i++) {  // bit range of every pid seen (the sort plans with it)
        k_or |= (uint32_t)tab[i];
        k_and &= (uint32_t)tab[i];
    }
    uint64_t ti = 0;
    for (uint64_t i = 0; i < P; i++) {
        if (i && (ppid[i] < ppid[i - 1] || (ppid[i] == ppid[i - 1] && parr[i] <= parr[i - 1])))
            return fail(c, CHRONOS_E_ARG, "checkpoint: pending entry %llu out of order", (unsigned long long)i);
        if (parr[i] >= MAX_ARRIVAL) return fail(c, CHRONOS_E_ARG, "checkpoint: arrival index of entry %llu too large", (unsigned long long)i);
        while (ti < K && (uint32_t)tab[ti] < ppid[i]) ti++;  // both lists ascend: one merge walk
        if (ti == K || (uint32_t)tab[ti] != ppid[i])
            return fail(c, CHRONOS_E_ARG, "checkpoint: pending pid %u is not among the PID entries", ppid[i]);
    }
    int rc;
    for (BatchSet &f : c->set)  // nothing may be in flight while the state is replaced
        if (f.state == BatchSet::SUBMITTED && (rc = complete_batch(c, f))) return rc;
    if ((rc = chronos_reset(c))) return rc;
    CK(cudaDeviceSynchronize());
    BatchSet &b = c->set[c->last];
    if (P + c->max_batch > c->cap_M || P > b.cap_M) {  // state larger than the pending budget: grow once
        const uint64_t want = std::min<uint64_t>(MAX_GROUPED, P + c->max_batch);
        if ((rc = reserve_shared(c, c->max_batch, want))) return rc;
        for (BatchSet &s : c->set)
            if ((rc = reserve_set(c, s, c->max_batch, want))) return rc;
    }
    if ((rc = ensure_pid_table(c, K, c->max_batch + P))) return rc;
    cudaStream_t st = c->stream;
    if (K) {
        if ((rc = ensure(c, c->pack, K * 8))) return rc;
        CK(cudaMemcpyAsync(c->pack.p, tab.data(), K * 8, cudaMemcpyHostToDevice, st));
        pt_load_kernel<<<(unsigned)std::min<uint64_t>((K + 255) / 256, 2048), 256, 0, st>>>(
            ptr<uint64_t>(c->pid_tab), c->pt_cap - 1, c->pt_epoch, ptr<uint64_t>(c->pack), (uint32_t)K);
        CK(cudaGetLastError());
        const uint32_t k32 = (uint32_t)K;
        CK(cudaMemcpyAsync(c->pt_count.p, &k32, 4, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
    }
    if (P) {
        const uint64_t off = b.pend_cap - P;  // right-aligned, as window_reverse leaves them
        CK(cudaMemcpyAsync(ptr<uint32_t>(b.pend_key) + off, ppid.data(), P * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(ptr<uint64_t>(b.pend_arrival) + off, parr.data(), P * 8, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(ptr<uint8_t>(b.pend_meta) + off, in + L.off_meta, P, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(b.arena.p, in + L.off_rec, P * CH_REC_BYTES, cudaMemcpyHostToDevice, st));
        iota_u32_kernel<<<(unsigned)std::min<uint64_t>((P + 255) / 256, 2048), 256, 0, st>>>(ptr<uint32_t>(b.pend_slot) + off,
                                                                                           (uint32_t)P);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(st));
    }
    c->n_pending = P;
    c->n_pids_seen = K;
    return set_carry(c, (uint32_t)P, k_or, k_and);
}

int chronos_get_device_view(chronos_ctx *c, chronos_device_view *out) {
    BatchSet *bp;
    int rc = need_fetched(c, &bp);
    if (rc) return rc;
    if (!out) return CHRONOS_E_ARG;
    BatchSet &b = *bp;
    const uint64_t P = b.pending_after, pc = b.pend_cap;
    out->arena = b.arena.p;
    out->windows = ptr<chronos_window>(b.windows);
    out->survivor_idx = ptr<uint32_t>(b.surv_src);
    out->survivor_meta = ptr<uint8_t>(b.surv_meta);
    out->pending_pid = P ? ptr<uint32_t>(b.pend_key) + (pc - P) : nullptr;
    out->pending_slot = P ? ptr<uint32_t>(b.pend_slot) + (pc - P) : nullptr;
    out->pending_arrival = P ? ptr<uint64_t>(b.pend_arrival) + (pc - P) : nullptr;
    return CHRONOS_OK;
}

int chronos_last_timing(chronos_ctx *c, float *total_ms, float *filter_ms) {
    BatchSet *bp;
    int rc = need_fetched(c, &bp);
    if (rc) return rc;
    if (total_ms) *total_ms = bp->t_total;
    if (filter_ms) *filter_ms = bp->t_filter_kernel;
    return CHRONOS_OK;
}

int chronos_last_stage_times(chronos_ctx *c, float *ms, uint32_t cap) {
    BatchSet *bp;
    int rc = need_fetched(c, &bp);
    if (rc) return rc;
    if (!ms) return CHRONOS_E_ARG;
    for (uint32_t i = 0; i < cap && i < 6; i++) ms[i] = bp->stage_ms[i];
    return CHRONOS_OK;
}

int chronos_set_host_trim(chronos_ctx *c, int threads) {
    if (!c || threads < 0 || threads > 256) return CHRONOS_E_ARG;
    CK(cudaSetDevice(c->device));
    CK(cudaDeviceSynchronize());  // (nothing may be using the worker buffers while their number changes)
    if (threads < (int)c->trim.size()) {
        for (size_t t = (size_t)threads; t < c->trim.size(); t++) {
            auto &w = c->trim[t];
            for (int k = 0; k < 2; k++) {
                if (w.pinned[k]) cudaFreeHost(w.pinned[k]);
                if (w.dev[k]) cudaFree(w.dev[k]);
                if (w.done[k]) cudaEventDestroy(w.done[k]);
            }
            if (w.stream) cudaStreamDestroy(w.stream);
        }
        c->trim.resize((size_t)threads);
    }
    c->trim_threads = threads;
    return threads ? trim_pools(c) : CHRONOS_OK;
}

int chronos_last_upload_bytes(chronos_ctx *c, uint64_t *bytes) {
    if (!c || !bytes) return CHRONOS_E_ARG;
    *bytes = c->upload_bytes;
    return CHRONOS_OK;
}

int chronos_set_overlap(chronos_ctx *c, int on) {
    if (!c) return CHRONOS_E_ARG;
    c->overlap = on != 0;
    return CHRONOS_OK;
}

int chronos_reset(chronos_ctx *c) {
    if (!c) return CHRONOS_E_ARG;
    CK(cudaSetDevice(c->device));
    // Takes effect in submit order and waits for nothing: the resets below are queued on the main
    // stream behind the window stage of every batch submitted so far. Those batches finish against
    // the old state and stay fetchable; batches submitted from now on start from an empty one.
    c->gen++;
    c->n_pending = 0;
    c->n_pids_seen = 0;
    c->x_sent = false;
    // slots written under an older epoch count as empty: a reset costs nothing on the device
    c->pt_epoch = (c->pt_epoch + 1) & 0xFFu;
    if (c->pt_epoch == 0) {  // tags wrapped: really clear the table once every 255 resets
        c->pt_epoch = 1;
        if (c->pt_cap) CK(cudaMemsetAsync(c->pid_tab.p, 0, (size_t)c->pt_cap * 8, c->stream));
    }
    CK(cudaMemsetAsync(c->pt_count.p, 0, 256, c->stream));
    return set_carry(c, 0u, 0u, 0xffffffffu);
}

int chronos_last_launches(chronos_ctx *c, uint64_t *launches) {
    if (!c || !launches) return CHRONOS_E_ARG;
    const int k = (c->fetched >= 0 && c->set[c->fetched].state == BatchSet::FETCHED) ? c->fetched : c->last;
    *launches = c->set[k].launches;
    return CHRONOS_OK;
}

void *chronos_stream(chronos_ctx *c) { return c ? (void *)c->stream : nullptr; }

uint32_t chronos_owner_of(uint32_t pid, uint32_t world) { return world ? hash32(pid) % world : 0; }

// Stage 1 of the sharded path: filter the slice, partition the survivors by owner rank and
// deliver them (pk.direct: into the owners' windows; else into the caller's send buffers):
// filter -> compact_scan -> xchg_hist -> scan -> xchg_pack, all on the main stream, in the
// buffer set the NEXT batch will use. The per-PID state of the context is not touched.
// (Cutting the slice into pieces so that the delivery of piece k overlaps the filter of piece
//  k+1 was built and measured on 2xB200, 12.5 M events per rank: stage 1 got SLOWER, 0.80 ->
//  1.15 ms with four pieces -- the filter loses its ~19 single-warp CTAs per SM to the pack
//  kernel -- so it was removed again.)
static int stage1(chronos_ctx *c, const void *records, uint64_t n, uint64_t base, XchgPackParams &pk) {
    int rc;
    if ((rc = check_records_arg(c, records, n))) return rc;
    if (n > c->max_batch) return fail(c, CHRONOS_E_CAPACITY, "batch of %llu events > max_batch_events %llu",
                                      (unsigned long long)n, (unsigned long long)c->max_batch);
    if (base >= MAX_ARRIVAL || base + n > MAX_ARRIVAL) return fail(c, CHRONOS_E_ARG, "arrival indices must stay below 2^48");
    BatchSet &b = c->set[c->last ^ 1];
    if (b.state == BatchSet::SUBMITTED && (rc = complete_batch(c, b))) return rc;
    // (the newest batch may still be in flight: stage 1 is queued on the main stream behind its tail,
    //  it only touches the other buffer set and the stage-1 temporaries)
    if (c->fetched == (c->last ^ 1)) c->fetched = -1;
    b.state = BatchSet::FREE;
    cudaStream_t st = c->stream;
    CK(cudaStreamWaitEvent(st, b.ev_packed, 0));
    const uint8_t *d_recs = nullptr;
    if ((rc = stage_records(c, b, records, n, st, &d_recs))) return rc;
    c->launches = 0;
    c->scan_slot = 0;
    c->n_tev = 0;
    DeviceScalars *ds = dscal(b);
    init_batch_kernel<<<1, 256, 0, st>>>(ds, ptr<uint32_t>(c->tickets_f), ptr<uint32_t>(c->tickets_t), 1);
    set_u32_kernel<<<1, 1, 0, st>>>(&ds->xchg_err, 0u);
    c->launches += 2;
    mark(c, "stage 1 start");
    b.filter_timed = false;
    if (n && (rc = run_filter_stage(c, b, d_recs, n, st))) return rc;
    b.x_filter_timed = n != 0;
    mark(c, "filter + compact");

    const uint32_t W = pk.world;
    const uint32_t T = (uint32_t)std::max<uint64_t>((n + XCHG_TILE - 1) / XCHG_TILE, 1);  // upper bound of the tile count
    const uint64_t table_len = (uint64_t)2 * W * T + 1;
    if ((rc = ensure(c, c->owner_table, (2ull * XCHG_MAX_WORLD * ((c->max_batch + XCHG_TILE - 1) / XCHG_TILE + 1) + 1) * 4))) return rc;
    uint32_t *table = ptr<uint32_t>(c->owner_table);
    const unsigned xblocks = std::min<uint32_t>(T, (uint32_t)c->sm_count * 4);  // persistent CTAs
    xchg_hist<<<xblocks, XCHG_THREADS, 0, st>>>(ptr<uint32_t>(b.surv_pid), ptr<uint8_t>(b.surv_ns), &ds->surv_count, W, T, table);
    if ((rc = dbg_sync(c, "xchg_hist"))) return rc;
    if ((rc = launch_scan(c, st, table, table_len))) return rc;
    if ((rc = dbg_sync(c, "owner table scan"))) return rc;
    mark(c, "owner histogram + scan");
    pk.pid = ptr<uint32_t>(b.surv_pid);
    pk.surv_src = ptr<uint32_t>(b.surv_src);
    pk.surv_meta = ptr<uint8_t>(b.surv_meta);
    pk.surv_ns = ptr<uint8_t>(b.surv_ns);
    pk.n_ptr = &ds->surv_count;
    pk.batch = d_recs;
    pk.base_arrival = base;
    pk.table = table;
    pk.T = T;
    pk.totals = ptr<uint32_t>(c->x_totals);
    pk.err = &ds->xchg_err;
    pk.done = ptr<uint32_t>(c->tickets_f) + 58;
    CK(cudaEventRecord(b.ev_prepack, st));
    xchg_pack<<<xblocks, XCHG_THREADS, XCHG_PACK_SMEM, st>>>(pk);
    CK(cudaGetLastError());
    CK(cudaEventRecord(b.ev_s1end, st));
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
    for (uint32_t d = 0; d < world; d++) entry_counts_host[d] = slot_counts_host[d] = 0;
    if (!n) return CHRONOS_OK;
    int rc;
    XchgPackParams pk;
    memset(&pk, 0, sizeof(pk));
    pk.world = world;
    pk.direct = false;
    pk.send_slots = static_cast<uint4 *>(send_slots);
    pk.send_hdr = send_hdr;
    pk.cap_entries = cap_entries;
    pk.cap_slots = cap_slots;
    if ((rc = stage1(c, records, n, base_arrival_index, pk))) return rc;
    BatchSet &b = c->set[c->last ^ 1];
    uint32_t hc[2 * XCHG_MAX_WORLD];
    CK(cudaMemcpyAsync(hc, c->x_totals.p, sizeof(hc), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(b.h_scalars, dscal(b), sizeof(DeviceScalars), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    uint64_t tot_e = 0, tot_s = 0;
    for (uint32_t d = 0; d < world; d++) {
        entry_counts_host[d] = hc[d];
        slot_counts_host[d] = hc[XCHG_MAX_WORLD + d];
        tot_e += hc[d];
        tot_s += hc[XCHG_MAX_WORLD + d];
    }
    cudaEventElapsedTime(&b.t_filter_kernel, b.ev_f0, b.ev_f1);
    cudaGetLastError();
    c->set[c->last].launches = c->launches;  // (chronos_last_launches right after this call)
    if (b.h_scalars->xchg_err & XERR_OVERFLOW)
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
    for (BatchSet &b : c->set)
        if ((rc = reserve_transport(c, b))) return rc;
    if ((rc = ensure(c, c->owner_table, (2ull * XCHG_MAX_WORLD * ((c->max_batch + XCHG_TILE - 1) / XCHG_TILE + 1) + 1) * 4))) return rc;
    CK(cudaMemsetAsync(c->x_window.p, 0, 2 * c->x_L.half_bytes, c->stream));
    CK(cudaStreamSynchronize(c->stream));  // peers may write as soon as they hold the address
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
    int rc;
    CK(cudaEventRecord(c->set[c->last ^ 1].ev_x0, c->stream));
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
    BatchSet *bp;
    cudaStream_t s1;
    int rc = begin_batch(c, n_max, 0, true, &bp, &s1, true);
    if (rc) return rc;
    BatchSet &b = *bp, &prev = c->set[c->last];
    c->launches += stage1_launches;
    const uint8_t *hb = ptr<uint8_t>(c->x_window) + (size_t)(c->x_epoch & 1u) * L.half_bytes;
    b.slots = reinterpret_cast<const uint4 *>(hb + L.off_slots);
    b.explicit_arrival = true;
    b.exchanged = true;
    b.x_epoch = c->x_epoch;
    c->x_sent = false;
    DeviceScalars *ds = dscal(b);
    XchgUnpackParams u;
    memset(&u, 0, sizeof(u));
    u.window = ptr<uint8_t>(c->x_window);
    u.L = L;
    u.epoch = c->x_epoch;
    u.timeout_ns = c->x_timeout_ns;
    u.arrival = ptr<uint64_t>(b.surv_arrival);
    u.meta = ptr<uint8_t>(b.surv_meta);
    u.ns = ptr<uint8_t>(b.surv_ns);
    u.slot_off = ptr<uint32_t>(b.slot_off);
    u.key = nullptr;
    u.pid_copy = ptr<uint32_t>(b.surv_pid);
    u.src = ptr<uint32_t>(b.surv_src);
    u.n_out = &ds->surv_count;
    u.k_or = &ds->key_or;
    u.k_and = &ds->key_and;
    u.err = &ds->xchg_err;
    u.n_cap = (uint32_t)n_max;
    // (the survivor scalars of this set were used by stage 1: back to neutral before the receive)
    init_batch_kernel<<<1, 256, 0, s1>>>(ds, ptr<uint32_t>(c->tickets_f), ptr<uint32_t>(c->tickets_t), 1);
    xchg_unpack<<<(unsigned)std::min<uint64_t>((n_max + 255) / 256, (uint64_t)c->sm_count * 4), 256, 0, s1>>>(u);
    CK(cudaGetLastError());
    c->launches += 2;
    if ((rc = join_stage1(c, b, s1))) return rc;
    if ((rc = run_window_stage(c, b, prev))) return rc;
    commit_batch(c, b);
    return CHRONOS_OK;
}

int chronos_exchange_set_timeout(chronos_ctx *c, double seconds) {
    if (!c || !(seconds > 0)) return CHRONOS_E_ARG;
    c->x_timeout_ns = (uint64_t)(seconds * 1e9);
    return CHRONOS_OK;
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
    l2_flush_kernel<<<device_sm_count(device) * 8, 256>>>(static_cast<uint4 *>(buf[device]), bytes / 16, v++);
    return cudaDeviceSynchronize() == cudaSuccess ? CHRONOS_OK : CHRONOS_E_CUDA;
}

int chronos_device_synchronize(int device) {
    if (cudaSetDevice(device) != cudaSuccess) return CHRONOS_E_CUDA;
    return cudaDeviceSynchronize() == cudaSuccess ? CHRONOS_OK : CHRONOS_E_CUDA;
}

int chronos_device_mem_info(int device, uint64_t *free_bytes, uint64_t *total_bytes) {
    if (cudaSetDevice(device) != cudaSuccess) return CHRONOS_E_CUDA;
    size_t f = 0, t = 0;
    if (cudaMemGetInfo(&f, &t) != cudaSuccess) return CHRONOS_E_CUDA;
    if (free_bytes) *free_bytes = f;
    if (total_bytes) *total_bytes = t;
    return CHRONOS_OK;
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
