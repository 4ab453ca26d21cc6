/*
 * chronos.h -- C ABI of the B200-native CHRONOS telemetry hot path.
 *
 * The reference (Riyaz246/Project-CHRONOS, /root/reference/chronos_sensor.py) has no
 * FFI of its own: its hot path is the eBPF openat filter (chronos_sensor.py:28-47,
 * :74-92) plus the Python perf-buffer callback print_event (:124-157) working on
 * `struct data_t` records (:18-23) and the module-global process_memory (:105).  This
 * header is the boundary a maintainer binds instead (ctypes stub in INTEGRATION.md):
 * raw 288-byte records in, surviving-event set + flushed context windows + pending
 * per-PID state out.  Plain C types only, int status returns, no exceptions; one
 * context per GPU, calls on a context externally serialised (the reference is a
 * single-threaded poll loop, :161-162).  There is no CPU fallback: every entry point
 * fails with CHRONOS_E_CUDA when no sm_100-class device is usable.
 */
#ifndef CHRONOS_B200_H
#define CHRONOS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* sizeof(struct data_t), chronos_sensor.py:18-23: u32 pid @0, comm[16] @4, argv[256] @20,
 * type[10] @276, 2 bytes tail padding. */
#define CHRONOS_RECORD_BYTES 288u

#define CHRONOS_OK 0
#define CHRONOS_E_ARG (-1)      /* bad pointer / size / alignment */
#define CHRONOS_E_CUDA (-2)     /* CUDA runtime error, or no usable device */
#define CHRONOS_E_NOMEM (-3)    /* device or pinned allocation failed */
#define CHRONOS_E_CAPACITY (-4) /* batch larger than the context was created for */
#define CHRONOS_E_STATE (-5)    /* call out of order (e.g. fetch with nothing submitted) */

/* per-event class bits kept with every surviving event */
#define CHRONOS_META_TRIG 1u  /* chronos_sensor.py:141 trigger needles matched */
#define CHRONOS_META_CURL 2u  /* extension: "curl" in event_str */
#define CHRONOS_META_CHMOD 4u /* extension: "chmod" in event_str */
#define CHRONOS_META_EXEC 8u  /* extension: type field is exactly "EXEC" (:55) */

typedef struct chronos_ctx chronos_ctx;

/* One flushed context window = one call of analyze_sequence(pid, history)
 * (chronos_sensor.py:142-147, :157).  Its events are records
 * [start, start+len) of the grouped-records arena, already in arrival order. */
typedef struct chronos_window {
    uint64_t trigger_idx; /* arrival index of the event that flushed it; windows are sorted by this */
    uint64_t start;       /* first record of the window in the arena */
    uint32_t len;         /* number of events (>= 2, :142) */
    uint32_t pid;
    uint32_t flags;       /* bit0: extension, a curl->chmod->EXEC chain completed inside this window */
    uint32_t reserved;
} chronos_window;

typedef struct chronos_counts {
    uint64_t n_events;    /* raw records in the submitted batch */
    uint64_t n_survivors; /* events appended to process_memory (:138) in this batch */
    uint64_t n_windows;   /* windows flushed by this batch */
    uint64_t n_arena;     /* records in the arena = carried-in pending + n_survivors */
    uint64_t n_pending;   /* events left in process_memory after this batch (sum of list lengths) */
    uint64_t n_chains;    /* extension: curl->chmod->EXEC chains completed in this batch */
    uint64_t n_pids_seen; /* keys of process_memory so far (PIDs with >= 1 surviving event) */
    uint64_t reserved;
} chronos_counts;

/* Device-side view of the last batch's results (pointers owned by the context, valid
 * until the next submit).  Lets a caller that already lives on the GPU skip the copies. */
typedef struct chronos_device_view {
    const void *arena;              /* n_arena * 288 bytes, grouped by pid, arrival order inside */
    const chronos_window *windows;  /* n_windows, ascending trigger_idx */
    const uint32_t *survivor_idx;   /* n_survivors, batch-local arrival index, ascending */
    const uint8_t *survivor_meta;   /* n_survivors, CHRONOS_META_* */
    const uint32_t *pending_pid;    /* n_pending, ascending pid, arrival order inside a pid */
    const uint32_t *pending_slot;   /* n_pending, arena record index */
    const uint64_t *pending_arrival;/* n_pending, global arrival index */
} chronos_device_view;

/* Library / device probe.  chronos_sensor.py has no counterpart (it has no native
 * library); returns CHRONOS_OK when `device` is a usable sm_100-class GPU. */
int chronos_device_ok(int device);

/* Replaces: module start-up + `process_memory = defaultdict(list)`
 * (chronos_sensor.py:100-105).  Allocates ALL device memory the context needs for batches of
 * up to max_batch_events records (< 2^27) on top of up to max_pending_events events pending in
 * process_memory: two buffer sets (arena, pending lists, survivor and window arrays), the
 * sort / scan temporaries and the per-PID table.  A submit allocates nothing, with two
 * documented exceptions: the first submit from HOST memory creates the upload staging buffer
 * (max_batch_events x 288 bytes per buffer set), and a context whose pending lists outgrow
 * max_pending_events is re-sized once by the submit that finds out (it waits for the GPU, the
 * state is kept).  process_memory itself is unbounded in the reference (:105); the one hard
 * limit here is pending + one batch < 2^27 events.  chronos_create uses
 * max_pending_events = max(max_batch_events, 65536). */
int chronos_create(int device, uint64_t max_batch_events, chronos_ctx **out);
int chronos_create_sized(int device, uint64_t max_batch_events, uint64_t max_pending_events, chronos_ctx **out);

/* Replaces: process exit (chronos_sensor.py:163). */
int chronos_destroy(chronos_ctx *ctx);

/* Replaces: restarting the sensor, i.e. a fresh `process_memory` (chronos_sensor.py:105).
 * Takes effect in submit order and waits for nothing: batches submitted before the call finish
 * against the old state and stay fetchable, batches submitted after it start from an empty
 * process_memory. Buffers stay allocated. */
int chronos_reset(chronos_ctx *ctx);
/* Host-side trimming of uploads. threads > 0: a submit from HOST memory of at least 131072 records
 * lets `threads` host threads cut every record, losslessly, to the 16-byte chunks that hold
 * anything (the NUL padding of the path field is most of a record) and copies only those; a
 * kernel rebuilds the 288-byte records on the GPU, byte for byte, before the filter runs. Every
 * decision about an event stays on the GPU. It pays when the host has cores to spare: 16 threads
 * get through 73 GB/s of records on the bench box, PCIe moves 55 GB/s.
 * Allocates, per thread, two pinned and two device buffers of 9.6 MB at this call. threads = 0
 * (default) switches it off. No reference counterpart. */
int chronos_set_host_trim(chronos_ctx *ctx, int threads);
/* Bytes the last submit from host memory sent over PCIe (n x 288, or less with host trimming). */
int chronos_last_upload_bytes(chronos_ctx *ctx, uint64_t *bytes);
/* on = 0: a batch starts on the GPU only when the previous batch has ended, even when it was
 * submitted earlier (the host still enqueues ahead); on = 1 (default): its filter stage may run
 * beside the previous batch's window stage. No reference counterpart. */
int chronos_set_overlap(chronos_ctx *ctx, int on);

/* Human-readable text for the last non-OK status on this context (or a static string
 * for ctx == NULL).  No reference counterpart. */
const char *chronos_last_error(const chronos_ctx *ctx);

/* Replaces: one perf_buffer_poll() worth of print_event callbacks
 * (chronos_sensor.py:124-142, :157, :160-162) preceded by the openat hook's filter
 * (:74-92).  `records` = n*288 bytes in arrival order, host or device memory (detected;
 * device pointers must be 16-byte aligned); record i has global arrival index
 * base_arrival_index + i (all arrival indices must stay below 2^48).  Work is only
 * queued; per-PID state carries over from earlier submits.
 * TWO batches may be in flight: a submit while the previous batch is unfetched starts
 * its filter stage beside that batch's window stage (the stages of one batch never
 * overlap, and window stages run in submit order), and chronos_fetch then returns the
 * batches in submit order.  A third submit drops the results of the oldest unfetched
 * batch (its effect on the state is kept).  `records` in device memory must stay
 * valid until the batch has been fetched.  A submit that is refused (CHRONOS_E_ARG,
 * CHRONOS_E_CAPACITY, CHRONOS_E_NOMEM) leaves the context exactly as it was. */
int chronos_submit(chronos_ctx *ctx, const void *records, uint64_t n, uint64_t base_arrival_index);

/* Waits for the oldest batch not fetched yet and returns its counts; the copy calls below
 * then refer to that batch until the next chronos_fetch. */
int chronos_fetch(chronos_ctx *ctx, chronos_counts *out);

/* Copies to host memory, after chronos_fetch.  cap_* are capacities in elements; the
 * call fails with CHRONOS_E_ARG when the result does not fit. */
int chronos_copy_survivors(chronos_ctx *ctx, uint64_t *global_idx, uint8_t *meta /* may be NULL */, uint64_t cap);
int chronos_copy_windows(chronos_ctx *ctx, chronos_window *dst, uint64_t cap);
int chronos_copy_arena(chronos_ctx *ctx, void *dst, uint64_t cap_records);
/* Only the records that belong to flushed windows, window after window in
 * trigger_idx order (window w starts at the running sum of len); the form the
 * analyze_sequence caller needs (chronos_sensor.py:147). */
int chronos_copy_window_records(chronos_ctx *ctx, void *dst, uint64_t cap_records);
int chronos_copy_chains(chronos_ctx *ctx, uint32_t *pid, uint64_t *global_idx, uint64_t cap);
/* One 64-bit digest per window of the fetched batch (same order as chronos_copy_windows),
 * computed on the GPU over the window's records in canonical form: bytes behind a field's
 * first NUL -- which the reference never reads, chronos_sensor.py:125-130 -- count as zero.
 * Equal digests = the same (pid, history) would reach analyze_sequence (:147). Used to compare
 * the output of a sharded multi-GPU run with a single-GPU run without moving the records. */
int chronos_window_digests(chronos_ctx *ctx, uint64_t *dst, uint64_t cap);

/* Replaces: reading process_memory (chronos_sensor.py:105) -- the pending per-PID
 * lists after the last fetched batch, ascending pid, arrival order inside a pid.
 * records (may be NULL) receives the 288-byte events themselves. */
int chronos_state(chronos_ctx *ctx, uint32_t *pid, uint64_t *global_idx, void *records, uint64_t cap);
/* The keys of process_memory (PIDs that ever had a surviving event, including those
 * whose list is currently []), unordered. */
int chronos_state_pids(chronos_ctx *ctx, uint32_t *pid, uint64_t cap);

int chronos_get_device_view(chronos_ctx *ctx, chronos_device_view *out);

/* Timing of the last fetched batch as measured with CUDA events on the context's
 * stream: the whole submit and the filter kernel alone (events right around its
 * launch), in milliseconds. */
int chronos_last_timing(chronos_ctx *ctx, float *total_ms, float *filter_ms);
/* Stage boundaries of the last fetched batch (CUDA events on the context's streams), in
 * milliseconds: [0] filter kernel, [1] cross-tile compaction (sharded path: + owner histogram and
 * scan), [2] sharded path only: pack into the owners' windows, [3] sharded path only: wait for the
 * peers + unpack, [4] init + sort, [5] window kernels + record gather. cap <= 6 values are written. */
int chronos_last_stage_times(chronos_ctx *ctx, float *ms, uint32_t cap);
/* Number of kernel launches the last submit issued (bench.py's gpu_launches). */
int chronos_last_launches(chronos_ctx *ctx, uint64_t *launches);
/* The CUDA stream the context launches on (cudaStream_t as void*), for callers that
 * want to order their own work after it. */
void *chronos_stream(chronos_ctx *ctx);

/* ---- multi-GPU building blocks (PID-hash sharding) ------------------- */

/* owner rank of a pid: hash32(pid) % world. Exposed so host code and tests agree. */
uint32_t chronos_owner_of(uint32_t pid, uint32_t world);

/* Stage 1 on the rank that holds a contiguous slice of the batch: filter it, then pack the
 * survivors per owner rank (stable: arrival order kept inside each destination) into
 * caller-provided DEVICE buffers in TRANSPORT FORM:
 *   send_hdr   one u64 per surviving event: arrival index (48 bits) | META bits << 48 |
 *              slot count << 56
 *   send_slots the event's record as 3..18 slots of 16 bytes: chunk 0 (pid, comm[0:12]),
 *              chunk 1 (comm[12:16], argv[0:12]), chunk 17 (argv[252:256], type) and then as
 *              many of the chunks 2..16 as argv reaches into. Bytes behind a field's NUL
 *              are not transported (a live sensor zero-fills them, chronos_sensor.py:52/:69;
 *              the receiver restores zeros), which cuts the NVLink volume about 3.5x.
 * entry_counts_host[world] / slot_counts_host[world] receive the per-destination counts
 * (this call waits for the GPU). Destination d's data starts at the running sums.
 * world <= CHRONOS_EXCHANGE_MAX_WORLD. */
int chronos_filter_partition(chronos_ctx *ctx, const void *records, uint64_t n, uint64_t base_arrival_index,
                             uint32_t world, void *send_slots, uint64_t cap_slots, uint64_t *send_hdr,
                             uint64_t cap_entries, uint64_t *entry_counts_host, uint64_t *slot_counts_host);

/* Stage 2 on the owner rank: run grouping + window logic on `n_entries` already-filtered
 * events in transport form (device pointers, concatenated in source-rank order = global
 * arrival order). Results are fetched with the calls above. */
int chronos_submit_survivors(chronos_ctx *ctx, const void *slots, const uint64_t *hdr, uint64_t n_entries,
                             uint64_t n_slots);

/* ---- checkpoint of the per-PID state -----------------------------------
 *
 * Everything a context carries from one batch to the next -- the pending per-PID lists with
 * their records (process_memory, chronos_sensor.py:105), the set of PIDs seen (its keys,
 * including lists emptied at :157) and the chain labeller's per-PID state -- as one
 * little-endian blob in HOST memory: a 64-byte header ("CHRNCKP1", version, counts), the PID
 * entries sorted by pid, then pid / arrival index / META bits / 288-byte record of every
 * pending entry. A context that loads the blob continues exactly where the saved one was
 * (same windows, same final state). The reference has no counterpart: its state dies with
 * the process. Not allowed between chronos_submit and chronos_fetch. */
int chronos_checkpoint_bytes(chronos_ctx *ctx, uint64_t *bytes);
int chronos_checkpoint_save(chronos_ctx *ctx, void *dst, uint64_t cap_bytes);
/* Replaces the context's state (an implicit chronos_reset first). Malformed or truncated
 * input: CHRONOS_E_ARG and the state is left reset or untouched, never half loaded. */
int chronos_checkpoint_load(chronos_ctx *ctx, const void *src, uint64_t bytes);

/* ---- asynchronous download of the results ------------------------------------------------
 *
 * The chronos_copy_* calls block. A replay loop that wants the PCIe link busy in both
 * directions (results of batch k going down while batch k+1 comes up) uses these two
 * instead: chronos_download_begin gathers survivors, windows and the windows' records of the
 * FETCHED batch into one device staging buffer and starts ONE device-to-host copy on a
 * separate stream; it returns at once and the context is free for the next chronos_submit.
 * host_dst should be pinned memory (chronos_host_alloc_pinned); with pageable memory the copy
 * still works but is staged by the driver. Two downloads may be in flight; a third call waits
 * for the oldest. The data is complete after chronos_download_wait. No reference counterpart
 * (the reference prints each window as it flushes, chronos_sensor.py:143-155). */
typedef struct chronos_download {
    uint64_t n_survivors, n_windows, n_window_records;
    uint64_t off_survivor_idx;    /* u64[n_survivors]  global arrival index, ascending */
    uint64_t off_survivor_meta;   /* u8[n_survivors]   CHRONOS_META_* bits */
    uint64_t off_windows;         /* chronos_window[n_windows], ascending trigger arrival */
    uint64_t off_window_records;  /* 288 B x n_window_records, window after window */
    uint64_t total_bytes;         /* what host_dst must hold */
} chronos_download;
/* cap_bytes too small: CHRONOS_E_ARG, and *out still tells the size needed. */
int chronos_download_begin(chronos_ctx *ctx, void *host_dst, uint64_t cap_bytes, chronos_download *out);
int chronos_download_wait(chronos_ctx *ctx);

/* ---- direct exchange: survivors go straight into the owner's memory over NVLink -------------
 *
 * The two calls above leave the all-to-all to the host (torch.distributed / NCCL). On one
 * NVSwitch box the repartition can skip that: every rank owns a RECEIVE WINDOW in its HBM,
 * peers map it (CUDA IPC between processes, plain pointers inside one process), and the
 * pack kernel of stage 1 writes each survivor into its owner's window while it computes the
 * partition; the last thread block to finish publishes the per-source counts and raises a
 * flag the owner's stage 2 waits on. No staging buffer, no count exchange, no host round
 * trip: chronos_filter_exchange + chronos_submit_exchanged only enqueue work; the first wait
 * is chronos_fetch. No reference counterpart (the reference is one process on one host).
 *
 * Protocol, identical on every rank and every step:
 *   chronos_exchange_create   once; all ranks with the same world and capacities
 *   (exchange the window pointers or IPC handles, e.g. with an all-gather)
 *   chronos_exchange_connect  once
 *   per step: chronos_filter_exchange(slice) ; chronos_submit_exchanged() ; chronos_fetch(...)
 * Every rank must take part in every step (an empty slice is fine): stage 2 waits for the
 * flag of every source rank and gives up with CHRONOS_E_STATE after the timeout (20 s unless
 * chronos_exchange_set_timeout changed it). A (source, owner) pair that exceeds the capacities
 * makes both ends report CHRONOS_E_CAPACITY at fetch; nothing is truncated silently.
 * Ranks must not destroy their context while peers may still send (barrier first). */
#define CHRONOS_IPC_HANDLE_BYTES 64
#define CHRONOS_EXCHANGE_MAX_WORLD 16

/* Allocate this rank's receive window: room for `cap_entries_per_peer` survivors and
 * `cap_slots_per_peer` 16-byte transport slots (0 = 8 per entry) from EACH source rank,
 * twice (two steps in flight). world * cap_entries_per_peer must not exceed the context's
 * max_batch_events. window_out (optional) receives the device address, ipc_handle_out
 * (optional, CHRONOS_IPC_HANDLE_BYTES) the handle another process opens it with. */
int chronos_exchange_create(chronos_ctx *ctx, uint32_t world, uint32_t rank, uint64_t cap_entries_per_peer,
                            uint64_t cap_slots_per_peer, void **window_out, uint8_t *ipc_handle_out);
/* Addresses of the peers' windows: peer_windows[world] (device pointers valid in THIS process
 * and reachable from this GPU; entries may be NULL) and/or peer_ipc_handles
 * [world * CHRONOS_IPC_HANDLE_BYTES] for ranks in other processes. The own rank is skipped. */
int chronos_exchange_connect(chronos_ctx *ctx, void *const *peer_windows, const uint8_t *peer_ipc_handles);
/* Stage 1: filter the slice (host or device pointer, as chronos_submit) and deliver every
 * survivor to owner = chronos_owner_of(pid, world). Asynchronous. */
int chronos_filter_exchange(chronos_ctx *ctx, const void *records, uint64_t n, uint64_t base_arrival_index);
/* Stage 2: wait (on the GPU) for all sources of this step, then grouping + window logic on
 * what arrived, in source-rank order = global arrival order. Asynchronous; results with
 * chronos_fetch and the copy calls. */
int chronos_submit_exchanged(chronos_ctx *ctx);
int chronos_exchange_set_timeout(chronos_ctx *ctx, double seconds);

/* ---- allocation-free filtering ------------------------------------------------------------ */

/* Largest batch chronos_filter_noalloc accepts. Its scratch is a fixed array inside the
 * library image (about 2.3 MiB of device memory held by every GPU that loads the library,
 * until the process exits), sized for this many events: four 64K-event micro-batches. */
#define CHRONOS_NOALLOC_MAX_EVENTS 262144u

/* No counterpart in the reference. The filter stage alone (openat noise filter, UTF-8
 * gate, comm ignore-list, trigger bits: the same decisions chronos_submit takes), for
 * callers that must not allocate GPU memory when they call us: a sensor thread on a
 * fixed latency budget (cudaMalloc and cudaFree take a device-wide lock and may wait for
 * running work), a process whose GPU memory is wholly owned by another framework's pool,
 * or a plug-in host that forbids allocation after start-up. The call makes no cudaMalloc,
 * cudaFree or other allocating call; the caller passes every input and output:
 *   records        n*288 bytes, DEVICE memory, 16-byte aligned, arrival order
 *   survivor_idx   DEVICE, room for n u32: batch-local index of each surviving event, ascending
 *   survivor_pid   DEVICE, room for n u32: its pid
 *   survivor_meta  DEVICE, room for n u8 : its CHRONOS_META_* bits
 *   count          DEVICE, one u32       : number of survivors
 * Work is queued on `stream` (a cudaStream_t, NULL for the default stream) and not waited
 * for. n above CHRONOS_NOALLOC_MAX_EVENTS is refused with CHRONOS_E_CAPACITY; errbuf
 * (may be NULL) receives a readable message for any failure.
 * Every call on a GPU shares the one scratch array, which still holds the previous
 * call's data; the parts that must start at zero are cleared at the head of each call.
 * Two calls whose GPU work overlaps in time (different streams, no ordering between
 * them) therefore race: keep the calls of one GPU on one stream or order them with events.
 * The first call on each device also looks up the scratch array's device address and the
 * launch geometry; make that first call before starting any stream capture. */
int chronos_filter_noalloc(int device, void *stream, const void *records, uint64_t n, uint32_t *survivor_idx,
                           uint32_t *survivor_pid, uint8_t *survivor_meta, uint32_t *count, char *errbuf,
                           uint64_t errbuf_len);

/* ---- synthetic trace (no reference counterpart: the reference ships no trace) ------ */

typedef struct chronos_tracegen_cfg {
    uint64_t seed;
    uint32_t n_pids;
    uint32_t pid_base;
    uint32_t pid_stride;
    uint32_t noise_permille;
    uint32_t exec_permille;
    uint32_t adv_permille;
    uint32_t long_permille;
    uint32_t chain_period;
    uint32_t chain_pid_base;
    uint32_t use_cdf;
} chronos_tracegen_cfg;

/* Fill dst (host memory) with records [first, first+n) of the trace. cdf: n_pids u64
 * thresholds (host) when cfg->use_cdf, else NULL. */
int chronos_tracegen_host(const chronos_tracegen_cfg *cfg, const uint64_t *cdf, uint64_t first, uint64_t n, void *dst);
/* Same bytes, generated by a CUDA kernel into device memory (dst and cdf are device
 * pointers on `device`). */
int chronos_tracegen_device(int device, const chronos_tracegen_cfg *cfg, const uint64_t *cdf_dev, uint64_t first,
                            uint64_t n, void *dst_dev);

/* Plain device/pinned memory helpers so that a ctypes-only caller needs no other CUDA
 * binding (no reference counterpart). */
int chronos_dev_alloc(int device, uint64_t bytes, void **out);
int chronos_dev_free(int device, void *p);
int chronos_host_alloc_pinned(uint64_t bytes, void **out);
int chronos_host_free_pinned(void *p);
int chronos_memcpy_h2d(int device, void *dst_dev, const void *src_host, uint64_t bytes);
int chronos_memcpy_d2h(int device, void *dst_host, const void *src_dev, uint64_t bytes);
/* Writes `bytes` of a scratch device buffer, to evict the 126 MB L2 between timed runs. */
int chronos_flush_l2(int device);
/* cudaDeviceSynchronize for ctypes-only callers. */
int chronos_device_synchronize(int device);
/* cudaMemGetInfo for ctypes-only callers (tests use it to show that submits allocate nothing). */
int chronos_device_mem_info(int device, uint64_t *free_bytes, uint64_t *total_bytes);

#ifdef __cplusplus
}
#endif
#endif /* CHRONOS_B200_H */
