// common.cuh -- shared device helpers (sm_100a): mbarrier + bulk-copy PTX, SWAR byte
// tricks, single-pass ("decoupled look-back") tile-status words.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define CH_REC_BYTES 288
#define CH_REC_WORDS 72
#define CH_REC_CHUNKS 18  // 16-byte chunks per record

// Bounds checks of every index a kernel computes before it writes through it. They cost
// registers and branches, so they are compiled in only by tools/checked_build.sh
// (-DCHRONOS_CHECKED: the -m gpu suite is then run against that library; a violated check traps
// the kernel and the next API call reports cudaErrorAssert). compute-sanitizer is not available
// on this pool (profiles/r02_sanitizer_unavailable.txt); this is what stands in for memcheck.
#ifdef CHRONOS_CHECKED
#include <assert.h>
#define CH_ASSERT(x) assert(x)
#else
#define CH_ASSERT(x) ((void)0)
#endif

namespace chronos {

// ---------------------------------------------------------------------------------
// mbarrier / TMA bulk copy (1-D cp.async.bulk: global -> shared, completion by bytes)
// ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// L2 eviction policy for data that is read exactly once (the raw batch): evict first, so that a
// multi-GB stream does not push the lines a kernel writes out of the 126 MB L2.
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// Loads of records or slots that lie SPARSELY in a large array (survivors in the batch, tile-local
// slots): without a qualifier a miss fills the whole 128-byte line from DRAM; `.L2::64B` limits the
// fill to the 64-byte half that is asked for (ncu: the record gather read 474 MB for 343 MB of
// records before, 399 MB with it).
__device__ __forceinline__ uint32_t ld_sparse_u32(const uint32_t *p) {
    uint32_t v;
    asm("ld.global.L2::64B.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint4 ld_sparse_nc_v4(const uint4 *p) {
    uint4 v;
    asm("ld.global.nc.L2::64B.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
// 16 bytes global -> shared memory without passing through registers (LDGSTS)
__device__ __forceinline__ void cp_async16(void *dst_smem, const void *src, uint32_t src_bytes) {
    // src_bytes = 16: plain copy; 0: the 16 bytes are zero-filled, nothing is read
    // (.L2::64B: a record is 4.5 such halves; see ld_sparse_u32 in common.cuh)
    asm volatile("cp.async.cg.shared.global.L2::64B [%0], [%1], 16, %2;" ::"r"(smem_u32(dst_smem)), "l"(src), "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// 1-D bulk copy through the TMA unit with an L2 cache hint; size and both addresses must be multiples of 16.
__device__ __forceinline__ void bulk_g2s_hint(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar,
                                              uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
// 1-D bulk copy through the TMA unit; size and both addresses must be multiples of 16.
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---------------------------------------------------------------------------------
// SWAR helpers on little-endian 4-byte words
// ---------------------------------------------------------------------------------
// 0x80 in every byte of x that is zero, 0 elsewhere (exact, no false positives).
__device__ __forceinline__ uint32_t zero_bytes(uint32_t x) {
    return ~(((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x | 0x7f7f7f7fu);
}
// index (0..3) of the lowest zero byte given z = zero_bytes(x) != 0
__device__ __forceinline__ int first_zero_index(uint32_t z) { return (__ffs((int)z) - 1) >> 3; }
// keep the low k bytes (k in 0..3) of x
__device__ __forceinline__ uint32_t keep_low_bytes(uint32_t x, int k) { return x & ((1u << (8 * k)) - 1u); }

__host__ __device__ constexpr uint32_t lit4(char a, char b, char c, char d) {
    return (uint32_t)(uint8_t)a | ((uint32_t)(uint8_t)b << 8) | ((uint32_t)(uint8_t)c << 16) |
           ((uint32_t)(uint8_t)d << 24);
}
__host__ __device__ constexpr uint32_t lit3(char a, char b, char c) { return lit4(a, b, c, 0); }
__host__ __device__ constexpr uint32_t lit2(char a, char b) { return lit4(a, b, 0, 0); }

// ---------------------------------------------------------------------------------
// look-back status words: [63:62] status, payload below
// ---------------------------------------------------------------------------------
constexpr uint64_t ST_NONE = 0, ST_AGG = 1, ST_PREFIX = 2;
__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(uint64_t *p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t *p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// Device-resident sizes of the batch in flight, written by plan_kernel once the survivor
// count is known ON THE DEVICE. Every kernel after the filter reads its extents from here
// and is launched with an upper-bound grid, so a submit needs no host round trip.
struct Plan {
    uint32_t S, P, M;          // survivors of this batch, carried-in pending entries, M = P + S
    uint32_t n_passes;         // radix passes actually needed (pids differ only in some bits)
    uint32_t shift[4], mask[4];
    uint32_t tile_keys, n_tiles, table_len;  // radix tiling of M
    uint32_t tiles_m, tiles_s;               // window tiling of M and of S
    uint32_t overflow;         // set when M exceeds what the buffers were sized for
    uint32_t key_or, key_and;  // over every pid seen since the last reset
    uint32_t cap_m, cap_s, cap_w;  // what the buffers of this batch hold: grouped entries, survivors, windows
    const uint32_t *skey;      // where the sorted (pid, value) arrays ended up
    const uint32_t *sval;
};

// What one batch hands to the next ON THE DEVICE (so that a submit never needs the host to know
// it): the number of pending entries it left in process_memory (chronos_sensor.py:105) and
// the OR / AND over every pid seen since the last reset (the bit range the sort covers).
struct Carry {
    uint32_t P, key_or, key_and, pad;
};

// owner rank of a pid for PID-hash sharding (also exported through the C ABI)
__host__ __device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16;
    x *= 0x7feb352dU;
    x ^= x >> 15;
    x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}

}  // namespace chronos
