// Host-side cost of shrinking a batch before it crosses PCIe: every 288-byte
// record is cut losslessly to the 16-byte chunks that hold anything (chunks 0, 1, 17 and the
// argv chunks up to the last non-zero one), T threads, 4 M synthetic records with 20..80-byte
// paths. Prints the rate at which the CPU gets through the INPUT bytes and the size ratio.
//   g++ -O3 -mavx2 -pthread -o trim_probe tools/trim_probe.cpp && ./trim_probe <threads>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <chrono>
#include <thread>
#include <vector>
#include <immintrin.h>
// lossless zero-run trim of one 288-byte record: chunks 0,1,17 always; chunks 2..1+k where k covers the last nonzero byte of [32,272)
static inline uint32_t trim_one(const uint8_t *r, uint8_t *out) {
    // find last nonzero 16-byte chunk among chunks 2..16
    int last = 1;
    for (int ch = 16; ch >= 2; ch--) {
        __m128i v = _mm_loadu_si128((const __m128i *)(r + ch * 16));
        if (_mm_movemask_epi8(_mm_cmpeq_epi8(v, _mm_setzero_si128())) != 0xFFFF) { last = ch; break; }
    }
    memcpy(out, r, 32);
    memcpy(out + 32, r + 272, 16);
    int extra = last - 1;
    memcpy(out + 48, r + 32, (size_t)extra * 16);
    return 3 + extra;
}
// the same with non-temporal stores for the output (no read-for-ownership of the staging lines)
static inline uint32_t trim_one_nt(const uint8_t *rb, uint8_t *out) {
    const __m128i *r = (const __m128i *)rb;
    const __m128i zero = _mm_setzero_si128();
    int last = 1;
    for (int ch = 16; ch >= 2; ch--) {
        if (_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128(r + ch), zero)) != 0xFFFF) { last = ch; break; }
    }
    __m128i *o = (__m128i *)out;
    _mm_stream_si128(o, _mm_loadu_si128(r));
    _mm_stream_si128(o + 1, _mm_loadu_si128(r + 1));
    _mm_stream_si128(o + 2, _mm_loadu_si128(r + 17));
    for (int k = 2; k <= last; k++) _mm_stream_si128(o + 1 + k, _mm_loadu_si128(r + k));
    return 2 + last;
}
int main(int argc, char **argv) {
    const bool nt = argc > 2 && atoi(argv[2]) == 1;
    size_t n = 4000000; int T = argc > 1 ? atoi(argv[1]) : 1;
    std::vector<uint8_t> in(n * 288, 0); uint8_t *outp = (uint8_t *)aligned_alloc(4096, n * 288); memset(outp, 1, n * 288);
    for (size_t i = 0; i < n; i++) { uint8_t *r = &in[i * 288]; memcpy(r, &i, 4); memcpy(r + 4, "bash", 4); int len = 20 + (i * 7) % 60; memset(r + 20, 'a', len); memcpy(r + 276, "OPEN", 4); }
    std::vector<uint8_t> ns(n);
    for (int rep = 0; rep < 3; rep++) {
        auto t0 = std::chrono::steady_clock::now();
        std::vector<std::thread> th; std::vector<size_t> slots(T);
        for (int t = 0; t < T; t++) th.emplace_back([&, t] {
            size_t lo = n * t / T, hi = n * (t + 1) / T; uint8_t *o = outp + lo * 288; size_t s = 0;
            for (size_t i = lo; i < hi; i++) { uint32_t k = nt ? trim_one_nt(&in[i * 288], o + s * 16) : trim_one(&in[i * 288], o + s * 16); ns[i] = (uint8_t)k; s += k; }
            if (nt) _mm_sfence();
            slots[t] = s; });
        for (auto &x : th) x.join();
        double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        size_t tot = 0; for (auto s : slots) tot += s;
        printf("%s T=%d %.1f ms  in %.1f GB/s  out ratio %.3f\n", nt ? "nt-stores" : "memcpy   ", T, dt * 1e3, n * 288 / dt / 1e9, tot * 16.0 / (n * 288));
    }
}
