/*
 * ref_filter_wrap.c -- glue around the reference's OWN eBPF helper functions.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/README.md). oracle/build_oracle.py
 * concatenates, in a temporary directory outside the repository,
 *     (1) lines 28-47 of /root/reference/chronos_sensor.py, read where they lie
 *         (`ends_with` and `starts_with`, plain C inside the BPF program string), and
 *     (2) this file,
 * and compiles the result into oracle/_ref/libchronos_ref_filter.so. No reference
 * source is copied into the repository; only the shared object is kept (git-ignored).
 *
 * The only thing restated here is the decision table of the openat hook's three `if`
 * blocks (chronos_sensor.py:76-80, :83-87, :90-92); the string helpers it calls are the
 * reference's code.
 */
#include <stdint.h>
#include <string.h>

static char *const k_prefix_a[] = {"/lib", "/usr/lib", "/usr/share", "/etc/ssl", "/etc/fonts", "/etc/host", 0};
static char *const k_suffix[] = {".so", ".cache", ".mo", ".conf", ".crt", ".curlrc", 0};
static char *const k_prefix_b[] = {"/dev/", "/proc/", 0};

/* 1 = the openat hook returns before perf_submit (event never reaches user space). */
int chronos_ref_openat_drop(char *argv) {
    for (int i = 0; k_prefix_a[i]; i++) if (starts_with(argv, k_prefix_a[i])) return 1; /* :76-80 */
    for (int i = 0; k_suffix[i]; i++) if (ends_with(argv, k_suffix[i])) return 1;       /* :83-87 */
    for (int i = 0; k_prefix_b[i]; i++) if (starts_with(argv, k_prefix_b[i])) return 1; /* :90-92 */
    return 0;
}

/* Raw-replay stage A over a batch of 288-byte records: drop[i] = 1 iff record i is an
 * OPEN record (type field reads "OPEN") that the hook would have dropped. */
void chronos_ref_filter_batch(const uint8_t *recs, uint64_t n, uint8_t *drop) {
    for (uint64_t i = 0; i < n; i++) {
        const uint8_t *r = recs + i * 288;
        char argv[256];
        memcpy(argv, r + 20, 256);
        int is_open = memcmp(r + 276, "OPEN", 5) == 0;
        drop[i] = (uint8_t)(is_open && chronos_ref_openat_drop(argv));
    }
}
