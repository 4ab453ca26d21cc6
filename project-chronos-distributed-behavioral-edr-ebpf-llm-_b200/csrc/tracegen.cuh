// tracegen.cuh -- deterministic synthetic execve/openat trace, one 288-byte
// `struct data_t` (reference: chronos_sensor.py:18-23) per event.
//
// The reference ships no trace, only attack_chain.sh:6-14 and one screenshot, so the
// benchmark workloads of BASELINE.json are synthesised. Record i is a pure function of
// (cfg, i): a counter-based splitmix64 stream, integer arithmetic only, so the host
// twin and the CUDA twin (same source, compiled twice) are bit-identical and a rank can
// generate any slice of a trace without the rest.
//
// Record-shape conventions follow the hooks: EXEC carries argv[0] and the caller's
// pre-exec comm (chronos_sensor.py:50-64), OPEN carries the file name (:67-72); fields
// are zero-filled after the string (:52, :69) except for the adversarial class, which
// also exercises garbage after the NUL, unterminated fields and broken UTF-8.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define CH_HD __host__ __device__ __forceinline__
#else
#define CH_HD inline
#endif

#include "chronos.h"  // struct chronos_tracegen_cfg

namespace chronos_gen {

CH_HD uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

struct Rng {
    uint64_t s;
    CH_HD uint64_t next() {
        s += 0x9E3779B97F4A7C15ULL;
        return mix64(s);
    }
    CH_HD uint32_t below(uint32_t n) { return (uint32_t)(((next() >> 32) * (uint64_t)n) >> 32); }
};

// All string tables live in one blob so host and device share one initialiser.
// Entries are NUL-separated; OFFS[] index the starts.
#define CHRONOS_GEN_BLOB                                                                     \
    /* 0..23 comm names: survivors, trigger-bearing, ignore-list hits */                     \
    "bash\0sshd\0systemd\0cron\0curl\0cat\0chmod\0ls\0nc\0sync\0attack_chain.sh\0"           \
    "python3\0node\0code\0git\0chrome\0ollama\0vmtoolsd\0dockerd\0containerd-shim\0"         \
    "nginx\0postgres\0unicode_start\0tar\0"                                                  \
    /* 24..35 noise directory prefixes (openat hook drops by prefix) */                      \
    "/lib/x86_64-linux-gnu/\0/usr/lib/x86_64-linux-gnu/\0/usr/share/locale/\0"               \
    "/usr/share/zoneinfo/\0/etc/ssl/certs/\0/etc/fonts/conf.d/\0/etc/host\0/proc/\0"         \
    "/dev/\0/lib64/\0/usr/libexec/\0/usr/share/icons/\0"                                     \
    /* 36..41 noise suffixes (openat hook drops by suffix) */                                \
    ".so\0.cache\0.mo\0.conf\0.crt\0.curlrc\0"                                               \
    /* 42..53 surviving directory prefixes */                                                \
    "/tmp/\0/home/kali/\0/var/log/\0/etc/\0/home/user/projects/\0/sys/class/net/\0"          \
    "/run/user/1000/\0/opt/app/data/\0relative/lib/\0/var/lib/dpkg/\0/srv/www/\0/mnt/data/\0" \
    /* 54..69 name words */                                                                  \
    "malware\0payload\0report\0session\0index\0libc\0license\0application\0locale\0"         \
    "history\0backup\0socket\0notes\0cache\0config\0data\0"                                  \
    /* 70..77 surviving extensions */                                                        \
    ".bin\0.txt\0.log\0.json\0.so.1\0.tmp\0\0.db\0"                                          \
    /* 78..89 argv[0] of EXEC events */                                                      \
    "curl\0chmod\0cat\0bash\0ls\0sh\0/usr/bin/env\0nc\0tar\0./payload\0/tmp/malware.bin\0sync\0" \
    /* 90..101 exact surviving/near-miss paths */                                            \
    "/etc/localtime\0/home/kali/.config/curlrc\0/tmp/malware.bin\0/etc/passwd\0/dev\0/li\0"  \
    "/usr/li\0/etc/hos\0so\0/proc\0/tmp/x.so.1\0.s\0"

static const char g_blob_host[] = CHRONOS_GEN_BLOB;
#if defined(__CUDACC__)
__device__ const char g_blob_dev[] = CHRONOS_GEN_BLOB;
#endif
// (both are always declared; only the choice below depends on the compilation pass)
#if defined(__CUDA_ARCH__)
#define CHRONOS_GEN_BLOB_PTR g_blob_dev
#else
#define CHRONOS_GEN_BLOB_PTR g_blob_host
#endif

struct Out {
    uint8_t *p;
    int len, cap;
    CH_HD void put(uint8_t c) {
        if (len < cap) p[len] = c;
        len++;
    }
    CH_HD void puts_idx(int idx) {  // idx-th NUL-separated entry of the blob
        const char *b = CHRONOS_GEN_BLOB_PTR;
        int k = 0;
        while (k < idx) {
            while (*b) b++;
            b++;
            k++;
        }
        while (*b) put((uint8_t)*b++);
    }
    CH_HD void put_num(uint32_t v) {
        char t[10];
        int n = 0;
        do {
            t[n++] = (char)('0' + v % 10);
            v /= 10;
        } while (v);
        while (n) put((uint8_t)t[--n]);
    }
    CH_HD void put_alnum(Rng &r, int n) {
        for (int i = 0; i < n; i++) {
            uint32_t x = r.below(36);
            put((uint8_t)(x < 26 ? 'a' + x : '0' + (x - 26)));
        }
    }
};

CH_HD uint32_t pid_of(const chronos_tracegen_cfg &c, uint32_t k) {
    return c.pid_base + (uint32_t)(((uint64_t)k * c.pid_stride) % 4194000ULL);
}

CH_HD uint32_t pick_pid_index(const chronos_tracegen_cfg &c, Rng &r, const uint64_t *cdf) {
    if (c.use_cdf && cdf) {
        uint64_t u = r.next();
        uint32_t lo = 0, hi = c.n_pids - 1;  // first k with cdf[k] > u; cdf[n-1] == UINT64_MAX
        while (lo < hi) {
            uint32_t mid = lo + (hi - lo) / 2;
            if (cdf[mid] > u) hi = mid;
            else lo = mid + 1;
        }
        return lo;
    }
    return r.below(c.n_pids);
}

// Write one record (288 bytes at `rec`, any alignment).
CH_HD void gen_record(const chronos_tracegen_cfg &c, uint64_t i, const uint64_t *cdf, uint8_t *rec) {
    for (int k = 0; k < 288; k++) rec[k] = 0;
    Rng r;
    r.s = mix64(c.seed ^ mix64(i + 0x632BE59BD9B4E019ULL));
    Out comm = {rec + 4, 0, 15};   // <=15 chars + NUL (bpf_get_current_comm)
    Out argv = {rec + 20, 0, 255}; // <=255 chars + NUL (bpf_probe_read_user_str)
    Out type = {rec + 276, 0, 9};
    uint32_t pid;

    // planted dropper chain (config 4): three EXEC events of one dedicated PID
    if (c.chain_period) {
        uint32_t ph = (uint32_t)(i % c.chain_period);
        uint32_t cid = (uint32_t)(i / c.chain_period);
        int step = ph == 7 ? 0 : (ph == c.chain_period / 2 ? 1 : (ph == c.chain_period - 3 ? 2 : -1));
        if (step >= 0 && c.chain_period >= 32) {
            pid = c.chain_pid_base + cid;
            rec[0] = (uint8_t)pid; rec[1] = (uint8_t)(pid >> 8); rec[2] = (uint8_t)(pid >> 16); rec[3] = (uint8_t)(pid >> 24);
            comm.puts_idx(0);  // bash
            type.put('E'); type.put('X'); type.put('E'); type.put('C');
            if (step == 0) argv.puts_idx(78);       // curl
            else if (step == 1) argv.puts_idx(79);  // chmod
            else { argv.puts_idx(42); argv.put('.'); argv.put_num(cid); argv.put('/'); argv.puts_idx(55); }  // /tmp/.<c>/payload
            return;
        }
    }

    pid = pid_of(c, pick_pid_index(c, r, cdf));
    rec[0] = (uint8_t)pid; rec[1] = (uint8_t)(pid >> 8); rec[2] = (uint8_t)(pid >> 16); rec[3] = (uint8_t)(pid >> 24);

    uint32_t cls = r.below(1000);
    if (cls < c.adv_permille) {
        // ---- adversarial edge cases -------------------------------------------------
        uint32_t a = r.below(16);
        bool is_exec = (r.below(4) == 0);
        comm.puts_idx((int)r.below(12));
        if (is_exec) { type.put('E'); type.put('X'); type.put('E'); type.put('C'); }
        else { type.put('O'); type.put('P'); type.put('E'); type.put('N'); }
        switch (a) {
        case 0: break;  // empty argv (screenshot window 3: "[OPEN] chmod -> ")
        case 1:         // 255-byte path ending in ".so" 
            argv.puts_idx(42);
            argv.put_alnum(r, 255 - 5 - 3);
            argv.puts_idx(36);
            break;
        case 2:  // unterminated 256-byte argv (replay only; no NUL in the field)
            argv.cap = 256;
            argv.puts_idx(43);
            argv.put_alnum(r, 256 - 11);
            break;
        case 3: argv.puts_idx(94 + (int)r.below(8)); break;  // /dev /li /usr/li /etc/hos so /proc /tmp/x.so.1 .s
        case 4:  // invalid UTF-8 in argv: stray continuation / overlong / surrogate / >U+10FFFF
        {
            argv.puts_idx(42);
            uint32_t v = r.below(6);
            if (v == 0) argv.put(0x80);
            else if (v == 1) { argv.put(0xC0); argv.put(0xAF); }
            else if (v == 2) { argv.put(0xED); argv.put(0xA0); argv.put(0x80); }
            else if (v == 3) { argv.put(0xF4); argv.put(0x90); argv.put(0x80); argv.put(0x80); }
            else if (v == 4) { argv.put(0xE2); argv.put(0x82); }  // truncated tail
            else argv.put(0xFF);
            argv.put_alnum(r, (int)r.below(6));
            break;
        }
        case 5:  // valid multi-byte UTF-8 (2, 3 and 4 byte forms)
            argv.puts_idx(43);
            argv.put(0xC3); argv.put(0xA9);
            argv.put(0xE2); argv.put(0x82); argv.put(0xAC);
            argv.put(0xF0); argv.put(0x9F); argv.put(0x98); argv.put(0x80);
            argv.puts_idx(71);
            break;
        case 6:  // 255-byte cut splitting a code point (truncated by the 256-byte copy)
            argv.puts_idx(42);
            argv.put_alnum(r, 255 - 5 - 2);
            argv.put(0xE2); argv.put(0x82); argv.put(0xAC);  // third byte falls off the end
            break;
        case 7:  // invalid UTF-8 in comm
            comm.len = 0; comm.put('s'); comm.put('h'); comm.put(0xC3);
            argv.puts_idx(92);
            break;
        case 8:  // 15-char comm + junk after every NUL
            comm.len = 0; comm.puts_idx(19);
            argv.puts_idx(90);
            for (int k = argv.len + 1; k < 256; k++) rec[20 + k] = (uint8_t)('a' + (k % 23));
            rec[20 + 255] = 'o'; rec[20 + 254] = 's'; rec[20 + 253] = '.';
            for (int k = comm.len + 1; k < 16; k++) rec[4 + k] = 'c';
            rec[276 + 5] = 'c'; rec[276 + 6] = 'a'; rec[276 + 7] = 't';
            break;
        case 9:  // unterminated 16-byte comm carrying an ignore needle across the end
            comm.cap = 16; comm.len = 0;
            for (int k = 0; k < 12; k++) comm.put('x');
            comm.put('n'); comm.put('o'); comm.put('d'); comm.put('e');
            argv.puts_idx(92);
            break;
        case 10:  // needle-bearing / odd type fields (replay may carry any <=10 bytes)
            type.len = 0;
            for (int k = 0; k < 10; k++) rec[276 + k] = 0;
            if (r.below(2)) { type.put('c'); type.put('a'); type.put('t'); }
            else { type.cap = 10; for (int k = 0; k < 10; k++) type.put((uint8_t)("OPENOPENnc"[k])); }
            argv.puts_idx(27); argv.puts_idx(58);  // noise-looking path that is NOT filtered (type != OPEN)
            break;
        case 11: argv.puts_idx(91); break;  // .../.config/curlrc: no dot, survives
        case 12:  // broad trigger hits: license (nc), application (cat), sync
            argv.puts_idx(42 + (int)r.below(4)); argv.puts_idx(60 + (int)r.below(2));
            break;
        case 13:  // suffix exactly equal to the whole string / one short
            argv.puts_idx(36 + (int)r.below(6));
            break;
        case 14:  // trigger straddling a 16-byte chunk boundary of the record
            argv.puts_idx(42);
            argv.put_alnum(r, 3 + (int)r.below(20));
            argv.puts_idx(78 + (int)r.below(4));
            argv.put_alnum(r, (int)r.below(5));
            break;
        default:  // near-miss type: "OPEN" + extra char => not path-filtered
            type.put('X');
            argv.puts_idx(24); argv.puts_idx(59); argv.puts_idx(36);
            break;
        }
        return;
    }

    bool is_exec = r.below(1000) < c.exec_permille;
    if (is_exec) {
        // comm = the parent's name before exec replaces it; argv = argv[0]
        type.put('E'); type.put('X'); type.put('E'); type.put('C');
        uint32_t w = r.below(16);
        comm.puts_idx(w < 6 ? 0 : (w < 9 ? 10 : (int)r.below(24)));
        argv.puts_idx(78 + (int)r.below(12));
        return;
    }

    type.put('O'); type.put('P'); type.put('E'); type.put('N');
    comm.puts_idx((int)r.below(24));
    bool noise = r.below(1000) < c.noise_permille;
    bool deep = r.below(1000) < c.long_permille;
    if (noise) {
        uint32_t kind = r.below(16);
        if (kind < 11) {  // dropped by prefix
            argv.puts_idx(24 + (int)r.below(12));
            if (deep) {
                int target = 60 + (int)r.below(190);
                while (argv.len < target) { argv.put_alnum(r, 2 + (int)r.below(9)); argv.put('/'); }
            } else if (r.below(2)) { argv.put_alnum(r, 3 + (int)r.below(10)); argv.put('/'); }
            argv.puts_idx(54 + (int)r.below(16));
            uint32_t e = r.below(4);
            if (e == 0) { argv.puts_idx(36); argv.put('.'); argv.put_num(r.below(10)); }
            else if (e == 1) argv.puts_idx(36 + (int)r.below(6));
            else if (e == 2) argv.put_num(r.below(100000));
        } else {  // dropped by suffix only: surviving prefix + noise suffix
            argv.puts_idx(42 + (int)r.below(12));
            if (deep) {
                int target = 60 + (int)r.below(185);
                while (argv.len < target) { argv.put_alnum(r, 2 + (int)r.below(9)); argv.put('/'); }
            }
            argv.puts_idx(54 + (int)r.below(16));
            if (r.below(2)) argv.put_alnum(r, 1 + (int)r.below(6));
            argv.len = argv.len > 248 ? 248 : argv.len;
            argv.puts_idx(36 + (int)r.below(6));
        }
    } else {
        uint32_t kind = r.below(16);
        if (kind == 0) argv.puts_idx(90 + (int)r.below(4));  // /etc/localtime, curlrc, malware.bin, passwd
        else {
            argv.puts_idx(42 + (int)r.below(12));
            if (deep) {
                int target = 60 + (int)r.below(190);
                while (argv.len < target) { argv.put_alnum(r, 2 + (int)r.below(9)); argv.put('/'); }
            } else if (r.below(2)) { argv.put_alnum(r, 2 + (int)r.below(8)); argv.put('/'); }
            argv.puts_idx(54 + (int)r.below(16));
            if (r.below(3) == 0) { argv.put('_'); argv.put_num(r.below(1000)); }
            argv.puts_idx(70 + (int)r.below(8));
        }
    }
    // a live sensor never delivers more than 255 chars: anything longer was cut by `cap`
}

}  // namespace chronos_gen
