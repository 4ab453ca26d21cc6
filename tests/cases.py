"""Hand-written record sets shared by the golden generator, the oracle tests and the GPU
parity tests. Every case mirrors an edge the reference's code has."""
from oracle.pyoracle import make_record as R


def screenshot_trace():
    """Arrival order that reproduces the four windows of the reference's screenshot
    (README.md:43-47), with the libc/locale/ssl noise the openat hook drops in between."""
    return [
        R(2769, b"attack_chain.sh", b"/tmp/malware.bin", b"OPEN"),
        R(2769, b"attack_chain.sh", b"curl", b"EXEC"),
        R(2769, b"curl", b"/etc/ld.so.cache", b"OPEN"),
        R(2769, b"curl", b"/lib/x86_64-linux-gnu/libcurl.so.4", b"OPEN"),
        R(2780, b"attack_chain.sh", b"/dev/null", b"OPEN"),
        R(2780, b"attack_chain.sh", b"cat", b"EXEC"),
        R(2780, b"cat", b"/etc/ld.so.cache", b"OPEN"),
        R(2780, b"cat", b"/usr/lib/locale/locale-archive", b"OPEN"),
        R(2780, b"cat", b"/tmp/malware.bin", b"OPEN"),
        R(2779, b"attack_chain.sh", b"chmod", b"EXEC"),
        R(2779, b"chmod", b"/lib/x86_64-linux-gnu/libc.so.6", b"OPEN"),
        R(2779, b"chmod", b"", b"OPEN"),
        R(2769, b"curl", b"/etc/ssl/openssl.cnf", b"OPEN"),
        R(2769, b"curl", b"/etc/localtime", b"OPEN"),
        R(2769, b"curl", b"/usr/share/locale/en/LC_MESSAGES/curl.mo", b"OPEN"),
        R(2769, b"curl", b"/home/kali/.config/curlrc", b"OPEN"),
        R(2769, b"curl", b"/home/kali/.curlrc", b"OPEN"),
        R(2769, b"curl", b"/etc/hosts", b"OPEN"),
    ]


SCREENSHOT_WINDOWS = [
    (2769, ["[OPEN] attack_chain.sh -> /tmp/malware.bin", "[EXEC] attack_chain.sh -> curl"]),
    (2780, ["[EXEC] attack_chain.sh -> cat", "[OPEN] cat -> /tmp/malware.bin"]),
    (2779, ["[EXEC] attack_chain.sh -> chmod", "[OPEN] chmod -> "]),
    (2769, ["[OPEN] curl -> /etc/localtime", "[OPEN] curl -> /home/kali/.config/curlrc"]),
]


def adversarial_trace():
    recs = []
    pid = [100]

    def add(comm, argv, etype=b"OPEN", p=None, junk=b""):
        recs.append(R(pid[0] if p is None else p, comm, argv, etype, junk=junk))

    # --- anchored prefixes (chronos_sensor.py:76-78, :90) and their near misses
    for a in [b"/lib", b"/lib64/ld.so", b"/library", b"/usr/lib", b"/usr/libexec/x", b"/usr/share/a", b"/etc/ssl/c",
              b"/etc/fonts/f", b"/etc/host", b"/etc/hosts", b"/etc/hostname", b"/dev/null", b"/proc/1/stat",
              b"/dev", b"/proc", b"/li", b"/usr/li", b"/usr/shar", b"/etc/ss", b"/etc/font", b"/etc/hos",
              b"relative/lib/x", b"", b"/", b"/usr", b"/etc", b"x/dev/null", b"/Lib/x", b"/usr/share"]:
        add(b"sh", a)
    # --- suffix table (:83-85): exact, too short, near misses, max length
    for a in [b".so", b"so", b"o", b".s", b"/tmp/x.so", b"/tmp/x.so.1", b"/tmp/.cache", b".cache", b"cache",
              b"/a/b.mo", b".mo", b"mo", b"/x/y.conf", b".conf", b"conf", b"x.conf ", b"/c/d.crt", b".crt",
              b"/home/kali/.curlrc", b".curlrc", b"/home/kali/.config/curlrc", b"curlrc", b"/tmp/a.SO",
              b"/tmp/" + b"a" * 247 + b".so", b"/tmp/" + b"b" * 250, b"/tmp/" + b"c" * 248 + b".so",
              b"/tmp/" + b"d" * 251, b"/tmp/" + b"e" * 248 + b".mo"]:
        add(b"ls", a)
    # --- EXEC records are never path-filtered (:50-64)
    for a in [b"/lib/x", b"/usr/bin/x.so", b"/dev/null", b"curl", b"", b"/proc/self/exe"]:
        add(b"bash", a, b"EXEC")
    # --- type variants seen only in replays
    for t in [b"OPENX", b"open", b"", b"cat", b"OPENOPENnc", b"EXEC2", b"OPEN\xc3\xa9", b"\xff", b"nc"]:
        add(b"sh", b"/lib/not-filtered-unless-OPEN", t)
    # --- comm ignore-list (:134-135): substrings, case, 15/16-byte fields
    for c in [b"python3", b"vscode", b"gitlab-runner", b"unicode_start", b"nodejs", b"chrome", b"ollama",
              b"vmtoolsd", b"Node", b"Python", b"gi", b"ollam", b"xxxxxxxxxxxxnode", b"xxxxxxxxxxxxxnod",
              b"containerd-shim", b"mygit", b"pytho", b"vmtool", b"chrom", b"cod", b"digit"]:
        add(c, b"/tmp/file")
    # --- trigger needles (:141), incl. the broad ones
    for c, a in [(b"sync", b"/tmp/a"), (b"ls", b"/tmp/license"), (b"locate", b"/tmp/a"),
                 (b"ls", b"/etc/application"), (b"bash", b"/tmp/a"), (b"ls", b"chmod"), (b"ls", b"chmo"),
                 (b"ls", b"/x/curl"), (b"ls", b"cur"), (b"cat", b""), (b"ls", b"/tmp/ca"), (b"ls", b"n"),
                 (b"ls", b"/tmp/" + b"z" * 240 + b"nc"), (b"ls", b"/tmp/" + b"z" * 11 + b"bash"),
                 (b"ls", b"/tmp/" + b"z" * 9 + b"chmod"), (b"ls", b"/tmp/" + b"z" * 10 + b"chmod"),
                 (b"ls", b"/tmp/" + b"z" * 11 + b"chmod"), (b"ls", b"/tmp/" + b"z" * 12 + b"chmod")]:
        add(c, a)
    # --- UTF-8 gate (:128-131): valid multi-byte, every class of invalid, in each field
    good = [b"/tmp/\xc3\xa9", b"/tmp/\xe2\x82\xac", b"/tmp/\xf0\x9f\x98\x80", b"/tmp/\xed\x9f\xbf", b"/tmp/\xf4\x8f\xbf\xbf"]
    bad = [b"/tmp/\x80", b"/tmp/\xc0\xaf", b"/tmp/\xc1\xbf", b"/tmp/\xed\xa0\x80", b"/tmp/\xf4\x90\x80\x80",
           b"/tmp/\xe2\x82", b"/tmp/\xff", b"/tmp/\xe0\x9f\xbf", b"/tmp/\xf0\x8f\xbf\xbf", b"/tmp/\xf5\x80\x80\x80",
           b"/tmp/" + b"q" * 248 + b"\xe2\x82\xac"]
    for a in good + bad:
        add(b"sh", a)
    add(b"sh\xc3", b"/tmp/x")
    add(b"caf\xc3\xa9", b"/tmp/x")
    add(b"\xe2\x82\xacuro", b"/tmp/x", b"EXEC")
    # --- junk after the NUL must be invisible (replays; a live sensor zero-fills, :52/:69)
    add(b"sh", b"/tmp/clean", junk=b"cat.so")
    add(b"sh", b"/tmp/x.so", junk=b"zz")
    add(b"sh", b"", junk=b"/lib/curl")
    # --- flush automaton (:141-142, :157) on one PID: N, T flush, T, T flush, T, N, N, T flush
    p = 4242
    for a in [b"/tmp/n1", b"/tmp/cat1", b"/tmp/cat2", b"/tmp/cat3", b"/tmp/cat4", b"/tmp/n2", b"/tmp/n3", b"/tmp/cat5"]:
        add(b"sh", a, p=p)
    # trigger as the very first event of a PID, then a non-trigger: no window
    add(b"sh", b"/tmp/cat", p=5151)
    add(b"sh", b"/tmp/plain", p=5151)
    # interleaved PIDs, pid 0 and the largest pid
    for k in range(12):
        add(b"sh", b"/tmp/cat" if k % 3 == 2 else b"/tmp/f", p=[0, 0xFFFFFFFF, 7][k % 3])
    return recs
