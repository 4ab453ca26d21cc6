"""CPU: properties that pin the oracle's building blocks (CPython's own decoder, the
reference's own C helpers) and, where /root/reference is present, a differential test
of the whole path against the UNMODIFIED reference on adversarial random records."""
import random

import pytest

from oracle import coracle, pyoracle, ref_shim

ALPHABET = [b"c", b"u", b"r", b"l", b"h", b"m", b"o", b"d", b"b", b"a", b"s", b"n", b"t", b"/", b".", b"i", b"g",
            b"e", b"p", b"y", b"v", b"x", b" ", b"-", b">", b"[", b"]", b"'", b"\"", b"\\", b"\xc3\xa9",
            b"\xe2\x82\xac", b"\xf0\x9f\x98\x80", b"\x80", b"\xc0", b"\xed\xa0\x80", b"\xff", b"\x01"]
WORDS = [b"curl", b"chmod", b"bash", b"nc", b"cat", b"node", b"code", b"ollama", b"python", b"chrome", b"vmtools",
         b"git", b"/lib", b"/usr/lib", b"/usr/share", b"/etc/ssl", b"/etc/fonts", b"/etc/host", b"/dev/", b"/proc/",
         b".so", b".cache", b".mo", b".conf", b".crt", b".curlrc", b"/tmp/", b"/home/kali/"]


def rand_bytes(rng, maxlen):
    n = rng.choice([0, 1, 2, 3, 5, 8, 13, 15, 16, 17, 40, 100, 254, 255, 256])
    out = b""
    while len(out) < min(n, maxlen):
        out += rng.choice(WORDS) if rng.random() < 0.4 else rng.choice(ALPHABET)
    return out[:maxlen] if rng.random() < 0.8 else out


def rand_record(rng):
    pid = rng.choice([1, 2, 3, 7, 0, 0xFFFFFFFF, rng.randrange(1 << 32)])
    comm = rand_bytes(rng, 16)
    argv = rand_bytes(rng, 256)
    etype = rng.choice([b"OPEN", b"OPEN", b"OPEN", b"EXEC", b"EXEC", b"", b"OPENX", rand_bytes(rng, 10)])
    junk = rng.choice([b"", b"", b"cat", b"\xff.so"])
    return pyoracle.make_record(pid, comm, argv, etype, junk=junk)


def test_utf8_validity_matches_cpython_exhaustively_on_short_sequences():
    def py_ok(b):
        try:
            b.decode("utf-8")
            return True
        except UnicodeDecodeError:
            return False

    interesting = [0x00 + 1, 0x41, 0x7F, 0x80, 0x8F, 0x90, 0x9F, 0xA0, 0xBF, 0xC0, 0xC1, 0xC2, 0xDF, 0xE0, 0xE1, 0xEC,
                   0xED, 0xEE, 0xEF, 0xF0, 0xF1, 0xF3, 0xF4, 0xF5, 0xFF]
    for a in interesting:
        assert coracle.utf8_valid(bytes([a])) == py_ok(bytes([a]))
        for b in interesting:
            assert coracle.utf8_valid(bytes([a, b])) == py_ok(bytes([a, b]))
            for c in interesting:
                s = bytes([a, b, c])
                assert coracle.utf8_valid(s) == py_ok(s), s
                for d in (0x41, 0x80, 0x8F, 0x90, 0xBF, 0xC2):
                    s4 = bytes([a, b, c, d])
                    assert coracle.utf8_valid(s4) == py_ok(s4), s4
    assert coracle.utf8_valid(b"")


def test_c_oracle_equals_python_port_on_random_adversarial_records():
    rng = random.Random(1234)
    for trial in range(40):
        data = b"".join(rand_record(rng) for _ in range(rng.choice([1, 5, 50, 300])))
        out = coracle.COracle().feed(data, 10)
        s = pyoracle.PySensor().feed(data, 10)
        assert out["survivors"].tolist() == s.survivors
        assert out["win_trigger"].tolist() == s.trigger_idx
        assert out["win_pid"].tolist() == [p for p, _ in s.windows]
        assert [int(b - a) for a, b in zip(out["win_off"][:-1], out["win_off"][1:])] == [len(h) for _, h in s.windows]


def test_openat_filter_equals_reference_c_helpers():
    """oracle/_ref = the reference's own starts_with / ends_with (chronos_sensor.py:28-47)."""
    from oracle import build_oracle

    if build_oracle.build_ref() is None:
        pytest.skip("oracle/_ref not built (no reference on this machine and no prebuilt library)")
    rng = random.Random(99)
    recs = [pyoracle.make_record(1, b"sh", rand_bytes(rng, 256), b"OPEN", junk=rng.choice([b"", b".so"]))
            for _ in range(3000)]
    drop = ref_shim.ref_filter_drop(b"".join(recs))
    for r, d in zip(recs, drop):
        assert coracle.openat_noise(r[20:276]) == bool(d)
        assert pyoracle.openat_noise(r[20:276]) == bool(d)


@pytest.mark.skipif(not ref_shim.available(), reason="the unmodified reference is only present in the build container")
def test_whole_path_equals_unmodified_reference_on_random_records():
    rng = random.Random(2024)
    for trial in range(6):
        data = b"".join(rand_record(rng) for _ in range(400))
        ref = ref_shim.run_reference(data)
        out = coracle.COracle().feed(data, 0)
        s = pyoracle.PySensor().feed(data, 0)
        assert out["survivors"].tolist() == ref["survivors"] == s.survivors
        assert [(p, h) for p, h in s.windows] == ref["windows"]
        assert dict(s.process_memory) == ref["process_memory"]
        assert out["win_pid"].tolist() == [p for p, _ in ref["windows"]]
        for hist, prompt in zip([h for _, h in ref["windows"]], ref["prompts"]):
            assert pyoracle.prompt_of(hist) == prompt
