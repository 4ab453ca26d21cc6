"""-m gpu: the CUDA path, called through the C ABI, against the oracle (bit-exact)."""
import numpy as np
import pytest

from helpers import Trace, compare_batch
from oracle import coracle
from oracle.pyoracle import make_record as R

pytestmark = pytest.mark.gpu


@pytest.fixture()
def ctx():
    import chronos_b200

    c = chronos_b200.Context(device=0, max_batch_events=1 << 21)
    yield c
    c.close()


def screenshot_trace():
    import cases

    return b"".join(cases.screenshot_trace())


def test_screenshot_kat(ctx):
    """The reference's only recorded output: four windows (README screenshot)."""
    buf = screenshot_trace()
    res, _ = compare_batch(ctx, coracle.COracle(), Trace(), buf)
    assert res.histories() == [
        (2769, ["[OPEN] attack_chain.sh -> /tmp/malware.bin", "[EXEC] attack_chain.sh -> curl"]),
        (2780, ["[EXEC] attack_chain.sh -> cat", "[OPEN] cat -> /tmp/malware.bin"]),
        (2779, ["[EXEC] attack_chain.sh -> chmod", "[OPEN] chmod -> "]),
        (2769, ["[OPEN] curl -> /etc/localtime", "[OPEN] curl -> /home/kali/.config/curlrc"]),
    ]


def test_empty_and_tiny_batches(ctx):
    eng, tr = coracle.COracle(), Trace()
    compare_batch(ctx, eng, tr, b"", 0)
    compare_batch(ctx, eng, tr, R(7, b"sh", b"/tmp/a", b"OPEN"), 0)
    compare_batch(ctx, eng, tr, R(7, b"sh", b"/tmp/cat", b"OPEN"), 1)
    compare_batch(ctx, eng, tr, R(7, b"sh", b"/lib/x", b"OPEN"), 2)
    compare_batch(ctx, eng, tr, b"", 3)


@pytest.mark.parametrize("name,n", [("attack_replay_10k", 10_000), ("uniform_10m_4k", 200_000),
                                    ("skewed_zipf_chains", 200_000), ("sharded_100m_1m", 150_000)])
def test_workload_parity(ctx, name, n):
    from chronos_b200 import tracegen

    rows = tracegen.generate_host(tracegen.WORKLOADS[name], n=n)
    compare_batch(ctx, coracle.COracle(), Trace(), rows)


def test_streaming_state_carries_over(ctx):
    """Config 5 in small: micro-batches with persistent per-PID state == one-shot oracle."""
    from chronos_b200 import tracegen

    spec = dict(tracegen.WORKLOADS["stream_64k"])
    spec["n_pids"] = 256
    rows = tracegen.generate_host(spec, n=40_000)
    eng, tr = coracle.COracle(), Trace()
    base = 0
    for size in (1, 7, 4096, 128, 10_000, 3, 25_765):
        compare_batch(ctx, eng, tr, rows[base:base + size], base)
        base += size
    assert base == 40_000


def test_random_pids_and_single_pid(ctx):
    rng = np.random.default_rng(7)
    recs = []
    comms = [b"sh", b"bash", b"sshd", b"cat", b"cron"]
    paths = [b"/tmp/a", b"/tmp/nc", b"/home/u/x.txt", b"/lib/q", b"/var/log/cat.log", b"", b"/etc/passwd"]
    for i in range(30_000):
        pid = int(rng.integers(0, 2**32)) if i % 3 else int(rng.integers(0, 50))
        recs.append(R(pid, comms[rng.integers(len(comms))], paths[rng.integers(len(paths))],
                      b"EXEC" if rng.integers(8) == 0 else b"OPEN"))
    compare_batch(ctx, coracle.COracle(), Trace(), b"".join(recs))
    one = b"".join(R(4242, comms[i % 5], paths[(i * 3) % 7], b"OPEN") for i in range(5000))
    ctx.reset()  # fresh process_memory, like the fresh oracle below
    compare_batch(ctx, coracle.COracle(), Trace(), one)


def test_tracegen_device_twin_matches_host():
    from chronos_b200 import tracegen

    for name in ("uniform_10m_4k", "skewed_zipf_chains"):
        spec = tracegen.WORKLOADS[name]
        host = tracegen.generate_host(spec, first=12345, n=50_000)
        dev = tracegen.generate_device(spec, device=0, first=12345, n=50_000)
        got = dev.download().reshape(-1, 288)
        dev.free()
        np.testing.assert_array_equal(got, host)


@pytest.mark.parametrize("name", ["screenshot_kat.json", "adversarial.json", "config1_trace.json"])
def test_golden_vectors_from_the_unmodified_reference(ctx, name):
    """CUDA path vs what chronos_sensor.py itself produced (tests/golden, made in the build
    container by running the unmodified script)."""
    from test_oracle_golden import fixture_records

    data, ref = fixture_records(name)
    ctx.submit(np.frombuffer(data, dtype=np.uint8), 0)
    res = ctx.fetch(records=True)
    assert res.survivors.tolist() == ref["survivors"]
    assert [[p, h] for p, h in res.histories()] == ref["windows"]
    pm = ctx.process_memory(res.counts["n_pending"], res.counts["n_pids_seen"])
    expect = {int(k): v for k, v in ref["process_memory"].items()}  # including the [] of flushed PIDs (:157)
    assert pm == expect


def test_random_adversarial_records(ctx):
    """Needle fragments, separators, multi-byte and broken UTF-8, 15/16/255/256-byte fields,
    junk after the NUL, odd type fields: differential against the oracle."""
    import random

    from test_oracle_pinning import rand_record

    rng = random.Random(77)
    eng, tr = coracle.COracle(), Trace()
    base = 0
    for n in (1, 33, 4000, 20_000):
        data = b"".join(rand_record(rng) for _ in range(n))
        compare_batch(ctx, eng, tr, data, base)
        base += n


def test_sharded_stages_emulated_on_one_gpu():
    """chronos_filter_partition + chronos_submit_survivors (the two CUDA stages of the
    multi-GPU path) with 3 emulated ranks on one device: the all-to-all is replaced by
    host-side routing of the packed send buffers; merged output must equal the oracle."""
    import ctypes
    from ctypes import c_void_p

    import chronos_b200
    from chronos_b200 import _ffi, sharded, tracegen
    from helpers import canonical

    L = _ffi.lib()
    world = 3
    spec = dict(tracegen.WORKLOADS["skewed_zipf_chains"])
    rows = tracegen.generate_host(spec, n=90_000)
    n = len(rows)
    bounds = [r * n // world for r in range(world + 1)]
    ctxs = [chronos_b200.Context(device=0, max_batch_events=n) for _ in range(world)]
    sent = []  # per source rank: (slots[*,16], hdr[cnt], entry counts, slot counts)
    for r in range(world):
        sl = rows[bounds[r]:bounds[r + 1]]
        cap = len(sl)
        d_in = tracegen.DeviceBuffer(0, cap * 288)
        d_in.upload(sl)
        d_slots, d_hdr = tracegen.DeviceBuffer(0, cap * 18 * 16), tracegen.DeviceBuffer(0, cap * 8)
        ec, sc = (ctypes.c_uint64 * world)(), (ctypes.c_uint64 * world)()
        rc = L.chronos_filter_partition(ctxs[r]._h, c_void_p(d_in.ptr), cap, bounds[r], world, c_void_p(d_slots.ptr),
                                        cap * 18, c_void_p(d_hdr.ptr), cap, ec, sc)
        ctxs[r]._check(rc)
        ec, sc = list(ec), list(sc)
        sent.append((d_slots.download(sum(sc) * 16).reshape(-1, 16), d_hdr.download(sum(ec) * 8).view(np.uint64), ec, sc))
        for b in (d_in, d_slots, d_hdr):
            b.free()
    # survivors of all ranks = oracle survivors; transport is much smaller than 288 B per survivor
    keep, meta = coracle.classify(rows)
    all_arr = np.sort(np.concatenate([s[1] & np.uint64(sharded.HDR_ARRIVAL_MASK) for s in sent]))
    np.testing.assert_array_equal(all_arr, np.nonzero(keep)[0].astype(np.uint64))
    assert sum(sum(s[3]) for s in sent) * 16 < 0.6 * 288 * len(all_arr)
    merged = []
    for d in range(world):  # route: destination d receives its block from every source, in source order
        slots, hdrs = [], []
        for r in range(world):
            sl, hd, ec, sc = sent[r]
            lo_e, lo_s = sum(ec[:d]), sum(sc[:d])
            slots.append(sl[lo_s:lo_s + sc[d]]); hdrs.append(hd[lo_e:lo_e + ec[d]])
        sl, hd = np.concatenate(slots), np.concatenate(hdrs)
        arr = (hd & np.uint64(sharded.HDR_ARRIVAL_MASK)).astype(np.int64)
        assert np.all(np.diff(arr) > 0)
        np.testing.assert_array_equal((hd >> np.uint64(48)) & np.uint64(0x0F), meta[arr] & 0x0F)
        k = len(arr)
        b_slots, b_hdr = tracegen.DeviceBuffer(0, max(len(sl), 1) * 16), tracegen.DeviceBuffer(0, max(k, 1) * 8)
        if k:
            b_slots.upload(sl); b_hdr.upload(hd)
        ctxs[d].submit_survivors_ptr(b_slots.ptr, b_hdr.ptr, k, len(sl))
        res = ctxs[d].fetch(records=True)
        np.testing.assert_array_equal(res.survivors.astype(np.int64), arr)
        if k:
            pid = res.window_records[:, :4].copy().view("<u4").reshape(-1)
            assert all(chronos_b200.owner_of(int(p), world) == d for p in pid[:200])
        o = 0
        for w in res.windows:
            ln = int(w["len"])
            merged.append((int(w["trigger_idx"]), int(w["pid"]), res.window_records[o:o + ln].copy()))
            o += ln
        for b in (b_slots, b_hdr):
            b.free()
    merged.sort(key=lambda t: t[0])
    ref = coracle.COracle().feed(rows, 0)
    assert [m[0] for m in merged] == ref["win_trigger"].tolist()
    assert [m[1] for m in merged] == ref["win_pid"].tolist()
    # the transport form drops whole chunks behind argv's NUL; junk inside a transported chunk
    # survives. Both are invisible to the reference (fields are read up to the NUL, :125-130),
    # so records are compared in canonical form: bytes behind each field's NUL zeroed.
    exp = canonical(rows[ref["members"].astype(np.int64)])
    np.testing.assert_array_equal(canonical(np.concatenate([m[2] for m in merged])), exp)
    for c in ctxs:
        c.close()


def test_sensor_drop_in_reproduces_reference_stdout_and_prompts():
    """GpuSensor (print_event / process_memory / analyze_sequence surface) against what the
    unmodified chronos_sensor.py printed and POSTed for the same records (golden fixtures;
    the stub LLM answers like oracle/ref_shim.py's)."""
    import ctypes

    from chronos_b200 import sensor
    from test_oracle_golden import fixture_records, load

    for name in ("screenshot_kat.json", "adversarial.json"):
        data, ref = fixture_records(name)
        lines, posts = [], []

        class Resp:
            def json(self):
                return {"response": '{"risk_score": 0, "verdict": "SAFE", "reason": "stub"}'}

        def post(url, json=None, timeout=None):
            posts.append(json["prompt"])
            return Resp()

        out = lambda s: lines.append(s)  # noqa: E731
        s = sensor.GpuSensor(sensor.default_analyze_sequence("http://x:11434/api/generate", post=post, out=out),
                             device=0, batch_events=7, out=out)
        buf = ctypes.create_string_buffer(data, len(data))
        for i in range(len(data) // 288):  # the way BCC would call it
            s.print_event(0, ctypes.c_void_p(ctypes.addressof(buf) + 288 * i), 288)
        pm = s.process_memory
        assert posts == ref["prompts"]
        assert pm == {int(k): v for k, v in ref["process_memory"].items()}
        # everything the unmodified script printed after its banner (:143-155), byte for byte
        assert "\n".join(lines) + "\n" == ref["stdout_after_banner"]
        s.close()


def test_async_download_equals_blocking_copies(ctx):
    """chronos_download_begin/_wait deliver exactly what the blocking copy calls return, also
    with two downloads in flight while the context already works on the next batch."""
    import ctypes

    from chronos_b200 import _ffi, tracegen

    L = _ffi.lib()
    rows = tracegen.generate_host(tracegen.WORKLOADS["stream_64k"], n=60_000)
    thirds = [rows[0:20_000], rows[20_000:40_000], rows[40_000:60_000]]
    cap = 20_000 * (288 + 9 + 32) * 2
    bufs = []
    for _ in range(3):
        p = ctypes.c_void_p()
        assert L.chronos_host_alloc_pinned(cap, ctypes.byref(p)) == 0
        bufs.append(p)
    ctx.reset()
    expect, layouts = [], []
    for k, part in enumerate(thirds):
        ctx.submit(part, k * 20_000)
        res = ctx.fetch(records=True)                        # blocking reference copies
        expect.append(res)
        layouts.append(ctx.download_begin(bufs[k].value, cap))   # and the same batch asynchronously
    ctx.download_wait()
    for k in range(3):
        arr = np.ctypeslib.as_array((ctypes.c_uint8 * cap).from_address(bufs[k].value))
        idx, meta, win, rec = ctx.parse_download(arr, layouts[k])
        np.testing.assert_array_equal(idx, expect[k].survivors)
        np.testing.assert_array_equal(meta, expect[k].survivor_meta)
        assert win.tobytes() == expect[k].windows.tobytes()
        np.testing.assert_array_equal(rec, expect[k].window_records)
        assert layouts[k]["n_windows"] > 100
    with pytest.raises(_ffi.ChronosError):
        ctx.download_begin(bufs[0].value, 16)                # too small: reported, nothing written
    for p in bufs:
        L.chronos_host_free_pinned(p)


def test_one_pid_across_many_tiles_and_extreme_pids(ctx):
    """A single PID whose list spans hundreds of scan tiles (look-back carries one open group
    across all of them), runs of consecutive triggers, and the pid values 0 and 2^32-1."""
    rng = np.random.default_rng(11)
    comms = [b"sh", b"bash", b"sshd", b"cat", b"cron", b"vim"]
    paths = [b"/tmp/a", b"/tmp/nc", b"/home/u/x.txt", b"/var/log/cat.log", b"", b"/etc/passwd", b"/opt/tool"]
    n = 300_000
    one = b"".join(R(77, comms[int(rng.integers(len(comms)))], paths[int(rng.integers(len(paths)))],
                     b"EXEC" if rng.integers(16) == 0 else b"OPEN") for _ in range(n))
    ctx.reset()
    eng, tr = coracle.COracle(), Trace()
    compare_batch(ctx, eng, tr, one, 0)
    # long stretch without any trigger, then all triggers: the list only grows, then flushes pairwise
    quiet = b"".join(R(77, b"vim", b"/home/u/notes", b"OPEN") for _ in range(50_000))
    loud = b"".join(R(77, b"bash", b"/tmp/x", b"OPEN") for _ in range(50_001))
    compare_batch(ctx, eng, tr, quiet, n)
    compare_batch(ctx, eng, tr, loud, n + 50_000)
    ext = b"".join(R(pid, comms[i % 6], paths[i % 7], b"OPEN") for i in range(4000) for pid in (0, 0xFFFFFFFF, 1, 0x80000000))
    compare_batch(ctx, eng, tr, ext, n + 100_001)
