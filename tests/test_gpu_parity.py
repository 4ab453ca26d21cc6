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
    pm = ctx.process_memory(res.counts["n_pending"])
    expect = {int(k): v for k, v in ref["process_memory"].items() if v}
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
