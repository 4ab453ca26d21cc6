"""CPU, world_size 2 over gloo: the host logic of the PID-hash sharded path
(chronos_b200.sharded.exchange + owner function + merge by trigger index). The two CUDA
stages are stood in for by the oracle here (test infrastructure only); the property
checked is the one SURVEY 8e demands: sharding is unobservable, i.e. the merged result
of the ranks equals the single-process result."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

REC = 288


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _trace():
    import cases
    from chronos_b200 import tracegen

    spec = dict(tracegen.WORKLOADS["attack_replay_10k"])
    spec["n_pids"] = 300
    rows = tracegen.generate_host(spec, n=6000)
    adv = np.frombuffer(b"".join(cases.adversarial_trace()), dtype=np.uint8).reshape(-1, REC)
    return np.concatenate([rows[:3000], adv, rows[3000:]])


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from chronos_b200 import sharded
        from oracle import coracle

        rows = _trace()
        n = len(rows)
        lo, hi = rank * n // world, (rank + 1) * n // world
        mine = rows[lo:hi]
        # ---- stage 1 stand-in: classify the slice, pack survivors per owner (stable) ----
        keep, meta = coracle.classify(mine)
        idx = np.nonzero(keep)[0]
        pid = mine[idx, :4].copy().view("<u4").reshape(-1)
        owner = sharded.owner_of_np(pid, world)
        order = np.argsort(owner, kind="stable")
        counts = np.bincount(owner, minlength=world)
        send_records = torch.from_numpy(np.ascontiguousarray(mine[idx][order]).reshape(-1))
        # header words as on the GPU path; the payload unit here is a whole 288-byte record
        hdr = sharded.pack_hdr(lo + idx[order], meta[idx][order], np.full(len(idx), 18))
        send_hdr = torch.from_numpy(hdr.view(np.int64).copy())
        cap = n
        recv_records = torch.empty(cap * REC, dtype=torch.uint8)
        recv_hdr = torch.empty(cap, dtype=torch.int64)
        n_recv, n_units = sharded.exchange(send_records, send_hdr, counts, counts, REC, recv_records, recv_hdr)
        assert n_units == n_recv
        got = recv_records[: n_recv * REC].numpy().reshape(-1, REC)
        rh = recv_hdr[:n_recv].numpy().view(np.uint64)
        arr = (rh & np.uint64(sharded.HDR_ARRIVAL_MASK)).astype(np.int64)
        # every received event belongs here, and arrives in global arrival order
        rp = got[:, :4].copy().view("<u4").reshape(-1)
        assert np.all(sharded.owner_of_np(rp, world) == rank)
        assert np.all(np.diff(arr) > 0)
        assert np.array_equal(got, rows[arr])
        # ---- stage 2 stand-in: this rank's process_memory over what it received ----------
        eng = coracle.COracle()
        wins = []
        for k in range(n_recv):
            o = eng.feed(got[k], int(arr[k]))
            for w in range(len(o["win_pid"])):
                a, b = int(o["win_off"][w]), int(o["win_off"][w + 1])
                wins.append((int(o["win_trigger"][w]), int(o["win_pid"][w]), o["members"][a:b].tolist()))
        pids, offs, mem = eng.state()
        state = {int(p): mem[int(offs[k]):int(offs[k + 1])].tolist() for k, p in enumerate(pids)}
        gathered = [None] * world
        dist.all_gather_object(gathered, (wins, state, arr.tolist()))
        if rank == 0:
            ref = coracle.COracle()
            o = ref.feed(rows, 0)
            exp = [(int(o["win_trigger"][w]), int(o["win_pid"][w]),
                    o["members"][int(o["win_off"][w]):int(o["win_off"][w + 1])].tolist())
                   for w in range(len(o["win_pid"]))]
            merged = sorted(w for g in gathered for w in g[0])  # merge by trigger arrival index
            assert merged == exp
            assert sorted(a for g in gathered for a in g[2]) == o["survivors"].tolist()
            rp_, ro, rm = ref.state()
            exp_state = {int(p): rm[int(ro[k]):int(ro[k + 1])].tolist() for k, p in enumerate(rp_)}
            got_state = {}
            for g in gathered:
                assert not (set(g[1]) & set(got_state)), "a PID lives on two ranks"
                got_state.update(g[1])
            assert got_state == exp_state
            open(os.path.join(out_dir, "ok"), "w").write(f"{len(exp)} windows")
    finally:
        dist.destroy_process_group()


def test_two_rank_sharding_is_unobservable(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok").exists()


def test_numpy_owner_function_matches_the_library():
    import chronos_b200
    from chronos_b200 import sharded

    rng = np.random.default_rng(5)
    pids = np.concatenate([rng.integers(0, 2**32, 500, dtype=np.uint64).astype(np.uint32),
                           np.array([0, 1, 0xFFFFFFFF, 2769], dtype=np.uint32)])
    for world in (1, 2, 3, 8):
        exp = [chronos_b200.owner_of(int(p), world) for p in pids]
        assert sharded.owner_of_np(pids, world).tolist() == exp
