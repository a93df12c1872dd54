"""Multi-GPU host logic on CPU: stream partition + output gather over gloo, world_size 2
(SURVEY.md 8(e)).  Each rank stands in for a GPU by producing its shard's per-frame rows
with the CPU oracle; rank 0 must see exactly what a single process tracking all streams sees."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from boxmot_b200 import sharding
from conftest import OCSORT_YAML

N_STREAMS, N_FRAMES, K, ROWS = 5, 12, 12, 32


def test_shard_range_covers_every_stream_once():
    for n in (0, 1, 5, 8, 1024, 8192, 1023):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                lo, hi = sharding.shard_range(n, r, world)
                assert 0 <= lo <= hi <= n
                seen += list(range(lo, hi))
                for s in range(lo, hi):
                    assert sharding.owner_of(s, n, world) == r
            assert seen == list(range(n))
            sizes = [np.subtract(*sharding.shard_range(n, r, world)[::-1]) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(8, 2, 2)


def test_compact_rows_is_ragged_view_of_padded_output():
    rng = np.random.default_rng(0)
    out = rng.random((4, 6, 8))
    nout = np.array([0, 6, 2, 1], np.int32)
    rows, cnt = sharding.compact_rows(out, nout)
    assert rows.shape == (9, 9) and cnt.sum() == 9
    assert np.array_equal(rows[:, 0], [1] * 6 + [2] * 2 + [3])
    assert np.array_equal(rows[6:8, 1:], out[2, :2])


def _track_streams(lo, hi):
    """Per-frame padded outputs [F][S_local, ROWS, 8] / counts of streams lo..hi-1."""
    from boxmot_b200.synth import make_stream
    from oracle.ocsort import OcSortOracle
    outs, cnts = [], []
    trackers = [OcSortOracle(**OCSORT_YAML) for _ in range(lo, hi)]
    frames = [make_stream(s, K, N_FRAMES) for s in range(lo, hi)]
    for f in range(N_FRAMES):
        out = np.zeros((hi - lo, ROWS, 8))
        n = np.zeros((hi - lo,), np.int32)
        for i, t in enumerate(trackers):
            rows = np.asarray(t.update(frames[i][f]), dtype=float).reshape(-1, 8)
            out[i, : len(rows)] = rows
            n[i] = len(rows)
        outs.append(out)
        cnts.append(n)
    return outs, cnts


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = sharding.shard_range(N_STREAMS, rank, world)
        outs, cnts = _track_streams(lo, hi)
        got = []
        for f in range(N_FRAMES):
            g = sharding.gather_rows(torch.from_numpy(outs[f]), torch.from_numpy(cnts[f]), N_STREAMS)
            assert (g is None) == (rank != 0)
            got.append(g)
        t, u = sharding.reduce_timing(10.0 + rank, 100 * (rank + 1))
        assert t == 10.0 + world - 1 and u == 100 * world * (world + 1) / 2
        if rank == 0:
            q.put(got)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.timeout(300)
def test_two_rank_gather_equals_single_process():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=240)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    outs, cnts = _track_streams(0, N_STREAMS)
    assert sum(int(c.sum()) for c in cnts) > 50          # the case is not vacuous
    for f in range(N_FRAMES):
        assert len(got[f]) == N_STREAMS
        for s in range(N_STREAMS):
            assert np.array_equal(got[f][s], outs[f][s, : cnts[f][s]]), (f, s)
