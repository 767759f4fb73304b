"""Utterance sharding across ranks: host logic on CPU, world_size 2 over gloo.

The reference splits ``wav.scp`` into N contiguous parts and merges per-process result files
(``/root/reference/infer_kaldidata.py:493-544,1013-1062``); no tensor ever crosses ranks."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

from funasr_b200.sharding import gather_results, shard_contiguous, shard_lpt, utterance_cost


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_contiguous_split_is_the_reference_scheme():
    # infer_kaldidata.py:511-542: base = n // N, the first n % N parts get one more, order preserved
    parts = shard_contiguous(10, 4)
    assert parts == [[0, 1, 2], [3, 4, 5], [6, 7], [8, 9]]
    assert shard_contiguous(2, 4) == [[0], [1], [], []]
    assert sum(shard_contiguous(1237, 8), []) == list(range(1237))


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_lpt_is_a_partition_and_balances_varlen(world):
    rng = np.random.default_rng(0)
    T = [int(-(-(1 + (int(s * 16000) - 400) // 160) // 6)) for s in rng.uniform(1.0, 30.0, 512)]
    parts = shard_lpt(T, world)
    assert sorted(sum(parts, [])) == list(range(len(T)))
    assert all(p == sorted(p) for p in parts)
    load = [sum(utterance_cost(T[i]) for i in p) for p in parts]
    assert max(load) <= 1.02 * (sum(load) / world)  # within 2 % of perfect balance on 512 utterances
    if world > 1:
        naive = [sum(utterance_cost(T[i]) for i in p) for p in shard_contiguous(len(T), world)]
        assert max(load) <= max(naive)
    assert shard_lpt(T, world) == parts  # deterministic: every rank computes the same partition without talking


def test_lpt_more_ranks_than_utterances():
    parts = shard_lpt([100, 50], 4)
    assert sorted(sum(parts, [])) == [0, 1] and sum(1 for p in parts if p) == 2


def _worker(rank, world, port, n_items, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        T = [17 + (7 * i) % 480 for i in range(n_items)]
        mine = shard_lpt(T, world)[rank]
        # stand-in for the per-utterance result of the hot path: (rows, checksum) derived from the utterance only
        local = {f"utt{i:04d}": (T[i], (T[i] * 2654435761) % 9973) for i in mine}
        merged = gather_results(local)
        q.put((rank, mine, list(merged.items())))
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo_shard_and_merge():
    world, n = 2, 37
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    got.sort()
    (_, mine0, merged0), (_, mine1, merged1) = got
    assert sorted(mine0 + mine1) == list(range(n)) and not set(mine0) & set(mine1)
    assert merged0 == merged1  # every rank ends with the same merged view
    keys = [k for k, _ in merged0]
    assert keys == sorted(keys) and len(keys) == n  # ordered by utterance id like the reference's merged files
    T = [17 + (7 * i) % 480 for i in range(n)]
    assert all(v == (T[int(k[3:])], (T[int(k[3:])] * 2654435761) % 9973) for k, v in merged0)


def test_bind_to_gpu_numa_is_best_effort():
    """No GPU / no NVML here: the helper must change nothing and say so (None); with NVML it returns the cores it bound to."""
    import os

    from funasr_b200.sharding import bind_to_gpu_numa

    before = os.sched_getaffinity(0)
    got = bind_to_gpu_numa(0)
    after = os.sched_getaffinity(0)
    assert got is None and after == before or set(got) == after and after <= before


def test_pipeline_piece_boundaries():
    """Host logic of AudioPipeline.run_pinned's overlapped schedule: contiguous pieces of near-equal cost, every piece non-empty,
    boundaries strictly increasing from 0 to B, never more pieces than asked for (or than utterances)."""
    import numpy as np

    from funasr_b200.pipeline import AudioPipeline as P

    cost = lambda T: 0.4657152 * T + 1.51552e-4 * T.astype(np.float64) ** 2
    rng = np.random.default_rng(5)
    for B in (2, 3, 10, 64, 360):
        for T in (np.full(B, 167), rng.integers(1, 500, B)):
            for pieces in (2, 3, 4, 8, 1000):
                cuts = P._cuts(T, pieces)
                assert cuts[0] == 0 and cuts[-1] == B and all(a < b for a, b in zip(cuts[:-1], cuts[1:]))
                assert len(cuts) - 1 <= min(pieces, B)
                if B >= 8 * pieces:  # enough utterances: every piece within one utterance's cost (each side) of the equal share
                    c = cost(T)
                    share, worst = c.sum() / pieces, c.max()
                    for a, b in zip(cuts[:-1], cuts[1:]):
                        assert abs(c[a:b].sum() - share) <= 2 * worst + 1e-9
    assert P._cuts(np.full(360, 167), 2) == [0, P._cut(np.full(360, 167)), 360]
