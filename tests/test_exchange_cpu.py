"""world_size-2 (and 4) gloo tests of the multi-GPU step: contiguous candidate shards, one 16-byte
all-gather per rank, identical deterministic argmax on every rank.  The per-shard
results come from the oracle here; on the GPU box the same exchange runs over NCCL in bench.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pointcloud_processor_b200 import exchange


def test_shard_ranges_partition_the_candidates():
    for C in (0, 1, 7, 232, 1930, 14656):
        for world in (1, 2, 4, 8):
            r = [exchange.shard_range(C, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == C
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [e - b for b, e in r]
            assert max(sizes) - min(sizes) <= 1


def test_pick_best_is_first_maximum():
    pairs = torch.stack([exchange.pack_pair(3.5, 40), exchange.pack_pair(7.25, 90), exchange.pack_pair(7.25, 61),
                         exchange.pack_pair(-np.inf, -1)])
    assert exchange.pick_best(pairs) == (7.25, 61)
    assert exchange.pick_best(torch.stack([exchange.pack_pair(-np.inf, -1)] * 2)) == (-np.inf, -1)
    s, i = exchange.pick_best(torch.stack([exchange.pack_pair(0.1 + 0.2, 5)]))
    assert s == 0.1 + 0.2 and i == 5          # the double survives the int64 round trip bit for bit


def _worker(rank, world, port, scores, out_q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        b, e = exchange.shard_range(len(scores), rank, world)
        local = scores[b:e]
        if len(local):
            k = int(np.argmax(local))          # numpy argmax = first maximum, like the kernel
            pair = exchange.pack_pair(float(local[k]), b + k)
        else:
            pair = exchange.pack_pair(-np.inf, -1)
        gathered = exchange.all_gather_best(pair)
        out_q.put((rank, exchange.pick_best(gathered)))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_argmax_over_gloo(world, scene_cache):
    from oracle import binding as O
    s = scene_cache("tiny")
    ref = O.sweep(O.Index(s.terrain, O.GRID), None, s.cells, s.params, s.fixed, s.candidates)
    scores = ref["scores"].copy()
    scores[3] = scores[ref["best_index"]]      # plant a tie on a lower index and on another shard
    expect = (float(scores.max()), int(np.argmax(scores)))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, scores, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert len(got) == world and all(g[1] == expect for g in got), (got, expect)
