"""CPU tests of the N > 1 host logic with the gloo backend, world_size 2."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gym_air_hockey_b200.parallel import all_reduce_mean, reduce_statistics, shard_range, statistics_dict


def test_shard_range_partitions():
    for total in (1, 7, 4096, 1048576, 65536 + 3):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, results):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # statistics: rank r contributes (r+1) * [0..31]
        stats = torch.arange(32, dtype=torch.int64) * (rank + 1)
        total = reduce_statistics(stats)
        d = statistics_dict(total)
        # gradients: every rank holds the flat gradient vector rank + 1 (the trainer's layout: one tensor, one collective)
        flat = torch.full((22,), float(rank + 1))
        all_reduce_mean(flat)
        results[rank] = (total.tolist(), d['steps'], d['goals_top'], [flat[:15].unique().tolist(), flat[15:].unique().tolist()])
    finally:
        dist.destroy_process_group()


def test_statistics_and_gradient_allreduce_gloo():
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    results = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, world, port, results)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    for r in range(world):
        total, steps, goals_top, grads = results[r]
        assert total == [3 * k for k in range(32)]        # (1 + 2) * k
        assert steps == 0 and goals_top == 6
        assert grads == [[1.5], [1.5]]                    # mean of 1 and 2
