"""CPU: the multi-GPU host logic (environment sharding + rollout-end all-gather) with world_size-2 gloo."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from modularevogym_b200.sharding import gather_env_rows, owner_of, shard_range


def test_shard_range_partitions():
    for n in (0, 1, 7, 4096, 65536, 65537):
        for world in (1, 2, 3, 8):
            blocks = [shard_range(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            for (a, b), (c, d) in zip(blocks, blocks[1:]):
                assert b == c
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
            for e in (0, n // 2, n - 1):
                if 0 <= e < n:
                    r = owner_of(e, n, world)
                    assert blocks[r][0] <= e < blocks[r][1]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_total, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_range(n_total, rank, world)
    # per-env "reward" rows = (global env id, 10 * id): every rank must end up with all of them in order
    ids = torch.arange(lo, hi, dtype=torch.float64)
    local = torch.stack([ids, 10.0 * ids], dim=1)
    full = gather_env_rows(local, n_total)
    want = torch.stack([torch.arange(n_total, dtype=torch.float64), 10.0 * torch.arange(n_total, dtype=torch.float64)], dim=1)
    q.put((rank, bool(torch.equal(full, want))))
    dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [8, 9])
def test_gather_env_rows_gloo_world2(n_total):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_total, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(results) == [(0, True), (1, True)]
