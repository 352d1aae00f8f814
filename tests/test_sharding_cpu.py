"""World-size-2 (and 3, ragged) gloo tests of the episode sharding / all-gather host logic."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from robomanipbaselines_b200.common.sharding import (AsyncEpisodeGather, episode_block, episode_seed, gather_episodes,
                                                     gather_results)


def test_episode_blocks_partition():
    for total in (0, 1, 7, 256, 2048, 4096):
        for world in (1, 2, 3, 4, 8):
            blocks = [episode_block(total, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == total
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
    assert episode_seed(7, 12) == 19
    with pytest.raises(ValueError):
        episode_block(8, 2, 2)


def _worker(rank, world, port, total, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    a, b = episode_block(total, rank, world)
    idx = torch.arange(a, b, dtype=torch.float64)
    local = torch.stack([idx * 10 + d for d in range(7)], dim=1)           # action[e, d] = 10 e + d
    full = gather_episodes(local, total)
    s, r, t = gather_results((idx % 2 == 0).to(torch.uint8), idx.float() * 0.5, idx.float() + 1, total)
    ok = (full.shape == (total, 7)
          and torch.equal(full[:, 0], torch.arange(total, dtype=torch.float64) * 10)
          and torch.equal(full[:, 6], torch.arange(total, dtype=torch.float64) * 10 + 6)
          and torch.equal(s, (torch.arange(total) % 2 == 0).to(torch.uint8))
          and torch.equal(r, torch.arange(total).float() * 0.5) and torch.equal(t, torch.arange(total).float() + 1))
    if total % world == 0:      # the pipelined gather: two submits outstanding, results come back in submit order
        ag = AsyncEpisodeGather(total, local.shape, local.dtype, "cpu", depth=2)
        ag.submit(local)
        ag.submit(local + 1000)
        r0 = ag.result().clone()
        ag.submit(local + 2000)
        r1, r2 = ag.result().clone(), ag.drain().clone()
        ok = ok and torch.equal(r0, full) and torch.equal(r1, full + 1000) and torch.equal(r2, full + 2000)
        try:
            ag.result()
            ok = False
        except RuntimeError:
            pass
    q.put((rank, bool(ok)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,total", [(2, 8), (2, 7), (3, 10)])
def test_gather_in_global_episode_order(world, total):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(r, True) for r in range(world)]
