"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: shard arithmetic, seeded-noise sharding that
reproduces a single-rank run bit for bit, ragged gather, cond/uncond pair groups."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from stablediffusion_b200.distributed import gather_batches, make_pair_group, shard_list, shard_range, sharded_noise


def test_shard_range_covers_everything_once():
    for n in (0, 1, 7, 8, 9, 64):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    assert shard_list(list("abcde"), 1, 2) == ["d", "e"]
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        B, seed = 5, 1234
        full = torch.randn((B, 4, 8, 8), generator=torch.Generator().manual_seed(seed))
        mine = sharded_noise((B, 4, 8, 8), seed, rank, world, "cpu")
        lo, hi = shard_range(B, rank, world)
        ok_noise = torch.equal(mine, full[lo:hi])
        # each rank "generates" from its shard; the ragged gather must equal the single-rank result
        gathered = gather_batches(mine * 2.0)
        ok_gather = torch.equal(gathered, full * 2.0)
        grp, pr = make_pair_group()
        eps_local = torch.full((3, 4), float(pr))
        eps_full = torch.empty(6, 4)
        dist.all_gather_into_tensor(eps_full, eps_local, group=grp)   # the per-step exchange of the pair split
        ok_pair = torch.equal(eps_full[:3], torch.zeros(3, 4)) and torch.equal(eps_full[3:], torch.ones(3, 4)) and pr == rank % 2
        q.put((rank, ok_noise, ok_gather, ok_pair))
    finally:
        dist.destroy_process_group()


def test_two_rank_data_parallel_and_pair_split_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, a, b, c in res:
        assert a and b and c, (rank, a, b, c)
