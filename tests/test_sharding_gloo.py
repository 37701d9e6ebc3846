"""N>1 host path on CPU: world_size-2 (and 3, ragged) gloo processes run the product's
sharding helpers; the gathered result must equal the single-process result."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_sampler(y, vals, mask, *, seed, sample_offset, x_T=None):
    """Stands in for the GPU sampler: a deterministic function of the GLOBAL sample index,
    the seed and the shard's own conditions (what Philox keyed by global index gives)."""
    n = y.shape[0]
    idx = torch.arange(sample_offset, sample_offset + n, dtype=torch.float32)
    base = idx[:, None] * 1000.0 + seed + y[:, None].float() * 10 + vals.sum(1, keepdim=True) + mask.sum(1, keepdim=True)
    out = base[:, :, None].repeat(1, 3, 4)
    if x_T is not None:
        out = out + x_T.reshape(n, -1)[:, :1, None]
    return out


def _worker(rank, world, port, n_total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from dmb200.sharding import sample_sharded, shard_range
        g = torch.Generator().manual_seed(5)
        y = torch.randint(1, 4, (n_total,), generator=g)
        vals = torch.rand((n_total, 12), generator=g)
        mask = (torch.rand((n_total, 12), generator=g) > 0.5).float()
        x_T = torch.randn((n_total, 4, 28, 28), generator=g)
        full = sample_sharded(_fake_sampler, y, vals, mask, seed=77, x_T=x_T)
        b, e = shard_range(n_total, rank, world)
        q.put((rank, full, (b, e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_total", [(2, 8), (2, 7), (3, 10)])
def test_sharded_equals_single_process(world, n_total):
    import sys
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    g = torch.Generator().manual_seed(5)
    y = torch.randint(1, 4, (n_total,), generator=g)
    vals = torch.rand((n_total, 12), generator=g)
    mask = (torch.rand((n_total, 12), generator=g) > 0.5).float()
    x_T = torch.randn((n_total, 4, 28, 28), generator=g)
    single = _fake_sampler(y, vals, mask, seed=77, sample_offset=0, x_T=x_T)
    covered = []
    for rank, full, (b, e) in sorted(got, key=lambda t: t[0]):
        assert torch.equal(full, single), f"rank {rank}"
        covered += list(range(b, e))
    assert covered == list(range(n_total))


def test_shard_range_properties():
    from dmb200.sharding import shard_range
    for n in (0, 1, 5, 4096, 4097):
        for w in (1, 2, 3, 8):
            rs = [shard_range(n, r, w) for r in range(w)]
            assert rs[0][0] == 0 and rs[-1][1] == n
            assert all(rs[i][1] == rs[i + 1][0] for i in range(w - 1))
            sizes = [e - b for b, e in rs]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def test_gather_without_process_group_is_identity():
    from dmb200.sharding import gather_shards
    x = torch.arange(12).reshape(4, 3)
    assert gather_shards(x, 4) is x
    with pytest.raises(ValueError):
        gather_shards(x, 5)
