"""world_size-2 gloo test of the multi-GPU host logic (sharding, cost-term all-gather, global selection,
winner broadcast) on CPU tensors; the oracle's selection stands in for the CUDA kernel."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _oracle_select(terms):
    """cost.py:56-88 over gathered terms (free, len, smooth) -> vector with [0] = best global index or -1."""
    free = terms[:, 0] != 0
    out = torch.zeros(8)
    if not bool(free.any()):
        out[0] = -1
        return out
    idx = torch.nonzero(free).flatten()
    l, s = terms[idx, 1], terms[idx, 2]
    ln = (l - l.min()) / (l.max() - l.min())
    sn = (s - s.min()) / (s.max() - s.min())
    c = 0.1 * sn + 0.9 * ln
    out[0] = float(idx[torch.argmin(c)])
    return out


def _worker(rank, world, port, total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ramp_b200 import sharding
        g = torch.Generator().manual_seed(7)
        x_all = torch.randn(total, 8, 4, generator=g)
        terms_all = torch.rand(total, 3, generator=g)
        terms_all[:, 0] = (terms_all[:, 0] > 0.4).float()
        lo, hi = sharding.shard_bounds(total, world, rank)
        traj, best, _ = sharding.select_global(x_all[lo:hi], terms_all[lo:hi], _oracle_select, total=total)
        gathered = sharding.gather_terms(terms_all[lo:hi], total=total)
        ok = torch.equal(gathered, terms_all)
        expect = int(_oracle_select(terms_all)[0])
        ok = ok and best == expect and torch.equal(traj, x_all[expect])
        # no survivor anywhere -> None on every rank, no broadcast
        none_terms = terms_all.clone()
        none_terms[:, 0] = 0
        t2, b2, _ = sharding.select_global(x_all[lo:hi], none_terms[lo:hi], _oracle_select, total=total)
        ok = ok and t2 is None and b2 == -1
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total", [16, 13])
def test_select_global_two_ranks(total):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, total, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]


def test_shard_bounds_cover_batch():
    from ramp_b200 import sharding
    for total in (0, 1, 7, 8192, 65536):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_bounds(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1
            for i in range(0, total, max(1, total // 5)):
                r, loc = sharding.owner_of(i, total, world)
                assert spans[r][0] + loc == i
