"""Batch sharding across the GPUs of one box.

Trajectories are independent through the whole reverse-diffusion loop, so the batch is split
contiguously over ranks with no data-path collective.  The one exchange is best-of-N selection
(cost.py:56-88 normalises by the GLOBAL min/max over collision-free survivors): an all-gather of
three fp32 per trajectory (free, length, smoothness), the same selection kernel on every rank, and
a broadcast of the winning [H, D] trajectory from its owner.  Host logic only — the arithmetic is
`ops.cost_terms` / `ops.select_best` (CUDA); tests drive this module over gloo with a checker in
place of the kernel."""
import torch
import torch.distributed as dist


def shard_bounds(total, world, rank):
    """Contiguous shard [lo, hi) of `total` trajectories; the first total % world ranks get one more."""
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_sizes(total, world):
    return [shard_bounds(total, world, r)[1] - shard_bounds(total, world, r)[0] for r in range(world)]


def gather_terms(terms, total=None, group=None):
    """terms [B_local, 3] -> [B_total, 3] in global trajectory order on every rank."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return terms
    world = dist.get_world_size(group)
    if total is None or total % world == 0:
        out = torch.empty((world * terms.shape[0],) + tuple(terms.shape[1:]), device=terms.device, dtype=terms.dtype)
        dist.all_gather_into_tensor(out, terms.contiguous(), group=group)
        return out
    sizes = shard_sizes(total, world)
    pad = max(sizes)
    buf = torch.zeros((pad,) + tuple(terms.shape[1:]), device=terms.device, dtype=terms.dtype)
    buf[:terms.shape[0]] = terms
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf, group=group)
    return torch.cat([p[:n] for p, n in zip(parts, sizes)], dim=0)


def owner_of(global_index, total, world):
    """(rank, local index) of global trajectory `global_index`."""
    for r in range(world):
        lo, hi = shard_bounds(total, world, r)
        if lo <= global_index < hi:
            return r, global_index - lo
    raise IndexError(global_index)


def select_global(x_local, terms_local, select_fn, total=None, group=None):
    """Best-of-N over all ranks.  select_fn(terms_all) -> tensor whose [0] is the best global index or -1
    (ops.select_best).  Returns (best trajectory [H, D] on every rank or None, global index, selection vector)."""
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank(group) if world > 1 else 0
    total = total if total is not None else world * x_local.shape[0]
    terms_all = gather_terms(terms_local, total, group)
    sel = select_fn(terms_all)
    best = int(sel[0])                      # host read: the replan loop branches on "no survivor" (cost.py:59-61)
    if best < 0:
        return None, -1, sel
    owner, local = owner_of(best, total, world)
    traj = x_local[local].clone() if rank == owner else torch.empty_like(x_local[0])
    if world > 1:
        dist.broadcast(traj, src=dist.get_global_rank(group, owner) if group is not None else owner, group=group)
    return traj, best, sel
