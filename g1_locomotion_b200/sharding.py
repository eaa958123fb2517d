"""Multi-GPU plumbing for the batched MPC path: one process per GPU, contiguous batch shards, one gather.

The problems are independent (the reference constructs a fresh solver on every call, g1_mpc/src/mpc.cpp:380),
so there is no data-path collective: rank r owns problems [lo_r, hi_r) and the only inter-GPU traffic is the
collection of results.  `torch.distributed` is plumbing here: NCCL over NVLink on the GPU box,
gloo in the CPU tests (tests/test_sharding_cpu.py).
"""
import torch
import torch.distributed as dist


def shard_range(total, rank, world):
    """Contiguous shard [lo, hi) of `total` problems owned by `rank`; sizes differ by at most one."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, rem = divmod(int(total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_sizes(total, world):
    return [shard_range(total, r, world)[1] - shard_range(total, r, world)[0] for r in range(world)]


def wbid_payload(u_opt, x_opt):
    """What the whole-body layer consumes from one MPC solve: first-step GRFs u_opt[:,0,:] and the next
    state x_opt[:,1,:] (g1_mujoco_sim/src/ros_run_simulation.py:409, wbid.py:232-297) -> [B, nu + 13]."""
    return torch.cat([u_opt[:, 0, :], x_opt[:, 1, :]], dim=1).contiguous()


def gather_results(local, total, group=None):
    """All-gather per-rank result rows `local` [b_r, ...] (b_r = shard size of this rank) into [total, ...] in
    global problem order on every rank.  Ragged shards are padded to the largest shard for the collective."""
    if not (dist.is_available() and dist.is_initialized()):
        if local.shape[0] != total:
            raise ValueError("single process: local batch must be the whole batch")
        return local
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    sizes = shard_sizes(total, world)
    if local.shape[0] != sizes[rank]:
        raise ValueError(f"rank {rank}: expected {sizes[rank]} rows, got {local.shape[0]}")
    mx = max(sizes)
    pad = local
    if local.shape[0] < mx:
        pad = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        pad[: local.shape[0]] = local
    pad = pad.contiguous()
    out = torch.empty((world * mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    if dist.get_backend(group) == "nccl":
        dist.all_gather_into_tensor(out, pad, group=group)
    else:
        parts = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(parts, pad, group=group)
        out = torch.cat(parts, 0)
    if all(s == mx for s in sizes):
        return out
    return torch.cat([out[r * mx: r * mx + sizes[r]] for r in range(world)], 0)


def solve_sharded(solve_fn, make_inputs, total, group=None, payload=wbid_payload):
    """Rank-local driver: build this rank's shard with `make_inputs(lo, hi)` (a tuple of tensors), solve it with
    `solve_fn(*inputs) -> (u_opt, x_opt, info)`, gather `payload(u_opt, x_opt)` from every rank.

    Returns (gathered [total, ...], u_opt_local, x_opt_local, info_local)."""
    if dist.is_available() and dist.is_initialized():
        world, rank = dist.get_world_size(group), dist.get_rank(group)
    else:
        world, rank = 1, 0
    lo, hi = shard_range(total, rank, world)
    ins = make_inputs(lo, hi)
    u, x, info = solve_fn(*ins)
    return gather_results(payload(u, x), total, group), u, x, info
