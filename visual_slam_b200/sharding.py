"""Host-side sharding of independent units (frame pairs, sequences) over ranks and the gather of the
fixed-shape per-pair results.  Backend-agnostic: NCCL on the GPU box, gloo in CPU tests."""
import torch
import torch.distributed as dist


def shard_range(n_units, rank, world):
    """Contiguous block of units for `rank`: sizes differ by at most one."""
    base, rem = divmod(n_units, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_sequences(n_seq, rank, world):
    """Round-robin assignment of whole sequences (C5: 8 sequences over 1/2/4/8 ranks)."""
    return [s for s in range(n_seq) if s % world == rank]


def gather_results(local, n_units, world=None):
    """local: tensor [n_local, F] of this rank's results (contiguous shard from shard_range).  Returns
    the [n_units, F] tensor of all ranks in unit order on every rank (one all_gather of padded shards)."""
    if world is None:
        world = dist.get_world_size() if dist.is_initialized() else 1
    if world == 1:
        return local
    cap = (n_units + world - 1) // world
    buf = torch.zeros((cap,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    buf[: local.shape[0]] = local
    allb = torch.empty((world, cap) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    if local.is_cuda:
        dist.all_gather_into_tensor(allb, buf)   # NCCL over NVLink
    else:
        dist.all_gather(list(allb.unbind(0)), buf)  # gloo (CPU tests)
    parts = []
    for r in range(world):
        lo, hi = shard_range(n_units, r, world)
        parts.append(allb[r, : hi - lo])
    return torch.cat(parts, dim=0)
