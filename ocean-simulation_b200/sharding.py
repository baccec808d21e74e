"""Multi-GPU sharding of independent ocean states: replicas of independent units, no
collective on the data path.  State s is owned by rank s % world (round robin); each rank runs its
shard through an ensemble runner on its own GPU; the only communication is the optional final gather
of the per-state results (torch.distributed all_gather over NCCL on GPUs, gloo in the CPU tests)."""
import numpy as np


def shard_indices(n_states, rank, world):
    """Indices of the states owned by `rank` (round robin, as BASELINE config 5 prescribes)."""
    return np.arange(rank, n_states, world, dtype=np.int64)


def run_sharded(run_states, oceans, seeds, rank=0, world=1, gather=None):
    """run_states(oceans_shard, seeds_shard) -> uint64 [n_local][3] checksums (OceanEnsemble.run on a GPU).
    Returns the instance-ordered [n][3] array on every rank when `gather` is a callable
    all_gather(list_of_arrays_per_rank <- local_array); without it only the local rows are filled."""
    oceans = np.asarray(oceans, np.float32).reshape(-1, 6)
    seeds = np.asarray(seeds, np.uint64)
    n = oceans.shape[0]
    mine = shard_indices(n, rank, world)
    local = np.asarray(run_states(oceans[mine], seeds[mine]), np.uint64).reshape(len(mine), 3)
    out = np.zeros((n, 3), np.uint64)
    if world == 1 or gather is None:
        out[mine] = local
        return out
    # ragged shards: pad to the longest so that a fixed-size all_gather can carry them
    longest = (n + world - 1) // world
    padded = np.zeros((longest, 3), np.uint64)
    padded[:len(mine)] = local
    for r, part in enumerate(gather(padded)):
        idx = shard_indices(n, r, world)
        out[idx] = np.asarray(part, np.uint64).reshape(longest, 3)[:len(idx)]
    return out


def torch_all_gather(device=None):
    """all_gather over the default torch.distributed group (NCCL on GPUs, gloo on CPU)."""
    import torch
    import torch.distributed as dist

    def gather(local):
        # uint64 has no collective support everywhere: ship the bit pattern as int64
        t = torch.from_numpy(local.view(np.int64).copy())
        if device is not None:
            t = t.to(device)
        parts = [torch.empty_like(t) for _ in range(dist.get_world_size())]
        dist.all_gather(parts, t)
        return [p.cpu().numpy().view(np.uint64) for p in parts]

    return gather
