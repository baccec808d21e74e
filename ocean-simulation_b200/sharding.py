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


class _DeviceArray:
    """A library-owned device array (ocean_get_output) as a __cuda_array_interface__ object: torch.as_tensor(...) then
    views the library's memory without a copy, so NCCL can send it straight over NVLink."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(int(s) for s in shape), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 3, "strides": None}


def device_view(body, which):
    """torch view (no copy) of one output array of a WaterBody: [C][N][N][4] float32 on the body's device.
    The caller orders its use after the frame (body.Sync(), or work on the body's stream)."""
    import torch
    ptr, nbytes = body.device_pointer(which)
    C, N = len(body.cascades), body.texturesSize
    assert nbytes == C * N * N * 16
    return torch.as_tensor(_DeviceArray(ptr, (C, N, N, 4)), device=torch.device("cuda", body.device))


def all_gather_arrays(local, world=None):
    """The optional final gather of SURVEY 8(e): every rank contributes one tensor of the same shape (an output array of
    the ocean it owns: device_view(...) on GPUs -> NCCL over NVLink / NVSwitch; a CPU tensor under gloo) and receives
    [world][...] with rank r's array at index r.  Off the hot path: 16 MiB per 4 x 512^2 array per rank."""
    import torch
    import torch.distributed as dist
    world = world or dist.get_world_size()
    out = torch.empty((world,) + tuple(local.shape), dtype=local.dtype, device=local.device)
    dist.all_gather([out[r] for r in range(world)], local.contiguous())  # (views of `out`: no extra copy; works under gloo too)
    return out
