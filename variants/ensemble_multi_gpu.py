"""BASELINE config 5 across GPUs: 4096 independent ocean states (3 cascades x 512^2; init + one frame at t = 1 s + checksums),
sharded round robin over the ranks (ocean-simulation_b200/sharding.py), no collective on the data path; the per-state
checksums are gathered at the end (NCCL all_gather) and must not depend on the number of GPUs.

    python variants/ensemble_multi_gpu.py                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 variants/ensemble_multi_gpu.py
"""
import hashlib, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oceanpkg
pkg = oceanpkg.load()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cfg = pkg.scene.config(5)
n = int(os.environ.get("STATES", 4096))
batch = int(os.environ.get("BATCH", 32))
oceans, seeds = pkg.scene.ensemble_states(n)
ens = pkg.OceanEnsemble(cfg["N"], cfg["cascades"], batch=batch, device=local)
mine = pkg.sharding.shard_indices(n, rank, world)
ens.run(oceans[mine][:2 * batch], seeds[mine][:2 * batch], time=cfg["time"])  # warm-up
def barrier():
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
times = []
for rep in range(3):
    barrier()
    t0 = time.perf_counter()
    sums = pkg.sharding.run_sharded(lambda o, s: ens.run(o, s, time=cfg["time"]), oceans, seeds, rank, world, gather=None)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1: dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    times.append(float(dt.item()))
gathered = pkg.sharding.run_sharded(lambda o, s: sums[mine], oceans, seeds, rank, world,
                                    gather=pkg.sharding.torch_all_gather(torch.device("cuda", local)) if world > 1 else None)
if rank == 0:
    best = min(times)
    print(json.dumps({"config": "5: %d states x 3 cascades x 512^2, batch %d" % (n, batch), "n_gpus": world,
                      "ocean_instances_per_s": n / best, "cascade_frames_per_s": 3 * n / best, "seconds": best,
                      "us_per_instance_per_gpu": best / (n / world) * 1e6,
                      "checksum_of_checksums": hashlib.sha256(np.ascontiguousarray(gathered).tobytes()).hexdigest()[:16]}), flush=True)
ens.close()
if world > 1: dist.destroy_process_group()
