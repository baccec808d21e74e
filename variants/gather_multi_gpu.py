"""The optional final gather of SURVEY 8(e) over NCCL: every rank steps its own 4 x 512^2 ocean and all ranks gather the
displacement arrays (16 MiB each) with all_gather_into_tensor on views of the library's own device memory.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 variants/gather_multi_gpu.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import oceanpkg
pkg = oceanpkg.load(); L = pkg._lib
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cfg = pkg.scene.config(3); o = cfg["ocean"]
b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=cfg["N"],
                  cascades=cfg["cascades"], seed=100 + rank, device=local).Awake()
for f in range(3): b.CalculateWavesTexturesAtTime(f / 60.0)
b.Sync()
view = pkg.sharding.device_view(b, L.DISPLACEMENT)
out = pkg.sharding.all_gather_arrays(view)
torch.cuda.synchronize()
assert np.array_equal(out[rank].cpu().numpy(), b.displacementsTextures)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
dist.barrier(); e0.record()
for _ in range(20): out = pkg.sharding.all_gather_arrays(view)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 20 * 1e3
sums = [float(out[r].abs().sum()) for r in range(world)]
if rank == 0:
    print(json.dumps({"n_gpus": world, "array_MiB": view.numel() * 4 / 2 ** 20, "all_gather_us": us,
                      "bus_GBps_per_rank": (world - 1) * view.numel() * 4 / us / 1e3, "distinct_oceans": len(set(sums))}))
b.OnDisable(); dist.barrier(); dist.destroy_process_group()
