"""Multi-GPU path on CPU: world_size-2 gloo run of the state sharding + final gather.  The
per-rank runner is the CPU oracle here (test infrastructure standing in for OceanEnsemble.run on a GPU);
what is under test is the host-side partitioning, ragged-shard padding and instance-ordered gather."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import numpy as np
sys.path.insert(0, %r)
import torch.distributed as dist
import oceanpkg
from oracle import pyoracle as po
pkg = oceanpkg.load()
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%%s" %% sys.argv[1], rank=int(sys.argv[2]), world_size=int(sys.argv[3]))
rank, world = dist.get_rank(), dist.get_world_size()
N, casc = 32, pkg.scene.SCENE_CASCADES[:3]
oceans, seeds = pkg.scene.ensemble_states(7)

def run_states(oc, sd):
    out = []
    for o, s in zip(oc, sd):
        b = po.OracleB(N, tuple(float(v) for v in o), casc, po.fill_noise(int(s), N))
        b.init_spectrum(); b.step(1.0)
        out.append([int(b.get(w).view(np.uint32).astype(np.uint64).sum()) for w in (po.DISP, po.DERIV, po.TURB)])
        b.close()
    return np.array(out, np.uint64).reshape(-1, 3)

got = pkg.sharding.run_sharded(run_states, oceans, seeds, rank, world, gather=pkg.sharding.torch_all_gather())
np.save(sys.argv[4] + ".rank%%d.npy" %% rank, got)
# the optional final gather of an output ARRAY (SURVEY 8e): rank r contributes the displacement of state r
import torch
b = po.OracleB(N, tuple(float(v) for v in oceans[rank]), casc, po.fill_noise(int(seeds[rank]), N))
b.init_spectrum(); b.step(1.0)
arrays = pkg.sharding.all_gather_arrays(torch.from_numpy(b.get(po.DISP)))
np.save(sys.argv[4] + ".arrays%%d.npy" %% rank, arrays.numpy())
mine = pkg.sharding.shard_indices(7, rank, world)
print("rank", rank, "owns", mine.tolist())
dist.barrier(); dist.destroy_process_group()
''' % ROOT


def test_round_robin_shards_cover_everything_once():
    import oceanpkg
    sh = oceanpkg.load().sharding
    for n, world in ((4096, 8), (7, 2), (3, 4), (0, 2)):
        parts = [sh.shard_indices(n, r, world) for r in range(world)]
        assert sorted(np.concatenate(parts).tolist()) == list(range(n))
        assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def test_world_size_2_gloo_gather_equals_serial(tmp_path, po, pkg):
    port = str(29500 + os.getpid() % 2000)
    prefix = str(tmp_path / "out")
    procs = [subprocess.Popen([sys.executable, "-c", WORKER, port, str(r), "2", prefix], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=300)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    a, b = np.load(prefix + ".rank0.npy"), np.load(prefix + ".rank1.npy")
    assert np.array_equal(a, b) and a.shape == (7, 3)
    # serial run (world = 1) through the same helper
    N, casc = 32, pkg.scene.SCENE_CASCADES[:3]
    oceans, seeds = pkg.scene.ensemble_states(7)

    def run_states(oc, sd):
        out = []
        for o, s in zip(oc, sd):
            ob = po.OracleB(N, tuple(float(v) for v in o), casc, po.fill_noise(int(s), N))
            ob.init_spectrum()
            ob.step(1.0)
            out.append([int(ob.get(w).view(np.uint32).astype(np.uint64).sum()) for w in (po.DISP, po.DERIV, po.TURB)])
        return np.array(out, np.uint64).reshape(-1, 3)

    serial = pkg.sharding.run_sharded(run_states, oceans, seeds)
    assert np.array_equal(serial, a)
    assert len(set(a[:, 0].tolist())) == 7  # seven different oceans
    # gathered output arrays: [world][C][N][N][4], identical on both ranks, rank r's ocean at index r
    g0, g1 = np.load(prefix + ".arrays0.npy"), np.load(prefix + ".arrays1.npy")
    assert g0.shape == (2, 3, N, N, 4) and np.array_equal(g0, g1)
    for r in range(2):
        ob = po.OracleB(N, tuple(float(v) for v in oceans[r]), casc, po.fill_noise(int(seeds[r]), N))
        ob.init_spectrum()
        ob.step(1.0)
        assert np.array_equal(g0[r], ob.get(po.DISP))
