"""BASELINE config 5 (ensemble of independent oceans, 3 cascades x 512^2): ocean-instances/s for a few pool sizes."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oceanpkg
pkg = oceanpkg.load()
cfg = pkg.scene.config(5)
n = int(os.environ.get("STATES", 512))
oceans, seeds = pkg.scene.ensemble_states(n)
for batch in (4, 8, 16, 32, 64):
    e = pkg.OceanEnsemble(cfg["N"], cfg["cascades"], batch=batch)
    e.run(oceans[:batch * 2], seeds[:batch * 2], time=cfg["time"])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sums = e.run(oceans, seeds, time=cfg["time"])
    dt = time.perf_counter() - t0
    print("batch %3d: %d states in %.1f ms = %.0f ocean-instances/s (%.1f us per instance), checksum[0]=%s" % (batch, n, dt * 1e3, n / dt, dt / n * 1e6, sums[0].tolist()), flush=True)
    e.close()
