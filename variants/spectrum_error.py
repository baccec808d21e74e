"""rel-L2 of h0 / waveData against Oracle B for the five BASELINE parameter sets (tolerance 1e-6), and the config-5 rate."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, oceanpkg
from oracle import pyoracle as po
pkg = oceanpkg.load()
rel = lambda a, w: float(np.linalg.norm((a.astype(np.float64) - w).ravel()) / np.linalg.norm(w.astype(np.float64).ravel()))
worst = 0.0
cases = [(n, pkg.scene.config(n)["ocean"], pkg.scene.config(n)["cascades"]) for n in (1, 2, 3, 4)]
oceans, seeds = pkg.scene.ensemble_states(6)
cases += [("5/%d" % i, tuple(float(v) for v in oceans[i]), pkg.scene.config(5)["cascades"]) for i in range(6)]
for name, o, casc in cases:
    N = 256
    noise = po.fill_noise(11, N)
    b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=N, cascades=casc, noise=noise).Awake()
    ob = po.OracleB(N, o, casc, noise, threads=0); ob.init_spectrum()
    e = max(rel(b.initialSpectrumTextures[c], ob.get(po.H0)[c]) for c in range(len(casc)))
    worst = max(worst, e)
    print("config %s: h0 rel-L2 (worst cascade) %.3g" % (name, e), flush=True)
    b.OnDisable(); ob.close()
print("worst %.3g (tolerance 1e-6)" % worst)
cfg = pkg.scene.config(5); n = 1024
oceans, seeds = pkg.scene.ensemble_states(n)
e = pkg.OceanEnsemble(cfg["N"], cfg["cascades"], batch=16)
e.run(oceans[:32], seeds[:32], time=cfg["time"]); torch.cuda.synchronize()
best = min((lambda t0: (e.run(oceans, seeds, time=cfg["time"]), time.perf_counter() - t0)[1])(time.perf_counter()) for _ in range(3))
print("config 5: %.2f us per instance = %.0f ocean-instances/s" % (best / n * 1e6, n / best))
