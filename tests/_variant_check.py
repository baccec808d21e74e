"""Child process of test_gpu_variants.py: the frame kernels selected by the environment (OCEAN_ROWS_PAIR,
OCEAN_FUSED, OCEAN_NO_PDL ...) against the CPU oracle on config 3 at N = 256 and 512.  Prints one JSON line."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oceanpkg  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

pkg = oceanpkg.load()
cfg = pkg.scene.config(3)
o = cfg["ocean"]
out = {}
for N in (256, 512):
    noise = po.fill_noise(7, N)
    body = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5],
                         texturesSize=N, cascades=cfg["cascades"], noise=noise).Awake()
    ob = po.OracleB(N, o, cfg["cascades"], noise, threads=os.cpu_count() or 1)
    ob.init_spectrum()
    for f in range(3):
        body.CalculateWavesTexturesAtTime(0.3 + f / 60.0)
        ob.step(0.3 + f / 60.0)
    worst = 0.0
    for got, which in ((body.displacementsTextures, po.DISP), (body.derivativesTextures, po.DERIV), (body.turbulenceTextures, po.TURB)):
        want = ob.get(which).astype(np.float64)
        worst = max(worst, float(np.linalg.norm((got - want).ravel()) / np.linalg.norm(want.ravel())))
    out[str(N)] = worst
    out["describe"] = body.describe()
    body.OnDisable()
print(json.dumps(out))
