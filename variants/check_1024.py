"""Config 4 (4 x 1024^2) with whatever frame path the environment selects, against Oracle B: prints rel-L2 of the three arrays."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, oceanpkg
from oracle import pyoracle as po
pkg = oceanpkg.load(); cfg = pkg.scene.config(4); o = cfg["ocean"]; N = cfg["N"]
noise = po.fill_noise(cfg["seed"], N)
b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=N, cascades=cfg["cascades"], noise=noise).Awake()
ob = po.OracleB(N, o, cfg["cascades"], noise, threads=0); ob.init_spectrum()
for f in range(3):
    ob.step(f / 60.0); b.CalculateWavesTexturesAtTime(f / 60.0)
rel = lambda a, w: float(np.linalg.norm((a - w).ravel()) / np.linalg.norm(w.ravel()))
print("rel-L2 disp %.3g deriv %.3g turb %.3g | %s" % (rel(b.displacementsTextures, ob.get(po.DISP)), rel(b.derivativesTextures, ob.get(po.DERIV)),
      rel(b.turbulenceTextures, ob.get(po.TURB)), b.describe()[:60]))
