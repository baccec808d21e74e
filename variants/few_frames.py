import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oceanpkg
pkg = oceanpkg.load(); cfg = pkg.scene.config(3); o = cfg["ocean"]
b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=cfg["N"], cascades=cfg["cascades"], seed=3).Awake()
for f in range(6): b.CalculateWavesTexturesAtTime(f / 60)
b.Sync(); print("ok")
