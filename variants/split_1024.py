import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, oceanpkg
pkg = oceanpkg.load()
for n in (3, 4):
    cfg = pkg.scene.config(n); o = cfg["ocean"]
    b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=cfg["N"], cascades=cfg["cascades"], seed=n).Awake()
    st = torch.cuda.ExternalStream(b.stream)
    for f in range(10): b.CalculateWavesTexturesAtTime(f / 60)
    b.Sync()
    K = 100
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(K)]
    with torch.cuda.stream(st):
        for f in range(K):
            evs[f][0].record(); b.StepPartial(f / 60, 1); evs[f][1].record(); b.StepPartial(f / 60, 2); evs[f][2].record()
    torch.cuda.synchronize()
    print("config %d: rows %.1f us, cols %.1f us (each bracketed by events)" % (n, 1e3 * np.median([e[0].elapsed_time(e[1]) for e in evs]), 1e3 * np.median([e[1].elapsed_time(e[2]) for e in evs])))
    b.OnDisable()
