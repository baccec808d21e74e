"""Host cost of issuing a frame (ocean_step through ctypes): issue time of K frames (K small enough that the launch
queue never fills) against the time until the GPU has finished them."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, oceanpkg
pkg = oceanpkg.load()
cfg = pkg.scene.config(3); o = cfg["ocean"]; N = cfg["N"]
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
bodies = []
for i in range(4):
    b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=N, cascades=cfg["cascades"], seed=3 + i).Awake()
    b.SetStream(st.cuda_stream); bodies.append(b)
for f in range(20): bodies[f % 4].CalculateWavesTexturesAtTime(f / 60.0)
torch.cuda.synchronize()
for K in (50, 100, 200, 400):
    res = []
    for rep in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for f in range(K): bodies[f % 4].CalculateWavesTexturesAtTime(f / 60.0)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        res.append(((t1 - t0) / K * 1e6, (t2 - t0) / K * 1e6))
    res.sort()
    print("K=%d: host issue %.2f us/frame, until GPU done %.2f us/frame (median of 5)" % (K, res[2][0], res[2][1]), flush=True)
# raw ctypes loop without the Python wrapper
lib = pkg._lib.load()
hs = [b._h for b in bodies]
import ctypes
K = 200
torch.cuda.synchronize()
t0 = time.perf_counter()
for f in range(K): lib.ocean_step(hs[f % 4], ctypes.c_float(f / 60.0))
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("raw ctypes K=200: host issue %.2f us/frame, until GPU done %.2f" % ((t1 - t0) / K * 1e6, (t2 - t0) / K * 1e6))
