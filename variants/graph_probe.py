"""Would a CUDA graph of K frames beat K stream launches?  Captures K ocean_step calls (rotating 4 oceans on one stream) and
replays the graph ONCE (the hand-off counter targets are baked into the kernel parameters, so only the first replay is valid)."""
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
K = 400
for f in range(40): bodies[f % 4].CalculateWavesTexturesAtTime(f / 60.0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for f in range(K): bodies[f % 4].CalculateWavesTexturesAtTime(f / 60.0)
e1.record(st); torch.cuda.synchronize()
print("stream launches: %.2f us/frame" % (e0.elapsed_time(e1) / K * 1e3))
ref = bodies[(K - 1) % 4].displacementsTextures.copy()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g, stream=st, capture_error_mode="relaxed"):
    for f in range(K): bodies[f % 4].CalculateWavesTexturesAtTime(f / 60.0)
torch.cuda.synchronize()
e0.record(st); g.replay(); e1.record(st); torch.cuda.synchronize()
print("graph replay:    %.2f us/frame" % (e0.elapsed_time(e1) / K * 1e3))
got = bodies[(K - 1) % 4].displacementsTextures
print("same displacement as the stream run:", bool(np.array_equal(ref, got)))
