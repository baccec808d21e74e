"""Frames of 4 oceans issued round robin on S streams (S = 1: the bench's arrangement; S = 2, 4: oceans i % S share a stream and a
workspace).  Device time from a joined start to a joined end, us per frame (aggregate).   python variants/two_streams.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, oceanpkg
pkg = oceanpkg.load(); cfg = pkg.scene.config(3); o = cfg["ocean"]; N = cfg["N"]
def run(S, K=800):
    streams = [torch.cuda.Stream() for _ in range(S)]
    bodies = []
    for i in range(4):
        b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=N, cascades=cfg["cascades"], seed=3 + i).Awake()
        b.SetStream(streams[i % S].cuda_stream); bodies.append(b)
    for f in range(40): bodies[f % 4].CalculateWavesTexturesAtTime(f / 60.0)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        for f in range(12): bodies[f % 4].CalculateWavesTexturesAtTime(f / 60.0)
        start = torch.cuda.Event(enable_timing=True); ends = [torch.cuda.Event(enable_timing=True) for _ in range(S)]
        # joined start: every stream waits for the others' priming frames, then the start event is recorded on stream 0
        joins = [torch.cuda.Event() for _ in range(S)]
        for s, j in zip(streams, joins): j.record(s)
        for s in streams:
            for j in joins: s.wait_event(j)
        start.record(streams[0])
        for s in streams[1:]: s.wait_event(start)
        for f in range(K): bodies[f % 4].CalculateWavesTexturesAtTime(f / 60.0)
        for s, e in zip(streams, ends): e.record(s)
        torch.cuda.synchronize()
        best = min(best, max(start.elapsed_time(e) for e in ends) / K)
    for b in bodies: b.OnDisable()
    return 1e3 * best
for S in (1, 2, 4, 1):
    print("%d stream(s): %.2f us per frame" % (S, run(S)), flush=True)
