"""us/frame and fraction of the measured HBM peak for BASELINE configs 1-4 (single ocean, back-to-back frames)."""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, oceanpkg
pkg = oceanpkg.load(); L = pkg._lib
peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
for n in (1, 2, 3, 4):
    cfg = pkg.scene.config(n); o = cfg["ocean"]; N = cfg["N"]; C = len(cfg["cascades"])
    b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=N, cascades=cfg["cascades"], seed=n).Awake()
    st = torch.cuda.ExternalStream(b.stream)
    K = 400
    for f in range(20): b.CalculateWavesTexturesAtTime(f / 60)
    b.Sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        e0.record()
        for f in range(K): b.CalculateWavesTexturesAtTime(f / 60)
        e1.record()
    e1.synchronize()
    us = e0.elapsed_time(e1) / K * 1e3
    bytes_ = 96 * C * N * N
    line = "config %d: %d x %d^2: %.2f us/frame, %.0f GB/s algorithmic = %.1f%% of %.0f GB/s" % (n, C, N, us, bytes_ / us / 1e3, 100 * bytes_ / us / 1e3 / peak, peak)
    if "query_points" in cfg:
        q = cfg["query_points"]
        pts = (torch.rand(q, 2, device="cuda") - 0.5) * N
        h = torch.empty(q, device="cuda")
        with torch.cuda.stream(st):
            e0.record()
            for f in range(100):
                b.CalculateWavesTexturesAtTime(f / 60)
                L.check(b._lib.ocean_query_heights(b._h, ctypes.c_void_p(pts.data_ptr()), q, L.QUERY_BILINEAR_ALL_CASCADES, ctypes.c_void_p(h.data_ptr()), 1, 1))
            e1.record()
        e1.synchronize()
        line += "; with %d bilinear all-cascade queries per frame: %.2f us/frame" % (q, e0.elapsed_time(e1) / 100 * 1e3)
    print(line, flush=True)
    print("   ", b.describe(), flush=True)
    b.OnDisable()
