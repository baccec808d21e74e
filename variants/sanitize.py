"""Small tour of every kernel of liboceanfft.so for compute-sanitizer (memcheck / racecheck / synccheck):
   compute-sanitizer --tool memcheck python variants/sanitize.py
Sizes are small (the tools slow kernels down 10-100x); OCEAN_FUSED in the environment picks the frame path."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oceanpkg
pkg = oceanpkg.load()
L = pkg._lib
cas = pkg.scene.SCENE_CASCADES


def tour(N, C, frames=3):
    b = pkg.WaterBody(windSpeed=10.0, windDirection=(1.0, -1.0), gravity=9.81, fetch=1e5, depth=20.0,
                      texturesSize=N, cascades=cas[:C], seed=7).Awake()
    for f in range(frames):
        b.CalculateWavesTexturesAtTime(f / 60.0)
    b.Sync()
    for f in range(frames):
        b.Update((frames + f) / 60.0)
    b.Sync()
    d = b.displacementsTextures
    assert np.isfinite(d).all() and np.abs(d).max() > 0
    pts = (np.random.default_rng(1).uniform(-N / 2, N / 2, (257, 2))).astype(np.float32)
    for mode in (L.QUERY_NEAREST_REF, L.QUERY_BILINEAR, L.QUERY_BILINEAR_ALL_CASCADES, L.QUERY_DISPLACED):
        h = b.GetWaterHeights(pts, mode)
        assert np.isfinite(h).all()
    b.GenerateMips()
    out = b.SampleSurface(pts, lod=1.3)
    assert all(np.isfinite(o).all() for o in out)
    b.SetReadbackMode(L.READBACK_HEIGHT)
    b.Update(1.0)
    b.WaitReadback()
    boxes = np.zeros((64, 6), np.float32); boxes[:, :3] = pts[:64].mean() ; boxes[:, 0] = pts[:64, 0]; boxes[:, 2] = pts[:64, 1]; boxes[:, 3:] = 1.0
    ans = np.zeros((64, 4), np.float32)
    b.UpdateBuoyancy(1.1, boxes, ans)
    b.Sync()
    assert np.isfinite(ans).all()
    b.OnDisable()
    print("tour N=%d C=%d ok" % (N, C), flush=True)


SIZES = [int(a) for a in sys.argv[1:]] or [64, 256, 512]
if 64 in SIZES: tour(64, 2)        # generic row / column kernels
if 256 in SIZES: tour(256, 3)      # fused frame kernel (or rows<256> + cols_tma<256> with OCEAN_FUSED=0)
if 512 in SIZES: tour(512, 1)      # rows<512> + cols_tma<512> (or a fused kernel with OCEAN_FUSED=1 / 2)
oceans, seeds = pkg.scene.ensemble_states(3)
for fused in ("0", "1"):
    os.environ["OCEAN_ENSEMBLE_FUSED_INIT"] = fused
    e = pkg.OceanEnsemble(128, cas[:3], batch=2)
    s = e.run(oceans, seeds, time=1.0)
    assert s.shape == (3, 3)
    e.close()
    print("ensemble fused_init=%s ok" % fused, flush=True)
print("sanitize tour complete")
