"""What would a DEFERRED column pass buy?  Stream order today: R(f) C(f) R(f+1) C(f+1): C(f) waits for the slowest row CTA
of its own frame.  Deferred: R(f+1) C(f) R(f+2) C(f+1): the column pass of a frame is enqueued after the NEXT frame's row
pass, so its rows are long finished.  Emulated here with two oceans on one stream and the x256 build (column pass does not
wait on the row counters; results are wrong, only the timing matters):  loop { X.rows ; Y.cols }  with X, Y alternating.
    python variants/deferred_probe.py variants/lib_x256_cols_no_await.so"""
import os, sys
HERE = os.path.dirname(os.path.abspath(__file__)); ROOT = os.path.dirname(HERE); sys.path.insert(0, ROOT)
if len(sys.argv) > 1: os.environ["OCEANFFT_LIB"] = os.path.abspath(sys.argv[1])
os.environ["OCEAN_FUSED"] = "0"
import torch, oceanpkg
pkg = oceanpkg.load()
cfg = pkg.scene.config(3); o = cfg["ocean"]; N = cfg["N"]
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
bodies = []
for i in range(4):
    b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=N, cascades=cfg["cascades"], seed=3 + i).Awake()
    b.SetStream(st.cuda_stream); bodies.append(b)
def run(K, deferred):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if deferred:
        bodies[0].StepPartial(0.0, 1)
        for f in range(1, 13):
            bodies[f % 4].StepPartial(f / 60.0, 1); bodies[(f - 1) % 4].StepPartial(f / 60.0, 2)
        e0.record(st)
        for f in range(13, 13 + K):
            bodies[f % 4].StepPartial(f / 60.0, 1); bodies[(f - 1) % 4].StepPartial(f / 60.0, 2)
        e1.record(st)
        bodies[(12 + K) % 4].StepPartial(0.0, 2)
    else:
        for f in range(12): bodies[f % 4].CalculateWavesTexturesAtTime(f / 60.0)
        e0.record(st)
        for f in range(K): bodies[f % 4].CalculateWavesTexturesAtTime(f / 60.0)
        e1.record(st)
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / K
for rep in range(3):
    print("in order  R(f) C(f):        %.2f us/frame" % run(400, False))
    print("deferred  R(f+1) C(f):      %.2f us/frame" % run(400, True), flush=True)
