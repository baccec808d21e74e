"""Times rows / cols / frame for each lib_*.so variant on config 3 (4 x 512^2), rotating 4 instances."""
import glob, json, os, subprocess, sys
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CHILD = r'''
import os, sys, json
sys.path.insert(0, %r)
import numpy as np, torch, oceanpkg
pkg = oceanpkg.load()
cfg = pkg.scene.config(3); o = cfg["ocean"]; N = cfg["N"]
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
bodies = []
for i in range(4):
    b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=N, cascades=cfg["cascades"], seed=3 + i).Awake()
    b.SetStream(st.cuda_stream); bodies.append(b)
K = 200
for f in range(20): bodies[f %% 4].CalculateWavesTexturesAtTime(f / 60.0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for f in range(K): bodies[f %% 4].CalculateWavesTexturesAtTime(f / 60.0)
e1.record(st); torch.cuda.synchronize()
frame = e0.elapsed_time(e1) / K
evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(K)]
for f in range(K):
    b = bodies[f %% 4]
    evs[f][0].record(st); b.StepPartial(f / 60.0, 1); evs[f][1].record(st); b.StepPartial(f / 60.0, 2); evs[f][2].record(st)
torch.cuda.synchronize()
rows = float(np.median([e[0].elapsed_time(e[1]) for e in evs])); cols = float(np.median([e[1].elapsed_time(e[2]) for e in evs]))
chk = float(np.abs(bodies[0].displacementsTextures).sum())
print(json.dumps({"frame_us": 1e3 * frame, "rows_us": 1e3 * rows, "cols_us": 1e3 * cols, "chk": chk}))
''' % ROOT
for lib in sorted(glob.glob(os.path.join(HERE, "lib_*.so"))):
    env = dict(os.environ, OCEANFFT_LIB=lib)
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=300)
    print(os.path.basename(lib), r.stdout.strip() or r.stderr[-500:], flush=True)
