"""Times the frame (config 3, 4 x 512^2, 4 rotating oceans; or --n256: config 2, 3 x 256^2) for each (library, environment)
combination given as  name:LIB:VAR=VAL,VAR=VAL ...  and prints a digest of the three output arrays after a fixed frame
sequence, so that variants which must be bit-identical can be checked in the same run.
    python variants/time_modes.py two:: fused1::OCEAN_FUSED=1 fused2::OCEAN_FUSED=2 [--n256] [--frames K]"""
import hashlib, json, os, subprocess, sys
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CHILD = r'''
import os, sys, json, hashlib
sys.path.insert(0, %r)
import numpy as np, torch, oceanpkg
pkg = oceanpkg.load()
cfgno, K = int(sys.argv[1]), int(sys.argv[2])
cfg = pkg.scene.config(cfgno); o = cfg["ocean"]; N = cfg["N"]
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
bodies = []
for i in range(4):
    b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=N, cascades=cfg["cascades"], seed=3 + i).Awake()
    b.SetStream(st.cuda_stream); bodies.append(b)
for f in range(24): bodies[f %% 4].CalculateWavesTexturesAtTime(f / 60.0)
torch.cuda.synchronize()
h = hashlib.sha256()
for b in bodies[:2]:
    for arr in (b.displacementsTextures, b.derivativesTextures, b.turbulenceTextures): h.update(arr.tobytes())
best = 1e9
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for f in range(12): bodies[f %% 4].CalculateWavesTexturesAtTime(f / 60.0)
    e0.record(st)
    for f in range(K): bodies[f %% 4].CalculateWavesTexturesAtTime(f / 60.0)
    e1.record(st); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / K)
print(json.dumps({"frame_us": round(1e3 * best, 3), "digest": h.hexdigest()[:12], "finite": bool(np.isfinite(bodies[0].turbulenceTextures).all())}))
''' % ROOT
args = [a for a in sys.argv[1:] if not a.startswith("--")]
for flag in ("--frames", "--cfg"):
    if flag in sys.argv: args.remove(sys.argv[sys.argv.index(flag) + 1])
cfgno = 2 if "--n256" in sys.argv else (int(sys.argv[sys.argv.index("--cfg") + 1]) if "--cfg" in sys.argv else 3)
K = int(sys.argv[sys.argv.index("--frames") + 1]) if "--frames" in sys.argv else 400
for spec in args:
    name, lib, envs = (spec.split(":") + ["", ""])[:3]
    env = dict(os.environ)
    if lib: env["OCEANFFT_LIB"] = lib if os.path.isabs(lib) else os.path.join(ROOT, lib)
    for kv in filter(None, envs.split(",")):
        k, v = kv.split("="); env[k] = v
    try:
        r = subprocess.run([sys.executable, "-c", CHILD, str(cfgno), str(K)], env=env, capture_output=True, text=True, timeout=240)
        print("%-28s %s" % (name, r.stdout.strip() or ("rc=%d " % r.returncode) + r.stderr[-600:]), flush=True)
    except subprocess.TimeoutExpired:
        print("%-28s TIMEOUT" % name, flush=True)
