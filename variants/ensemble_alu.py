"""Config 5 rate as a function of OCEAN_ENSEMBLE_ALU_CTAS_PER_SM (grid cap of the ALU-bound noise / spectrum kernels) and batch.
    python variants/ensemble_alu.py"""
import hashlib, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import os, sys, time, hashlib, json
sys.path.insert(0, %r)
import numpy as np, torch, oceanpkg
pkg = oceanpkg.load(); cfg = pkg.scene.config(5)
n, batch = int(sys.argv[1]), int(sys.argv[2])
oceans, seeds = pkg.scene.ensemble_states(n)
e = pkg.OceanEnsemble(cfg["N"], cfg["cascades"], batch=batch)
e.run(oceans[:batch * 2], seeds[:batch * 2], time=cfg["time"]); torch.cuda.synchronize()
best = 1e9
for rep in range(3):
    t0 = time.perf_counter(); sums = e.run(oceans, seeds, time=cfg["time"]); best = min(best, time.perf_counter() - t0)
print(json.dumps({"us_per_instance": round(best / n * 1e6, 2), "instances_per_s": round(n / best), "digest": hashlib.sha256(np.ascontiguousarray(sums).tobytes()).hexdigest()[:12]}))
''' % ROOT
for batch in (16, 32):
    for k in ("0", "2", "3", "4", "6", "8", "12"):
        env = dict(os.environ, OCEAN_ENSEMBLE_ALU_CTAS_PER_SM=k)
        r = subprocess.run([sys.executable, "-c", CHILD, "1024", str(batch)], env=env, capture_output=True, text=True, timeout=300)
        print("batch %2d alu_ctas_per_sm %2s: %s" % (batch, k, r.stdout.strip() or r.stderr[-400:]), flush=True)
