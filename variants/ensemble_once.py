"""One small ensemble run (for ncu): 32 states in waves of 16."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oceanpkg
pkg = oceanpkg.load(); cfg = pkg.scene.config(5)
oceans, seeds = pkg.scene.ensemble_states(32)
e = pkg.OceanEnsemble(cfg["N"], cfg["cascades"], batch=16)
print(e.run(oceans, seeds, time=cfg["time"])[0].tolist()); e.close()
