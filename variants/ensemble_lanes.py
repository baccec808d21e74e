"""Config 5 with one lane and with two (OCEAN_ENSEMBLE_LANES): ocean-instances/s and identical checksums."""
import os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for lanes in ("1", "2"):
    env = dict(os.environ, OCEAN_ENSEMBLE_LANES=lanes)
    print("lanes =", lanes, flush=True)
    subprocess.run([sys.executable, os.path.join(root, "variants", "ensemble_rate.py")], env=env, check=True)
