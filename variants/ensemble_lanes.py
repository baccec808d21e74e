"""Config 5 in its three forms: one lane / two lanes (OCEAN_ENSEMBLE_LANES) with the separate spectrum kernel,
and with the spectrum fused into the row pass (OCEAN_ENSEMBLE_FUSED_INIT=1, opt-in): ocean-instances/s."""
import os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for lanes, fused in (("1", "0"), ("2", "0"), ("1", "1"), ("2", "1")):
    env = dict(os.environ, OCEAN_ENSEMBLE_LANES=lanes, OCEAN_ENSEMBLE_FUSED_INIT=fused)
    print("lanes = %s, fused init = %s" % (lanes, fused), flush=True)
    subprocess.run([sys.executable, os.path.join(root, "variants", "ensemble_rate.py")], env=env, check=True)
