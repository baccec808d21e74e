"""config 3 and config 4 frame times for each lib_*.so variant."""
import glob, os, subprocess, sys
HERE = os.path.dirname(os.path.abspath(__file__))
for lib in sorted(glob.glob(os.path.join(HERE, "lib_*.so"))):
    r = subprocess.run([sys.executable, os.path.join(HERE, "config_rates.py")], env=dict(os.environ, OCEANFFT_LIB=lib), capture_output=True, text=True, timeout=300)
    print(os.path.basename(lib)); print("\n".join(l[:120] for l in r.stdout.splitlines() if l.startswith("config")) or r.stderr[-400:], flush=True)
