"""Builds tuning variants of liboceanfft.so (experiments only; not part of the product build).
Only ocean_frame.cu is recompiled per variant; the other objects come from the product build."""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oceanpkg
HERE = os.path.dirname(os.path.abspath(__file__))
VARIANTS = {
    "base": [],
    "x04_no_turb_load": ["-DOCEAN_EXP=4"],
    "x16_2cta": ["-DOCEAN_EXP=16", "-DOCEAN_COLS_TMA_CTAS_PER_SM=2"],
    "x16_3cta": ["-DOCEAN_EXP=16", "-DOCEAN_COLS_TMA_CTAS_PER_SM=3"],
}
if __name__ == "__main__":
    bm = oceanpkg.build_module()
    bm.build()
    names = sys.argv[1:] or list(VARIANTS)
    for n in names:
        obj = os.path.join(HERE, "frame_%s.o" % n)
        cmd = ["nvcc", "-ccbin", "g++"] + bm.ARCH + [f for f in bm.COMMON if f not in ("-Xptxas", "-v")] + VARIANTS[n] + ["-c", os.path.join(bm.CSRC, "ocean_frame.cu"), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode: print(n, "FAILED", r.stderr[-2000:]); continue
        info = re.findall(r"ocean_(rows|rows_seq|cols|cols_tma)_kernelILi512E.*?\n.*?(\d+) bytes spill stores.*?\n.*?Used (\d+) registers", r.stderr)
        others = [os.path.join(bm.OBJ, o) for o in ("ocean_init.o", "ocean_host.o", "ocean_surface.o")]
        subprocess.check_call(["nvcc", "-ccbin", "g++"] + bm.ARCH + ["-shared", "-o", os.path.join(HERE, "lib_%s.so" % n), obj] + others)
        os.remove(obj)
        print(n, info, flush=True)
