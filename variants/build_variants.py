"""Builds tuning variants of liboceanfft.so (experiments only; not part of the product build).
Only ocean_frame.cu is recompiled per variant; the other objects come from the product build."""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oceanpkg
HERE = os.path.dirname(os.path.abspath(__file__))
VARIANTS = {
    "base": [],
    # sensitivity builds (OCEAN_EXP bits, ocean_frame.cu): results are wrong on purpose, only the timing matters
    "x01_rows_l2_loads": ["-DOCEAN_EXP=1"],
    "x02_cols_no_store": ["-DOCEAN_EXP=2"],
    "x04_no_turb_load": ["-DOCEAN_EXP=4"],
    "x08_rows_no_store": ["-DOCEAN_EXP=8"],
    "x15_no_hbm": ["-DOCEAN_EXP=15"],
    "x32_no_sincos": ["-DOCEAN_EXP=32"],
    "x47_no_hbm_no_sincos": ["-DOCEAN_EXP=47"],
    "x64_counters_only": ["-DOCEAN_EXP=64"],
    "x128_return": ["-DOCEAN_EXP=128"],
    # alternatives of the same arithmetic
    "fast_sincos": ["-DOCEAN_ROWS_FAST_SINCOS=1"],
    "early_producer": ["-DOCEAN_COLS_EARLY_PRODUCER=1"],
    "rows_r2": ["-DOCEAN_ROWS512_R=2", "-DOCEAN_ROWS512_MINB=7"],
    # fused2 timing experiments (OCEAN_F2_EXP bits, ocean_fused2.cuh)
    "f2x1_fast_hot": ["-DOCEAN_F2_EXP=1"],
    "f2x2_no_foam_load": ["-DOCEAN_F2_EXP=2"],
    "f2x3_both": ["-DOCEAN_F2_EXP=3"],
    "x256_cols_no_await": ["-DOCEAN_EXP=256"],
    "hetero_g4": ["-DOCEAN_HETERO_G=4"],
    "hetero_g5": ["-DOCEAN_HETERO_G=5"],
    "hetero_g6": ["-DOCEAN_HETERO_G=6"],
    "hetero_g7": ["-DOCEAN_HETERO_G=7"],
    "rows_m16": ["-DOCEAN_ROWS512_MINB=16"],
    "cols_tw_global": ["-DOCEAN_COLS_TW_SMEM=0"],
    "rows_m16_cols_tw_global": ["-DOCEAN_ROWS512_MINB=16", "-DOCEAN_COLS_TW_SMEM=0"],
    "rows_recompute_hot": ["-DOCEAN_ROWS_RECOMPUTE_HOT=1"],
    "cols_reverse_1024": ["-DOCEAN_COLS_REVERSE_LAYERS=1"],
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
        info = re.findall(r"ocean_(rows|cols|cols_tma)_kernelILi512E.*?\n.*?(\d+) bytes spill stores.*?\n.*?Used (\d+) registers", r.stderr)
        others = [os.path.join(bm.OBJ, o) for o in ("ocean_init.o", "ocean_host.o", "ocean_surface.o")]
        subprocess.check_call(["nvcc", "-ccbin", "g++"] + bm.ARCH + ["-shared", "-o", os.path.join(HERE, "lib_%s.so" % n), obj] + others)
        os.remove(obj)
        print(n, info, flush=True)
