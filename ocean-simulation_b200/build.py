"""Builds liboceanfft.so in-tree with nvcc for sm_100a (B200).  No JIT cache, no torch extension:
the library is a plain C-ABI shared object (include/oceanfft.h) that links only the CUDA runtime.

    python ocean-simulation_b200/build.py [--force] [--verbose]
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "build")
LIB = os.path.join(HERE, "liboceanfft.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math", "-Xptxas", "-v"]
# per translation unit flags: the one-time spectrum kernels keep the reference's rounding
# (no FMA contraction); the frame kernels are free to contract.
UNITS = {
    "ocean_init.cu": ["-fmad=false"],
    "ocean_frame.cu": [],
    "ocean_surface.cu": ["-fmad=false"],  # sampler / shading arithmetic: bit-comparable with the fp32 oracle
    "ocean_host.cu": [],
}
HEADERS = ["ocean_internal.h", "ocean_fft.cuh", "ocean_tma.cuh", "ocean_fused.cuh", "ocean_evolve.cuh", "ocean_spectrum.cuh", "../../include/oceanfft.h", "../../include/ocean_noise.h"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False, defines=(), out=None):
    """Compile every CUDA translation unit for sm_100a and link liboceanfft.so.  Returns the path.
    `defines` / `out` build a tuning variant (e.g. ["-DOCEAN_COLS512_W=4"]) into another file."""
    global OBJ, LIB
    if out:
        OBJ = os.path.join(CSRC, "build_" + os.path.basename(out).replace(".so", ""))
        LIB = out
        force = True
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdrs = [os.path.join(CSRC, h) for h in HEADERS] + [os.path.abspath(__file__)]
    objs = []
    log = []
    for src, extra in UNITS.items():
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            # use the PATH gcc: the image's CXX=/opt/gcc/bin/g++ wrapper is not a full toolchain
            cmd = [nvcc, "-ccbin", "g++"] + ARCH + COMMON + extra + list(defines) + ["-c", s, "-o", o]
            r = subprocess.run(cmd, capture_output=True, text=True)
            log.append("$ " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
            if r.returncode != 0:
                raise RuntimeError("nvcc failed:\n" + log[-1])
    if force or _stale(LIB, objs):
        cmd = [nvcc, "-ccbin", "g++"] + ARCH + ["-shared", "-o", LIB] + objs
        r = subprocess.run(cmd, capture_output=True, text=True)
        log.append("$ " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + log[-1])
    if log:
        with open(os.path.join(OBJ, "build.log"), "w") as f:
            f.write("\n".join(log))
        if verbose:
            print("\n".join(log))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
