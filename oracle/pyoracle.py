"""ctypes bindings for the CPU oracles.  TEST INFRASTRUCTURE ONLY.

Oracle B (`OracleB`): oracle/ocean_oracle.hpp, the hand transcription (fp32, or fp64 arbiter).
Oracle A (`OracleA`): the reference's own compute-shader text compiled as C++ (oracle/_ref/),
available when `make -C oracle ref` has been run where /root/reference is visible.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package never does.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_B = os.path.join(HERE, "_build", "libocean_oracle.so")
LIB_A = os.path.join(HERE, "_ref", "libocean_ref.so")

# `which` numbering shared by both libraries
H0, WAVE, SPEC0, SPEC1, SPEC2, SPEC3, PING, DISP, DERIV, TURB, BFLY, NOISE = range(12)
ROWS0 = 12  # ROWS0..ROWS0+3: state after the horizontal passes (Oracle B only)

_fp = ctypes.POINTER(ctypes.c_float)
_dp = ctypes.POINTER(ctypes.c_double)


def build(ref=True, quiet=True):
    """Compile the checker libraries (gcc only).  Oracle A is skipped when the reference is absent."""
    targets = ["oracle"] + (["ref"] if ref else [])
    out = subprocess.run(["make", "-C", HERE] + targets, capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    if not quiet:
        print(out.stdout)


def have_ref():
    return os.path.exists(LIB_A)


_libB = None
_libA = None


def _load_b():
    global _libB
    if _libB is None:
        if not os.path.exists(LIB_B):
            build(ref=False)
        L = ctypes.CDLL(LIB_B)
        L.oracle_create.restype = ctypes.c_void_p
        L.oracle_create.argtypes = [ctypes.c_int, ctypes.c_int, _fp, _fp, _fp, ctypes.c_int]
        L.oracle_fill_noise.argtypes = [ctypes.c_uint64, ctypes.c_int, _fp]
        for name in ("oracle_destroy", "oracle_init_spectrum", "oracle_reset_foam"):
            getattr(L, name).argtypes = [ctypes.c_void_p]
        L.oracle_set_threads.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.oracle_get_threads.argtypes = [ctypes.c_void_p]
        for name in ("oracle_step", "oracle_evolve", "oracle_step_rows_only"):
            getattr(L, name).argtypes = [ctypes.c_void_p, ctypes.c_double]
        L.oracle_size.restype = ctypes.c_longlong
        L.oracle_size.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.oracle_get_f32.argtypes = [ctypes.c_void_p, ctypes.c_int, _fp]
        L.oracle_get_f64.argtypes = [ctypes.c_void_p, ctypes.c_int, _dp]
        L.oracle_set_turbulence.argtypes = [ctypes.c_void_p, _fp]
        L.oracle_query_heights.argtypes = [ctypes.c_void_p, _fp, ctypes.c_int, _fp]
        _libB = L
    return _libB


def _load_a():
    global _libA
    if _libA is None:
        if not os.path.exists(LIB_A):
            raise RuntimeError("Oracle A (oracle/_ref/libocean_ref.so) has not been built")
        L = ctypes.CDLL(LIB_A)
        L.ref_create.restype = ctypes.c_void_p
        L.ref_create.argtypes = [ctypes.c_int, ctypes.c_int, _fp, _fp, _fp]
        for name in ("ref_destroy", "ref_init_spectrum", "ref_reset_foam"):
            getattr(L, name).argtypes = [ctypes.c_void_p]
        L.ref_set_threads.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.ref_step.argtypes = [ctypes.c_void_p, ctypes.c_float]
        L.ref_size.restype = ctypes.c_longlong
        L.ref_size.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.ref_get_f32.argtypes = [ctypes.c_void_p, ctypes.c_int, _fp]
        _libA = L
    return _libA


def fill_noise(seed, N):
    """N x N x 2 float32 noise texture, [y][x][ch], from include/ocean_noise.h."""
    out = np.empty((N, N, 2), np.float32)
    _load_b().oracle_fill_noise(ctypes.c_uint64(seed), N, out.ctypes.data_as(_fp))
    return out


def _shape(N, C, which, size):
    if which in (BFLY,):
        return (N, size // (4 * N), 4)
    if which == NOISE:
        return (N, N, 2)
    ch = size // (C * N * N)
    return (C, N, N, ch)


def _pack(params, cascades, noise):
    p = np.ascontiguousarray(np.asarray(params, np.float32).reshape(6))
    c = np.ascontiguousarray(np.asarray(cascades, np.float32).reshape(-1, 5))
    n = np.ascontiguousarray(noise, dtype=np.float32)
    return p, c, n


class OracleB:
    """params = (windSpeed, windDirX, windDirY, gravity, fetch, depth);
    cascades = C x (wavelength, cutoffLow, cutoffHigh, swell, fade); noise = [N][N][2]."""

    def __init__(self, N, params, cascades, noise, double=False, threads=1):
        self.L = _load_b()
        p, c, n = _pack(params, cascades, noise)
        assert n.shape == (N, N, 2)
        self.N, self.C, self.double = N, c.shape[0], double
        self.h = self.L.oracle_create(N, self.C, p.ctypes.data_as(_fp), c.ctypes.data_as(_fp), n.ctypes.data_as(_fp), int(double))
        self.set_threads(threads)

    def set_threads(self, n):
        self.L.oracle_set_threads(self.h, n)
        return self.L.oracle_get_threads(self.h)

    def init_spectrum(self):
        self.L.oracle_init_spectrum(self.h)

    def step(self, t):
        self.L.oracle_step(self.h, float(t))

    def evolve(self, t):
        self.L.oracle_evolve(self.h, float(t))

    def step_rows_only(self, t):
        self.L.oracle_step_rows_only(self.h, float(t))

    def reset_foam(self):
        self.L.oracle_reset_foam(self.h)

    def set_turbulence(self, arr):
        a = np.ascontiguousarray(arr, np.float32)
        self.L.oracle_set_turbulence(self.h, a.ctypes.data_as(_fp))

    def get(self, which, dtype=None):
        size = self.L.oracle_size(self.h, which)
        if size < 0:
            raise ValueError(which)
        dtype = dtype or (np.float64 if self.double else np.float32)
        out = np.empty(size, dtype)
        if dtype == np.float64:
            self.L.oracle_get_f64(self.h, which, out.ctypes.data_as(_dp))
        else:
            self.L.oracle_get_f32(self.h, which, out.ctypes.data_as(_fp))
        w = which if which < ROWS0 else SPEC0
        return out.reshape(_shape(self.N, self.C, w, size))

    def query_heights(self, xz):
        xz = np.ascontiguousarray(xz, np.float32).reshape(-1, 2)
        out = np.empty(xz.shape[0], np.float32)
        self.L.oracle_query_heights(self.h, xz.ctypes.data_as(_fp), xz.shape[0], out.ctypes.data_as(_fp))
        return out

    def close(self):
        if self.h:
            self.L.oracle_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class OracleA:
    """The reference's own shader code on the CPU.  One live instance per process at a time
    (the shader translation units keep HLSL-style global uniforms)."""

    def __init__(self, N, params, cascades, noise, threads=1):
        self.L = _load_a()
        p, c, n = _pack(params, cascades, noise)
        assert n.shape == (N, N, 2)
        self.N, self.C = N, c.shape[0]
        self.h = self.L.ref_create(N, self.C, p.ctypes.data_as(_fp), c.ctypes.data_as(_fp), n.ctypes.data_as(_fp))
        self.L.ref_set_threads(self.h, threads)

    def set_threads(self, n):
        self.L.ref_set_threads(self.h, n)

    def init_spectrum(self):
        self.L.ref_init_spectrum(self.h)

    def step(self, t):
        self.L.ref_step(self.h, float(t))

    def reset_foam(self):
        self.L.ref_reset_foam(self.h)

    def get(self, which):
        size = self.L.ref_size(self.h, which)
        if size < 0:
            raise ValueError(which)
        out = np.empty(size, np.float32)
        self.L.ref_get_f32(self.h, which, out.ctypes.data_as(_fp))
        return out.reshape(_shape(self.N, self.C, which, size))

    def close(self):
        if self.h:
            self.L.ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---------------------------------------------------------------------------------------------
# Oracle A for the array consumers (oracle/surface_ref_host.cpp + the reference's own lines of
# Water.shader / WaterBody.cs / BuoyantObject.cs, built by `make -C oracle ref`).
LIB_S = os.path.join(HERE, "_ref", "libsurface_ref.so")
_libS = None


def have_surface_ref():
    return os.path.exists(LIB_S)


def _load_s():
    global _libS
    if _libS is None:
        if not os.path.exists(LIB_S):
            raise RuntimeError("surface Oracle A (oracle/_ref/libsurface_ref.so) has not been built")
        L = ctypes.CDLL(LIB_S)
        i, f = ctypes.c_int, ctypes.c_float
        L.surface_ref_domain.argtypes = [_fp, i, i, _fp, _fp, i, f, _fp]
        L.surface_ref_fragment.argtypes = [_fp, _fp, _fp, _fp, i, i, _fp, _fp, i, f, _fp, _fp]
        L.surface_ref_get_water_height.argtypes = [_fp, i, _fp, i, _fp]
        L.surface_ref_buoyancy.argtypes = [_fp, i, _fp, i, f, f, _fp]
        _libS = L
    return _libS


def _f32(a):
    a = np.ascontiguousarray(a, np.float32)
    return a, a.ctypes.data_as(_fp)


def pack_mips(levels_per_layer):
    """[layer][level0, level1, ...] -> the packed chain layout of the library: [layer][level 1 | level 2 | ...]."""
    return np.concatenate([np.concatenate([lv.reshape(-1, 4) for lv in levels[1:]]) for levels in levels_per_layer]).astype(np.float32)


def surface_ref_domain(disp, wavelengths, pts, lod=0.0):
    """Water.shader:322-326 (reference text) -> displacement [n][3]."""
    d, dp = _f32(disp)
    w, wp = _f32(wavelengths)
    p, pp = _f32(np.asarray(pts, np.float32).reshape(-1, 2))
    out = np.empty((p.shape[0], 3), np.float32)
    _load_s().surface_ref_domain(dp, d.shape[1], d.shape[0], wp, pp, p.shape[0], float(lod), out.ctypes.data_as(_fp))
    return out


def surface_ref_fragment(deriv, deriv_mips, turb, turb_mips, wavelengths, pts, lod=0.0):
    """Water.shader:337-347 (reference text) -> (normalOS [n][3], turbulence sum [n]).  *_mips: packed chains or None."""
    d, dp = _f32(deriv)
    t, tp = _f32(turb)
    dm = tm = None
    dmp = tmp = None
    if deriv_mips is not None:
        dm, dmp = _f32(deriv_mips)
        tm, tmp = _f32(turb_mips)
    w, wp = _f32(wavelengths)
    p, pp = _f32(np.asarray(pts, np.float32).reshape(-1, 2))
    normal, foam = np.empty((p.shape[0], 3), np.float32), np.empty(p.shape[0], np.float32)
    _load_s().surface_ref_fragment(dp, dmp, tp, tmp, d.shape[1], d.shape[0], wp, pp, p.shape[0], float(lod),
                                   normal.ctypes.data_as(_fp), foam.ctypes.data_as(_fp))
    return normal, foam


def surface_ref_get_water_height(layer0, pts):
    """WaterBody.GetWaterHeight WB:195-209 (reference text).  layer0 [N][N][4] or None (no readback yet)."""
    p, pp = _f32(np.asarray(pts, np.float32).reshape(-1, 2))
    out = np.empty(p.shape[0], np.float32)
    if layer0 is None:
        _load_s().surface_ref_get_water_height(None, 0, pp, p.shape[0], out.ctypes.data_as(_fp))
    else:
        a, ap = _f32(layer0)
        _load_s().surface_ref_get_water_height(ap, a.shape[0], pp, p.shape[0], out.ctypes.data_as(_fp))
    return out


def surface_ref_buoyancy(layer0, boxes, density=1.0, gravity_y=-9.81):
    """BuoyantObject.FixedUpdate BO:17-37 (reference text) over WaterBody.GetWaterHeight (reference text):
    [n][2] = (waterHeight, y of the buoyant acceleration it applied)."""
    a, ap = _f32(layer0)
    b, bp = _f32(np.asarray(boxes, np.float32).reshape(-1, 6))
    out = np.empty((b.shape[0], 2), np.float32)
    _load_s().surface_ref_buoyancy(ap, a.shape[0], bp, b.shape[0], float(density), float(gravity_y), out.ctypes.data_as(_fp))
    return out
