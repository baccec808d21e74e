"""C-ABI and host-logic tests that need no GPU: the library loads, exports exactly what include/oceanfft.h
declares, validates arguments before touching CUDA, and fails loudly (no CPU fallback) without a device."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "oceanfft.h")


def header_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ocean_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(pkg):
    L = pkg._lib
    lib = L.load()
    declared = header_symbols()
    assert declared == sorted(L.ABI_SYMBOLS), "header and binding list differ"
    for name in declared:
        assert hasattr(lib, name), name
    out = subprocess.run(["nm", "-D", "--defined-only", L.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (ocean_[a-z_0-9]+)", out))
    assert set(declared) <= exported
    assert lib.ocean_abi_version() == 1


def test_struct_layouts_match_header(pkg):
    L = pkg._lib
    assert ctypes.sizeof(L.OceanParams) == 32 and ctypes.sizeof(L.OceanCascade) == 20
    assert [f[0] for f in L.OceanParams._fields_][:6] == ["windSpeed", "windDirX", "windDirY", "gravity", "fetch", "depth"]


def _create(L, lib, params, casc, n):
    h = ctypes.c_void_p()
    rc = lib.ocean_create(ctypes.byref(params), casc, n, None, 0, 0, ctypes.byref(h))
    if rc == L.OK:
        lib.ocean_destroy(h)
    return rc, lib.ocean_last_error().decode()


def test_argument_validation_precedes_cuda(pkg):
    """SURVEY 8b error contract: rejects bad N, C outside [1,5], zero wind, cutoffLow <= 0, non-positive U/F/g/D.
    These are checked before any CUDA call, so the status codes are the same with or without a GPU."""
    L = pkg._lib
    lib = L.load()
    good_c = (L.OceanCascade * 5)(*[L.OceanCascade(100.0, 1e-4, 10.0, 0.4, 0.1)] * 5)

    def P(**kw):
        base = dict(windSpeed=8, windDirX=1, windDirY=-1, gravity=9.81, fetch=5e4, depth=2560, textureSize=256, lambda_=1.0)
        base.update(kw)
        return L.OceanParams(*[base[k] for k in ("windSpeed", "windDirX", "windDirY", "gravity", "fetch", "depth", "textureSize", "lambda_")])

    for N in (0, 8, 100, 4096, -16):
        assert _create(L, lib, P(textureSize=N), good_c, 1)[0] == L.ERR_INVALID_SIZE
    for bad in (dict(windSpeed=0), dict(gravity=-9.81), dict(fetch=0), dict(depth=float("nan")), dict(windDirX=0, windDirY=0)):
        rc, msg = _create(L, lib, P(**bad), good_c, 1)
        assert rc == L.ERR_INVALID_PARAMS and msg
    for n in (0, 6, -1):
        assert _create(L, lib, P(), good_c, n)[0] == L.ERR_INVALID_CASCADES
    bad_c = (L.OceanCascade * 1)(L.OceanCascade(100.0, 0.0, 10.0, 0.4, 0.1))
    assert _create(L, lib, P(), bad_c, 1)[0] == L.ERR_INVALID_CASCADES
    bad_c = (L.OceanCascade * 1)(L.OceanCascade(-5.0, 1e-4, 10.0, 0.4, 0.1))
    assert _create(L, lib, P(), bad_c, 1)[0] == L.ERR_INVALID_CASCADES
    assert lib.ocean_create(None, good_c, 1, None, 0, 0, ctypes.byref(ctypes.c_void_p())) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_create(ctypes.byref(P()), good_c, 1, None, 0, 0, None) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_step(None, 0.0) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_generate_mips(None) == L.ERR_INVALID_ARGUMENT and lib.ocean_set_auto_mips(None, 1) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_mip_levels(None) == -1
    assert lib.ocean_sample_surface(None, None, 0, 0.0, None, None, None, 0) == L.ERR_INVALID_ARGUMENT
    e = ctypes.c_void_p()
    assert lib.ocean_ensemble_create(100, good_c, 3, 4, 0, ctypes.byref(e)) == L.ERR_INVALID_SIZE
    assert lib.ocean_ensemble_create(256, good_c, 3, 0, 0, ctypes.byref(e)) == L.ERR_INVALID_ARGUMENT


def test_no_cpu_fallback(pkg):
    """Without a CUDA device the product path fails loudly instead of computing on the host."""
    lib = pkg._lib.load()
    if lib.ocean_device_count() > 0:
        pytest.skip("a GPU is present")
    body = pkg.WaterBody(windSpeed=8, windDirection=(1, -1), fetch=5e4, depth=2560, texturesSize=64,
                         cascades=pkg.scene.SCENE_CASCADES[:1])
    with pytest.raises(pkg._lib.OceanError) as ei:
        body.Awake()
    assert ei.value.status == pkg._lib.ERR_CUDA and "no CPU path" in ei.value.message
    with pytest.raises(pkg._lib.OceanError):
        pkg.OceanEnsemble(64, pkg.scene.SCENE_CASCADES[:3], batch=2)


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under the package (or the C sources) may reference it."""
    pkgdir = os.path.join(ROOT, "ocean-simulation_b200")
    for dirpath, _, files in os.walk(pkgdir):
        if os.path.basename(dirpath).startswith("build"):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "pyoracle" not in text and "oracle/" not in text and "ocean_oracle" not in text, os.path.join(dirpath, f)


def test_waterbody_mirror_defaults_and_scene(pkg):
    """Field defaults are the reference's initialisers (WB:10-14, WB:29, WC:10-24); scene numbers are SC's."""
    b = pkg.WaterBody()
    assert (b.windSpeed, b.windDirection, b.gravity, b.fetch, b.depth, b.texturesSize) == (1.0, (1.0, 1.0), 9.81, 1.0, 4.0, 256)
    c = pkg.WaterCascade()
    assert (c.wavelength, c.cutoffHigh, c.cutoffLow, c.swell, c.fade) == (10.0, 5.0, 0.0001, 0.4, 0.1)
    assert pkg.scene.SCENE_OCEAN == (8.0, 1.0, -1.0, 9.81, 50000.0, 2560.0) and len(pkg.scene.SCENE_CASCADES) == 4
    for n, (N, C) in {1: (256, 1), 2: (256, 3), 3: (512, 4), 4: (1024, 4)}.items():
        cfg = pkg.scene.config(n)
        assert cfg["N"] == N and len(cfg["cascades"]) == C
    assert all(c[3] == 0.8 for c in pkg.scene.config(3)["cascades"]) and pkg.scene.config(3)["ocean"][5] == 20.0
    oceans, seeds = pkg.scene.ensemble_states(4096)
    assert oceans.shape == (4096, 6) and seeds.tolist()[:3] == [0, 1, 2]
    assert 3 <= oceans[:, 0].min() and oceans[:, 0].max() <= 20 and 1e4 <= oceans[:, 4].min() and oceans[:, 5].max() <= 3000
    with pytest.raises(RuntimeError):
        b.CalculateWavesTexturesAtTime(0.0)  # Awake() first


def _build_abi_driver(pkg):
    """gcc tests/abi_driver.c against include/oceanfft.h and the in-tree library (no CUDA headers needed)."""
    L = pkg._lib
    exe = os.path.join(ROOT, "tests", "_build", "abi_driver")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    libdir = os.path.dirname(L.LIB_PATH)
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-O1", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "abi_driver.c"), "-o", exe, "-L", libdir, "-l:liboceanfft.so",
           "-Wl,-rpath," + libdir, "-lm"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    return exe


def test_c_driver_compiles_against_the_header_and_links_every_symbol(pkg):
    """A C caller (tests/abi_driver.c) that includes include/oceanfft.h: catches header / library drift that
    the hand-written ctypes signatures cannot.  Without a device it runs the error paths only."""
    exe = _build_abi_driver(pkg)
    src = open(os.path.join(ROOT, "tests", "abi_driver.c")).read()
    for name in header_symbols():
        assert "(const void*)%s" % name in src, "abi_driver.c does not reference %s" % name
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "symbols linked" in r.stdout


@pytest.mark.gpu
def test_c_driver_device_flow(pkg):
    """create -> init -> step -> copy -> query -> buoyancy -> update/readback -> destroy from plain C on cuda:0."""
    exe = _build_abi_driver(pkg)
    r = subprocess.run([exe, "--device"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "device flow ok" in r.stdout
