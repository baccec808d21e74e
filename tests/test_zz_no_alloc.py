"""ocean_buoyancy_noalloc -- the allocation-free variant of ocean_buoyancy (scratch in a static device array of the
library; no counterpart in the reference) -- against ocean_buoyancy itself,
through the C ABI.  The two run the same kernels in the same order, so the comparison is bit for bit, like every other
buoyancy test here (tests/test_gpu_surface.py).

This file is named to be collected LAST (a failure here must not hide other tests) and has
to pass both alone in a fresh process and after every other test file; it therefore builds its own ocean and relies on no
state left behind by other tests.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ocean(pkg):
    cfg = pkg.scene.config(2)  # scene ocean, 3 cascades, 256^2
    o = cfg["ocean"]
    body = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5],
                         texturesSize=cfg["N"], cascades=cfg["cascades"], seed=77).Awake()
    for f in range(3):
        body.CalculateWavesTexturesAtTime(f / 60.0)
    return body, cfg["N"]


def _boxes(rng, n, N):
    b = np.empty((n, 6), np.float32)
    b[:, 0] = rng.uniform(-N / 2 - 8, N / 2 + 8, n)
    b[:, 2] = rng.uniform(-N / 2 - 8, N / 2 + 8, n)
    b[:, 1] = rng.uniform(-2.0, 2.0, n)
    b[:, 3:] = rng.uniform(0.2, 3.0, (n, 3))
    return b


def test_noalloc_buoyancy_equals_buoyancy_on_two_calls_in_a_row(pkg):
    L = pkg._lib
    body, N = _ocean(pkg)
    rng = np.random.default_rng(2024)
    # second call: fewer boxes, other positions, another mode order -- the scratch still holds the first call's contents
    first, second = _boxes(rng, 40000, N), _boxes(rng, 1237, N)
    for boxes in (first, second):
        for mode in (L.QUERY_NEAREST_REF, L.QUERY_BILINEAR_ALL_CASCADES, L.QUERY_DISPLACED):
            want = body.Buoyancy(boxes, mode, density=0.9, gravityY=-9.81)
            got = body.BuoyancyNoAlloc(boxes, mode, density=0.9, gravityY=-9.81)
            assert got.shape == want.shape == (len(boxes), 4)
            assert np.isfinite(got).all()
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), "mode %d, n %d" % (mode, len(boxes))
        assert (want[:, 1] > 0).sum() > 100 and (want[:, 1] == 0).sum() > 100  # both branches of BO:21 were taken
    # the full capacity, and one box
    for n in (L.NOALLOC_MAX_BOXES, 1):
        boxes = _boxes(rng, n, N)
        assert np.array_equal(body.BuoyancyNoAlloc(boxes), body.Buoyancy(boxes))
    # a frame later the answers follow the new frame (the entry point reads the live arrays, not a cached copy)
    body.CalculateWavesTexturesAtTime(1.0)
    assert np.array_equal(body.BuoyancyNoAlloc(second, L.QUERY_DISPLACED), body.Buoyancy(second, L.QUERY_DISPLACED))
    body.OnDisable()


def test_noalloc_buoyancy_with_device_buffers(pkg):
    """Caller-owned device inputs and outputs (the graph-capture shape: nothing is staged, only points / heights use the
    static scratch); asynchronous on the handle's stream."""
    torch = pytest.importorskip("torch")
    L = pkg._lib
    lib = L.load()
    body, N = _ocean(pkg)
    rng = np.random.default_rng(7)
    for n in (5000, 300):  # two calls in a row
        boxes = _boxes(rng, n, N)
        d_boxes = torch.from_numpy(boxes).cuda()
        d_out = torch.full((n, 4), float("nan"), device="cuda")
        torch.cuda.synchronize()
        L.check(lib.ocean_buoyancy_noalloc(body._h, ctypes.c_void_p(d_boxes.data_ptr()), n, L.QUERY_BILINEAR_ALL_CASCADES, 1.0, -9.81,
                                           ctypes.c_void_p(d_out.data_ptr()), 1, 1))
        body.Sync()
        want = body.Buoyancy(boxes, L.QUERY_BILINEAR_ALL_CASCADES, density=1.0, gravityY=-9.81)
        assert np.array_equal(d_out.cpu().numpy().view(np.uint32), want.view(np.uint32))
    body.OnDisable()


def test_noalloc_buoyancy_rejects_what_does_not_fit(pkg):
    L = pkg._lib
    lib = L.load()
    body, N = _ocean(pkg)
    n = L.NOALLOC_MAX_BOXES + 1
    boxes = np.zeros((n, 6), np.float32)
    out = np.zeros((n, 4), np.float32)
    rc = lib.ocean_buoyancy_noalloc(body._h, boxes.ctypes.data, n, L.QUERY_NEAREST_REF, 1.0, -9.81, out.ctypes.data, 0, 0)
    assert rc == L.ERR_INVALID_ARGUMENT
    msg = lib.ocean_last_error().decode()
    assert "OCEAN_NOALLOC_MAX_BOXES" in msg and str(L.NOALLOC_MAX_BOXES) in msg
    assert not out.any()  # nothing was written
    assert lib.ocean_buoyancy_noalloc(body._h, boxes.ctypes.data, 8, L.QUERY_READBACK_REF, 1.0, -9.81, out.ctypes.data, 0, 0) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_buoyancy_noalloc(body._h, None, 8, L.QUERY_NEAREST_REF, 1.0, -9.81, out.ctypes.data, 0, 0) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_buoyancy_noalloc(body._h, boxes.ctypes.data, 0, L.QUERY_NEAREST_REF, 1.0, -9.81, out.ctypes.data, 0, 0) == L.OK
    body.OnDisable()


def test_noalloc_buoyancy_on_a_second_device(pkg):
    """The static scratch exists once per GPU and its device address is resolved per device (cudaGetSymbolAddress with that device
    current): an ocean on cuda:1 must get cuda:1's copy, whichever device is current in the calling thread.  Skipped on one GPU."""
    L = pkg._lib
    lib = L.load()
    if lib.ocean_device_count() < 2:
        pytest.skip("needs two GPUs")
    cfg = pkg.scene.config(2)
    o = cfg["ocean"]
    bodies = []
    for dev in (1, 0, 1):  # device 1 first: its address is resolved before device 0's
        b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=cfg["N"],
                          cascades=cfg["cascades"], seed=77, device=dev).Awake()
        for f in range(3):
            b.CalculateWavesTexturesAtTime(f / 60.0)
        bodies.append(b)
    rng = np.random.default_rng(5)
    boxes = _boxes(rng, 20000, cfg["N"])
    results = []
    for b in bodies:
        want = b.Buoyancy(boxes, L.QUERY_BILINEAR_ALL_CASCADES)
        got = b.BuoyancyNoAlloc(boxes, L.QUERY_BILINEAR_ALL_CASCADES)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        results.append(got)
    assert np.array_equal(results[0], results[1]) and np.array_equal(results[0], results[2])  # same seed, same ocean on both GPUs
    for b in bodies:
        b.OnDisable()
