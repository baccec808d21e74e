"""GPU parity of the array consumers: mip chain WB:191-192, Water.shader:322-347 surface
sampling, displaced height queries -- CUDA through the C ABI against oracle/surface_oracle.py.
The kernels are compiled without FMA contraction, so they are compared bit for bit where the inputs are
the library's own arrays, and at the field tolerance of SURVEY 8(c) end to end against Oracle B."""
import ctypes

import numpy as np
import pytest

from conftest import max_abs_rel
from test_gpu_parity import make_body

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def so():
    from oracle import surface_oracle
    return surface_oracle


def stepped_body(pkg, po, n_cfg=3, N=128, frames=3, **kw):
    cfg = pkg.scene.config(n_cfg)
    body, noise = make_body(pkg, po, cfg, N=N, **kw)
    for f in range(frames):
        body.CalculateWavesTexturesAtTime(0.5 + f / 60.0)
    return cfg, body, noise


def test_mip_chain_matches_the_box_filter_oracle(pkg, po, so):
    L = pkg._lib
    cfg, body, _ = stepped_body(pkg, po)
    body.GenerateMips()
    assert body.mipLevels == 7
    for which, arr in ((L.DERIVATIVES, body.derivativesTextures), (L.TURBULENCE, body.turbulenceTextures)):
        for c in range(arr.shape[0]):
            want = so.mip_chain(arr[c])
            for level in range(0, 8):
                got = body.Mip(which, level, c)
                assert got.shape == want[level].shape
                assert np.array_equal(got, want[level]), (which, c, level)
    body.OnDisable()


def test_auto_mips_follow_every_frame(pkg, po, so):
    L = pkg._lib
    cfg, body, _ = stepped_body(pkg, po, n_cfg=2, N=64, frames=1)
    body.GenerateMips(auto=True)
    before = body.Mip(L.TURBULENCE, 2, 1)
    body.CalculateWavesTexturesAtTime(0.9)
    after = body.Mip(L.TURBULENCE, 2, 1)
    assert not np.array_equal(before, after)
    assert np.array_equal(after, so.mip_chain(body.turbulenceTextures[1])[2])
    assert np.array_equal(body.Mip(L.DERIVATIVES, 6, 0), so.mip_chain(body.derivativesTextures[0])[6])
    body.OnDisable()


def test_auto_mips_with_back_to_back_frames_and_readbacks(pkg, po, so):
    """200 frames enqueued without a host sync, chains rebuilt beside every next frame and a height readback per
    frame: nothing deadlocks, and the final chains belong to the final arrays."""
    L = pkg._lib
    cfg, body, _ = stepped_body(pkg, po, N=None, frames=1)  # 4 x 512^2: the headline kernels
    body.GenerateMips(auto=True)
    body.SetReadbackMode(L.READBACK_HEIGHT)
    for f in range(200):
        body.Update(0.7 + f / 60.0)
    body.Sync()
    frame, heights = body.WaitReadback()
    assert frame == body.frame - 1
    assert np.array_equal(heights, body.displacementsTextures[0][..., 1])
    for which, arr in ((L.DERIVATIVES, body.derivativesTextures), (L.TURBULENCE, body.turbulenceTextures)):
        want = so.mip_chain(arr[2])
        for level in (1, 4, 9):
            assert np.array_equal(body.Mip(which, level, 2), want[level])
    body.OnDisable()


@pytest.mark.parametrize("lod", [0.0, 1.0, 2.6, 40.0])
def test_surface_sample_bit_exact_on_the_library_arrays(pkg, po, so, lod):
    cfg, body, _ = stepped_body(pkg, po)
    rng = np.random.default_rng(50)
    pts = rng.uniform(-2000, 2000, (20000, 2)).astype(np.float32)
    pts[:4] = [[0, 0], [-0.0, 1530.0], [1e6, -1e6], [5.9765625, 5.9765625]]
    d, n, f = body.SampleSurface(pts, lod)
    wd, wn, wf = so.surface(body.displacementsTextures, body.derivativesTextures, body.turbulenceTextures,
                            body.wavelengths, pts, lod)
    assert np.array_equal(d, wd)
    assert np.array_equal(f, wf)
    assert np.array_equal(n, wn)
    assert np.allclose(np.linalg.norm(n.astype(np.float64), axis=1), 1.0, atol=1e-6)
    assert f.min() >= 0.0 and f.max() <= len(body.wavelengths)
    body.OnDisable()


def test_surface_sample_end_to_end_against_oracle_b(pkg, po, so):
    """Spectrum -> frame -> mips -> sampling on the GPU vs the same chain on the CPU oracles."""
    cfg, body, noise = stepped_body(pkg, po, frames=2)
    ob = po.OracleB(128, cfg["ocean"], cfg["cascades"], noise, threads=0)
    ob.init_spectrum()
    for f in range(2):
        ob.step(0.5 + f / 60.0)
    pts = np.random.default_rng(51).uniform(-800, 800, (5000, 2)).astype(np.float32)
    for lod in (0.0, 1.5):
        d, n, f = body.SampleSurface(pts, lod)
        wd, wn, wf = so.surface(ob.get(po.DISP), ob.get(po.DERIV), ob.get(po.TURB), body.wavelengths, pts, lod)
        assert max_abs_rel(d, wd) <= 1e-4
        assert np.abs(n - wn).max() <= 1e-4
        # foam saturates: a texel whose foam branch flipped (RF:31) moves the sum by O(J); allow a few
        bad = np.abs(f - wf) > 1e-4
        assert bad.mean() <= 1e-3
    body.OnDisable()


def test_surface_outputs_are_optional_and_device_buffers_work(pkg, po, so):
    torch = pytest.importorskip("torch")
    cfg, body, _ = stepped_body(pkg, po, n_cfg=2, N=64)
    pts = np.random.default_rng(52).uniform(-300, 300, (999, 2)).astype(np.float32)
    d, n, f = body.SampleSurface(pts, 0.0)
    d2, n2, f2 = body.SampleSurface(pts, 0.0, normal=False, foam=False)
    assert n2 is None and f2 is None and np.array_equal(d, d2)
    _, n3, _ = body.SampleSurface(pts, 0.0, displacement=False, foam=False)
    assert np.array_equal(n, n3)
    # device pointers, asynchronous on the handle's stream
    tp = torch.from_numpy(pts).cuda()
    tf = torch.empty(999, device="cuda")
    tn = torch.empty(999, 3, device="cuda")
    pkg._lib.check(body._lib.ocean_sample_surface(body._h, ctypes.c_void_p(tp.data_ptr()), 999, 0.0, None,
                                                  ctypes.c_void_p(tn.data_ptr()), ctypes.c_void_p(tf.data_ptr()), 1))
    body.Sync()
    assert np.array_equal(tf.cpu().numpy(), f) and np.array_equal(tn.cpu().numpy(), n)
    # empty input is a no-op
    e = body.SampleSurface(np.zeros((0, 2), np.float32))
    assert e[0].shape == (0, 3) and e[2].shape == (0,)
    body.OnDisable()


def test_surface_argument_errors(pkg, po):
    L = pkg._lib
    cfg, body, _ = stepped_body(pkg, po, n_cfg=1, N=64, frames=1)
    lib, h = body._lib, body._h
    pts = np.zeros((4, 2), np.float32)
    out = np.zeros((4, 3), np.float32)
    assert lib.ocean_sample_surface(h, pts.ctypes.data, 4, -1.0, out.ctypes.data, None, None, 0) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_sample_surface(h, pts.ctypes.data, 4, float("nan"), out.ctypes.data, None, None, 0) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_sample_surface(h, pts.ctypes.data, 4, 0.0, None, None, None, 0) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_sample_surface(h, None, 4, 0.0, out.ctypes.data, None, None, 0) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_sample_surface(h, pts.ctypes.data, -1, 0.0, out.ctypes.data, None, None, 0) == L.ERR_INVALID_ARGUMENT
    buf = np.zeros((64, 64, 4), np.float32)
    fp = buf.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
    assert lib.ocean_copy_mip(h, L.TURBULENCE, 1, 0, fp) == L.ERR_NOT_INITIALIZED    # no chain yet
    assert lib.ocean_copy_mip(h, L.TURBULENCE, 0, 0, fp) == L.OK                     # level 0 = the array
    assert np.array_equal(buf, body.turbulenceTextures[0])
    body.GenerateMips()
    assert lib.ocean_copy_mip(h, L.DISPLACEMENT, 1, 0, fp) == L.ERR_INVALID_ARGUMENT  # WB:227: no chain
    assert lib.ocean_copy_mip(h, L.TURBULENCE, 7, 0, fp) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_copy_mip(h, L.TURBULENCE, 1, 1, fp) == L.ERR_INVALID_ARGUMENT
    p, b = ctypes.c_void_p(), ctypes.c_size_t()
    assert lib.ocean_get_mips(h, L.DERIVATIVES, ctypes.byref(p), ctypes.byref(b)) == L.OK
    assert b.value == (64 * 64 - 1) // 3 * 16 and p.value
    body.OnDisable()


def test_displaced_query_matches_the_oracle_and_inverts_the_displacement(pkg, po, so):
    L = pkg._lib
    cfg, body, _ = stepped_body(pkg, po)
    pts = np.random.default_rng(53).uniform(-1500, 1500, (20000, 2)).astype(np.float32)
    got = body.GetWaterHeights(pts, L.QUERY_DISPLACED)
    disp = body.displacementsTextures
    want, (qx, qz) = so.displaced_heights(disp, body.wavelengths, pts, L.DISPLACED_ITERATIONS, return_q=True)
    assert np.array_equal(got, want)
    # the solved surface point lands over the query: |q + Dxz(q) - (x, z)| is small against the displacement
    d, _, _ = so.surface(disp, disp * 0, disp * 0, body.wavelengths, np.stack([qx, qz], 1), 0.0, mips=False)
    res = np.hypot(qx + d[:, 0] - pts[:, 0], qz + d[:, 2] - pts[:, 1])
    scale = np.hypot(d[:, 0], d[:, 2]).max()
    # (where the choppy surface folds over itself -- Jacobian < 0, the foam criterion RF:27 -- no unique
    # inverse exists and the iteration does not contract: bound the bulk, not the maximum)
    d0, _, _ = so.surface(disp, disp * 0, disp * 0, body.wavelengths, pts, 0.0, mips=False)
    res0 = np.hypot(d0[:, 0], d0[:, 2])  # residual of the un-inverted lookup
    assert np.median(res) <= 2e-3 * scale and np.median(res) <= 0.02 * np.median(res0)
    assert (res > 0.05 * scale).mean() <= 0.05
    # and it is not the plain lookup
    plain = body.GetWaterHeights(pts, L.QUERY_BILINEAR_ALL_CASCADES)
    assert np.abs(plain - got).max() > 1e-3 * np.abs(got).max()
    body.OnDisable()


def test_displaced_query_without_choppiness_is_the_plain_lookup(pkg, po):
    """lambda = 0: no horizontal displacement, so the inversion is the identity."""
    L = pkg._lib
    cfg, body, _ = stepped_body(pkg, po, n_cfg=2, N=64, choppiness=0.0)
    pts = np.random.default_rng(54).uniform(-200, 200, (3000, 2)).astype(np.float32)
    a = body.GetWaterHeights(pts, L.QUERY_DISPLACED)
    b = body.GetWaterHeights(pts, L.QUERY_BILINEAR_ALL_CASCADES)
    assert np.abs(a - b).max() <= 1e-5 * np.abs(b).max()
    body.OnDisable()


def test_surface_full_size_properties(pkg, po):
    """BASELINE config 3 at full size (4 x 512^2): size-independent properties of the consumers."""
    L = pkg._lib
    cfg, body, _ = stepped_body(pkg, po, N=None, frames=4)
    body.GenerateMips()
    N = cfg["N"]
    for which, arr in ((L.DERIVATIVES, body.derivativesTextures), (L.TURBULENCE, body.turbulenceTextures)):
        for c in (0, 3):
            mean0 = arr[c].astype(np.float64).reshape(-1, 4).mean(0)
            top = body.Mip(which, body.mipLevels, c)[0, 0].astype(np.float64)
            assert np.allclose(top, mean0, rtol=1e-4, atol=1e-6 * np.abs(arr[c]).max())  # the box filter keeps the mean
    # a regular grid of world positions: unit normals, foam in [0, C], texel-centre samples return texels
    wl0 = body.wavelengths[0]
    i = np.arange(N, dtype=np.float64)
    xs = ((i + 0.5) * wl0 / N)
    gx, gz = np.meshgrid(xs, xs[:64])
    pts = np.stack([gx.ravel(), gz.ravel()], 1).astype(np.float32)
    d, n, f = body.SampleSurface(pts, 0.0)
    assert np.allclose(np.linalg.norm(n.astype(np.float64), axis=1), 1.0, atol=1e-6)
    assert (n[:, 1] > 0).all() and f.min() >= 0 and f.max() <= 4
    body.OnDisable()


@pytest.mark.parametrize("mode_name", ["QUERY_NEAREST_REF", "QUERY_BILINEAR_ALL_CASCADES", "QUERY_DISPLACED"])
def test_buoyancy_matches_the_oracle_bit_for_bit(pkg, po, so, mode_name):
    """ocean_buoyancy = BuoyantObject.FixedUpdate BO:18-30 for a batch of boxes: the height is the query kernel's, the force
    epilogue is compared bit for bit with the fp32 oracle fed with those heights; boxes above, across and below the surface."""
    L = pkg._lib
    mode = getattr(L, mode_name)
    cfg, body, _ = stepped_body(pkg, po)
    N = body.texturesSize
    rng = np.random.default_rng(11)
    n = 4099
    boxes = np.empty((n, 6), np.float32)
    boxes[:, 0] = rng.uniform(-N, N, n)
    boxes[:, 2] = rng.uniform(-N, N, n)
    boxes[:, 1] = rng.uniform(-3.0, 3.0, n)          # around the surface
    boxes[:, 3:6] = rng.uniform(0.2, 4.0, (n, 3))
    boxes[0, 1], boxes[1, 1] = 1e4, -1e4             # far above, far below
    heights = body.GetWaterHeights(boxes[:, [0, 2]], mode)
    got = body.Buoyancy(boxes, mode, density=1.0, gravityY=-9.81)
    want = so.buoyancy(heights, boxes, density=1.0, gravity_y=-9.81)
    assert np.array_equal(got, want)
    assert got[0, 1] == 0 and got[0, 2] == 0 and got[0, 3] == 0
    assert got[1, 3] == np.float32(np.float32(boxes[1, 3] * boxes[1, 4]) * boxes[1, 5])  # fully submerged: the whole volume
    assert (got[:, 1] > 0).sum() > n // 4 and (got[:, 1] == 0).sum() > n // 4
    body.OnDisable()


def test_buoyancy_argument_errors(pkg, po):
    import ctypes
    L = pkg._lib
    lib = L.load()
    cfg, body, _ = stepped_body(pkg, po, N=64, frames=1)
    box = np.zeros((1, 6), np.float32)
    out = np.zeros((1, 4), np.float32)
    assert lib.ocean_buoyancy(body._h, box.ctypes.data, 1, L.QUERY_READBACK_REF, 1.0, -9.81, out.ctypes.data, 0, 0) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_buoyancy(body._h, None, 1, L.QUERY_NEAREST_REF, 1.0, -9.81, out.ctypes.data, 0, 0) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_buoyancy(body._h, box.ctypes.data, -1, L.QUERY_NEAREST_REF, 1.0, -9.81, out.ctypes.data, 0, 0) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_buoyancy(body._h, box.ctypes.data, 0, L.QUERY_NEAREST_REF, 1.0, -9.81, out.ctypes.data, 0, 0) == L.OK
    body.OnDisable()


@pytest.mark.parametrize("N", [64, 256, 512])
def test_auto_mips_belong_to_the_frame_that_is_observed(pkg, po, so, N):
    """ocean_set_auto_mips rebuilds both chains with every frame (WB:191-192).  Whatever the observer (copy_mip,
    sample_surface, an explicit generate, a sync) and however many frames ran in between, the chains seen are those of the
    CURRENT frame's arrays, bit for bit -- the contract any other placement of the chain kernels has to keep."""
    L = pkg._lib
    cfg, body, _ = stepped_body(pkg, po, n_cfg=2, N=N, frames=1)
    body.GenerateMips(auto=True)
    rng = np.random.default_rng(3)
    pts = rng.uniform(-500, 500, (2000, 2)).astype(np.float32)
    t = 1.0
    for burst, observer in ((1, "copy"), (3, "sample"), (2, "generate"), (5, "copy"), (1, "sample"), (4, "sync")):
        for _ in range(burst):
            t += 1.0 / 60.0
            body.CalculateWavesTexturesAtTime(t)
        if observer == "sample":
            d, n, f = body.SampleSurface(pts, 1.7)
            wd, wn, wf = so.surface(body.displacementsTextures, body.derivativesTextures, body.turbulenceTextures, body.wavelengths, pts, 1.7)
            assert np.array_equal(n, wn) and np.array_equal(f, wf)
        elif observer == "generate":
            body.GenerateMips()
        elif observer == "sync":
            body.Sync()
        for which, arr in ((L.DERIVATIVES, body.derivativesTextures), (L.TURBULENCE, body.turbulenceTextures)):
            want = so.mip_chain(arr[1])
            for level in (1, 3, body.mipLevels):
                assert np.array_equal(body.Mip(which, level, 1), want[level]), (observer, which, level)
    body.GenerateMips(auto=False)
    body.CalculateWavesTexturesAtTime(t + 1.0)  # chains are now stale by design: level 0 moved on, the chain did not
    assert not np.array_equal(body.Mip(L.TURBULENCE, 1, 1), so.mip_chain(body.turbulenceTextures[1])[1])
    body.OnDisable()
