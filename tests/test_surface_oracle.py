"""The surface-consumer oracle (oracle/surface_oracle.py) against closed forms: sampler conventions of
WaterBody.cs:104-128, the mip chain of WB:191-192 (box filter), Water.shader:322-347 arithmetic and the
displaced-height inversion.  CPU only."""
import numpy as np
import pytest

from oracle import surface_oracle as so

F = np.float32


def test_bilinear_sampler_conventions():
    n = 8
    rng = np.random.default_rng(0)
    img = rng.standard_normal((n, n, 4)).astype(F)
    # texel centres return the texel
    i = np.arange(n)
    u = ((i + 0.5) / n).astype(F)
    got = so.sample_bilinear(img, u, np.full(n, 2.5 / n, F))
    assert np.array_equal(got, img[2, i])
    # half way between two centres: the mean; across the edge: Repeat wrap
    got = so.sample_bilinear(img, np.array([3.0 / n, 0.0], F), np.array([4.5 / n, 4.5 / n], F))
    assert np.allclose(got[0], 0.5 * (img[4, 2] + img[4, 3]), atol=1e-6)
    assert np.allclose(got[1], 0.5 * (img[4, n - 1] + img[4, 0]), atol=1e-6)
    # uv outside [0, 1) wraps
    a = so.sample_lod([img], np.array([0.3], F), np.array([0.7], F), 0.0)
    b = so.sample_lod([img], np.array([-3.7], F), np.array([5.7], F), 0.0)
    assert np.allclose(a, b, atol=2e-5)


def test_mip_chain_is_the_box_filter():
    img = np.arange(4 * 4 * 4, dtype=F).reshape(4, 4, 4)
    lv = so.mip_chain(img)
    assert [l.shape[0] for l in lv] == [4, 2, 1]
    assert np.array_equal(lv[1][0, 0], (img[0, 0] + img[0, 1] + img[1, 0] + img[1, 1]) / 4)
    assert np.array_equal(lv[1][1, 0], (img[2, 0] + img[2, 1] + img[3, 0] + img[3, 1]) / 4)
    assert np.allclose(lv[2][0, 0], img.reshape(-1, 4).mean(0))
    # the mean is preserved level to level
    rng = np.random.default_rng(1)
    lv = so.mip_chain(rng.standard_normal((64, 64, 4)).astype(F))
    for a, b in zip(lv[:-1], lv[1:]):
        assert np.allclose(a.reshape(-1, 4).mean(0), b.reshape(-1, 4).mean(0), atol=1e-5)


def test_trilinear_blends_two_levels():
    rng = np.random.default_rng(2)
    lv = so.mip_chain(rng.standard_normal((16, 16, 4)).astype(F))
    u, v = rng.random(50).astype(F), rng.random(50).astype(F)
    s1, s2 = so.sample_bilinear(lv[1], u, v), so.sample_bilinear(lv[2], u, v)
    assert np.array_equal(so.sample_lod(lv, u, v, 1.0), s1)
    assert np.allclose(so.sample_lod(lv, u, v, 1.25), s1 + (s2 - s1) * 0.25, atol=1e-6)
    # beyond the last level: clamps to the 1x1 level; a texture without a chain ignores the LOD
    assert np.allclose(so.sample_lod(lv, u, v, 99.0), np.broadcast_to(lv[-1][0, 0], (50, 4)), atol=0)
    assert np.array_equal(so.sample_lod(lv[:1], u, v, 3.0), so.sample_bilinear(lv[0], u, v))


def test_surface_closed_forms():
    """Constant arrays: Water.shader:322-347 collapses to closed forms."""
    C, N = 3, 16
    disp = np.zeros((C, N, N, 4), F)
    deriv = np.zeros((C, N, N, 4), F)
    turb = np.zeros((C, N, N, 4), F)
    for c in range(C):
        disp[c] = [0.5 * (c + 1), 1.0 + c, -0.25 * (c + 1), 0.0]
        deriv[c] = [0.1, -0.2, 0.05, 0.025]
        turb[c] = [0.3, 1.7, -0.4][c]
    pts = np.random.default_rng(3).uniform(-500, 500, (64, 2)).astype(F)
    for lod in (0.0, 2.5):
        d, nrm, foam = so.surface(disp, deriv, turb, [100.0, 31.0, 7.0], pts, lod)
        assert np.allclose(d, [3.0, 6.0, -1.5], atol=1e-6)
        sx, sz = 0.3 / 1.15, -0.6 / 1.075
        want = np.array([-sx, 1.0, -sz]) / np.sqrt(sx * sx + 1 + sz * sz)
        assert np.allclose(nrm, want, atol=1e-6)
        assert np.allclose(np.linalg.norm(nrm, axis=1), 1.0, atol=1e-6)
        assert np.allclose(foam, (1 - 0.3) + (1 - 1.0) + (1 - 0.0), atol=1e-6)  # 1 - saturate(x)


def test_surface_sums_cascades_at_their_own_wavelength():
    C, N = 2, 32
    rng = np.random.default_rng(4)
    disp = rng.standard_normal((C, N, N, 4)).astype(F)
    zero = np.zeros_like(disp)
    wl = [64.0, 10.0]
    # a point on a texel centre of both cascades: x = (i + 0.5) * wl / N
    x = F(2.5 * 64.0 / N)          # texel 2 of cascade 0; 2.5*2 = 5 -> u1 = 0.5 -> fx = 15.5 (between 15 and 16)
    d, _, _ = so.surface(disp, zero, zero, wl, np.array([[x, x]], F), 0.0)
    c1 = 0.25 * (disp[1, 15, 15] + disp[1, 15, 16] + disp[1, 16, 15] + disp[1, 16, 16])
    assert np.allclose(d[0], (disp[0, 2, 2] + c1)[:3], atol=1e-5)


def test_displaced_heights_invert_a_gerstner_wave():
    """One trochoidal wave along x: Dx = -A sin(kx), Dy = A cos(kx).  The surface above X is at the q with
    q - A sin(kq) = X; the fixed-point iteration contracts with ratio A k."""
    N, wl, A = 256, 64.0, 1.5
    k = 2 * np.pi / wl * 2  # two periods per tile: A k = 0.29
    xs = (np.arange(N) + 0.5) * wl / N
    disp = np.zeros((1, N, N, 4), F)
    disp[0, :, :, 0] = (-A * np.sin(k * xs))[None, :]
    disp[0, :, :, 1] = (A * np.cos(k * xs))[None, :]
    X = np.random.default_rng(5).uniform(0, wl, (500, 2)).astype(F)
    h, (qx, qz) = so.displaced_heights(disp, [wl], X, iterations=12, return_q=True)
    # exact inversion by Newton in float64
    q = X[:, 0].astype(np.float64)
    for _ in range(50):
        q -= (q - A * np.sin(k * q) - X[:, 0]) / (1 - A * k * np.cos(k * q))
    assert np.abs(qx - q).max() < 2e-3            # bilinear interpolation error of the sampled sine
    assert np.abs(h - A * np.cos(k * q)).max() < 2e-3
    assert np.array_equal(qz, X[:, 1])            # no z displacement
    # the un-inverted lookup is visibly wrong, and 4 iterations (the library's setting) are within (A k)^5
    h0 = so.displaced_heights(disp, [wl], X, iterations=0)
    assert np.abs(h0 - A * np.cos(k * q)).max() > 0.3
    h4 = so.displaced_heights(disp, [wl], X, iterations=4)
    assert np.abs(h4 - A * np.cos(k * q)).max() < A * (A * k) ** 5 * 3 + 2e-3


def test_buoyancy_oracle_closed_forms():
    """BO:18-30: a 1 x 2 x 3 box whose centre is 0.5 below the water line is a quarter submerged (0.5 / sy), one far below is
    fully submerged (clamp), one above the water feels nothing."""
    from oracle import surface_oracle as so
    boxes = np.array([[0, -0.5, 0, 1, 2, 3], [5, -10.0, 5, 1, 2, 3], [0, 4.0, 0, 1, 2, 3]], np.float32)
    out = so.buoyancy(np.array([0.0, 1.0, 3.5], np.float32), boxes, density=2.0, gravity_y=-10.0)
    assert np.allclose(out[0], [0.0, 0.5, 10.0 * 2.0 * 0.25 * 6.0, 0.25 * 6.0])
    assert np.allclose(out[1], [1.0, 11.0, 10.0 * 2.0 * 6.0, 6.0])
    assert np.array_equal(out[2], np.array([3.5, 0.0, 0.0, 0.0], np.float32))
    assert out.dtype == np.float32


# ---- the pin: oracle/surface_oracle.py == the reference's own lines compiled here (oracle/_ref/libsurface_ref.so) ----
# Water.shader:322-326 / 337-347 (HLSL) and WaterBody.cs:195-209 / BuoyantObject.cs:17-37 (C#) are streamed through sed
# into g++ by `make -C oracle ref`; only the texture sampler underneath them is a restatement (Unity / driver behaviour).

def _pin():
    from oracle import pyoracle as po
    if not po.have_surface_ref():
        if not __import__("os").path.isdir("/root/reference"):
            pytest.skip("oracle/_ref/libsurface_ref.so was not built (no /root/reference here)")
        po.build(ref=True)
    return po


def _random_ocean_arrays(seed, C, N):
    rng = np.random.default_rng(seed)
    disp = rng.standard_normal((C, N, N, 4)).astype(F)
    disp[..., 3] = 0
    deriv = (0.3 * rng.standard_normal((C, N, N, 4))).astype(F)
    turb = np.repeat(rng.uniform(-0.5, 1.8, (C, N, N, 1)).astype(F), 4, axis=3)  # straddles both saturate() bounds
    return disp, deriv, turb


@pytest.mark.parametrize("lod", [0.0, 0.75, 2.0, 3.5, 99.0])
def test_surface_oracle_equals_the_reference_shader_lines(lod):
    po = _pin()
    C, N = 3, 32
    disp, deriv, turb = _random_ocean_arrays(7, C, N)
    wl = np.array([1530.0, 201.0, 34.0], F)
    rng = np.random.default_rng(8)
    pts = rng.uniform(-4000, 4000, (5000, 2)).astype(F)
    pts[:4] = [[0, 0], [-0.0, 1530.0], [100.5, -17.0], [1e6, -1e6]]
    want_d, want_n, want_f = so.surface(disp, deriv, turb, wl, pts, lod=lod, mips=True)
    dm = po.pack_mips([so.mip_chain(deriv[c]) for c in range(C)])
    tm = po.pack_mips([so.mip_chain(turb[c]) for c in range(C)])
    got_d = po.surface_ref_domain(disp, wl, pts, lod)
    got_n, got_f = po.surface_ref_fragment(deriv, dm, turb, tm, wl, pts, lod)
    assert np.array_equal(got_d.view(np.uint32), want_d.view(np.uint32))
    assert np.array_equal(got_n.view(np.uint32), want_n.view(np.uint32))
    assert np.array_equal(got_f.view(np.uint32), want_f.view(np.uint32))
    # a texture without a chain reads level 0 at every LOD
    want_d0, want_n0, want_f0 = so.surface(disp, deriv, turb, wl, pts, lod=lod, mips=False)
    got_n0, got_f0 = po.surface_ref_fragment(deriv, None, turb, None, wl, pts, lod)
    assert np.array_equal(got_n0.view(np.uint32), want_n0.view(np.uint32)) and np.array_equal(got_f0, want_f0)


def _numpy_get_water_height(layer0, pts):
    """WB:199-208 in numpy fp32 -- the restatement the GPU tests compare OCEAN_QUERY_NEAREST_REF with."""
    N = layer0.shape[0]
    lo, hi = F(-(N // 2)), F(N // 2)
    u = np.clip((pts[:, 0] - lo) / (hi - lo), F(0), F(1)).astype(F)
    v = np.clip((pts[:, 1] - lo) / (hi - lo), F(0), F(1)).astype(F)
    x = np.clip((u * F(N)).astype(np.int32), 0, N - 1)
    y = np.clip((v * F(N)).astype(np.int32), 0, N - 1)
    return layer0[y, x, 1]


def test_get_water_height_and_buoyancy_equal_the_reference_csharp_lines():
    po = _pin()
    N = 64
    disp, _, _ = _random_ocean_arrays(9, 1, N)
    rng = np.random.default_rng(10)
    pts = rng.uniform(-N / 2 - 6, N / 2 + 6, (20000, 2)).astype(F)
    pts[:8] = [[-N / 2, -N / 2], [N / 2, N / 2], [0, 0], [-1e9, 1e9], [N / 2 - 1e-3, 0], [0.5, -0.5], [-0.0, 0.0], [31.99999, -32.0]]
    got = po.surface_ref_get_water_height(disp[0], pts)
    assert np.array_equal(got, _numpy_get_water_height(disp[0], pts))
    # ... and Oracle B's transcription of the same method (oracle/ocean_oracle.hpp) agrees bit for bit
    assert np.array_equal(po.surface_ref_get_water_height(None, pts), np.zeros(len(pts), F))  # WB:197: no readback yet
    # BuoyantObject.FixedUpdate over it: boxes straddling the surface, zero and huge scales
    boxes = np.empty((len(pts), 6), F)
    boxes[:, 0], boxes[:, 2] = pts[:, 0], pts[:, 1]
    boxes[:, 1] = rng.uniform(-3, 3, len(pts))
    boxes[:, 3:] = rng.uniform(0.05, 4.0, (len(pts), 3))
    boxes[:3, 4] = [1e-30, 1e30, 0.5]
    for density, g in ((1.0, -1.0), (1.0, -9.81), (0.7, 3.5)):
        want = so.buoyancy(_numpy_get_water_height(disp[0], pts), boxes, density=density, gravity_y=g)
        got = po.surface_ref_buoyancy(disp[0], boxes, density=density, gravity_y=g)
        assert np.array_equal(got[:, 0].view(np.uint32), want[:, 0].view(np.uint32))  # waterHeight   BO:18
        assert np.array_equal(got[:, 1].view(np.uint32), want[:, 2].view(np.uint32))  # |g.y| * (density * submergedVolume)  BO:26-29
        assert (want[:, 1] > 0).sum() > 1000 and (want[:, 1] == 0).sum() > 1000
    # density = |g.y| = 1 makes the applied acceleration the submerged volume itself (BO:25)
    want = so.buoyancy(_numpy_get_water_height(disp[0], pts), boxes, density=1.0, gravity_y=-1.0)
    got = po.surface_ref_buoyancy(disp[0], boxes, density=1.0, gravity_y=-1.0)
    assert np.array_equal(got[:, 1].view(np.uint32), want[:, 3].view(np.uint32))


def test_oracle_b_query_equals_the_reference_get_water_height(po):
    """Oracle B's hand transcription of GetWaterHeight == WB:195-209 compiled from the reference text."""
    pin = _pin()
    import oceanpkg
    cfg = oceanpkg.load().scene.config(1)
    N = 64
    noise = po.fill_noise(1, N)
    ob = po.OracleB(N, cfg["ocean"], cfg["cascades"], noise)
    ob.init_spectrum()
    ob.step(1.0)
    rng = np.random.default_rng(40)
    pts = rng.uniform(-N / 2 - 4, N / 2 + 4, (5000, 2)).astype(F)
    assert np.array_equal(ob.query_heights(pts), pin.surface_ref_get_water_height(ob.get(po.DISP)[0], pts))
