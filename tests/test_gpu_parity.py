"""GPU parity: the CUDA path, called through the C ABI (ctypes -> liboceanfft.so), against the
CPU oracle on identical noise.  Tolerances are SURVEY.md 8(c):

  displacement / derivatives  rel-L2 <= 1e-5 and max-abs <= 1e-4 * max|field|, per channel
  turbulence                  rel-L2 <= 1e-5 after excluding branch-flip texels (RF:31), flips bounded
  h0 / waveData               rel-L2 <= 1e-6
  arbiter                     CUDA and the fp32 oracle both within 1e-5 rel-L2 of the fp64 oracle

Nothing here reads /root/reference (it does not exist on the GPU box).
"""
import numpy as np
import pytest

from conftest import max_abs_rel, rel_l2

pytestmark = pytest.mark.gpu

TOL_FIELD_L2 = 1e-5
TOL_FIELD_MAX = 1e-4
TOL_SPECTRUM_L2 = 1e-6


def make_body(pkg, po, cfg, N=None, noise=None, **kw):
    N = N or cfg["N"]
    if noise is None:
        noise = po.fill_noise(cfg["seed"], N)
    o = cfg["ocean"]
    body = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5],
                         texturesSize=N, cascades=cfg["cascades"], noise=noise, **kw)
    body.Awake()
    return body, noise


def check_fields(got, want, what, tol_l2=TOL_FIELD_L2, tol_max=TOL_FIELD_MAX, channels=4):
    for c in range(got.shape[0]):
        for ch in range(channels):
            g, w = got[c, ..., ch], want[c, ..., ch]
            if np.abs(w).max() == 0:
                assert np.abs(g).max() == 0, "%s cascade %d ch %d: expected all zero" % (what, c, ch)
                continue
            l2, mx = rel_l2(g, w), max_abs_rel(g, w)
            assert l2 <= tol_l2, "%s cascade %d ch %d rel-L2 %.3g" % (what, c, ch, l2)
            assert mx <= tol_max, "%s cascade %d ch %d max-abs/max %.3g" % (what, c, ch, mx)


def check_turbulence(got, want, jac_scale=1.0):
    """Foam: `if (foam < J) foam += J` can flip on 1-ulp differences; a flip changes the texel by O(J).
    Exclude texels whose difference is of that order, bound their count, and test the rest."""
    g, w = got[..., 0].astype(np.float64), want[..., 0].astype(np.float64)
    diff = np.abs(g - w)
    flips = diff > 1e-3 * max(np.abs(w).max(), 1e-30)
    frac = flips.mean()
    assert frac <= 1e-4, "too many foam branch flips: %.3g of texels" % frac
    keep = ~flips
    l2 = np.linalg.norm((g - w)[keep]) / np.linalg.norm(w[keep])
    assert l2 <= TOL_FIELD_L2, "turbulence rel-L2 %.3g (flips excluded: %d)" % (l2, flips.sum())
    # all four channels carry the same value (RF:32)
    for ch in range(1, 4):
        assert np.array_equal(got[..., ch], got[..., 0])
    return int(flips.sum())


@pytest.mark.parametrize("N", [16, 32, 64, 128, 256, 512, 1024])
def test_initial_spectrum_matches_oracle(pkg, po, N):
    cfg = pkg.scene.config(3)
    body, noise = make_body(pkg, po, cfg, N=N)
    ob = po.OracleB(N, cfg["ocean"], cfg["cascades"], noise)
    ob.init_spectrum()
    h0, wave = body.initialSpectrumTextures, body.wavesDataTextures
    rh0, rwave = ob.get(po.H0), ob.get(po.WAVE)
    # structure is exact: zero pattern (cut-offs), k vectors, Hermitian packing (IS:139-141)
    assert np.array_equal(h0 == 0, rh0 == 0)
    assert np.array_equal(wave[..., 0], rwave[..., 0]) and np.array_equal(wave[..., 2], rwave[..., 2])
    for c in range(h0.shape[0]):
        assert rel_l2(h0[c], rh0[c]) <= TOL_SPECTRUM_L2
        assert rel_l2(wave[c], rwave[c]) <= TOL_SPECTRUM_L2
    m = (N - np.arange(N)) % N
    mirrored = h0[:, m][:, :, m]
    assert np.array_equal(h0[..., 2], mirrored[..., 0]) and np.array_equal(h0[..., 3], -mirrored[..., 1])
    body.OnDisable()


@pytest.mark.parametrize("N", [16, 32, 64, 128, 256, 512, 1024, 2048])
def test_row_pass_matches_oracle(pkg, po, N):
    """Time evolution fused into the x-direction FFT == oracle's evolve + log2(N) horizontal passes."""
    cfg = pkg.scene.config(1)
    body, noise = make_body(pkg, po, cfg, N=N)
    ob = po.OracleB(N, cfg["ocean"], cfg["cascades"], noise, threads=0)
    ob.init_spectrum()
    # same spectrum on both sides so the comparison isolates the frame kernels
    t = 1.0
    ob.step_rows_only(t)
    body.CalculateWavesTexturesAtTime(t)
    rows = body.rowSpectra  # [2][C][N][N][4]
    for k in range(4):
        want = ob.get(po.ROWS0 + k)  # [C][N][N][2]
        # plane layout (re_first, re_second, im_first, im_second): ocean_fft.cuh packs two spectra SoA
        got = np.stack([rows[k // 2][..., k % 2], rows[k // 2][..., 2 + k % 2]], axis=-1)
        assert rel_l2(got, want) <= TOL_FIELD_L2, "spectrum %d after the row pass" % k
    body.OnDisable()


@pytest.mark.parametrize("N", [16, 32, 64, 128, 256, 512, 1024, 2048])
def test_frame_matches_oracle_every_size(pkg, po, N):
    cfg = pkg.scene.config(1)
    body, noise = make_body(pkg, po, cfg, N=N)
    ob = po.OracleB(N, cfg["ocean"], cfg["cascades"], noise, threads=0)
    ob.init_spectrum()
    for t in (0.0, 1.0, 7.25):
        ob.step(t)
        body.CalculateWavesTexturesAtTime(t)
    check_fields(body.displacementsTextures, ob.get(po.DISP), "displacement")
    check_fields(body.derivativesTextures, ob.get(po.DERIV), "derivatives")
    check_turbulence(body.turbulenceTextures, ob.get(po.TURB))
    assert np.all(body.displacementsTextures[..., 3] == 0)
    body.OnDisable()


def test_config1_with_fp64_arbiter(pkg, po):
    """BASELINE config 1: 1 cascade x 256^2, U10=10, fetch 100 km, depth 500 m, t = 1 s."""
    cfg = pkg.scene.config(1)
    body, noise = make_body(pkg, po, cfg)
    N = cfg["N"]
    o32 = po.OracleB(N, cfg["ocean"], cfg["cascades"], noise, threads=0)
    o64 = po.OracleB(N, cfg["ocean"], cfg["cascades"], noise, double=True, threads=0)
    for o in (o32, o64):
        o.init_spectrum()
        o.step(1.0)
    body.CalculateWavesTexturesAtTime(1.0)
    for which, arr in ((po.DISP, body.displacementsTextures), (po.DERIV, body.derivativesTextures)):
        want32, want64 = o32.get(which), o64.get(which)
        check_fields(arr, want32, "config1 array %d" % which)
        e_cuda, e_cpu = rel_l2(arr, want64), rel_l2(want32, want64)
        assert e_cuda <= 1e-5 and e_cpu <= 1e-5
        # the CUDA result must be no farther from fp64 than the fp32 oracle is, within a small factor
        assert e_cuda <= 4 * e_cpu + 1e-7, (e_cuda, e_cpu)
    body.OnDisable()


def test_config3_headline_shape(pkg, po):
    """BASELINE config 3: 4 cascades x 512^2, depth 20 m (TMA active), swell 0.8; a few frames."""
    cfg = pkg.scene.config(3)
    body, noise = make_body(pkg, po, cfg)
    ob = po.OracleB(cfg["N"], cfg["ocean"], cfg["cascades"], noise, threads=0)
    ob.init_spectrum()
    for c in range(4):
        assert rel_l2(body.initialSpectrumTextures[c], ob.get(po.H0)[c]) <= TOL_SPECTRUM_L2
    for f in range(3):
        ob.step(cfg["times"][f])
        body.CalculateWavesTexturesAtTime(cfg["times"][f])
    check_fields(body.displacementsTextures, ob.get(po.DISP), "displacement")
    check_fields(body.derivativesTextures, ob.get(po.DERIV), "derivatives")
    check_turbulence(body.turbulenceTextures, ob.get(po.TURB))
    body.OnDisable()


def test_config2_foam_accumulation(pkg, po):
    """BASELINE config 2 (scene defaults, 3 cascades x 256^2): foam recursion over many frames.
    The oracle is run for 60 frames here; test_config2_all_600_frames_against_the_oracle runs all 600."""
    cfg = pkg.scene.config(2)
    body, noise = make_body(pkg, po, cfg)
    ob = po.OracleB(cfg["N"], cfg["ocean"], cfg["cascades"], noise, threads=0)
    ob.init_spectrum()
    flips = 0
    for f in range(60):
        ob.step(cfg["times"][f])
        body.CalculateWavesTexturesAtTime(cfg["times"][f])
        if f % 20 == 19:
            flips = max(flips, check_turbulence(body.turbulenceTextures, ob.get(po.TURB)))
    check_fields(body.displacementsTextures, ob.get(po.DISP), "displacement")
    check_fields(body.derivativesTextures, ob.get(po.DERIV), "derivatives")
    assert body.frame == 60
    # 600-frame property: foam_n = e^-2 foam_{n-1} (+ J if below J)  =>  J <= foam <= J / (1 - e^-2) wherever J > 0
    for f in range(60, 600):
        body.CalculateWavesTexturesAtTime(cfg["times"][f])
    turb = body.turbulenceTextures[..., 0].astype(np.float64)
    assert np.isfinite(turb).all()
    assert turb.max() <= 50.0  # bounded: no blow-up over 600 frames
    body.OnDisable()


def test_foam_checkpoint_resume(pkg, po):
    """get/set of the turbulence array resumes a run bit-for-bit (checkpoint/resume)."""
    cfg = pkg.scene.config(2)
    a, noise = make_body(pkg, po, cfg, N=128)
    b, _ = make_body(pkg, po, cfg, N=128, noise=noise)
    for f in range(10):
        a.CalculateWavesTexturesAtTime(f / 60.0)
    snap = a.turbulenceTextures
    b.SetTurbulence(snap)
    for f in range(10, 15):
        a.CalculateWavesTexturesAtTime(f / 60.0)
        b.CalculateWavesTexturesAtTime(f / 60.0)
    assert np.array_equal(a.turbulenceTextures, b.turbulenceTextures)
    assert np.array_equal(a.displacementsTextures, b.displacementsTextures)
    a.ResetFoam()
    a.CalculateWavesTexturesAtTime(0.0)
    c, _ = make_body(pkg, po, cfg, N=128, noise=noise)
    c.CalculateWavesTexturesAtTime(0.0)
    assert np.array_equal(a.turbulenceTextures, c.turbulenceTextures)
    for x in (a, b, c):
        x.OnDisable()


def test_device_noise_is_bit_identical_to_host(pkg, po):
    """ocean_create(noise=NULL, seed) generates the texture on the GPU with include/ocean_noise.h."""
    for N, seed in ((64, 1), (256, 12345678901234)):
        body = pkg.WaterBody(windSpeed=8, windDirection=(1, -1), fetch=50000, depth=2560, texturesSize=N,
                             cascades=pkg.scene.SCENE_CASCADES[:1], seed=seed).Awake()
        assert np.array_equal(body.randomNoiseTexture.view(np.uint32), po.fill_noise(seed, N).view(np.uint32))
        body.OnDisable()


def test_linearity_and_single_mode(pkg, po):
    """Size-independent properties at full size: the frame is linear in the noise, and a spectrum with
    a single +-k pair gives an analytic cosine travelling at omega = sqrt(g k)."""
    cfg = pkg.scene.config(3)
    N = cfg["N"]
    n1, n2 = po.fill_noise(11, N), po.fill_noise(12, N)
    outs = []
    for nz in (n1, n2, n1 + n2):
        body, _ = make_body(pkg, po, cfg, noise=nz)
        body.CalculateWavesTexturesAtTime(2.5)
        outs.append((body.displacementsTextures, body.derivativesTextures))
        body.OnDisable()
    for k in range(2):
        assert rel_l2(outs[0][k] + outs[1][k], outs[2][k]) <= 1e-5

    # single mode: noise non-zero only at texel (N/2 + 20, N/2 - 12) -> k = (20, -12) dk, near the spectral peak
    cfg1 = dict(cfg, cascades=cfg["cascades"][:1])
    nz = np.zeros((N, N, 2), np.float32)
    nz[N // 2 - 12, N // 2 + 20] = (1.0, 0.0)
    body, _ = make_body(pkg, po, cfg1, noise=nz)
    h0 = body.initialSpectrumTextures[0]
    wv = body.wavesDataTextures[0]
    a = float(h0[N // 2 - 12, N // 2 + 20, 0])
    kx, kz, om = (float(v) for v in wv[N // 2 - 12, N // 2 + 20, [0, 2, 3]])
    assert a > 0
    t = 3.0
    body.CalculateWavesTexturesAtTime(t)
    Dy = body.displacementsTextures[0, ..., 1].astype(np.float64)
    L = cfg1["cascades"][0][0]
    xs = np.arange(N) * L / N
    X, Z = np.meshgrid(xs, xs)
    # h(k,t) = a e^{i w t} at +k and a e^{-i w t} at -k  =>  Dy = 2 a cos(k.x + w t) after the ifftshift sign
    want = 2 * a * np.cos(kx * X + kz * Z + om * t)
    assert np.abs(Dy - want).max() <= 2e-5 * 2 * a + 1e-12
    body.OnDisable()


def test_query_nearest_matches_reference_lookup(pkg, po):
    cfg = pkg.scene.config(1)
    body, noise = make_body(pkg, po, cfg)
    N = cfg["N"]
    body.CalculateWavesTexturesAtTime(1.0)
    rng = np.random.default_rng(40)
    pts = rng.uniform(-N / 2 - 10, N / 2 + 10, (100000, 2)).astype(np.float32)
    pts[:8] = [[-N / 2, -N / 2], [N / 2, N / 2], [0, 0], [-1e9, 1e9], [N / 2 - 1e-3, 0], [0.5, -0.5], [-0.0, 0.0], [127.99999, -128.0]]
    got = body.GetWaterHeights(pts, pkg._lib.QUERY_NEAREST_REF)
    disp = body.displacementsTextures[0]
    # WB:199-208 restated with numpy fp32
    lo, hi = np.float32(-(N // 2)), np.float32(N // 2)
    u = np.clip((pts[:, 0] - lo) / (hi - lo), np.float32(0), np.float32(1)).astype(np.float32)
    v = np.clip((pts[:, 1] - lo) / (hi - lo), np.float32(0), np.float32(1)).astype(np.float32)
    x = np.clip((u * np.float32(N)).astype(np.int32), 0, N - 1)
    y = np.clip((v * np.float32(N)).astype(np.int32), 0, N - 1)
    assert np.array_equal(got, disp[y, x, 1])
    # and against the oracle's transcription of GetWaterHeight on the oracle's own displacement
    ob = po.OracleB(N, cfg["ocean"], cfg["cascades"], noise, threads=0)
    ob.init_spectrum()
    ob.step(1.0)
    want = ob.query_heights(pts)
    assert np.abs(got - want).max() <= 1e-4 * np.abs(want).max()
    body.OnDisable()


def test_query_bilinear_modes(pkg, po):
    cfg = pkg.scene.config(3)
    body, _ = make_body(pkg, po, cfg, N=128)
    N = 128
    body.CalculateWavesTexturesAtTime(0.5)
    disp = body.displacementsTextures.astype(np.float64)
    wl = body.wavelengths
    rng = np.random.default_rng(41)
    pts = rng.uniform(-3000, 3000, (20000, 2)).astype(np.float32)

    def sample(c, p):
        u = p[:, 0].astype(np.float64) / wl[c]
        v = p[:, 1].astype(np.float64) / wl[c]
        fx, fy = (u - np.floor(u)) * N - 0.5, (v - np.floor(v)) * N - 0.5
        x0, y0 = np.floor(fx).astype(int), np.floor(fy).astype(int)
        wx, wy = fx - x0, fy - y0
        x1, y1 = (x0 + 1) % N, (y0 + 1) % N
        x0, y0 = x0 % N, y0 % N
        d = disp[c, ..., 1]
        return (d[y0, x0] * (1 - wx) + d[y0, x1] * wx) * (1 - wy) + (d[y1, x0] * (1 - wx) + d[y1, x1] * wx) * wy

    got1 = body.GetWaterHeights(pts, pkg._lib.QUERY_BILINEAR)
    gotA = body.GetWaterHeights(pts, pkg._lib.QUERY_BILINEAR_ALL_CASCADES)
    want1 = sample(0, pts)
    wantA = sum(sample(c, pts) for c in range(4))
    # fp32 uv arithmetic at |world| ~ 3000 limits agreement with the float64 restatement
    assert np.abs(got1 - want1).max() <= 2e-3 * np.abs(want1).max()
    assert np.abs(gotA - wantA).max() <= 2e-3 * np.abs(wantA).max()
    body.OnDisable()


def test_update_readback_and_get_water_height(pkg, po):
    """WB:284-297 + WB:195-209: 0 before the first readback, then layer 0 of a completed frame."""
    cfg = pkg.scene.config(1)
    body, _ = make_body(pkg, po, cfg, N=128)
    assert body.GetWaterHeight((3.0, 0.0, -7.0)) == 0.0
    body.Update(0.25)
    frame, data = body.WaitReadback()
    assert frame == 0
    assert np.array_equal(data, body.displacementsTextures[0])
    N = 128
    x, y = int((3.0 + 64) / 128 * N), int((-7.0 + 64) / 128 * N)
    assert body.GetWaterHeight((3.0, 99.0, -7.0)) == data[y, x, 1]
    for f in range(1, 6):
        body.Update(0.25 + f / 60.0)
    frame, data = body.WaitReadback()
    assert frame == 5
    assert np.array_equal(data, body.displacementsTextures[0])
    body.OnDisable()


@pytest.mark.parametrize("N", [128, 256, 512, 1024])  # plain-load columns, fused frame kernel, TMA columns (2 shapes)
def test_height_only_readback_answers_like_the_rgba_readback(pkg, po, N):
    """OCEAN_READBACK_HEIGHT ships only `.g` of layer 0 (what WB:206-208 reads); GetWaterHeight is unchanged.
    The height plane is written by whichever column kernel the size selects."""
    cfg = pkg.scene.config(2)
    body, _ = make_body(pkg, po, cfg, N=N)
    L = pkg._lib
    pts = [(3.0, 0.0, -7.0), (-63.5, 0.0, 63.9), (0.0, 0.0, 0.0), (500.0, 0.0, -500.0)]
    body.Update(0.5)
    frame, rgba = body.WaitReadback()
    assert frame == 0 and rgba.shape == (N, N, 4)
    want = [body.GetWaterHeight(p) for p in pts]
    body.SetReadbackMode(L.READBACK_HEIGHT)
    body.Update(0.5)  # same time: same displacement (only the foam recursion advances)
    frame, h = body.WaitReadback()
    assert frame == 1 and h.shape == (N, N)
    assert np.array_equal(h, rgba[..., 1]) and np.array_equal(h, body.displacementsTextures[0][..., 1])
    assert [body.GetWaterHeight(p) for p in pts] == want
    body.OnDisable()


def test_choppiness_lambda(pkg, po):
    """lambda scales the horizontal displacement and enters the Jacobian; lambda = 1 is the reference."""
    cfg = pkg.scene.config(1)
    a, noise = make_body(pkg, po, cfg, N=128)
    b, _ = make_body(pkg, po, cfg, N=128, noise=noise, choppiness=0.5)
    a.CalculateWavesTexturesAtTime(1.0)
    b.CalculateWavesTexturesAtTime(1.0)
    da, db = a.displacementsTextures, b.displacementsTextures
    assert np.array_equal(db[..., 1], da[..., 1])
    assert np.array_equal(db[..., 0], da[..., 0] * np.float32(0.5)) and np.array_equal(db[..., 2], da[..., 2] * np.float32(0.5))
    ga, gb = a.derivativesTextures, b.derivativesTextures
    assert np.array_equal(gb[..., :2], ga[..., :2])
    assert np.array_equal(gb[..., 2:], ga[..., 2:] * np.float32(0.5))
    a.OnDisable()
    b.OnDisable()


def test_reinit_after_parameter_change(pkg, po):
    """OnValidate path (WB:324-337): change wind/fetch/depth, recompute the spectrum, compare with the oracle."""
    cfg = pkg.scene.config(2)
    body, noise = make_body(pkg, po, cfg, N=128)
    body.windSpeed, body.windDirection, body.fetch, body.depth = 15.0, (0.3, 1.0), 2.0e5, 35.0
    body.OnValidate()
    ocean = (15.0, 0.3, 1.0, 9.81, 2.0e5, 35.0)
    ob = po.OracleB(128, ocean, cfg["cascades"], noise)
    ob.init_spectrum()
    assert rel_l2(body.initialSpectrumTextures, ob.get(po.H0)) <= TOL_SPECTRUM_L2
    ob.step(0.5)
    body.ResetFoam()
    body.CalculateWavesTexturesAtTime(0.5)
    check_fields(body.displacementsTextures, ob.get(po.DISP), "displacement")
    body.OnDisable()


def _ensemble_env(value, name="OCEAN_ENSEMBLE_FUSED_INIT"):
    """OCEAN_ENSEMBLE_FUSED_INIT / OCEAN_ENSEMBLE_LANES are read when an ensemble is created."""
    import contextlib, os

    @contextlib.contextmanager
    def cm():
        old = os.environ.get(name)
        if value is None:
            os.environ.pop(name, None)
        else:
            os.environ[name] = value
        try:
            yield
        finally:
            if old is None:
                os.environ.pop(name, None)
            else:
                os.environ[name] = old
    return cm()


def test_ensemble_lanes_do_not_change_results(pkg, po):
    """Waves alternate between two lanes (streams + pooled buffers) by default; one lane gives the same checksums,
    and so does a second run on the same runner (every wave restarts its lane's counters and buffers)."""
    cfg = pkg.scene.config(5)
    oceans, seeds = pkg.scene.ensemble_states(9)
    with _ensemble_env("1", "OCEAN_ENSEMBLE_LANES"):
        one = pkg.OceanEnsemble(256, cfg["cascades"], batch=2)
    with _ensemble_env(None, "OCEAN_ENSEMBLE_LANES"):
        two = pkg.OceanEnsemble(256, cfg["cascades"], batch=2)
    want = one.run(oceans, seeds, time=0.5)
    assert want.tolist() == two.run(oceans, seeds, time=0.5).tolist()
    assert want.tolist() == two.run(oceans, seeds, time=0.5).tolist()
    # 9 states in waves of 2: the last (ragged) wave ran on lane 0; its slot 0 holds state 8
    assert np.array_equal(one.copy_output(pkg._lib.TURBULENCE, 0), two.copy_output(pkg._lib.TURBULENCE, 0))
    one.close()
    two.close()


def test_ensemble_equals_individual_runs(pkg, po):
    """Ensemble runner (default: separate spectrum kernel; config 5 shape, reduced): every state == a stand-alone
    WaterBody with the same parameters and seed, bit for bit; checksums match the arrays; waves of `batch` cover
    ragged counts."""
    cfg = pkg.scene.config(5)
    N = 128
    oceans, seeds = pkg.scene.ensemble_states(7)
    with _ensemble_env(None):
        ens = pkg.OceanEnsemble(N, cfg["cascades"], batch=3)
        again = pkg.OceanEnsemble(N, cfg["cascades"], batch=4)
    sums = ens.run(oceans, seeds, time=1.0)
    assert sums.shape == (7, 3)
    last = ens.copy_output(pkg._lib.DISPLACEMENT, 0)  # last wave held state 6 in slot 0
    for i in (0, 3, 6):
        o = oceans[i]
        body = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5],
                             texturesSize=N, cascades=cfg["cascades"], seed=int(seeds[i])).Awake()
        body.CalculateWavesTexturesAtTime(1.0)
        arrs = (body.displacementsTextures, body.derivativesTextures, body.turbulenceTextures)
        for k in range(3):
            assert int(arrs[k].view(np.uint32).astype(np.uint64).sum()) == int(sums[i, k])
        if i == 6:
            assert np.array_equal(arrs[0], last)
        body.OnDisable()
    # and one state against the oracle
    o = oceans[3]
    ob = po.OracleB(N, tuple(float(v) for v in o), cfg["cascades"], po.fill_noise(int(seeds[3]), N))
    ob.init_spectrum()
    ob.step(1.0)
    again.run(oceans[:4], seeds[:4], time=1.0)
    check_fields(again.copy_output(pkg._lib.DISPLACEMENT, 3), ob.get(po.DISP), "ensemble displacement")
    assert sums[:4].tolist() == again.run(oceans[:4], seeds[:4], time=1.0).tolist()
    ens.close()
    again.close()


def test_fused_spectrum_request_falls_back_below_128(pkg, po):
    """ocean_rows_from_noise_kernel covers N >= 128; smaller ensembles ignore OCEAN_ENSEMBLE_FUSED_INIT and stay bit-identical
    to stand-alone oceans."""
    cfg = pkg.scene.config(5)
    oceans, seeds = pkg.scene.ensemble_states(3)
    with _ensemble_env("1"):
        ens = pkg.OceanEnsemble(64, cfg["cascades"], batch=2)
    sums = ens.run(oceans, seeds, time=1.0)
    o = oceans[2]
    body = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5],
                         texturesSize=64, cascades=cfg["cascades"], seed=int(seeds[2])).Awake()
    body.CalculateWavesTexturesAtTime(1.0)
    assert np.array_equal(ens.copy_output(pkg._lib.DISPLACEMENT, 0), body.displacementsTextures)
    assert int(body.derivativesTextures.view(np.uint32).astype(np.uint64).sum()) == int(sums[2, 1])
    body.OnDisable()
    ens.close()


@pytest.mark.parametrize("N", [128, 256, 512, 1024])
def test_ensemble_with_the_spectrum_fused_into_the_row_pass(pkg, po, N):
    """Opt-in ensemble path (ocean_rows_from_noise_kernel; OCEAN_ENSEMBLE_FUSED_INIT=1): h0 never reaches memory.  Every state
    is within the field tolerance of the oracle (all three arrays) and of a stand-alone WaterBody, the checksums
    are those of the arrays the ensemble holds, do not depend on the pool size, and the self-mirrored rows
    (0 and N/2) and ragged last waves are covered."""
    cfg = pkg.scene.config(5)
    n = 5 if N < 512 else 3
    oceans, seeds = pkg.scene.ensemble_states(n)
    with _ensemble_env("1"):
        ens = pkg.OceanEnsemble(N, cfg["cascades"], batch=2)
        again = pkg.OceanEnsemble(N, cfg["cascades"], batch=n)
    sums = ens.run(oceans, seeds, time=1.0)
    assert sums.tolist() == again.run(oceans, seeds, time=1.0).tolist()
    assert sums.tolist() == ens.run(oceans, seeds, time=1.0).tolist()
    ids = (pkg._lib.DISPLACEMENT, pkg._lib.DERIVATIVES, pkg._lib.TURBULENCE)
    for i in range(n):
        got = [again.copy_output(w, i) for w in ids]
        for k in range(3):
            assert int(got[k].view(np.uint32).astype(np.uint64).sum()) == int(sums[i, k])
        if i in (0, n - 1):
            o = oceans[i]
            ob = po.OracleB(N, tuple(float(v) for v in o), cfg["cascades"], po.fill_noise(int(seeds[i]), N), threads=8)
            ob.init_spectrum()
            ob.step(1.0)
            check_fields(got[0], ob.get(po.DISP), "fused-init ensemble displacement")
            check_fields(got[1], ob.get(po.DERIV), "fused-init ensemble derivatives")
            check_turbulence(got[2], ob.get(po.TURB))
            body = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5],
                                 texturesSize=N, cascades=cfg["cascades"], seed=int(seeds[i])).Awake()
            body.CalculateWavesTexturesAtTime(1.0)
            check_fields(got[0], body.displacementsTextures, "fused-init ensemble vs stand-alone ocean")
            body.OnDisable()
    # the last wave of `ens` was ragged (n odd, batch 2): slot 0 holds the last state
    assert np.array_equal(ens.copy_output(ids[0], 0), again.copy_output(ids[0], n - 1))
    ens.close()
    again.close()


def test_abi_flow_and_errors_on_device(pkg, po):
    """create -> init -> step x n -> read -> query -> destroy, as the C# shim would drive it."""
    import ctypes
    L = pkg._lib
    lib = L.load()
    assert lib.ocean_device_count() >= 1
    p = L.OceanParams(8, 1, -1, 9.81, 50000, 2560, 64, 1.0)
    casc = (L.OceanCascade * 2)(L.OceanCascade(1530, 1e-10, 1e12, 0.4, 0.1), L.OceanCascade(201, 1e-5, 1e6, 0.1, 0.1))
    h = ctypes.c_void_p()
    assert lib.ocean_create(ctypes.byref(p), casc, 2, None, 7, 0, ctypes.byref(h)) == L.OK
    assert lib.ocean_step(h, 0.0) == L.ERR_NOT_INITIALIZED
    assert b"ocean_init_spectrum" in lib.ocean_last_error()
    assert lib.ocean_init_spectrum(h) == L.OK
    before = lib.ocean_launch_count()
    for f in range(4):
        assert lib.ocean_step(h, f / 60.0) == L.OK
    assert lib.ocean_sync(h) == L.OK
    assert lib.ocean_launch_count() - before == 8  # two kernels per frame
    assert lib.ocean_frame_index(h) == 4
    ptr, nbytes = ctypes.c_void_p(), ctypes.c_size_t()
    assert lib.ocean_get_output(h, L.DISPLACEMENT, ctypes.byref(ptr), ctypes.byref(nbytes)) == L.OK
    assert ptr.value and nbytes.value == 2 * 64 * 64 * 16
    assert lib.ocean_get_output(h, 99, ctypes.byref(ptr), ctypes.byref(nbytes)) == L.ERR_INVALID_ARGUMENT
    out = np.empty((64, 64, 4), np.float32)
    assert lib.ocean_copy_output(h, L.DISPLACEMENT, 5, out.ctypes.data_as(ctypes.POINTER(ctypes.c_float))) == L.ERR_INVALID_ARGUMENT
    assert lib.ocean_create(ctypes.byref(p), casc, 2, None, 7, 99, ctypes.byref(ctypes.c_void_p())) == L.ERR_INVALID_ARGUMENT
    buf = ctypes.create_string_buffer(512)
    assert lib.ocean_describe(h, buf, 512) == L.OK and b"N=64" in buf.value
    lib.ocean_destroy(h)


@pytest.mark.parametrize("name", ["scene_n32", "shallow_n64"])
def test_cuda_matches_committed_golden_vectors(pkg, po, name):
    """tests/golden/*.npz hold outputs of the reference's own shader code (Oracle A, generated by
    tests/golden/make_golden.py); the CUDA path must reproduce them within the stated tolerances."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
    N = g["noise"].shape[0]
    o = g["ocean"]
    body = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=N,
                         cascades=[tuple(c) for c in g["cascades"]], noise=g["noise"]).Awake()
    for c in range(g["h0"].shape[0]):
        assert rel_l2(body.initialSpectrumTextures[c], g["h0"][c]) <= TOL_SPECTRUM_L2
        assert rel_l2(body.wavesDataTextures[c], g["wave"][c]) <= TOL_SPECTRUM_L2
    for t in g["times"]:
        body.CalculateWavesTexturesAtTime(float(t))
    check_fields(body.displacementsTextures, g["disp"], "golden displacement")
    check_fields(body.derivativesTextures, g["deriv"], "golden derivatives")
    check_turbulence(body.turbulenceTextures, g["turb"])
    body.OnDisable()


def test_sharded_ensemble_on_one_gpu(pkg, po):
    """The sharding helper driving the real ensemble runner: shards computed separately == one run."""
    cfg = pkg.scene.config(5)
    oceans, seeds = pkg.scene.ensemble_states(6)
    ens = pkg.OceanEnsemble(64, cfg["cascades"], batch=4)
    full = ens.run(oceans, seeds, time=1.0)
    parts = [pkg.sharding.run_sharded(lambda oc, sd: ens.run(oc, sd, time=1.0), oceans, seeds, r, 2) for r in range(2)]
    assert np.array_equal(parts[0] + parts[1], full)  # disjoint rows, zero elsewhere
    ens.close()


def test_config4_matches_oracle_with_queries(pkg, po):
    """BASELINE config 4 as specified: 4 cascades x 1024^2, scene ocean, seed 4 -- the regime where the
    row->column workspace (134 MB) exceeds L2 and four layers go through the 544-thread TMA column kernel
    and its per-layer counters -- three frames against Oracle B (WB:180-193, IFFT.cs:66-94), then the
    1 M query points of the config (seed 40) in NEAREST_REF: bit-exact against the numpy restatement of
    WB:199-208 on the library's own displacement, and within tolerance of OracleB.query_heights."""
    cfg = pkg.scene.config(4)
    N = cfg["N"]
    assert N == 1024 and len(cfg["cascades"]) == 4
    body, noise = make_body(pkg, po, cfg)
    ob = po.OracleB(N, cfg["ocean"], cfg["cascades"], noise, threads=0)
    ob.init_spectrum()
    for c in range(4):
        assert rel_l2(body.initialSpectrumTextures[c], ob.get(po.H0)[c]) <= TOL_SPECTRUM_L2
    for f in range(3):
        ob.step(cfg["times"][f])
        body.CalculateWavesTexturesAtTime(cfg["times"][f])
    disp = body.displacementsTextures
    check_fields(disp, ob.get(po.DISP), "config4 displacement")
    check_fields(body.derivativesTextures, ob.get(po.DERIV), "config4 derivatives")
    check_turbulence(body.turbulenceTextures, ob.get(po.TURB))
    rng = np.random.default_rng(cfg["query_seed"])
    pts = rng.uniform(-N / 2, N / 2, (cfg["query_points"], 2)).astype(np.float32)
    got = body.GetWaterHeights(pts, pkg._lib.QUERY_NEAREST_REF)
    lo, hi = np.float32(-(N // 2)), np.float32(N // 2)
    u = np.clip((pts[:, 0] - lo) / (hi - lo), np.float32(0), np.float32(1)).astype(np.float32)
    v = np.clip((pts[:, 1] - lo) / (hi - lo), np.float32(0), np.float32(1)).astype(np.float32)
    x = np.clip((u * np.float32(N)).astype(np.int32), 0, N - 1)
    y = np.clip((v * np.float32(N)).astype(np.int32), 0, N - 1)
    assert np.array_equal(got, disp[0][y, x, 1])
    want = ob.query_heights(pts)
    assert np.abs(got - want).max() <= 1e-4 * np.abs(want).max()
    body.OnDisable()


def test_config2_all_600_frames_against_the_oracle(pkg, po):
    """BASELINE config 2 in full: the foam recursion RF:28-32 over all 600 frames (t_f = f / 60) against
    Oracle B, turbulence checked with the branch-flip rule at frames 60 / 200 / 400 / 600 and the fields at
    the end.  A flipped texel differs by O(J) and decays by e^-2 per frame, so flips do not accumulate."""
    cfg = pkg.scene.config(2)
    body, noise = make_body(pkg, po, cfg)
    ob = po.OracleB(cfg["N"], cfg["ocean"], cfg["cascades"], noise, threads=0)
    ob.init_spectrum()
    flips = []
    for f in range(600):
        ob.step(cfg["times"][f])
        body.CalculateWavesTexturesAtTime(cfg["times"][f])
        if f + 1 in (60, 200, 400, 600):
            flips.append(check_turbulence(body.turbulenceTextures, ob.get(po.TURB)))
    check_fields(body.displacementsTextures, ob.get(po.DISP), "displacement after 600 frames")
    check_fields(body.derivativesTextures, ob.get(po.DERIV), "derivatives after 600 frames")
    assert body.frame == 600
    turb = body.turbulenceTextures[..., 0].astype(np.float64)
    assert np.isfinite(turb).all() and turb.max() <= 50.0
    body.OnDisable()


def test_default_ensemble_path_at_config5_size_matches_oracle(pkg, po):
    """The default (separate spectrum kernel, two lanes) ensemble path at the size config 5 runs:
    3 cascades x 512^2, three randomised states in waves of two, every array of two states against Oracle B."""
    cfg = pkg.scene.config(5)
    N = cfg["N"]
    assert N == 512 and len(cfg["cascades"]) == 3
    oceans, seeds = pkg.scene.ensemble_states(3, cfg["rng_seed"])
    with _ensemble_env(None):
        ens = pkg.OceanEnsemble(N, cfg["cascades"], batch=2)
        whole = pkg.OceanEnsemble(N, cfg["cascades"], batch=3)
    sums = ens.run(oceans, seeds, time=cfg["time"])
    assert sums.tolist() == whole.run(oceans, seeds, time=cfg["time"]).tolist()
    ids = (pkg._lib.DISPLACEMENT, pkg._lib.DERIVATIVES, pkg._lib.TURBULENCE)
    for i in (0, 2):
        got = [whole.copy_output(w, i) for w in ids]
        for k in range(3):
            assert int(got[k].view(np.uint32).astype(np.uint64).sum()) == int(sums[i, k])
        ob = po.OracleB(N, tuple(float(v) for v in oceans[i]), cfg["cascades"], po.fill_noise(int(seeds[i]), N), threads=0)
        ob.init_spectrum()
        ob.step(cfg["time"])
        check_fields(got[0], ob.get(po.DISP), "ensemble state %d displacement" % i)
        check_fields(got[1], ob.get(po.DERIV), "ensemble state %d derivatives" % i)
        check_turbulence(got[2], ob.get(po.TURB))
        ob.close()
    ens.close()
    whole.close()


def test_config4_opt_in_paths_match_the_default(pkg, po):
    """4 x 1024^2 through the two opt-in paths of round 2 -- the fused frame kernel with the layer-interleaved item order
    (OCEAN_FUSED_1024=1) and partial workspace residency (OCEAN_WS_RESIDENT=2) -- in child processes (the switches are read
    once): the digests of all three output arrays after three frames equal the default path's, bit for bit."""
    import json, os, subprocess, sys
    child = (
        "import os, sys, hashlib; sys.path.insert(0, %r); import oceanpkg; pkg = oceanpkg.load(); cfg = pkg.scene.config(4); o = cfg['ocean']\n"
        "b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5], texturesSize=cfg['N'], cascades=cfg['cascades'], seed=4).Awake()\n"
        "[b.CalculateWavesTexturesAtTime(f / 60.0) for f in range(3)]\n"
        "h = hashlib.sha256(); [h.update(a.tobytes()) for a in (b.displacementsTextures, b.derivativesTextures, b.turbulenceTextures)]; print(h.hexdigest())\n"
    ) % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    digests = {}
    for name, env in (("default", {}), ("fused", {"OCEAN_FUSED_1024": "1"}), ("resident2", {"OCEAN_WS_RESIDENT": "2"})):
        r = subprocess.run([sys.executable, "-c", child], env=dict(os.environ, **env), capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-1500:]
        digests[name] = r.stdout.strip().splitlines()[-1]
    assert digests["fused"] == digests["default"] and digests["resident2"] == digests["default"], digests
