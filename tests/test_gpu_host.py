"""GPU tests of the host-side state machine of liboceanfft.so (workspace sharing, staged parameter edits,
the readback ring across a foam reset, ensemble lambda), through the C ABI."""
import numpy as np
import pytest

from test_gpu_parity import check_fields, make_body

pytestmark = pytest.mark.gpu


def test_two_geometries_with_equal_bytes_on_one_stream(pkg, po):
    """512 x 1 and 256 x 4 have the same plane bytes; on one caller stream they must NOT share a workspace
    (its counters are one per layer and count rows of N texels).  Both oceans, interleaved frame by frame on
    the stream, must equal the same oceans run on private streams, bit for bit."""
    import torch
    cfg = pkg.scene.config(2)
    one = dict(cfg, cascades=cfg["cascades"][:1])
    four = dict(cfg, cascades=pkg.scene.SCENE_CASCADES)
    st = torch.cuda.Stream()
    a, na = make_body(pkg, po, one, N=512)
    b, nb = make_body(pkg, po, four, N=256)
    ra, _ = make_body(pkg, po, one, N=512, noise=na)
    rb, _ = make_body(pkg, po, four, N=256, noise=nb)
    a.SetStream(st.cuda_stream)
    b.SetStream(st.cuda_stream)
    for f in range(6):
        t = f / 60.0
        for x in (a, b, ra, rb):
            x.CalculateWavesTexturesAtTime(t)
    for x, r in ((a, ra), (b, rb)):
        assert np.array_equal(x.displacementsTextures, r.displacementsTextures)
        assert np.array_equal(x.derivativesTextures, r.derivativesTextures)
        assert np.array_equal(x.turbulenceTextures, r.turbulenceTextures)
    # same-geometry handles on the stream still share (and still agree)
    c, _ = make_body(pkg, po, four, N=256, noise=nb)
    c.SetStream(st.cuda_stream)
    for f in range(6):
        c.CalculateWavesTexturesAtTime(f / 60.0)
    assert np.array_equal(c.turbulenceTextures, rb.turbulenceTextures)
    for x in (a, b, c, ra, rb):
        x.OnDisable()


def test_set_params_is_staged_until_the_spectrum_is_recomputed(pkg, po):
    """ocean_set_params between two frames must not change the second one (new wavelengths with the old h0 /
    omega would be inconsistent); ocean_init_spectrum commits the edit and the oracle agrees afterwards."""
    cfg = pkg.scene.config(2)
    body, noise = make_body(pkg, po, cfg, N=128)
    twin, _ = make_body(pkg, po, cfg, N=128, noise=noise)
    body.CalculateWavesTexturesAtTime(0.0)
    twin.CalculateWavesTexturesAtTime(0.0)
    wl = body.wavelengths
    body.windSpeed, body.depth, body.choppiness = 15.0, 35.0, 0.5
    body.cascades = [pkg.WaterCascade.from_tuple((c[0] * 0.5, c[1], c[2], c[3], c[4])) for c in cfg["cascades"]]
    body.SetParams()
    body.CalculateWavesTexturesAtTime(0.1)
    twin.CalculateWavesTexturesAtTime(0.1)
    assert body.wavelengths == wl
    assert np.array_equal(body.displacementsTextures, twin.displacementsTextures)
    assert np.array_equal(body.turbulenceTextures, twin.turbulenceTextures)
    pts = np.array([[3.0, -7.0], [250.0, 90.0]], np.float32)
    assert np.array_equal(body.GetWaterHeights(pts, pkg._lib.QUERY_BILINEAR_ALL_CASCADES),
                          twin.GetWaterHeights(pts, pkg._lib.QUERY_BILINEAR_ALL_CASCADES))
    body.CalculateInitialSpectrumTextures()
    assert body.wavelengths == [np.float32(w * 0.5) for w in wl]
    casc = [(c[0] * 0.5, c[1], c[2], c[3], c[4]) for c in cfg["cascades"]]
    ocean = (15.0,) + tuple(cfg["ocean"][1:5]) + (35.0,)
    ob = po.OracleB(128, ocean, casc, noise)
    ob.init_spectrum()
    ob.step(0.2)
    body.ResetFoam()
    body.CalculateWavesTexturesAtTime(0.2)
    want = ob.get(po.DISP).copy()
    want[..., 0] *= np.float32(0.5)  # lambda = 0.5 scales the horizontal displacement (the oracle is lambda = 1)
    want[..., 2] *= np.float32(0.5)
    check_fields(body.displacementsTextures, want, "displacement after the committed edit")
    body.OnDisable()
    twin.OnDisable()


def test_reset_foam_restarts_the_readback_ring(pkg, po):
    """After ocean_reset_foam the frame index restarts; READBACK_REF must answer 0 (WB:197) until a
    post-reset readback completes, then serve it -- not a stale pre-reset layer with a larger frame index."""
    cfg = pkg.scene.config(1)
    body, _ = make_body(pkg, po, cfg, N=128)
    for f in range(5):
        body.Update(f / 60.0)
    frame, _ = body.WaitReadback()
    assert frame == 4
    assert body.GetWaterHeight((3.0, 0.0, -7.0)) != 0.0
    body.ResetFoam()
    assert body.GetWaterHeight((3.0, 0.0, -7.0)) == 0.0
    assert body.WaitReadback() == (-1, None)
    body.Update(2.0)
    frame, data = body.WaitReadback()
    assert frame == 0
    assert np.array_equal(data, body.displacementsTextures[0])
    for f in range(1, 5):  # every slot of the ring is usable again
        body.Update(2.0 + f / 60.0)
    frame, data = body.WaitReadback()
    assert frame == 4 and np.array_equal(data, body.displacementsTextures[0])
    body.OnDisable()


def test_ensemble_rejects_a_wave_with_mixed_lambda(pkg, po):
    cfg = pkg.scene.config(5)
    oceans, seeds = pkg.scene.ensemble_states(4)
    ens = pkg.OceanEnsemble(64, cfg["cascades"], batch=2)
    with pytest.raises(pkg._lib.OceanError) as e:
        ens.run(oceans, seeds, time=1.0, choppiness=[1.0, 0.5, 1.0, 1.0])
    assert e.value.status == pkg._lib.ERR_INVALID_PARAMS and "lambda" in e.value.message
    # one lambda per wave is fine, and is honoured per wave
    mixed = ens.run(oceans, seeds, time=1.0, choppiness=[1.0, 1.0, 0.5, 0.5])
    ones = ens.run(oceans, seeds, time=1.0, choppiness=1.0)
    halves = ens.run(oceans, seeds, time=1.0, choppiness=0.5)
    assert mixed[:2].tolist() == ones[:2].tolist() and mixed[2:].tolist() == halves[2:].tolist()
    assert mixed[2:].tolist() != ones[2:].tolist()
    ens.close()


@pytest.mark.parametrize("N", [128, 256, 512])  # plain-load columns, fused frame kernel, TMA columns
def test_update_buoyancy_answers_on_the_owning_gpu(pkg, po, N):
    """ocean_update_buoyancy: frame + BO:18-30 over WB:195-209 against the frame's height plane on the copy stream.
    Eight pipelined frames with a ring of host buffers must equal, bit for bit, a twin ocean stepped synchronously
    and asked through ocean_buoyancy(NEAREST_REF); READBACK_REF keeps serving the last shipped layer."""
    cfg = pkg.scene.config(2)
    body, noise = make_body(pkg, po, cfg, N=N)
    twin, _ = make_body(pkg, po, cfg, N=N, noise=noise)
    rng = np.random.default_rng(77)
    n = 4096
    boxes = np.empty((n, 6), np.float32)
    boxes[:, 0] = rng.uniform(-N / 2 - 5, N / 2 + 5, n)
    boxes[:, 2] = rng.uniform(-N / 2 - 5, N / 2 + 5, n)
    boxes[:, 1] = rng.uniform(-2, 2, n)
    boxes[:, 3:] = rng.uniform(0.2, 3.0, (n, 3))
    body.Update(0.0)
    twin.CalculateWavesTexturesAtTime(0.0)
    frame0, layer0 = body.WaitReadback()
    ring = [np.full((n, 4), np.nan, np.float32) for _ in range(3)]
    ins = [boxes.copy() for _ in range(3)]
    want = []
    for f in range(1, 9):
        k = f % 3
        if f > 3:  # the slot's previous answer (frame f - 3) is complete by the contract: check it before reuse
            body.Sync()
            assert np.array_equal(ring[k], want[f - 4]), "frame %d" % (f - 3)
        ins[k][:, 1] = boxes[:, 1] + np.float32(0.01 * f)
        body.UpdateBuoyancy(f / 60.0, ins[k], ring[k], density=0.9, gravityY=-9.81)
        twin.CalculateWavesTexturesAtTime(f / 60.0)
        want.append(twin.Buoyancy(ins[k], pkg._lib.QUERY_NEAREST_REF, density=0.9, gravityY=-9.81))
    body.Sync()
    for f in (6, 7, 8):
        assert np.array_equal(ring[f % 3], want[f - 1])
    assert (want[-1][:, 1] > 0).sum() > 100 and (want[-1][:, 1] == 0).sum() > 100
    assert np.array_equal(body.displacementsTextures, twin.displacementsTextures)
    # no layer was shipped since frame 0: the reference-latency lookup still answers from it
    frame, layer = body.WaitReadback()
    assert frame == frame0 and np.array_equal(layer, layer0)
    body.OnDisable()
    twin.OnDisable()


def test_both_device_noise_kernels_equal_the_host_generator(pkg, po):
    """The default noise kernel balances the WB:71-81 rejection loop over 8 draws per thread; OCEAN_NOISE_BALANCED=0 selects the
    one-draw-per-loop kernel.  Both must reproduce include/ocean_noise.h on the host bit for bit (child processes: the switch is
    read once), at a size with a single group per row block (16), a ragged last block (64 * 64 / 4 groups) and the headline size."""
    import hashlib, os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    child = (
        "import sys, hashlib; sys.path.insert(0, %r); import oceanpkg; pkg = oceanpkg.load()\n"
        "for N, seed in ((16, 5), (64, 1), (512, 12345678901234)):\n"
        "    b = pkg.WaterBody(windSpeed=8, windDirection=(1, -1), fetch=50000, depth=2560, texturesSize=N, cascades=pkg.scene.SCENE_CASCADES[:1], seed=seed).Awake()\n"
        "    print(N, hashlib.sha256(b.randomNoiseTexture.tobytes()).hexdigest()); b.OnDisable()\n"
    ) % root
    want = {N: hashlib.sha256(po.fill_noise(seed, N).tobytes()).hexdigest() for N, seed in ((16, 5), (64, 1), (512, 12345678901234))}
    for env in ({}, {"OCEAN_NOISE_BALANCED": "0"}):
        r = subprocess.run([sys.executable, "-c", child], env=dict(os.environ, **env), capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-1500:]
        got = {int(l.split()[0]): l.split()[1] for l in r.stdout.strip().splitlines()}
        assert got == want, (env, got, want)


def test_two_devices_in_one_process(pkg, po):
    """cudaFuncSetAttribute is per device: every frame path must configure its kernels on each GPU it runs on.  Oceans on cuda:1
    (created first) and cuda:0 in ONE process, at the sizes that select the fused frame kernel (256) and the two-kernel frame
    (512), plus an ensemble with the spectrum fused into the row pass: same bits on both GPUs.  Skipped on one GPU."""
    lib = pkg._lib.load()
    if lib.ocean_device_count() < 2:
        pytest.skip("needs two GPUs")
    cfg = pkg.scene.config(2)
    for N in (256, 512):
        outs = []
        for dev in (1, 0):
            b, _ = make_body(pkg, po, cfg, N=N, device=dev)
            for f in range(3):
                b.CalculateWavesTexturesAtTime(f / 60.0)
            outs.append((b.displacementsTextures, b.turbulenceTextures))
            b.OnDisable()
        assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
    oceans, seeds = pkg.scene.ensemble_states(3)
    import os
    old = os.environ.get("OCEAN_ENSEMBLE_FUSED_INIT")
    os.environ["OCEAN_ENSEMBLE_FUSED_INIT"] = "1"
    try:
        sums = []
        for dev in (1, 0):
            e = pkg.OceanEnsemble(256, pkg.scene.config(5)["cascades"], batch=2, device=dev)
            sums.append(e.run(oceans, seeds, time=1.0).tolist())
            e.close()
        assert sums[0] == sums[1]
    finally:
        if old is None:
            os.environ.pop("OCEAN_ENSEMBLE_FUSED_INIT", None)
        else:
            os.environ["OCEAN_ENSEMBLE_FUSED_INIT"] = old
