"""The in-kernel FFT plan (ocean-simulation_b200/fft_plan.py mirrors csrc/ocean_fft.cuh): exactness against
numpy for every supported size, and shared-memory bank conflicts of the layouts the kernels use."""
import importlib.util
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("fft_plan", os.path.join(ROOT, "ocean-simulation_b200", "fft_plan.py"))
plan = importlib.util.module_from_spec(spec)
spec.loader.exec_module(plan)


@pytest.mark.parametrize("N", [16, 32, 64, 128, 256, 512, 1024, 2048])
@pytest.mark.parametrize("dense0", [False, True])
def test_plan_is_unnormalised_inverse_dft(N, dense0):
    rng = np.random.default_rng(N)
    for W in (1, 2, 8):
        x = rng.standard_normal((W, N)) + 1j * rng.standard_normal((W, N))
        X, _ = plan.run_plan(x, W, dense0=dense0)
        ref = np.fft.ifft(x, axis=1) * N
        assert np.abs(X - ref).max() <= 1e-12 * np.abs(ref).max()


def test_stage_radices():
    assert plan.stage_radices(16) == [8, 2] and plan.stage_radices(256) == [8, 8, 4]
    assert plan.stage_radices(512) == [8, 8, 8] and plan.stage_radices(2048) == [8, 8, 8, 4]


def _worst(N, W, dense0):
    x = np.zeros((W, N), complex)
    _, conf = plan.run_plan(x, W, check_conflicts=True, dense0=dense0)
    return max(w for _, _, w in conf)


def test_headline_layouts_are_bank_conflict_free():
    # row kernel (one line per region == W = 1 model), every exchange swizzled
    for N in (512, 1024, 2048):
        assert _worst(N, 1, False) == 1
    # TMA column kernel at N = 512: tile of W columns lands dense, exchange 0 dense, exchange 1 swizzled
    for W in (2, 4, 8):
        assert _worst(512, W, True) == 1
    assert _worst(256, 8, True) == 1
    # generic column kernel (swizzled everywhere) at its default widths: at most 2-way
    assert _worst(1024, 2, False) <= 2 and _worst(2048, 2, False) <= 2 and _worst(128, 8, False) == 1


def test_column_tile_wavefronts_match_the_ncu_source_page():
    """The conflict model extended from the exchanges to every shared-memory instruction of a column tile.  The figures are
    those of the committed ncu capture, per tile (profiles/r2_ncu_smem_by_instruction.txt: 512 tiles per launch): everything
    is conflict-free except the scalar read of the incoming turbulence .x from the staged RGBA tile (4-way, by layout)."""
    got = plan.column_tile_wavefronts(512, 4)
    ncu_per_launch = {"LDS.128": (786432, 786432), "STS.128": (917504, 917504), "LDS.64": (114688, 114688), "LDS.32": (131072, 32768)}
    assert got == {k: (w // 512, i // 512) for k, (w, i) in ncu_per_launch.items()}
    excess = sum(w - i for w, i in got.values())
    assert excess == 192 == 8 * 8 * 3  # eight foam reads per thread, eight warps per tile, three extra wavefronts each
    for W in (2, 8):  # other tile widths: the transform itself stays conflict-free
        g = plan.column_tile_wavefronts(512, W)
        assert g["LDS.128"][0] == g["LDS.128"][1] and g["STS.128"][0] == g["STS.128"][1]
