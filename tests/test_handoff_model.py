"""The frame kernels' hand-off protocol, checked by exhaustive interleaving search on an abstract machine
(ocean-simulation_b200/handoff_model.py) -- the stand-in for compute-sanitizer's racecheck, which is closed on the GPU
pool (profiles/r2_sanitizer.txt).  CPU only."""
import pytest


@pytest.fixture(scope="module")
def hm(pkg):
    import importlib
    return importlib.import_module("ocean_simulation_b200.handoff_model")


def test_protocol_as_built_has_no_bad_read_and_no_deadlock(hm):
    base = hm.DEFAULT
    assert hm.explore(base) > 100
    # two persistent column CTAs sharing the tiles, and a tighter SM (column CTAs of the next frame wait for room)
    assert hm.explore(base._replace(col_ctas=2, capacity=4)) > 100
    assert hm.explore(base._replace(capacity=2)) > 50
    # two layers: columns of layer 0 start while layer 1's rows are still running
    assert hm.explore(base._replace(layers=2, rows=1, tiles=1, capacity=3)) > 100
    # the largest case: 2 frames x 2 layers x 2 rows x 2 tiles, two persistent column CTAs (~1.2e5 states)
    assert hm.explore(base._replace(layers=2, rows=2, tiles=2, col_ctas=2, capacity=5)) > 100000


def test_three_frames_chain(hm):
    assert hm.explore(hm.DEFAULT._replace(frames=3, rows=1, tiles=1, capacity=3)) > 100
    assert hm.explore(hm.DEFAULT._replace(frames=3, capacity=3)) > 1000
    assert hm.explore(hm.DEFAULT._replace(frames=3, layers=2, rows=1, tiles=1, col_ctas=2, capacity=4)) > 5000


@pytest.mark.parametrize("mutation,needle", [
    ("no_row_fence", "('ws'"),
    ("no_war_wait", "('ws'"),
    ("no_proxy_fence", "without a proxy fence"),
    ("no_col_fence", "('out'"),
    ("no_row_wait", "('ws'"),
])
def test_every_fence_and_wait_is_needed(hm, mutation, needle):
    with pytest.raises(hm.Violation) as e:
        hm.explore(hm.DEFAULT, mutate=mutation)
    assert needle in str(e.value)


def test_spinning_consumers_rely_on_the_programmatic_launch_rule(hm):
    """A column CTA spins for its layer's rows while holding its SM resources.  That is only safe because a dependent
    kernel's CTAs cannot start before EVERY CTA of the row pass has started (all of them are resident or finished)."""
    tight = hm.DEFAULT._replace(capacity=2)
    assert hm.explore(tight) > 50
    with pytest.raises(hm.Violation) as e:
        hm.explore(tight, mutate="no_launch_rule")
    assert "deadlock" in str(e.value)
