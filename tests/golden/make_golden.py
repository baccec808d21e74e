"""Generates tests/golden/*.npz from Oracle A -- the reference's own .compute shader text compiled as C++
(oracle/_ref, built by `make -C oracle ref` where /root/reference is visible) -- so the fixtures are
outputs of the reference itself run here.  Small cases only (N = 32, 64); larger configurations are
pinned by 64-bit checksums of the same run.

    python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po  # noqa: E402
import oceanpkg  # noqa: E402

scene = oceanpkg.load().scene


def checksum(a):
    return int(np.ascontiguousarray(a, np.float32).view(np.uint32).astype(np.uint64).sum())


def run(N, ocean, cascades, seed, times):
    noise = po.fill_noise(seed, N)
    o = po.OracleA(N, ocean, cascades, noise)
    o.init_spectrum()
    for t in times:
        o.step(t)
    out = {k: o.get(w) for k, w in (("h0", po.H0), ("wave", po.WAVE), ("disp", po.DISP), ("deriv", po.DERIV), ("turb", po.TURB))}
    out["bfly"] = o.get(po.BFLY)
    o.close()
    return noise, out


def main():
    po.build(ref=True)
    assert po.have_ref(), "Oracle A needs /root/reference"
    cases = {
        "scene_n32": dict(N=32, ocean=scene.SCENE_OCEAN, cascades=scene.SCENE_CASCADES[:3], seed=2, times=[0.0, 1 / 60, 2 / 60]),
        "shallow_n64": dict(N=64, ocean=scene.config(3)["ocean"], cascades=scene.config(3)["cascades"], seed=3, times=[0.5]),
    }
    sums = {}
    for name, c in cases.items():
        noise, out = run(**c)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), noise=noise, ocean=np.array(c["ocean"], np.float32),
                            cascades=np.array(c["cascades"], np.float32), times=np.array(c["times"], np.float64), **out)
        print(name, {k: v.shape for k, v in out.items()})
    # full-size configurations: checksums of the reference's outputs
    for n, frames in ((1, 1), (2, 5), (3, 2)):
        cfg = scene.config(n)
        times = cfg["times"][:frames] if n != 1 else cfg["times"]
        _, out = run(cfg["N"], cfg["ocean"], cfg["cascades"], cfg["seed"], times)
        sums["config%d" % n] = dict(frames=len(times), **{k: checksum(out[k]) for k in ("h0", "wave", "disp", "deriv", "turb")})
        print("config", n, sums["config%d" % n])
    with open(os.path.join(HERE, "checksums.json"), "w") as f:
        json.dump(sums, f, indent=1)


if __name__ == "__main__":
    main()
