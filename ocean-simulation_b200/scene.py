"""Parameter sets of the reference scene and the BASELINE.json configurations.

Numbers come from Assets/Scenes/Waves.unity (SC) in the reference:
  ocean   SC:1305-1314   windSpeed 8, windDirection (1,-1), gravity 9.81, fetch 50000, depth 2560, texturesSize 512
  C0      SC:1431-1435   C1  SC:470-474   C2  SC:1249-1253   C3  SC:1572-1576 (present in the scene, not wired)
"""

# (wavelength, cutoffLow, cutoffHigh, swell, fade)
SCENE_CASCADES = [
    (1530.0, 1e-10, 1e12, 0.4, 0.1),
    (1000.0, 1e-7, 1e7, 0.3, 0.2),
    (201.0, 1e-5, 1e6, 0.1, 0.1),
    (34.0, 1e-3, 10.0, 0.4, 0.1),
]

# (windSpeed, windDirX, windDirY, gravity, fetch, depth)
SCENE_OCEAN = (8.0, 1.0, -1.0, 9.81, 50000.0, 2560.0)


def config(n):
    """BASELINE.json configs[n-1] made concrete: dict(N, ocean, cascades, seed, times)."""
    if n == 1:  # single cascade 256^2, U10=10, fetch=100 km, depth=500 m, one frame
        return dict(N=256, ocean=(10.0, 1.0, -1.0, 9.81, 100000.0, 500.0), cascades=SCENE_CASCADES[:1], seed=1,
                    times=[1.0])
    if n == 2:  # reference default scene at 256^2, 3 cascades, 600 frames of foam accumulation
        return dict(N=256, ocean=SCENE_OCEAN, cascades=SCENE_CASCADES[:3], seed=2, times=[f / 60.0 for f in range(600)])
    if n == 3:  # 4 cascades x 512^2, shallow-water TMA (depth 20 m, swell 0.8) -- the headline
        casc = [(c[0], c[1], c[2], 0.8, c[4]) for c in SCENE_CASCADES]
        return dict(N=512, ocean=(10.0, 1.0, -1.0, 9.81, 100000.0, 20.0), cascades=casc, seed=3,
                    times=[f / 60.0 for f in range(110)])
    if n == 4:  # 4 cascades x 1024^2 + 1M buoyancy queries per frame
        return dict(N=1024, ocean=SCENE_OCEAN, cascades=SCENE_CASCADES, seed=4, times=[f / 60.0 for f in range(20)],
                    query_points=1 << 20, query_seed=40)
    if n == 5:  # ensemble of 4096 states x 3 cascades x 512^2
        return dict(N=512, cascades=SCENE_CASCADES[:3], states=4096, rng_seed=5, time=1.0)
    raise ValueError(n)


def ensemble_states(n_states, rng_seed=5):
    """Randomised ocean states of config 5: U10 ~ U[3,20], wind angle ~ U[0,2pi), fetch ~ logU[1e4,5e5],
    depth ~ logU[10,3000]; noise seed = state index."""
    import numpy as np
    rng = np.random.default_rng(rng_seed)
    U = rng.uniform(3.0, 20.0, n_states)
    ang = rng.uniform(0.0, 2.0 * np.pi, n_states)
    F = np.exp(rng.uniform(np.log(1e4), np.log(5e5), n_states))
    D = np.exp(rng.uniform(np.log(10.0), np.log(3000.0), n_states))
    oceans = np.stack([U, np.cos(ang), np.sin(ang), np.full(n_states, 9.81), F, D], axis=1).astype(np.float32)
    seeds = np.arange(n_states, dtype=np.uint64)
    return oceans, seeds
