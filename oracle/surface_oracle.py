"""CPU oracle for the consumers of the three output arrays.  TEST INFRASTRUCTURE ONLY: only tests/
and __graft_entry__.smoke() may import this module; the product package never does.

numpy fp32 restatement, operation for operation, of
  - the sampler the reference configures on its RenderTextures (WaterBody.cs:104-128: Repeat wrap,
    FilterMode.Trilinear): texel centres at (i + 0.5)/n, bilinear in a level, linear between levels;
  - derivativesTextures.GenerateMips() / turbulenceTextures.GenerateMips() (WaterBody.cs:191-192);
  - Water.shader Domain, lines 322-327 (displacement summed over cascades at worldUV / _Wavelengths[i]);
  - Water.shader Fragment, lines 337-347 (derivatives and turbulence summed over cascades, slope, normal);
  - the height lookup behind BuoyantObject.cs:18 -> WaterBody.GetWaterHeight (WaterBody.cs:195-209), extended
    to invert the horizontal displacement of Water.shader:327.

PIN: the arithmetic of Water.shader:322-326 / 337-347, WaterBody.GetWaterHeight (WaterBody.cs:195-209) and
BuoyantObject.FixedUpdate (BuoyantObject.cs:17-37) is pinned to the reference's own text: `make -C oracle ref`
streams those lines through sed into g++ (oracle/surface_ref_host.cpp supplies the types) and
tests/test_surface_oracle.py asserts this module equals that library bit for bit.  What stays UNPINNED, by
nature: the texture sampler underneath (Unity / driver behaviour: restated here and in surface_ref_host.cpp as
Repeat + trilinear with texel centres at (i + 0.5)/n) and Unity's mip filter, which is not part of the reference
tree, so the chain is DEFINED as the 2x2 box filter; the displaced-height inversion is this library's own
extension (checked against Newton in fp64).  The CUDA kernels (compiled without FMA contraction) are compared
with this module bit for bit.
"""
import numpy as np

F = np.float32


def mip_chain(level0):
    """level0 [n][n][ch] -> [level0, level1, ...] down to 1x1; dst = ((a + b) + (c + d)) * 0.25."""
    levels = [np.ascontiguousarray(level0, F)]
    while levels[-1].shape[0] > 1:
        s = levels[-1]
        a, b = s[0::2, 0::2], s[0::2, 1::2]
        c, d = s[1::2, 0::2], s[1::2, 1::2]
        levels.append((((a + b) + (c + d)) * F(0.25)).astype(F))
    return levels


def _lerp(a, b, t):
    return (a + (b - a) * t).astype(F)


def sample_bilinear(img, u, v):
    """img [n][n][ch] fp32; u, v fp32 arrays in [0, 1).  Repeat wrap, texel centres at (i + 0.5)/n."""
    n = img.shape[0]
    u, v = np.asarray(u, F), np.asarray(v, F)
    fx = (u * F(n) - F(0.5)).astype(F)
    fy = (v * F(n) - F(0.5)).astype(F)
    x0f, y0f = np.floor(fx), np.floor(fy)
    wx, wy = (fx - x0f).astype(F)[:, None], (fy - y0f).astype(F)[:, None]
    x0, y0 = x0f.astype(np.int64) & (n - 1), y0f.astype(np.int64) & (n - 1)
    x1, y1 = (x0 + 1) & (n - 1), (y0 + 1) & (n - 1)
    top = _lerp(img[y0, x0], img[y0, x1], wx)
    bot = _lerp(img[y1, x0], img[y1, x1], wx)
    return _lerp(top, bot, wy)


def sample_lod(levels, u, v, lod):
    """SAMPLE_TEXTURE2D_ARRAY_LOD on one layer.  levels: [level0, level1, ...] (a list of length 1 is a
    texture without a chain: every LOD reads level 0)."""
    u, v = np.asarray(u, F), np.asarray(v, F)
    u = (u - np.floor(u)).astype(F)
    v = (v - np.floor(v)).astype(F)
    top = len(levels) - 1
    lod = F(lod)
    if top == 0 or not lod > 0:
        return sample_bilinear(levels[0], u, v)
    lod = min(lod, F(top))
    l0 = int(lod)
    t = F(lod - F(l0))
    s0 = sample_bilinear(levels[l0], u, v)
    if t == 0 or l0 >= top:
        return s0
    return _lerp(s0, sample_bilinear(levels[l0 + 1], u, v), t)


def surface(disp, deriv, turb, wavelengths, world_xz, lod=0.0, mips=True):
    """disp/deriv/turb [C][N][N][4]; world_xz [n][2].  Returns (displacement [n][3], normal [n][3], foam [n]).
    Water.shader:322-327 and 337-347.  The displacement array has no chain (WaterBody.cs:227)."""
    p = np.asarray(world_xz, F).reshape(-1, 2)
    n = p.shape[0]
    d = np.zeros((n, 4), F)
    dv = np.zeros((n, 4), F)
    foam = np.zeros(n, F)
    for c in range(disp.shape[0]):
        wl = F(wavelengths[c])
        u, v = (p[:, 0] / wl).astype(F), (p[:, 1] / wl).astype(F)  # worldUV / _Wavelengths[i]
        d = (d + sample_lod([np.asarray(disp[c], F)], u, v, lod)).astype(F)
        dl = mip_chain(deriv[c]) if mips else [np.asarray(deriv[c], F)]
        tl = mip_chain(turb[c]) if mips else [np.asarray(turb[c], F)]
        dv = (dv + sample_lod(dl, u, v, lod)).astype(F)
        t = sample_lod(tl, u, v, lod)[:, 0]
        foam = (foam + (F(1) - np.clip(t, F(0), F(1)))).astype(F)  # turbulence += 1 - saturate(.x)
    sx = (dv[:, 0] / (F(1) + dv[:, 2])).astype(F)  # slope = (d.x / (1 + d.z), d.y / (1 + d.w))
    sz = (dv[:, 1] / (F(1) + dv[:, 3])).astype(F)
    inv = (F(1) / np.sqrt(((sx * sx).astype(F) + F(1)).astype(F) + (sz * sz).astype(F), dtype=F)).astype(F)
    normal = np.stack([(-sx * inv).astype(F), inv, (-sz * inv).astype(F)], axis=1)  # normalize(-slope.x, 1, -slope.y)
    return d[:, :3].copy(), normal, foam


def displaced_heights(disp, wavelengths, world_xz, iterations=4, return_q=False):
    """Height of the surface above (x, z): q <- (x, z) - Dxz(q), `iterations` times, then Dy(q)."""
    p = np.asarray(world_xz, F).reshape(-1, 2)
    qx, qz = p[:, 0].copy(), p[:, 1].copy()
    height = np.zeros(p.shape[0], F)
    for _ in range(iterations + 1):
        d = np.zeros((p.shape[0], 4), F)
        for c in range(disp.shape[0]):
            wl = F(wavelengths[c])
            d = (d + sample_lod([np.asarray(disp[c], F)], (qx / wl).astype(F), (qz / wl).astype(F), 0.0)).astype(F)
        height = d[:, 1].copy()
        lastq = (qx, qz)
        qx, qz = (p[:, 0] - d[:, 0]).astype(F), (p[:, 1] - d[:, 2]).astype(F)
    return (height, lastq) if return_q else height


def buoyancy(water_height, boxes, density=1.0, gravity_y=-9.81):
    """BuoyantObject.FixedUpdate BO:18-30 in fp32, operation for operation: water_height [n] (WaterBody.GetWaterHeight at each
    box position, BO:18), boxes [n][6] = (transform.position xyz, transform.localScale xyz) ->
    [n][4] = (waterHeight, heightSubmerged, upward acceleration |g.y| * density * submergedVolume, submergedVolume)."""
    f = np.float32
    wh = np.asarray(water_height, f)
    b = np.asarray(boxes, f).reshape(-1, 6)
    py, sx, sy, sz = b[:, 1], b[:, 3], b[:, 4], b[:, 5]
    sub = np.maximum(f(0), wh - py)                                     # BO:19
    with np.errstate(divide="ignore", invalid="ignore"):
        full = (sx * sy) * sz                                           # BO:24
        vol = np.clip(sub / sy, f(0), f(1)) * full                      # BO:25
        acc = f(abs(gravity_y)) * (f(density) * vol)                    # BO:26, BO:29
    wet = sub > 0                                                       # BO:21
    vol = np.where(wet, vol, f(0)).astype(f)
    acc = np.where(wet, acc, f(0)).astype(f)
    return np.stack([wh, sub.astype(f), acc, vol], axis=1)
