"""Host-side mirror of the reference's WaterBody / WaterCascade surface over liboceanfft.so.

Same member names, argument meaning and call order as Assets/Scripts/Water/WaterBody.cs (WB) and
WaterCascade.cs (WC) so that tests read like the reference's own usage:

    body = WaterBody(windSpeed=8, windDirection=(1, -1), fetch=50000, depth=2560,
                     texturesSize=512, cascades=[WaterCascade(1530, 1e12, 1e-10, 0.4, 0.1), ...])
    body.Awake()                                  # WB:211
    body.CalculateWavesTexturesAtTime(t)          # WB:180
    body.Update(t)                                # WB:284  (step + async readback of layer 0)
    h = body.GetWaterHeight((x, y, z))            # WB:195

All compute happens in the CUDA library; this file only marshals arguments (numpy <-> ctypes).
"""
import ctypes

import numpy as np

from . import _lib as L


class WaterCascade:
    """WC:10-24 with the reference's defaults."""

    def __init__(self, wavelength=10.0, cutoffHigh=5.0, cutoffLow=0.0001, swell=0.4, fade=0.1):
        self.wavelength = float(wavelength)
        self.cutoffHigh = float(cutoffHigh)
        self.cutoffLow = float(cutoffLow)
        self.swell = float(swell)
        self.fade = float(fade)

    @classmethod
    def from_tuple(cls, t):
        """(wavelength, cutoffLow, cutoffHigh, swell, fade) -- the order of scene.SCENE_CASCADES."""
        return cls(wavelength=t[0], cutoffLow=t[1], cutoffHigh=t[2], swell=t[3], fade=t[4])

    def _c(self):
        return L.OceanCascade(self.wavelength, self.cutoffLow, self.cutoffHigh, self.swell, self.fade)


def _cascade_array(cascades):
    cs = [c if isinstance(c, WaterCascade) else WaterCascade.from_tuple(c) for c in cascades]
    arr = (L.OceanCascade * max(len(cs), 1))()
    for i, c in enumerate(cs):
        arr[i] = c._c()
    return cs, arr


class WaterBody:
    """WB:7-56.  Defaults are the reference's field initialisers (WB:10-14, WB:29)."""

    def __init__(self, windSpeed=1.0, windDirection=(1.0, 1.0), gravity=9.81, fetch=1.0, depth=4.0,
                 texturesSize=256, cascades=(), noise=None, seed=0, device=0, choppiness=1.0):
        self.windSpeed = float(windSpeed)
        self.windDirection = (float(windDirection[0]), float(windDirection[1]))
        self.gravity = float(gravity)
        self.fetch = float(fetch)
        self.depth = float(depth)
        self.texturesSize = int(texturesSize)
        self.cascades = [c if isinstance(c, WaterCascade) else WaterCascade.from_tuple(c) for c in cascades]
        self.choppiness = float(choppiness)  # lambda; the reference is hard-wired to 1
        self.noise = noise                   # optional [N][N][2] float32 (WB:86-100), else generated from seed
        self.seed = int(seed)
        self.device = int(device)
        self._h = None
        self._lib = L.load()

    # -- helpers ---------------------------------------------------------------------------
    def _params(self):
        return L.OceanParams(self.windSpeed, self.windDirection[0], self.windDirection[1], self.gravity, self.fetch,
                             self.depth, self.texturesSize, self.choppiness)

    def _need(self):
        if self._h is None:
            raise RuntimeError("WaterBody.Awake() has not been called")
        return self._h

    # -- Unity messages --------------------------------------------------------------------
    def Awake(self):
        """WB:211-256: allocate, upload/generate noise, compute the initial spectrum."""
        if self._h is not None:
            return self
        p = self._params()
        _, arr = _cascade_array(self.cascades)
        noise_ptr = None
        if self.noise is not None:
            n = np.ascontiguousarray(self.noise, np.float32)
            if n.shape != (self.texturesSize, self.texturesSize, 2):
                raise ValueError("noise must be [N][N][2]")
            self._noise_keep = n
            noise_ptr = n.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
        h = ctypes.c_void_p()
        L.check(self._lib.ocean_create(ctypes.byref(p), arr, len(self.cascades), noise_ptr, self.seed, self.device,
                                       ctypes.byref(h)))
        self._h = h
        self.CalculateInitialSpectrumTextures()
        return self

    def OnDisable(self):
        """WB:300-309."""
        if self._h is not None:
            self._lib.ocean_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.OnDisable()
        except Exception:
            pass

    def __enter__(self):
        return self.Awake()

    def __exit__(self, *a):
        self.OnDisable()

    def SetParams(self):
        """Stages the current field values (ocean_set_params); they take effect at the next
        CalculateInitialSpectrumTextures -- frames issued in between keep the active set."""
        p = self._params()
        _, arr = _cascade_array(self.cascades)
        L.check(self._lib.ocean_set_params(self._need(), ctypes.byref(p), arr, len(self.cascades)))

    def OnValidate(self):
        """The reference's (commented-out) live parameter edit WB:324-337: push the current field
        values and recompute the initial spectrum."""
        self.SetParams()
        self.CalculateInitialSpectrumTextures()

    # -- the path --------------------------------------------------------------------------
    def CalculateInitialSpectrumTextures(self):
        """WB:171-178."""
        L.check(self._lib.ocean_init_spectrum(self._need()))

    def CalculateWavesTexturesAtTime(self, time):
        """WB:180-193.  Asynchronous; Sync() or any array read waits for it."""
        L.check(self._lib.ocean_step(self._need(), float(time)))

    def StepPartial(self, time, which):
        """Measurement hook: which = 1 row pass only, 2 column pass + assembly only, 0 both."""
        L.check(self._lib.ocean_step_partial(self._need(), float(time), int(which)))

    def Update(self, time):
        """WB:284-297: one frame + AsyncGPUReadback request of displacement layer 0."""
        L.check(self._lib.ocean_update(self._need(), float(time)))

    def Sync(self):
        L.check(self._lib.ocean_sync(self._need()))

    def GetWaterHeight(self, worldPosition):
        """WB:195-209: (x, y, z) -> height from the last completed readback; 0 until one exists."""
        pts = np.array([[worldPosition[0], worldPosition[2]]], np.float32)
        return float(self.GetWaterHeights(pts, L.QUERY_READBACK_REF)[0])

    def GetWaterHeights(self, worldXZ, mode=L.QUERY_NEAREST_REF):
        """Batched query: worldXZ [n][2] (x, z) -> heights [n]."""
        pts = np.ascontiguousarray(worldXZ, np.float32).reshape(-1, 2)
        out = np.empty(pts.shape[0], np.float32)
        L.check(self._lib.ocean_query_heights(self._need(), pts.ctypes.data, pts.shape[0], int(mode), out.ctypes.data, 0, 0))
        return out

    # -- consumers of the arrays -----------------------------------------------
    def GenerateMips(self, auto=None):
        """derivativesTextures.GenerateMips(); turbulenceTextures.GenerateMips()  (WB:191-192).
        auto=True/False switches the per-frame rebuild the reference does on or off."""
        if auto is not None:
            L.check(self._lib.ocean_set_auto_mips(self._need(), 1 if auto else 0))
        L.check(self._lib.ocean_generate_mips(self._need()))

    @property
    def mipLevels(self):
        return self._lib.ocean_mip_levels(self._need())

    def Mip(self, which, level, cascade):
        """One level of one layer of a mipped array: [(N>>level)][(N>>level)][4]."""
        n = self.texturesSize >> level
        out = np.empty((n, n, 4), np.float32)
        L.check(self._lib.ocean_copy_mip(self._need(), int(which), int(level), int(cascade),
                                         out.ctypes.data_as(ctypes.POINTER(ctypes.c_float))))
        return out

    def SampleSurface(self, worldXZ, lod=0.0, displacement=True, normal=True, foam=True):
        """Water.shader's Domain (WS:322-327) and Fragment (WS:337-347) arithmetic at world positions
        [n][2]: returns (displacement [n][3] | None, normal [n][3] | None, foam [n] | None)."""
        pts = np.ascontiguousarray(worldXZ, np.float32).reshape(-1, 2)
        n = pts.shape[0]
        d = np.empty((n, 3), np.float32) if displacement else None
        nr = np.empty((n, 3), np.float32) if normal else None
        f = np.empty(n, np.float32) if foam else None
        ptr = lambda a: a.ctypes.data if a is not None else None
        L.check(self._lib.ocean_sample_surface(self._need(), pts.ctypes.data, n, float(lod), ptr(d), ptr(nr), ptr(f), 0))
        return d, nr, f

    def Buoyancy(self, boxes, mode=L.QUERY_NEAREST_REF, density=1.0, gravityY=-9.81):
        """BuoyantObject.FixedUpdate BO:18-30 for a batch: boxes [n][6] = (position xyz, localScale xyz) ->
        [n][4] = (waterHeight, heightSubmerged, upward acceleration, submergedVolume).  density = WaterBody.density WB:18."""
        b = np.ascontiguousarray(boxes, np.float32).reshape(-1, 6)
        out = np.empty((b.shape[0], 4), np.float32)
        L.check(self._lib.ocean_buoyancy(self._need(), b.ctypes.data, b.shape[0], int(mode), float(density), float(gravityY),
                                         out.ctypes.data, 0, 0))
        return out

    def BuoyancyNoAlloc(self, boxes, mode=L.QUERY_NEAREST_REF, density=1.0, gravityY=-9.81):
        """Buoyancy() through ocean_buoyancy_noalloc: same results, no allocation inside the library (scratch in a static
        device array; at most L.NOALLOC_MAX_BOXES boxes per call; calls that overlap on the GPU race on that array)."""
        b = np.ascontiguousarray(boxes, np.float32).reshape(-1, 6)
        out = np.empty((b.shape[0], 4), np.float32)
        L.check(self._lib.ocean_buoyancy_noalloc(self._need(), b.ctypes.data, b.shape[0], int(mode), float(density), float(gravityY),
                                                 out.ctypes.data, 0, 0))
        return out

    def UpdateBuoyancy(self, time, boxes, out, density=1.0, gravityY=-9.81):
        """WaterBody.Update + BuoyantObject.FixedUpdate with the lookups answered on the owning GPU (ocean_update_buoyancy).
        boxes: float32 [n][6] and out: float32 [n][4], caller-owned (pinned for overlap) and untouched until Sync().
        Pass numpy arrays or (data_ptr, n) pairs of pinned torch tensors."""
        if isinstance(boxes, tuple):
            bptr, n = boxes
            optr = out[0] if isinstance(out, tuple) else out.ctypes.data
        else:
            assert boxes.dtype == np.float32 and boxes.flags.c_contiguous and out.dtype == np.float32 and out.flags.c_contiguous
            bptr, n, optr = boxes.ctypes.data, boxes.shape[0], out.ctypes.data
        L.check(self._lib.ocean_update_buoyancy(self._need(), float(time), bptr, int(n), float(density), float(gravityY), optr))

    @property
    def copyStream(self):
        s = ctypes.c_void_p()
        L.check(self._lib.ocean_get_copy_stream(self._need(), ctypes.byref(s)))
        return s.value

    def SetReadbackMode(self, mode):
        """L.READBACK_RGBA (the reference's whole-layer request, default) or L.READBACK_HEIGHT (packed .g only)."""
        L.check(self._lib.ocean_set_readback_mode(self._need(), int(mode)))

    def WaitReadback(self):
        """Blocks for the newest AsyncGPUReadback; returns (frame, array) or (-1, None).  The array is
        [N][N][4] in RGBA mode and [N][N] (heights) in HEIGHT mode."""
        fr = ctypes.c_int64(-1)
        p = self._lib.ocean_readback_ptr(self._need(), 1, ctypes.byref(fr))
        if not p:
            return -1, None
        N = self.texturesSize
        packed = self._lib.ocean_readback_mode_of_latest(self._need()) == L.READBACK_HEIGHT
        return fr.value, np.ctypeslib.as_array(p, shape=(N, N) if packed else (N, N, 4)).copy()

    # -- arrays ----------------------------------------------------------------------------
    def _array(self, which, cascade=-1):
        N, C = self.texturesSize, len(self.cascades)
        if which == L.NOISE:
            out = np.empty((N, N, 2), np.float32)
        elif which == L.ROW_SPECTRA:
            out = np.empty((2, C, N, N, 4) if cascade < 0 else (2, N, N, 4), np.float32)
        else:
            out = np.empty((C, N, N, 4) if cascade < 0 else (N, N, 4), np.float32)
        L.check(self._lib.ocean_copy_output(self._need(), which, cascade, out.ctypes.data_as(ctypes.POINTER(ctypes.c_float))))
        return out

    @property
    def displacementsTextures(self):   # WB:44
        return self._array(L.DISPLACEMENT)

    @property
    def derivativesTextures(self):     # WB:45
        return self._array(L.DERIVATIVES)

    @property
    def turbulenceTextures(self):      # WB:46
        return self._array(L.TURBULENCE)

    @property
    def initialSpectrumTextures(self):  # WB:37
        return self._array(L.INITIAL_SPECTRUM)

    @property
    def wavesDataTextures(self):       # WB:38
        return self._array(L.WAVES_DATA)

    @property
    def randomNoiseTexture(self):      # WB:34
        return self._array(L.NOISE)

    @property
    def rowSpectra(self):
        """Debug: state between the x and y passes: [2 planes][C][N][N][4]."""
        return self._array(L.ROW_SPECTRA)

    @property
    def wavelengths(self):             # WB:49, published as _Wavelengths WB:281
        out = (ctypes.c_float * L.OCEAN_MAX_CASCADES)()
        n = ctypes.c_int(0)
        L.check(self._lib.ocean_get_wavelengths(self._need(), out, ctypes.byref(n)))
        return [out[i] for i in range(n.value)]

    def device_pointer(self, which):
        p, b = ctypes.c_void_p(), ctypes.c_size_t()
        L.check(self._lib.ocean_get_output(self._need(), which, ctypes.byref(p), ctypes.byref(b)))
        return p.value, b.value

    def SetTurbulence(self, arr):
        a = np.ascontiguousarray(arr, np.float32)
        L.check(self._lib.ocean_set_turbulence(self._need(), a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))))

    def ResetFoam(self):
        L.check(self._lib.ocean_reset_foam(self._need()))

    @property
    def frame(self):
        return self._lib.ocean_frame_index(self._need())

    @property
    def stream(self):
        s = ctypes.c_void_p()
        L.check(self._lib.ocean_get_stream(self._need(), ctypes.byref(s)))
        return s.value

    def SetStream(self, stream):
        L.check(self._lib.ocean_set_stream(self._need(), ctypes.c_void_p(stream)))

    def describe(self):
        buf = ctypes.create_string_buffer(1024)
        L.check(self._lib.ocean_describe(self._need(), buf, 1024))
        return buf.value.decode()


class OceanEnsemble:
    """Pooled runner for many independent oceans (BASELINE config 5): ocean_ensemble_* ABI."""

    def __init__(self, texturesSize, cascades, batch=8, device=0):
        self._lib = L.load()
        self.texturesSize = int(texturesSize)
        self.cascades, arr = _cascade_array(cascades)
        self.batch = int(batch)
        h = ctypes.c_void_p()
        L.check(self._lib.ocean_ensemble_create(self.texturesSize, arr, len(self.cascades), self.batch, int(device), ctypes.byref(h)))
        self._h = h

    def run(self, oceans, seeds=None, time=1.0, choppiness=1.0):
        """oceans: [n][6] (windSpeed, windDirX, windDirY, gravity, fetch, depth).  Returns uint64 [n][3]
        checksums of (displacement, derivatives, turbulence)."""
        oceans = np.asarray(oceans, np.float32).reshape(-1, 6)
        n = oceans.shape[0]
        lam = np.broadcast_to(np.asarray(choppiness, np.float32), (n,))  # a scalar, or one lambda per state
        ps = (L.OceanParams * max(n, 1))()
        for i in range(n):
            o = oceans[i]
            ps[i] = L.OceanParams(o[0], o[1], o[2], o[3], o[4], o[5], self.texturesSize, float(lam[i]))
        sd = None
        if seeds is not None:
            s = np.ascontiguousarray(seeds, np.uint64)
            sd = s.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64))
        out = np.zeros((n, 3), np.uint64)
        L.check(self._lib.ocean_ensemble_run(self._h, ps, sd, n, float(time), out.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64))))
        return out

    def copy_output(self, which, slot):
        N, C = self.texturesSize, len(self.cascades)
        out = np.empty((C, N, N, 4), np.float32)
        L.check(self._lib.ocean_ensemble_copy_output(self._h, which, slot, out.ctypes.data_as(ctypes.POINTER(ctypes.c_float))))
        return out

    def close(self):
        if self._h is not None:
            self._lib.ocean_ensemble_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
