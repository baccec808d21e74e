"""ctypes binding of liboceanfft.so (include/oceanfft.h).  Loads the in-tree CUDA library or
fails loudly: there is no CPU or PyTorch fallback for any entry point."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("OCEANFFT_LIB") or os.path.join(HERE, "liboceanfft.so")  # env override: tuning variants

OCEAN_MAX_CASCADES = 5

# OceanStatus
OK, ERR_INVALID_ARGUMENT, ERR_INVALID_SIZE, ERR_INVALID_CASCADES, ERR_INVALID_PARAMS, ERR_CUDA, ERR_NOT_INITIALIZED = range(7)
# OceanArray
DISPLACEMENT, DERIVATIVES, TURBULENCE, INITIAL_SPECTRUM, WAVES_DATA, NOISE, ROW_SPECTRA = range(7)
# OceanReadbackMode
READBACK_RGBA, READBACK_HEIGHT = range(2)
# OceanQueryMode
QUERY_NEAREST_REF, QUERY_READBACK_REF, QUERY_BILINEAR, QUERY_BILINEAR_ALL_CASCADES, QUERY_DISPLACED = range(5)
DISPLACED_ITERATIONS = 4
NOALLOC_MAX_BOXES = 65536  # OCEAN_NOALLOC_MAX_BOXES

# every symbol include/oceanfft.h declares (tests check the library exports all of them)
ABI_SYMBOLS = [
    "ocean_abi_version", "ocean_last_error", "ocean_device_count", "ocean_create", "ocean_destroy",
    "ocean_set_params", "ocean_init_spectrum", "ocean_step", "ocean_step_partial", "ocean_sync", "ocean_update",
    "ocean_readback_ptr", "ocean_set_readback_mode", "ocean_readback_mode_of_latest", "ocean_get_output", "ocean_copy_output", "ocean_get_wavelengths",
    "ocean_set_turbulence", "ocean_reset_foam", "ocean_frame_index", "ocean_query_heights",
    "ocean_get_stream", "ocean_set_stream", "ocean_launch_count", "ocean_describe",
    "ocean_ensemble_create", "ocean_ensemble_destroy", "ocean_ensemble_run",
    "ocean_ensemble_get_output", "ocean_ensemble_copy_output",
    "ocean_generate_mips", "ocean_set_auto_mips", "ocean_mip_levels", "ocean_get_mips", "ocean_copy_mip",
    "ocean_sample_surface", "ocean_buoyancy", "ocean_update_buoyancy", "ocean_get_copy_stream",
    "ocean_buoyancy_noalloc",
]


class OceanParams(ctypes.Structure):
    _fields_ = [("windSpeed", ctypes.c_float), ("windDirX", ctypes.c_float), ("windDirY", ctypes.c_float),
                ("gravity", ctypes.c_float), ("fetch", ctypes.c_float), ("depth", ctypes.c_float),
                ("textureSize", ctypes.c_int32), ("lambda_", ctypes.c_float)]


class OceanCascade(ctypes.Structure):
    _fields_ = [("wavelength", ctypes.c_float), ("cutoffLow", ctypes.c_float), ("cutoffHigh", ctypes.c_float),
                ("swell", ctypes.c_float), ("fade", ctypes.c_float)]


class OceanBuoyantBox(ctypes.Structure):
    _fields_ = [("px", ctypes.c_float), ("py", ctypes.c_float), ("pz", ctypes.c_float),
                ("sx", ctypes.c_float), ("sy", ctypes.c_float), ("sz", ctypes.c_float)]


class OceanError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("oceanfft status %d: %s" % (status, message))
        self.status = status
        self.message = message


_lib = None


def load():
    """Returns the loaded library.  Raises if it has not been built (run build.py / __graft_entry__.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("liboceanfft.so is missing at %s: build it with `python ocean-simulation_b200/build.py` "
                          "(nvcc, sm_100a).  There is no fallback implementation." % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    vp, fp, ip = ctypes.c_void_p, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int)
    L.ocean_abi_version.restype = ctypes.c_int
    L.ocean_last_error.restype = ctypes.c_char_p
    L.ocean_device_count.restype = ctypes.c_int
    L.ocean_create.argtypes = [ctypes.POINTER(OceanParams), ctypes.POINTER(OceanCascade), ctypes.c_int, fp,
                               ctypes.c_uint64, ctypes.c_int, ctypes.POINTER(vp)]
    L.ocean_destroy.argtypes = [vp]
    L.ocean_destroy.restype = None
    L.ocean_set_params.argtypes = [vp, ctypes.POINTER(OceanParams), ctypes.POINTER(OceanCascade), ctypes.c_int]
    L.ocean_init_spectrum.argtypes = [vp]
    L.ocean_step.argtypes = [vp, ctypes.c_float]
    L.ocean_step_partial.argtypes = [vp, ctypes.c_float, ctypes.c_int]
    L.ocean_sync.argtypes = [vp]
    L.ocean_update.argtypes = [vp, ctypes.c_float]
    L.ocean_readback_ptr.argtypes = [vp, ctypes.c_int, ctypes.POINTER(ctypes.c_int64)]
    L.ocean_readback_ptr.restype = fp
    L.ocean_set_readback_mode.argtypes = [vp, ctypes.c_int]
    L.ocean_readback_mode_of_latest.argtypes = [vp]
    L.ocean_get_output.argtypes = [vp, ctypes.c_int, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
    L.ocean_copy_output.argtypes = [vp, ctypes.c_int, ctypes.c_int, fp]
    L.ocean_get_wavelengths.argtypes = [vp, fp, ip]
    L.ocean_set_turbulence.argtypes = [vp, fp]
    L.ocean_reset_foam.argtypes = [vp]
    L.ocean_frame_index.argtypes = [vp]
    L.ocean_frame_index.restype = ctypes.c_int64
    L.ocean_query_heights.argtypes = [vp, vp, ctypes.c_int, ctypes.c_int, vp, ctypes.c_int, ctypes.c_int]
    L.ocean_get_stream.argtypes = [vp, ctypes.POINTER(vp)]
    L.ocean_set_stream.argtypes = [vp, vp]
    L.ocean_launch_count.restype = ctypes.c_uint64
    L.ocean_describe.argtypes = [vp, ctypes.c_char_p, ctypes.c_size_t]
    L.ocean_ensemble_create.argtypes = [ctypes.c_int, ctypes.POINTER(OceanCascade), ctypes.c_int, ctypes.c_int,
                                        ctypes.c_int, ctypes.POINTER(vp)]
    L.ocean_ensemble_destroy.argtypes = [vp]
    L.ocean_ensemble_destroy.restype = None
    L.ocean_ensemble_run.argtypes = [vp, ctypes.POINTER(OceanParams), ctypes.POINTER(ctypes.c_uint64), ctypes.c_int,
                                     ctypes.c_float, ctypes.POINTER(ctypes.c_uint64)]
    L.ocean_ensemble_get_output.argtypes = [vp, ctypes.c_int, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
    L.ocean_ensemble_copy_output.argtypes = [vp, ctypes.c_int, ctypes.c_int, fp]
    L.ocean_generate_mips.argtypes = [vp]
    L.ocean_set_auto_mips.argtypes = [vp, ctypes.c_int]
    L.ocean_mip_levels.argtypes = [vp]
    L.ocean_get_mips.argtypes = [vp, ctypes.c_int, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
    L.ocean_copy_mip.argtypes = [vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, fp]
    L.ocean_sample_surface.argtypes = [vp, vp, ctypes.c_int, ctypes.c_float, vp, vp, vp, ctypes.c_int]
    L.ocean_buoyancy.argtypes = [vp, vp, ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_float, vp, ctypes.c_int, ctypes.c_int]
    L.ocean_update_buoyancy.argtypes = [vp, ctypes.c_float, vp, ctypes.c_int, ctypes.c_float, ctypes.c_float, vp]
    L.ocean_get_copy_stream.argtypes = [vp, ctypes.POINTER(vp)]
    L.ocean_buoyancy_noalloc.argtypes = [vp, vp, ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_float, vp, ctypes.c_int, ctypes.c_int]
    _lib = L
    return L


def check(status):
    if status != OK:
        raise OceanError(status, load().ocean_last_error().decode("utf-8", "replace"))
