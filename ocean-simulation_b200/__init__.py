"""ocean-simulation_b200: B200-native (sm_100a CUDA) spectral-ocean hot path behind WaterBody's surface.

The directory name carries a hyphen (it is the graft's package name), so import it through the
repo-root helper:  `import oceanpkg; pkg = oceanpkg.load()`  ->  pkg.WaterBody, pkg.WaterCascade.
"""
from . import _lib
from . import scene
from . import sharding
from .waterbody import OceanEnsemble, WaterBody, WaterCascade

__all__ = ["WaterBody", "WaterCascade", "OceanEnsemble", "scene", "sharding", "_lib"]
