"""B200-native Pixel simulator for Spatial Model Editor (one hot path).

``PixelSim`` is the drop-in for ``sme::simulate::PixelSim`` / ``CudaPixelSim``
behind the C-ABI of ``include/smeb200.h``; the rest of the package builds the
flat problem description the C-ABI consumes.
"""
from .problem import (RK101, RK212, RK323, RK435, DBL_MAX, STAGES, Compartment,
                      Membrane, Problem)
from .pixelsim import PixelSim, device_count, load_library

__all__ = ["RK101", "RK212", "RK323", "RK435", "DBL_MAX", "STAGES", "Compartment",
           "Membrane", "Problem", "PixelSim", "device_count", "load_library"]
