"""usgscsm_b200 -- B200-native batched sensor-model engine behind the CSM RasterGM interface.

The per-point camera math lives in hand-written sm_100a CUDA kernels (usgscsm_b200/csrc),
reached through the C-ABI of include/usgscsm_b200.h.  This package is the thin host mirror
of the reference's plugin interface; it has no CPU implementation of the math.
"""
from . import _capi, synth  # noqa: F401  (_capi raises ImportError when the .so is missing)
from .plugin import (Dem, EcefCoord, Error, GeoGrid, ImageCoord, RasterGM,  # noqa: F401
                     UsgsAstroPlugin, Warning, applyDistortion, groundToImageImages, imageToGroundGridImages,
                     imageToGroundImages, lagrangeInterp, removeDistortion)
