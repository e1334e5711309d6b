"""b200pt: B200-native wavefront path-tracing core behind pbrt-minus's integrator surface.

The product is libb200pt.so (hand-written CUDA for sm_100a, C-ABI in include/b200pt.h); this package is
the thin host-side mirror of the reference interface used by tests and bench.py.  There is no CPU
fallback: importing works anywhere, but every compute call needs the built library and a GPU.
"""
from . import api  # noqa: F401
from .api import B200ptError, load_library  # noqa: F401
from .scene import HLBVH, DeviceBuffer, Film, WavefrontPathIntegrator, sample_range  # noqa: F401
from . import pbrt_scene, spectral  # noqa: F401
from . import workloads  # noqa: F401
