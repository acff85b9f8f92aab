"""mitsuba-im_b200 — B200-native ray intersection + path tracing behind Mitsuba's plugin surface.

Layout: csrc/ (CUDA kernels + the C ABI of include/mtsb200.h -> lib/libmtsb200.so), plugin/ (the C++ Mitsuba
integrator plugin that adapts the C ABI to Integrator / ResponsiveIntegrator), binding.py (ctypes view of the C ABI
for tests and bench.py), scenes.py (procedural scenes of the BASELINE configs for both back ends).
The directory name contains a hyphen, so import it through __graft_entry__.load_package().
"""
from . import binding  # noqa: F401
from . import scenes   # noqa: F401
from . import multigpu  # noqa: F401
