"""rustraytracer_b200 — B200-native path-tracing core for RustRaytracer's hot path.

Only what the path needs: `csrc/` (hand-written sm_100a CUDA + the C ABI, built to librtcore.so) and
the host-side mirror of the reference's interface (`core`, `scenes`).  See DESIGN.md / INTEGRATION.md.
"""
from . import core, scenes  # noqa: F401
from .core import (Camera, Comm, Film, IntType, Objects, Samplers, get_intersection, make_rays, render_multi, render_samples, replicate,  # noqa: F401
                   tile_multithread, unoccluded)
