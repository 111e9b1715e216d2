"""B200-native path-tracing backend for the silvesthu/RayTracingWeekend color() hot path.

Public surface: `Builder` (reference constructors), `Scene` (device scene + render / intersect_batch),
`tonemap`, `write_ppm`, `build` (in-tree nvcc build).  See DESIGN.md and include/rtw_b200.h.
"""
from .api import Builder, Scene, SCENE_NAMES, tonemap, write_ppm, load_earth  # noqa: F401
from .build import build  # noqa: F401
