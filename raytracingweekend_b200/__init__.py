"""B200-native path-tracing backend for the silvesthu/RayTracingWeekend color() hot path.

Public surface: `Builder` (reference constructors), `Scene` (device scene + render / intersect_batch),
`tonemap`, `write_ppm[_binary]`, `write_png`, `decode_jpeg`, `build` (in-tree nvcc build).  See DESIGN.md and include/rtw_b200.h.
"""
from .api import (Builder, Scene, SCENE_NAMES, tonemap, write_ppm, write_ppm_binary, write_png, decode_jpeg,  # noqa: F401
                  load_earth)
from .build import build  # noqa: F401
