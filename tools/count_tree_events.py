"""Traversal events per path of the render kernel's OWN acceleration structures (trees built at upload, box records,
front-to-back culling), counted on the device by the diagnostics build of the library:

    python -c "from raytracingweekend_b200.build import build_count; build_count()"        # lib/librtw_b200_count.so
    RTW_B200_LIB=$PWD/raytracingweekend_b200/lib/librtw_b200_count.so python tools/count_tree_events.py <scene> <nx> <ny> <spp>

Prints one JSON line: {"per_path": {aabb, sphere_test, sphere_root, moving_center, rect_test, box_test, xf}, "rays_per_path"}.
bench.py prices these with the Appendix-D costs for roofline.kernel_tree (cfg1, cfg3)."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingweekend_b200 as r  # noqa: E402
from raytracingweekend_b200 import _lib  # noqa: E402

name, nx, ny, spp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
lib = _lib.load()
if not hasattr(lib, "rtw_debug_events"):
    raise SystemExit("this library was not built with -DRTW_COUNT_EVENTS (set RTW_B200_LIB to librtw_b200_count.so)")
lib.rtw_debug_events.argtypes = [C.POINTER(C.c_ulonglong), C.c_int, C.c_int]
sc = r.Scene.named(name, nx / ny)
ev = (C.c_ulonglong * 8)()
lib.rtw_debug_events(ev, 8, 1)
acc, st = sc.render(nx, ny, spp)
n = lib.rtw_debug_events(ev, 8, 1)
names = ["aabb", "sphere_test", "sphere_root", "moving_center", "rect_test", "box_test", "xf"]
print(json.dumps({"scene": name, "nx": nx, "ny": ny, "spp": spp, "rays_per_path": st["rays"] / st["paths"],
                  "per_path": {k: ev[i] / st["paths"] for i, k in enumerate(names[:n])}}))
