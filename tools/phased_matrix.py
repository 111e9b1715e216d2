"""Digest matrix of book2_final_scene renders over shapes / depths / flags: run once per RTW_PHASED setting and diff the outputs.
    RTW_PHASED=0 python tools/phased_matrix.py > a.txt; RTW_PHASED=1 python tools/phased_matrix.py > b.txt; diff a.txt b.txt"""
import hashlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingweekend_b200 as r
from raytracingweekend_b200 import _lib

for nx, ny in ((16, 16), (33, 17), (100, 60), (257, 129)):
    sc = r.Scene.named("book2_final_scene", nx / ny)
    for spp in (1, 5, 64):
        for depth in (1, 3, 50):
            for flags in (0, getattr(_lib, "RTW_FLAG_DEBUG_RAY", 2)):
                a, st = sc.render(nx, ny, spp, max_depth=depth, flags=flags, seed=5)
                print(nx, ny, spp, depth, flags, hashlib.sha256(a.tobytes()).hexdigest()[:16], st["rays"])
    a, st = sc.render(nx, ny, 7, seed=5, spp_begin=3)
    print(nx, ny, "spp_begin", hashlib.sha256(a.tobytes()).hexdigest()[:16], st["rays"])
    sc.close()
