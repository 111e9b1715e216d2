"""Where the end-to-end step of bench.py spends its time (N = 1): python tools/e2e_phases.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import raytracingweekend_b200 as rtw

NX = NY = 1024
SPP = 1024
host_t = torch.empty((NY, NX, 4), dtype=torch.float32).pin_memory()
host = host_t.numpy()
for it in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    b = rtw.Builder().named_scene("cornell_box_aluminum", 1.0)
    t1 = time.perf_counter()
    s = rtw.Scene(b, device=0)
    t2 = time.perf_counter()
    _, st = s.render(NX, NY, SPP, max_depth=100, seed=0, out=host)
    t3 = time.perf_counter()
    s.close(); b.close()
    t4 = time.perf_counter()
    print(f"iter {it}: build {1e3*(t1-t0):.2f} ms, Scene() {1e3*(t2-t1):.2f} ms, render {1e3*(t3-t2):.2f} ms (kernel {st['kernel_ms']:.2f}), close {1e3*(t4-t3):.2f} ms, total {1e3*(t4-t0):.2f}")
