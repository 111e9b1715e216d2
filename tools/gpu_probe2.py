import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as ol
import raytracingweekend_b200 as rtw
b = rtw.Builder().named_scene("cornell_box_aluminum", 1.0)
sc = rtw.Scene(b)
acc, st = sc.render(64, 64, 16384, seed=3)
np.save("gpurun_out/gpu_alu_64_16384.npy", acc)
print(st)
# smoke debug
cam = b.flatten().contents.camera
o, llc = np.array(cam.origin), np.array(cam.lower_left_corner)
h, v = np.array(cam.horizontal), np.array(cam.vertical)
m = 64
s, t = np.meshgrid((np.arange(m) + 0.5) / m, (np.arange(m) + 0.5) / m)
d = llc + s.reshape(-1, 1) * h + t.reshape(-1, 1) * v - o
d /= np.linalg.norm(d, axis=1, keepdims=True)
rays = np.concatenate([np.tile(o, (m * m, 1)), d, np.zeros((m * m, 1))], axis=1).astype(np.float32)
gpu = sc.intersect_batch(rays)
ref = ol.oracle_hits(b, rays.astype(np.float64))
bad = np.nonzero((gpu["prim"] >= 0) != (ref[:, 0] == 1))[0]
print("mismatch", len(bad))
for i in bad[:8]:
    print(i, rays[i], "gpu", gpu[i], "oracle", ref[i])
