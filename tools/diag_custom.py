import sys, os
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
import oracle_lib as ol
import raytracingweekend_b200 as rtw
import custom_scenes as cs

def check(name, b, nx=48, ny=32, spp=8):
    sc = rtw.Scene(b)
    acc, st = sc.render(nx, ny, spp, seed=3)
    osum, _, orays = ol.oracle_render(b, nx, ny, spp, mode="philox", seed=3)
    g = acc[..., :3].astype(np.float64)
    close = np.isclose(g, osum, rtol=2e-3, atol=1e-3 * spp).all(axis=2)
    print(f"{name:28s} close {close.mean():.4f} rays {st['rays']} vs {orays} mean {g.mean(axis=(0,1))/spp} vs {osum.mean(axis=(0,1))/spp}")
    sc.close()

for n, f in cs.CUSTOM.items():
    check(n, f(1.5))

# kitchen sink with one feature group at a time
def sub(which, aperture=0.0):
    b = rtw.Builder()
    white = b.constant_texture(0.73, 0.73, 0.73)
    green = b.constant_texture(0.12, 0.45, 0.15)
    objs = []
    if "ground_checker" in which: objs.append(b.xz_rect(-20, 20, -20, 20, -0.05, b.lambertian(b.checker_texture(green, b.noise_texture(2.0)))))
    else: objs.append(b.xz_rect(-20, 20, -20, 20, 0, b.lambertian(green)))
    glass = b.dielectric(1.5)
    if "hollow" in which: objs += [b.sphere((-2.2, 1, 0), 1.0, glass), b.sphere((-2.2, 1, 0), -0.9, glass)]
    if "metal" in which: objs += [b.sphere((2.2, 1, 0), 1.0, b.metal((0.8, 0.6, 0.2), 0.3)), b.sphere((0, 1, 2.5), 1.0, b.metal((0.9, 0.9, 0.9), 0.0))]
    if "chain" in which:
        box = b.box((0, 0, 0), (1, 2, 1), b.lambertian(white))
        objs.append(b.translate(b.rotate_y(b.translate(b.rotate_y(box, 20.0), (0.5, 0, 0)), -35.0), (-0.5, 0, -1.5)))
    if "fliprect" in which: objs.append(b.flip_normals(b.rotate_y(b.xy_rect(-1, 1, 0, 1.5, 4.0, b.lambertian(green)), 10.0)))
    if "medium" in which:
        fog_box = b.translate(b.rotate_y(b.box((0, 0, 0), (1.5, 1.5, 1.5), b.lambertian(white)), 30.0), (3.0, 0, 2.0))
        objs.append(b.constant_medium(fog_box, 0.8, b.isotropic(b.constant_texture(0.9, 0.9, 0.9))))
    if "moving" in which: objs.append(b.moving_sphere((0, 0.4, -2), 0.4, b.lambertian(white), (0, 0.8, -2), 0.0, 1.0))
    b.set_world(b.hittable_list(objs)); b.set_lights(None); b.set_background_type(1)
    b.camera((0, 2.5, -10), (0, 1, 0), (0, 1, 0), 35.0, 1.5, aperture, 10.0, 0.0, 1.0)
    return b

for w in (["plain"], ["ground_checker"], ["hollow"], ["metal"], ["chain"], ["fliprect"], ["medium"], ["moving"]):
    check("+".join(w), sub(w))
check("plain+aperture", sub(["plain"], 0.1))
check("all-no-aperture", sub(["ground_checker","hollow","metal","chain","fliprect","medium","moving"], 0.0))
