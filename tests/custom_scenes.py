"""Small scenes composed through the builder API (one call per reference constructor) that combine features the nine
named scenes do not: nested transform chains, flipped rotated rects, a hollow glass sphere (negative radius), a medium
inside a rotated box, bvh_node with 1, 2 and 3 objects and under a transform, a light list with every light type
(xz_rect, sphere, and one with the base-class defaults), no lights at all, an empty world."""
import raytracingweekend_b200 as rtw


def kitchen_sink(aspect):
    b = rtw.Builder()
    white = b.constant_texture(0.73, 0.73, 0.73)
    green = b.constant_texture(0.12, 0.45, 0.15)
    ground = b.lambertian(b.checker_texture(green, b.noise_texture(2.0)))
    # the checkered ground is NOT at y = 0: there sin(10 y) is a rounding-noise sign in the reference (p.y = o.y + t d.y) and
    # exactly 0 on the device, which snaps p onto the rect's plane — an ill-conditioned texture, not a parity case
    objs = [b.xz_rect(-20, 20, -20, 20, -0.05, ground)]
    glass = b.dielectric(1.5)
    objs += [b.sphere((-2.2, 1, 0), 1.0, glass), b.sphere((-2.2, 1, 0), -0.9, glass)]  # hollow: negative radius (Q8)
    objs += [b.sphere((2.2, 1, 0), 1.0, b.metal((0.8, 0.6, 0.2), 0.3)), b.sphere((0, 1, 2.5), 1.0, b.metal((0.9, 0.9, 0.9), 0.0))]
    box = b.box((0, 0, 0), (1, 2, 1), b.lambertian(white))
    objs.append(b.translate(b.rotate_y(b.translate(b.rotate_y(box, 20.0), (0.5, 0, 0)), -35.0), (-0.5, 0, -1.5)))
    objs.append(b.flip_normals(b.rotate_y(b.xy_rect(-1, 1, 0, 1.5, 4.0, b.lambertian(green)), 10.0)))
    fog_box = b.translate(b.rotate_y(b.box((0, 0, 0), (1.5, 1.5, 1.5), b.lambertian(white)), 30.0), (3.0, 0, 2.0))
    objs.append(b.constant_medium(fog_box, 0.8, b.isotropic(b.constant_texture(0.9, 0.9, 0.9))))
    objs.append(b.moving_sphere((0, 0.4, -2), 0.4, b.lambertian(white), (0, 0.8, -2), 0.0, 1.0))
    b.set_world(b.hittable_list(objs))
    b.set_lights(None)
    b.set_background_type(1)  # gradient
    b.camera((0, 2.5, -10), (0, 1, 0), (0, 1, 0), 35.0, aspect, 0.1, 10.0, 0.0, 1.0)
    return b


def every_light_type(aspect):
    b = rtw.Builder()
    white = b.lambertian(b.constant_texture(0.73, 0.73, 0.73))
    red = b.lambertian(b.constant_texture(0.65, 0.05, 0.05))
    emit = b.diffuse_light(b.constant_texture(7, 7, 7))
    rect_light = b.xz_rect(-1, 1, -1, 1, 3.9, emit)
    sphere_light = b.sphere((2.5, 1.0, 0.5), 0.5, emit)
    odd_light = b.xy_rect(-3, -2, 0.5, 1.5, 2.9, emit)  # hittable.h:36-37 defaults: pdf 0, direction (1,0,0)
    objs = [b.xz_rect(-4, 4, -4, 4, 0, white), b.flip_normals(b.xz_rect(-4, 4, -4, 4, 4, white)),
            b.flip_normals(b.xy_rect(-4, 4, 0, 4, 3, white)), b.yz_rect(0, 4, -4, 4, -4, red),
            b.flip_normals(b.yz_rect(0, 4, -4, 4, 4, red)), b.flip_normals(rect_light), sphere_light, odd_light,
            b.sphere((-1, 0.7, 0), 0.7, b.dielectric(1.5)), b.sphere((0.8, 0.5, -0.5), 0.5, white)]
    b.set_world(b.hittable_list(objs))
    b.set_lights(b.hittable_list([rect_light, sphere_light, odd_light]))
    b.set_background_type(0)
    b.camera((0, 2, -9), (0, 2, 0), (0, 1, 0), 40.0, aspect, 0.0, 10.0, 0.0, 1.0)
    return b


def small_trees(aspect):
    b = rtw.Builder()
    white = b.lambertian(b.constant_texture(0.73, 0.73, 0.73))
    blue = b.lambertian(b.constant_texture(0.2, 0.3, 0.8))
    one = b.bvh_node([b.sphere((-3, 0.5, 0), 0.5, blue)])
    two = b.bvh_node([b.sphere((-1.5, 0.5, 0), 0.5, blue), b.box((-1.2, 0, 1), (-0.4, 0.8, 1.8), white)])
    three = b.bvh_node([b.sphere((0.5, 0.5, 0), 0.5, b.metal((0.7, 0.7, 0.7), 0.1)), b.sphere((1.2, 0.3, -0.8), 0.3, b.dielectric(1.5)),
                        b.moving_sphere((1.8, 0.4, 0.5), 0.4, blue, (1.8, 0.7, 0.5), 0.0, 1.0)])
    grid = b.bvh_node([b.sphere((0.45 * i, 0.2, 0.45 * j), 0.2, white) for i in range(4) for j in range(4)])
    moved = b.translate(b.rotate_y(grid, 25.0), (2.5, 0, 1.0))
    b.set_world(b.hittable_list([b.xz_rect(-10, 10, -10, 10, 0, white), one, two, three, moved]))
    b.set_lights(None)
    b.set_background_type(1)
    b.camera((0, 3, -8), (0, 0.5, 0), (0, 1, 0), 35.0, aspect, 0.0, 10.0, 0.0, 1.0)
    return b


def empty_world(aspect):
    b = rtw.Builder()
    b.set_world(b.hittable_list([]))
    b.set_lights(None)
    b.set_background_type(1)
    b.camera((0, 0, 0), (0, 0, -1), (0, 1, 0), 60.0, aspect, 0.0, 1.0, 0.0, 1.0)
    return b


CUSTOM = {"kitchen_sink": kitchen_sink, "every_light_type": every_light_type, "small_trees": small_trees, "empty_world": empty_world}
