"""Host logic: the builder/flattener reproduces the reference's object graph (topology, order, tables)."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle_lib as ol
import raytracingweekend_b200 as rtw
from conftest import GOLDEN


def builder_for(name, aspect=1.5):
    earth = rtw.load_earth() if name in ol.EARTH_SCENES else None
    return rtw.Builder().named_scene(name, aspect, earth)


def test_earth_fixture_hash():
    """the 196 608-byte decode of the reference's earth.jpg is pinned by hash."""
    a = rtw.load_earth()
    assert a.shape == (256, 256, 3)
    assert hashlib.sha256(a.tobytes()).hexdigest() == "da3cdb69263fdf2cf28811cab55d915e2ed89ad1a492d35f11e6087513f87e39"


@pytest.mark.parametrize("scene", rtw.SCENE_NAMES)
def test_native_scene_flattens_like_the_reference_objects(scene):
    """rtw_build_named_scene + rtw_builder_flatten == adapter walk over the reference's own objects, byte for byte
    (primitive order, flips, transform chains, bvh_node topology from std::sort + minstd, Perlin tables, camera)."""
    sha = json.load(open(os.path.join(GOLDEN, "flat_sha256.json")))
    blob = builder_for(scene).serialize()
    assert hashlib.sha256(blob).hexdigest() == sha[scene]
    if ol.have_ref():
        assert blob == ol.ref_flat(scene, 1.5)


def test_scene_shapes():
    b1 = builder_for("cornell_box_aluminum", 1.0)  # keep the builder alive: the flat view is owned by it
    fs = b1.flatten().contents
    assert (fs.n_prims, fs.n_lights, fs.n_materials, fs.n_items, fs.n_xfs) == (13, 2, 6, 1, 2)
    assert [fs.prims[i].type for i in range(7)] == [3, 4, 4, 3, 3, 2, 0]  # light, green, red, ceiling, floor, back, glass
    assert [fs.prims[i].flip for i in range(7)] == [0, 1, 0, 1, 0, 1, 0]
    assert [fs.prims[i].flip for i in range(7, 13)] == [0, 1, 0, 1, 0, 1]  # box: min-side rects flipped
    assert fs.lights[0].type == 1 and fs.lights[1].type == 2  # xz_rect light + the glass sphere (Q18)
    b2 = builder_for("book2_final_scene", 1.0)
    fs = b2.flatten().contents
    assert fs.n_nodes == 511 + 1023 and fs.n_leaves == 400 + 1000 and fs.n_prims == 2400 + 1000 + 9
    kinds = [fs.items[i].type for i in range(fs.n_items)]
    assert kinds == [1, 0, 2, 2, 0, 1]  # bvh(boxes), prims, medium, fog, prims, bvh(spheres)
    assert fs.items[2].multiplicity == 2 and fs.items[3].multiplicity == 2
    assert fs.items[5].xf >= 0 and fs.xfs[fs.items[5].xf].kind == 1  # translate outermost, then rotate_y
    b3 = builder_for("random_balls_scene")
    fs = b3.flatten().contents
    assert fs.n_prims == 485 and sum(fs.prims[i].type == 1 for i in range(485)) > 300  # mostly moving spheres


def test_perlin_tables_match_reference():
    g = np.load(os.path.join(GOLDEN, "funcs.npz"))["perlin"]
    bl = builder_for("light_sample")
    fs = bl.flatten().contents
    rv = np.ctypeslib.as_array(fs.perlin_ranvec, (768,))
    pm = np.ctypeslib.as_array(fs.perlin_perm, (768,))
    assert np.array_equal(rv, g[:768]) and np.array_equal(pm.astype(np.float64), g[768:])
    assert np.array_equal(pm[:256], pm[256:512]) and np.array_equal(pm[:256], pm[512:])


def test_bvh_node_semantics():
    b = rtw.Builder()
    m = b.lambertian(b.constant_texture(0.5, 0.5, 0.5))
    one = b.bvh_node([b.sphere((0, 0, 0), 1, m)])
    ok, mn, mx = b.bounding_box(one)
    assert ok and mn == [-1, -1, -1] and mx == [1, 1, 1]
    # rect boxes are padded by the float literal 0.0001f (hittable.h:169)
    ok, mn, mx = b.bounding_box(b.xy_rect(0, 1, 0, 2, 5, m))
    pad = float(np.float32(0.0001))
    assert mn == [0, 0, 5 - pad] and mx == [1, 2, 5 + pad]
    # hittable_list::bounding_box returns true without writing the box (Q6)
    ok, mn, mx = b.bounding_box(b.hittable_list([b.sphere((3, 3, 3), 1, m)]))
    assert ok and mn == [0, 0, 0] and mx == [0, 0, 0]
    # moving sphere box ignores t0,t1 and unions both end positions (Q36)
    ok, mn, mx = b.bounding_box(b.moving_sphere((0, 0, 0), 1, m, (0, 2, 0), 0, 1))
    assert mn == [-1, -1, -1] and mx == [1, 3, 1]
    # n == 1 node keeps left == right
    b.set_world(b.hittable_list([one]))
    b.camera((0, 0, 5), (0, 0, 0), (0, 1, 0), 40, 1, 0, 10, 0, 1)
    fs = b.flatten().contents
    assert fs.n_nodes == 1 and fs.nodes[0].left == fs.nodes[0].right == ~0


def test_nested_lists_inline_and_multiply_medium_draws():
    b = rtw.Builder()
    iso = b.isotropic(b.constant_texture(1, 1, 1))
    m = b.lambertian(b.constant_texture(0.5, 0.5, 0.5))
    med = b.constant_medium(b.box((0, 0, 0), (1, 1, 1), m), 0.5, iso)
    inner = b.hittable_list([b.sphere((5, 0, 0), 1, m), med])
    b.set_world(b.hittable_list([b.sphere((0, 5, 0), 1, m), b.translate(inner, (1, 2, 3))]))
    b.camera((0, 0, 5), (0, 0, 0), (0, 1, 0), 40, 1, 0, 10, 0, 1)
    fs = b.flatten().contents
    assert [fs.items[i].type for i in range(fs.n_items)] == [0, 2]
    assert fs.items[0].count == 2 and fs.prims[1].xf >= 0 and fs.prims[0].xf == -1
    assert fs.items[1].multiplicity == 4 and fs.items[1].count == 6 and fs.items[1].xf >= 0
