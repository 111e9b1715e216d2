"""Pins the CPU oracle (oracle/rtw_oracle.cpp) to the reference, bit for bit in float64:
  * against committed golden vectors generated from the unmodified reference (tests/golden/make_golden.py);
  * against the reference itself (oracle/_ref/ref_harness) when it has been built in this tree.
Also restates the reference's own unit tests (CppTest/unittest1.cpp) as known-answer tests."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle_lib as ol
import raytracingweekend_b200 as rtw
from conftest import GOLDEN

SCENES = rtw.SCENE_NAMES


def builder_for(name, aspect):
    earth = rtw.load_earth() if name in ol.EARTH_SCENES else None
    return rtw.Builder().named_scene(name, aspect, earth)


def same(a, b):
    return (a == b) | (np.isnan(a) & np.isnan(b))


def test_reference_unit_tests_as_kats():
    """CppTest/unittest1.cpp:20-40 (_vec3), :72-109 (_aabb) — evaluated by the reference (kat.json) and the oracle."""
    ref = json.load(open(os.path.join(GOLDEN, "kat.json")))
    assert ref["dot"] == 2.0 and ref["cross"] == [0.0, 0.0, 1.0] and ref["point_at_parameter_3"] == [7.0, 7.0, 7.0]
    assert ref["aabb_hit"] == [1, 0, 1, 1, 0]
    assert ref["surrounding_min"] == [0.0, 0.0, 0.0] and ref["surrounding_max"] == [4.0, 4.0, 4.0]
    assert ref["sizeof_vec3"] == 24  # _vec3::_size (:15-18) is stale in the reference: vec3 is 3 doubles
    out = np.zeros(22)
    o = ol.oracle()
    o.oracle_kat.argtypes = [C.c_void_p]
    o.oracle_kat(out.ctypes.data)
    assert out[0] == ref["dot"] and list(out[1:4]) == ref["cross"] and list(out[4:7]) == ref["point_at_parameter_3"]
    assert list(out[7:12].astype(int)) == ref["aabb_hit"]
    assert list(out[12:15]) == ref["surrounding_min"] and list(out[15:18]) == ref["surrounding_max"] and out[18] == 24


def test_per_function_sequences_match_golden():
    g = np.load(os.path.join(GOLDEN, "funcs.npz"))
    mine = ol.oracle_funcs(g["perlin"])
    assert mine.shape == g["funcs"].shape
    assert same(mine, g["funcs"]).all(), np.nonzero(~same(mine, g["funcs"]))[0][:8]


@pytest.mark.parametrize("scene", SCENES)
def test_hits_match_golden_bit_for_bit(scene):
    g = np.load(os.path.join(GOLDEN, f"hits_{scene}.npz"))
    b = builder_for(scene, float(g["aspect"]))
    ref, mine = g["hits"], ol.oracle_hits(b, g["rays"])
    fs = b.flatten().contents
    iso = [i for i in range(fs.n_materials) if fs.materials[i].type == 4]
    stochastic = (ref[:, 0] == 1) & np.isin(ref[:, 10], iso)  # the reference scattered inside a constant_medium
    hit = ref[:, 0] == 1
    ok = (ref[:, 0] == mine[:, 0]) & (~hit | same(ref[:, 1:11], mine[:, 1:11]).all(axis=1))
    assert (ok | stochastic).all(), f"{(~(ok | stochastic)).sum()} rays differ"
    assert (hit & ~stochastic).sum() > 500


@pytest.mark.parametrize("scene", SCENES)
def test_serial_render_matches_golden_bit_for_bit(scene):
    g = np.load(os.path.join(GOLDEN, f"serial_{scene}.npz"))
    nx, ny, spp = int(g["nx"]), int(g["ny"]), int(g["spp"])
    b = builder_for(scene, nx / ny)
    s, q, rays = ol.oracle_render(b, nx, ny, spp, mode="minstd")
    assert same(s, g["sum"]).all() and same(q, g["sumsq"]).all()
    assert rays >= nx * ny * spp


def test_medium_density_doubling_kat():
    """a constant_medium inside a hittable_list behaves with twice its density (two passes)."""
    b = rtw.Builder()
    white = b.isotropic(b.constant_texture(1, 1, 1))
    bnd = b.sphere((0, 0, 0), 50, b.dielectric(1.5))
    b.set_world(b.hittable_list([b.constant_medium(bnd, 0.02, white)]))
    b.camera((0, 0, -200), (0, 0, 0), (0, 1, 0), 0.001, 1.0, 0.0, 10.0, 0.0, 1.0)
    b.set_background_type(1)
    fs = b.flatten().contents
    assert fs.items[0].multiplicity == 2
    # transmittance along the diameter: exp(-2*rho*L) = exp(-4) = 0.0183; isotropic white + gradient background:
    # estimate P(no scatter) as the fraction of 1-bounce paths, via rays/path in both RNG modes
    for mode in ("minstd", "philox"):
        s, q, rays = ol.oracle_render(b, 64, 64, 8, max_depth=1, mode=mode, threads=1)
        frac_bg = (s.sum(axis=2) > 0).mean()  # with max_depth 1 only unscattered paths see the background
        # P(pixel has >= 1 of 8 samples unscattered) = 1-(1-T)^8 with T = 0.0183 -> 0.137
        assert abs(frac_bg - (1 - (1 - np.exp(-4.0)) ** 8)) < 0.03, (mode, frac_bg)


@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref/ref_harness not built (needs /root/reference)")
class TestLiveReference:
    def test_funcs_live(self):
        assert same(ol.oracle_funcs(ol.ref_perlin()), ol.ref_funcs()).all()

    @pytest.mark.parametrize("scene", ["cornell_box_aluminum", "book2_final_scene", "random_balls_scene"])
    def test_hits_live(self, scene):
        rng = np.random.default_rng(5)
        b = builder_for(scene, 1.25)
        rays = np.concatenate([rng.normal(size=(3000, 3)) * 300 + 250, rng.normal(size=(3000, 3)), rng.random((3000, 1))], axis=1)
        ref, mine = ol.ref_hits(scene, 1.25, rays), ol.oracle_hits(b, rays)
        fs = b.flatten().contents
        iso = [i for i in range(fs.n_materials) if fs.materials[i].type == 4]
        stochastic = (ref[:, 0] == 1) & np.isin(ref[:, 10], iso)
        hit = ref[:, 0] == 1
        ok = (ref[:, 0] == mine[:, 0]) & (~hit | same(ref[:, 1:11], mine[:, 1:11]).all(axis=1))
        assert (ok | stochastic).all()

    def test_serial_render_live(self):
        b = builder_for("cornell_box_scene", 1.0)
        rs, rq, _ = ol.ref_render("cornell_box_scene", 20, 20, 3, threads=1)
        s, q, _ = ol.oracle_render(b, 20, 20, 3, mode="minstd")
        assert same(rs, s).all() and same(rq, q).all()

    def test_bvh_overlay_is_the_two_token_patch(self):
        diff = open(os.path.join(os.path.dirname(ol.REF_HARNESS), "bvh_overlay.diff")).read()
        changed = [l for l in diff.splitlines() if l.startswith(">")]
        assert len(changed) == 2 and "right = new bvh_node" in changed[0] and "right->hit" in changed[1]
