"""The C-ABI library loads, exports every symbol include/rtw_b200.h declares, and fails loudly (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

import raytracingweekend_b200 as rtw
from raytracingweekend_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "rtw_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rtw_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound():
    lib = _lib.load()
    names = header_functions()
    assert len(names) >= 45
    for n in names:
        assert hasattr(lib, n), f"{n} declared in rtw_b200.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature"
    assert lib.rtw_abi_version() == 1


def test_host_library_has_no_cuda_dependency():
    host = ctypes.CDLL(os.path.join(ROOT, "raytracingweekend_b200", "lib", "librtw_b200_host.so"))
    assert hasattr(host, "rtw_builder_flatten") and not hasattr(host, "rtw_render")


def test_struct_sizes_match_header():
    # sizes implied by the header's field lists (natural alignment)
    assert ctypes.sizeof(_lib.FlatPrim) == 16 + 72
    assert ctypes.sizeof(_lib.FlatXf) == 8 + 24
    assert ctypes.sizeof(_lib.FlatNode) == 48 + 8
    assert ctypes.sizeof(_lib.FlatItem) == 32 + 8
    assert ctypes.sizeof(_lib.FlatCamera) == 24 * 8
    assert ctypes.sizeof(_lib.RenderParams) == 32
    assert ctypes.sizeof(_lib.Ray) == 32 and ctypes.sizeof(_lib.Hit) == 32
    assert rtw.api.HIT_DTYPE.itemsize == 32


def test_error_behaviour_invalid_ids():
    b = rtw.Builder()
    with pytest.raises(_lib.RtwError):
        b.sphere((0, 0, 0), 1.0, 12345)  # not a material id
    t = b.constant_texture(1, 1, 1)
    with pytest.raises(_lib.RtwError):
        b.sphere((0, 0, 0), 1.0, t)  # a texture is not a material
    with pytest.raises(_lib.RtwError):
        b.flatten()  # no world
    m = b.lambertian(t)
    s = b.sphere((0, 0, 0), 1.0, m)
    b.set_world(b.hittable_list([s]))
    with pytest.raises(_lib.RtwError):
        b.flatten()  # no camera
    with pytest.raises(_lib.RtwError):
        b.named_scene("no_such_scene", 1.0)


def test_unsupported_graphs_are_reported_at_flatten():
    b = rtw.Builder()
    m = b.lambertian(b.constant_texture(0.5, 0.5, 0.5))
    ms = b.moving_sphere((0, 0, 0), 1.0, m, (0, 1, 0), 0.0, 1.0)
    b.set_world(b.hittable_list([ms]))
    b.set_lights(b.hittable_list([ms]))  # moving-sphere light is rejected
    b.camera((0, 0, 5), (0, 0, 0), (0, 1, 0), 40, 1.0, 0.0, 10.0, 0.0, 1.0)
    with pytest.raises(_lib.RtwError, match="moving_sphere"):
        b.flatten()


def test_no_cpu_fallback_without_gpu():
    if _lib.load().rtw_device_count() > 0:
        pytest.skip("a B200 is present")
    b = rtw.Builder().named_scene("cornell_box_scene", 1.0)
    with pytest.raises(_lib.RtwError, match="no CPU fallback"):
        rtw.Scene(b)


def test_tonemap_and_ppm_writer(tmp_path):
    # RayTracingWeekend.cpp:241-244 and :252-275
    acc = np.zeros((2, 3, 4), dtype=np.float32)
    acc[..., 3] = 4.0
    acc[0, 0, :3] = (1.0, 4.0, 16.0)   # means 0.25, 1, 4 -> sqrt 0.5, 1, clamp 1
    acc[1, 2, :3] = (0.04, 0.0, 400.0)
    c = rtw.tonemap(acc)
    assert np.allclose(c[0, 0], (0.5, 1.0, 1.0)) and np.allclose(c[1, 2], (0.1, 0.0, 1.0))
    p = tmp_path / "x.ppm"
    rtw.write_ppm(str(p), c)
    lines = p.read_text().split("\n")
    assert lines[0] == "P3" and lines[1] == "3 2" and lines[2] == "255"
    # top row first: j = ny-1 = 1, then j = 0
    assert lines[3 + 2] == f"{int(np.float64(np.float32(255.99)) * c[1, 2, 0])} 0 255"
    assert lines[3 + 3] == "127 255 255"
    assert len([l for l in lines if l]) == 3 + 6


def test_multi_gpu_entry_points_argument_handling():
    """rtw_comm_* / rtw_shard_range / rtw_reduce_accumulators without a GPU: ids, ranges, no-op cases and refusals."""
    import ctypes as C

    lib = _lib.load()
    b, n = C.c_int(), C.c_int()
    for spp, world in ((1024, 8), (100, 8), (5, 3), (0, 2), (7, 16)):
        got = []
        for r in range(world):
            assert lib.rtw_shard_range(spp, r, world, C.byref(b), C.byref(n)) == _lib.RTW_OK
            got.append((b.value, n.value))
        assert got[0][0] == 0 and sum(c for _, c in got) == spp
        assert all(b0 + c0 == b1 for (b0, c0), (b1, _) in zip(got, got[1:]))
    assert lib.rtw_shard_range(8, 2, 2, C.byref(b), C.byref(n)) != _lib.RTW_OK  # rank out of range
    assert lib.rtw_shard_range(8, 0, 0, C.byref(b), C.byref(n)) != _lib.RTW_OK
    # a single rank (or no communicator at all) reduces to a no-op; NULL handles are harmless
    assert lib.rtw_reduce_accumulators(None, None, 4, 4, 0, None) == _lib.RTW_OK
    assert lib.rtw_comm_rank(None) == 0 and lib.rtw_comm_world(None) == 1
    lib.rtw_comm_destroy(None)
    assert not lib.rtw_comm_create(None, 0, 1, 0)       # no id
    assert not lib.rtw_comm_adopt(None, 0, 2, 0)        # no communicator
    ident = (C.c_uint8 * _lib.RTW_COMM_ID_BYTES)()
    if lib.rtw_comm_unique_id(ident) == _lib.RTW_OK:    # libnccl present: two draws differ
        other = (C.c_uint8 * _lib.RTW_COMM_ID_BYTES)()
        assert lib.rtw_comm_unique_id(other) == _lib.RTW_OK and bytes(ident) != bytes(other)
        assert not lib.rtw_comm_create(ident, 3, 2, 0)  # rank >= world
    else:
        assert b"nccl" in lib.rtw_last_error().lower()
