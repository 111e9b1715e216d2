"""rtw_render_device_noalloc (include/rtw_b200.h): the wavefront render on the library's static device scratch — the entry
point for callers that cannot allocate (no counterpart in the reference).

Collected last on purpose.  Through the C ABI: its output must equal the existing entry
point's (rtw_render_device with RTW_FLAG_WAVEFRONT: same kernels, same Philox keys — identical ray counts, sums equal up to
the order of the float atomics, the tolerance of test_wavefront_pipeline_traces_the_same_paths_as_the_megakernel) on two calls
in a row that SHARE the scratch array: a big pool first, then a small pool on another image and seed, then the first again, so
stale contents of the previous call would show."""
import numpy as np
import pytest

import raytracingweekend_b200 as rtw
from raytracingweekend_b200 import _lib

pytestmark = pytest.mark.gpu


def _existing(sc, torch, nx, ny, spp, seed):
    acc = torch.zeros((ny, nx, 4), dtype=torch.float32, device="cuda:0")
    sc.render_device(acc.data_ptr(), nx, ny, spp, seed=seed, flags=_lib.RTW_FLAG_WAVEFRONT, stream=torch.cuda.current_stream().cuda_stream)
    st = sc.finish()
    return acc.cpu().numpy(), st


def _check(new, st_new, ref, st_ref, spp):
    assert st_new["rays"] == st_ref["rays"] and st_new["paths"] == st_ref["paths"] and st_new["kernel_launches"] > 2
    assert np.all(new[..., 3] == spp)
    fin = np.isfinite(ref).all(axis=2)
    assert np.array_equal(fin, np.isfinite(new).all(axis=2))
    close = np.isclose(new, ref, rtol=1e-4, atol=1e-5 * spp).all(axis=2) | ~fin
    assert close.mean() > 0.995, f"{(~close).sum()} pixels differ"
    assert np.allclose(new[fin], ref[fin], rtol=2e-2, atol=1e-3 * spp)


def test_noalloc_entry_point_matches_the_existing_one_on_consecutive_calls():
    import torch

    sc = rtw.Scene.named("cornell_box_aluminum", 1.0)
    stream = torch.cuda.current_stream().cuda_stream
    cases = [(96, 96, 16, 21, 0), (37, 21, 5, 4, 512), (96, 96, 16, 21, 0)]  # nx, ny, spp, seed, pool_slots
    for nx, ny, spp, seed, pool in cases:
        ref, st_ref = _existing(sc, torch, nx, ny, spp, seed)
        acc = torch.full((ny, nx, 4), 123.0, dtype=torch.float32, device="cuda:0")  # the call overwrites, it does not accumulate
        torch.cuda.synchronize()
        st = sc.render_device_noalloc(acc.data_ptr(), nx, ny, spp, seed=seed, pool_slots=pool, stream=stream)
        _check(acc.cpu().numpy(), st, ref, st_ref, spp)
    # a second scene on the same GPU shares the same array
    sc2 = rtw.Scene.named("two_perlin_spheres_earth", 1.0)
    ref, st_ref = _existing(sc2, torch, 40, 40, 8, 3)
    acc = torch.empty((40, 40, 4), dtype=torch.float32, device="cuda:0")
    st = sc2.render_device_noalloc(acc.data_ptr(), 40, 40, 8, seed=3, stream=stream)
    _check(acc.cpu().numpy(), st, ref, st_ref, 8)
    # sizes beyond the static array are refused with a clear error, nothing is rendered
    with pytest.raises(_lib.RtwError, match="RTW_NOALLOC_MAX_POOL_SLOTS"):
        sc.render_device_noalloc(acc.data_ptr(), 40, 40, 8, pool_slots=_lib.RTW_NOALLOC_MAX_POOL_SLOTS + 1)
    with pytest.raises(_lib.RtwError):
        sc.render_device_noalloc(0, 40, 40, 8)
    sc.close()
    sc2.close()
