"""Known-answer tests of the RNG and of the samplers that replace the reference's rejection loops.

* Philox4x32-10 against the published Random123 known-answer vectors (Random123 `kat_vectors`, "philox4x32 10" lines) —
  an anchor neither the oracle nor the CUDA code was written from: a plain-Python restatement of the algorithm, the oracle
  (oracle_philox) and the device function (rtw_philox4x32_10_batch) must all reproduce them.
* The closed-form expectations of the reference's Monte-Carlo scratchpad (MonteCarlo/MonteCarlo.cpp:151-183: the integral of
  cos^2 over the sphere is 4 pi / 3; :234-287: the cosine-weighted estimate of the integral of cos^3 over the hemisphere is
  pi / 2) and the moments of the uniform ball / disk, applied to the Philox-mode BOUNDED samplers (uniform ball by cbrt,
  polar lens disk, cosine lobe), which replace utility.h:27-35 and camera.h:61-69 — on the oracle here, on the device with
  -m gpu.  Tolerances are 5 standard errors of the estimator at the sample count used.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol

M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
# (counter[4], key[2]) -> output[4]
RANDOM123_KAT = [
    ((0x00000000, 0x00000000, 0x00000000, 0x00000000), (0x00000000, 0x00000000), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff), (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]
N = 1 << 20


def philox_py(c, k):
    c, k = list(c), list(k)
    for _ in range(10):
        p0, p1 = M0 * c[0], M1 * c[2]
        c = [(p1 >> 32) ^ c[1] ^ k[0], p1 & 0xffffffff, (p0 >> 32) ^ c[3] ^ k[1], p0 & 0xffffffff]
        k = [(k[0] + W0) & 0xffffffff, (k[1] + W1) & 0xffffffff]
    return tuple(c)


def oracle_philox(c, k):
    o = ol.oracle()
    cc, kk, out = (C.c_uint32 * 4)(*c), (C.c_uint32 * 2)(*k), (C.c_uint32 * 4)()
    o.oracle_philox(cc, kk, out)
    return tuple(out)


def oracle_samples(which, n, seed=1):
    o = ol.oracle()
    o.oracle_sampler_batch.argtypes = [C.c_int, C.c_int64, C.c_uint64, C.c_void_p]
    out = np.zeros((n, 3), dtype=np.float64)
    o.oracle_sampler_batch(which, n, seed, out.ctypes.data)
    return out


def check_sampler_moments(ball, disk, cosine):
    """Closed forms, each within 5 standard errors."""
    n = len(ball)

    def near(x, mean, sd, what):
        est, tol = float(np.mean(x)), 5.0 * sd / np.sqrt(n)
        assert abs(est - mean) < tol, f"{what}: {est} vs {mean} (tolerance {tol})"
        return est

    # uniform in the unit ball: |p| <= 1, E r = 3/4, E r^2 = 3/5, E p = 0, and its direction is uniform on the sphere:
    # MonteCarlo.cpp:151-183 — integral of cos^2(theta) over the sphere = E[cos^2 / (1 / 4 pi)] = 4 pi / 3
    r = np.linalg.norm(ball, axis=1)
    assert r.max() <= 1.0 + 1e-6
    near(r, 0.75, np.sqrt(3 / 5 - 9 / 16), "ball E r")
    near(r * r, 0.6, np.sqrt(3 / 7 - 9 / 25), "ball E r^2")
    for ax in range(3):
        near(ball[:, ax], 0.0, np.sqrt(0.2), f"ball E p[{ax}]")
    d = ball / r[:, None]
    near(d[:, 2] ** 2 * 4 * np.pi, 4 * np.pi / 3, 4 * np.pi * np.sqrt(1 / 5 - 1 / 9), "sphere integral of cos^2")
    # uniform in the unit disk: E r^2 = 1/2, E x = E y = 0, E x^2 = 1/4
    assert np.all(disk[:, 2] == 0) and np.linalg.norm(disk[:, :2], axis=1).max() <= 1.0 + 1e-6
    near((disk[:, :2] ** 2).sum(axis=1), 0.5, np.sqrt(1 / 3 - 1 / 4), "disk E r^2")
    near(disk[:, 0], 0.0, 0.5, "disk E x")
    near(disk[:, 1], 0.0, 0.5, "disk E y")
    near(disk[:, 0] ** 2, 0.25, np.sqrt(1 / 8 - 1 / 16), "disk E x^2")
    # cosine-weighted direction about +z: unit length, z >= 0, pdf = cos / pi:
    # MonteCarlo.cpp:234-287 — integral of cos^3 over the hemisphere = E[cos^3 / (cos / pi)] = pi E[cos^2] = pi / 2
    assert np.allclose(np.linalg.norm(cosine, axis=1), 1.0, atol=2e-6) and cosine[:, 2].min() >= 0.0
    near(np.pi * cosine[:, 2] ** 2, np.pi / 2, np.pi * np.sqrt(1 / 3 - 1 / 4), "hemisphere integral of cos^3")
    near(cosine[:, 2], 2 / 3, np.sqrt(1 / 2 - 4 / 9), "cosine E z")
    near(cosine[:, 0], 0.0, 0.5, "cosine E x")


def test_philox_matches_the_published_random123_vectors():
    for c, k, want in RANDOM123_KAT:
        assert philox_py(c, k) == want
        assert oracle_philox(c, k) == want


def test_oracle_philox_equals_the_python_restatement_on_random_counters():
    rng = np.random.default_rng(5)
    for _ in range(200):
        c = tuple(int(x) for x in rng.integers(0, 1 << 32, size=4))
        k = tuple(int(x) for x in rng.integers(0, 1 << 32, size=2))
        assert oracle_philox(c, k) == philox_py(c, k)


def test_oracle_bounded_samplers_meet_the_monte_carlo_closed_forms():
    check_sampler_moments(oracle_samples(0, N), oracle_samples(1, N), oracle_samples(2, N))


@pytest.mark.gpu
def test_device_philox_matches_random123_and_the_oracle():
    from raytracingweekend_b200 import _lib

    lib = _lib.load()
    for c, k, want in RANDOM123_KAT:
        cc = np.array([c], dtype=np.uint32)
        out = np.zeros((1, 4), dtype=np.uint32)
        _lib.check(lib.rtw_philox4x32_10_batch(0, cc.ctypes.data, 1, k[0] | (k[1] << 32), out.ctypes.data), "philox batch")
        assert tuple(int(x) for x in out[0]) == want
    rng = np.random.default_rng(6)
    cs = rng.integers(0, 1 << 32, size=(4096, 4), dtype=np.uint64).astype(np.uint32)
    key = (0x9abcdef0 << 32) | 0x12345678
    out = np.zeros_like(cs)
    _lib.check(lib.rtw_philox4x32_10_batch(0, cs.ctypes.data, len(cs), key, out.ctypes.data), "philox batch")
    for i in range(0, len(cs), 97):
        assert tuple(int(x) for x in out[i]) == philox_py([int(x) for x in cs[i]], (0x12345678, 0x9abcdef0))


@pytest.mark.gpu
def test_device_bounded_samplers_meet_the_closed_forms_and_follow_the_oracle():
    from raytracingweekend_b200 import _lib

    lib = _lib.load()
    got = []
    for which in range(3):
        out = np.zeros((N, 3), dtype=np.float32)
        _lib.check(lib.rtw_sampler_batch(0, which, N, 1, out.ctypes.data), "sampler batch")
        got.append(out.astype(np.float64))
    check_sampler_moments(*got)
    # same Philox counters, same uniforms: the FP32 samples follow the FP64 oracle's sample for sample
    for which, g in enumerate(got):
        o = oracle_samples(which, 4096)
        err = np.abs(g[:4096] - o).max()
        print(f"sampler {which}: max |device - oracle| = {err:.2e}")
        assert err < 2e-5
