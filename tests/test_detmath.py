"""include/mbots_detmath.h against float64 libm (CPU only). The header DEFINES the simulator's sin/cos/atan2
(DEVIATIONS.md D2); these tests keep a bug there from hiding as a common-mode error of oracle and kernels."""
import ctypes

import numpy as np

from oracle_binding import oracle_lib

PF = ctypes.POINTER(ctypes.c_float)


def _ulp_err(got, want64):
    want32 = want64.astype(np.float32)
    ulp = np.spacing(np.abs(want32)).astype(np.float64)
    ulp = np.maximum(ulp, np.float64(2.0 ** -149))
    return np.abs(got.astype(np.float64) - want64) / ulp


def _sincos(x):
    lib = oracle_lib(False)
    lib.mbo_detmath_sincos.argtypes = [PF, PF, PF, ctypes.c_int64]
    x = np.ascontiguousarray(x, dtype=np.float32)
    s, c = np.empty_like(x), np.empty_like(x)
    lib.mbo_detmath_sincos(x.ctypes.data_as(PF), s.ctypes.data_as(PF), c.ctypes.data_as(PF), len(x))
    return s, c


def _atan2(y, x):
    lib = oracle_lib(False)
    lib.mbo_detmath_atan2.argtypes = [PF, PF, PF, ctypes.c_int64]
    y = np.ascontiguousarray(y, dtype=np.float32)
    x = np.ascontiguousarray(x, dtype=np.float32)
    o = np.empty_like(x)
    lib.mbo_detmath_atan2(y.ctypes.data_as(PF), x.ctypes.data_as(PF), o.ctypes.data_as(PF), len(x))
    return o


def test_sincos_accuracy_over_the_simulator_range():
    # ray angles are yaw + [-pi/4, 5pi/4]: |x| < 8
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-8, 8, 400000), np.linspace(-8, 8, 100001)]).astype(np.float32)
    s, c = _sincos(x)
    x64 = x.astype(np.float64)
    # absolute error bound near zeros of the function, ulp bound elsewhere
    es, ec = np.abs(s - np.sin(x64)), np.abs(c - np.cos(x64))
    assert es.max() < 1.2e-7 and ec.max() < 1.2e-7, (es.max(), ec.max())
    big = np.abs(np.sin(x64)) > 0.1
    assert _ulp_err(s[big], np.sin(x64[big])).max() <= 2.0
    big = np.abs(np.cos(x64)) > 0.1
    assert _ulp_err(c[big], np.cos(x64[big])).max() <= 2.0
    # unit length to rounding
    assert np.abs(s.astype(np.float64) ** 2 + c.astype(np.float64) ** 2 - 1).max() < 3e-7


def test_sincos_exact_points():
    s, c = _sincos(np.array([0.0, -0.0], dtype=np.float32))
    assert s[0] == 0 and c[0] == 1 and c[1] == 1
    s, c = _sincos(np.array([np.pi / 2, np.pi, -np.pi / 2], dtype=np.float32))
    assert abs(s[0] - 1) < 1e-7 and abs(c[1] + 1) < 1e-7 and abs(s[2] + 1) < 1e-7


def test_atan2_accuracy():
    rng = np.random.default_rng(1)
    ang = rng.uniform(-np.pi, np.pi, 400000)
    r = rng.uniform(0.5, 1.5, ang.size)
    y, x = (r * np.sin(ang)).astype(np.float32), (r * np.cos(ang)).astype(np.float32)
    a = _atan2(y, x)
    want = np.arctan2(y.astype(np.float64), x.astype(np.float64))
    assert np.abs(a - want).max() < 4e-7
    ok = np.abs(want) > 0.05
    assert _ulp_err(a[ok], want[ok]).max() <= 3.0
    assert a.min() >= -np.float32(np.pi) and a.max() <= np.float32(np.pi)


def test_atan2_axes_and_identity_heading():
    # an agent that never turned has rot = (1, 0): forward = (1, 0) -> yaw exactly 0
    a = _atan2(np.array([0, 0, 1, -1, 0.0], dtype=np.float32), np.array([1, -1, 0, 0, 2.0], dtype=np.float32))
    assert a[0] == 0 and a[4] == 0
    assert a[1] == np.float32(np.pi)
    assert a[2] == np.float32(np.pi / 2) and a[3] == -np.float32(np.pi / 2)
