"""include/esb_math.h (shared by the CUDA kernels and the oracle) against libm: the deterministic
replacements must stay within a few ulp, far inside the model's own sensitivity."""
import ctypes as C

import numpy as np

import oracle_lib as ol


def _f():
    L = ol.lib()
    L.orc_math.restype = C.c_double
    L.orc_math.argtypes = [C.c_int, C.c_double, C.c_double]
    L.orc_rand_u01.restype = C.c_double
    L.orc_rand_u01.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64]
    return L


def _ulps(a, b):
    return np.abs(a - b) / np.maximum(np.spacing(np.abs(b)), 1e-300)


def test_sin_cos_log_exp_pow():
    L = _f()
    rng = np.random.default_rng(0)
    x = rng.uniform(-1e4, 1e4, 4000)
    s = np.array([L.orc_math(0, v, 0) for v in x]); c = np.array([L.orc_math(1, v, 0) for v in x])
    assert np.abs(s - np.sin(x)).max() < 5e-16 and np.abs(c - np.cos(x)).max() < 5e-16
    p = 10.0 ** rng.uniform(-12, 12, 4000)
    lg = np.array([L.orc_math(2, v, 0) for v in p])
    assert _ulps(lg, np.log(p)).max() <= 2
    e = rng.uniform(-300, 300, 4000)
    ex = np.array([L.orc_math(3, v, 0) for v in e])
    assert _ulps(ex, np.exp(e)).max() <= 2
    b = rng.uniform(0.01, 10, 4000); y = rng.uniform(-4, 4, 4000)
    pw = np.array([L.orc_math(4, u, v) for u, v in zip(b, y)])
    assert (np.abs(pw - b ** y) / (b ** y)).max() < 5e-15
    assert L.orc_math(4, 0.0, 2.0) == 0.0 and L.orc_math(4, 0.0, 0.0) == 1.0


def test_noise_uniform_and_stateless():
    L = _f()
    u = np.array([L.orc_rand_u01(3, 17, k) for k in range(20000)])
    assert 0 <= u.min() and u.max() < 1 and abs(u.mean() - 0.5) < 0.01 and abs(u.var() - 1 / 12) < 0.005
    assert L.orc_rand_u01(3, 17, 5) == u[5] and L.orc_rand_u01(4, 17, 5) != u[5]
