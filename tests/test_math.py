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


def test_math_on_the_argument_ranges_the_gas_model_produces():
    """esb_math.h is the one piece of code the product and the oracle share, so bit-exact GPU-vs-oracle tests cannot see an error
    in it: this test holds it to libm on the arguments the physics actually feeds it.
      pow(ratio, 1/gamma), ratio = p_down / p_up in (0, 1]                   GasSystem::flowRate
      pow(1 + (gamma-1)/2 M^2, gamma/(gamma-1)), Mach 0 .. 1                  GasSystem::dynamicPressure
      pow(2/(gamma+1), ...) once per load; pow(throttle, gamma) in [0,1]      choked-flow constants, throttle linkage
      exp(-(v/v_st)^2), exp(-2 min(v/v_c, 300))                              ring friction
      sin / cos of body angles (|theta| up to thousands of radians after minutes of running)"""
    L = _f()
    rng = np.random.default_rng(7)
    g = 1.4
    ratio = np.concatenate([rng.uniform(1e-6, 1.0, 6000), 1.0 - 10.0 ** rng.uniform(-12, -1, 2000), [1.0, 0.5282817877171742]])
    got = np.array([L.orc_math(4, v, 1.0 / g) for v in ratio])
    assert _ulps(got, ratio ** (1.0 / g)).max() <= 8      # pow = exp(y log x) without extended precision: measured 4 ulp here
    mach = np.concatenate([rng.uniform(0.0, 1.0, 6000), 10.0 ** rng.uniform(-9, -2, 2000)])
    x = 1.0 + (g - 1.0) / 2.0 * mach * mach
    got = np.array([L.orc_math(4, v, g / (g - 1.0)) for v in x])
    assert _ulps(got, x ** (g / (g - 1.0))).max() <= 2
    assert L.orc_math(4, 1.0, g / (g - 1.0)) == 1.0                         # no flow: exactly no dynamic pressure
    thr = rng.uniform(0.0, 1.0, 3000)
    for gamma in (0.5, 1.0, 2.0):
        got = np.array([L.orc_math(4, v, gamma) for v in thr])
        assert _ulps(got, thr ** gamma).max() <= 8
    q = rng.uniform(0.0, 30.0, 4000)
    got = np.array([L.orc_math(3, -(v * v), 0) for v in q])
    ref = np.exp(-(q * q))
    ok = ref > 1e-300
    assert _ulps(got[ok], ref[ok]).max() <= 2 and (got[~ok] <= 1e-300).all()
    z = -2.0 * np.minimum(rng.uniform(0.0, 400.0, 4000), 300.0)
    got = np.array([L.orc_math(3, v, 0) for v in z])
    assert _ulps(got, np.exp(z)).max() <= 2
    th = np.concatenate([rng.uniform(-6e4, 6e4, 6000), rng.uniform(-7, 7, 2000)])
    s = np.array([L.orc_math(0, v, 0) for v in th]); c = np.array([L.orc_math(1, v, 0) for v in th])
    assert np.abs(s - np.sin(th)).max() < 5e-16 and np.abs(c - np.cos(th)).max() < 5e-16
    assert np.abs(s * s + c * c - 1.0).max() < 1e-15
