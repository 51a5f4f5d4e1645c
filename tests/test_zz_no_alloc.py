"""EsbCheckFastArithNoAlloc: the allocation-free variant of EsbCheckFastArith (device staging in a fixed array inside the library,
no cudaMalloc) must return exactly what the allocating entry point returns - results bit for bit, the same flags - on calls in a
row that share the array, including a short call after a long one (stale contents of the array must not leak into it).
Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _esb():
    import engine_sim_cli_b200 as esb
    return esb


def _same(x, y):
    return np.array_equal(x.view(np.uint64), y.view(np.uint64))


def _operands(rng, n):
    # magnitudes of the gas model (pressures, volumes, mole counts, temperatures) plus the values FastAr flags
    a = rng.uniform(-1.0, 1.0, n) * 10.0 ** rng.uniform(-12, 12, n)
    b = rng.uniform(0.1, 1.0, n) * 10.0 ** rng.uniform(-12, 12, n) * rng.choice([-1.0, 1.0], n)
    special = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 5e-324, 1e-310, 1e300, -1e300])
    k = min(len(special), n)
    a[:k] = special[:k]
    b[n - k:] = special[:k]
    return a, b


def test_no_alloc_matches_allocating_entry_point_on_calls_in_a_row():
    esb = _esb()
    rng = np.random.default_rng(20260)
    # long call first (fills the whole array), then shorter ones, then the long one again
    for call, n in enumerate((esb.NOALLOC_MAX_N, 1000, 37, 1, esb.NOALLOC_MAX_N)):
        a, b = _operands(rng, n)
        for op in (0, 1):
            args = (op, a, b) if op == 0 else (op, np.abs(a))
            want, want_flag = esb.check_fast_arith(*args)
            got, got_flag = esb.check_fast_arith_noalloc(*args)
            assert _same(got, want), f"call {call}, op {op}, n {n}: results differ at {np.nonzero(got.view(np.uint64) != want.view(np.uint64))[0][:8]}"
            assert np.array_equal(got_flag, want_flag), f"call {call}, op {op}, n {n}: flags differ"
            # the same call again straight away: same answer from the shared array
            again, again_flag = esb.check_fast_arith_noalloc(*args)
            assert _same(again, want) and np.array_equal(again_flag, want_flag)
    # unflagged quotients / roots are the IEEE ones, so the two entry points agree with the host as well
    a = rng.uniform(1.0, 2.0, 4096)
    b = rng.uniform(1.0, 2.0, 4096)
    q, f = esb.check_fast_arith_noalloc(0, a, b)
    assert not f.any() and _same(q, a / b)
    r, f = esb.check_fast_arith_noalloc(1, a)
    assert not f.any() and _same(r, np.sqrt(a))


def test_no_alloc_rejects_sizes_beyond_the_array():
    esb = _esb()
    a = np.ones(esb.NOALLOC_MAX_N + 1)
    with pytest.raises(esb.EngineSimError) as e:
        esb.check_fast_arith_noalloc(0, a, a)
    assert "ESB_NOALLOC_MAX_N" in str(e.value) and str(esb.NOALLOC_MAX_N) in str(e.value)
    # and the next call of a legal size works
    out, flagged = esb.check_fast_arith_noalloc(0, a[:8], a[:8])
    assert (out == 1.0).all() and not flagged.any()
