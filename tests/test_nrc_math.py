"""include/nrc_math.h (the normative transcendental definitions shared by oracle and kernels)
against float64 libm: a few ulp is what GLSL asks of these builtins."""
import numpy as np

from oracle import oracle as O


def _ulp_err(f, ref):
    ref32 = ref.astype(np.float32)
    ulp = np.spacing(np.abs(ref32)).astype(np.float64)
    return float(np.max(np.abs(f.astype(np.float64) - ref) / np.maximum(ulp, 1e-45)))


def test_sin_cos_accuracy():
    x = np.random.default_rng(0).uniform(-100, 100, 400000).astype(np.float32)
    x64 = x.astype(np.float64)
    assert np.max(np.abs(O.math_eval("sin", x) - np.sin(x64))) < 2e-7
    assert np.max(np.abs(O.math_eval("cos", x) - np.cos(x64))) < 2e-7
    assert _ulp_err(O.math_eval("sin", x), np.sin(x64)) < 4.0


def test_exp_accuracy_and_limits():
    x = np.random.default_rng(1).uniform(-100, 20, 400000).astype(np.float32)
    assert _ulp_err(O.math_eval("exp", x), np.exp(x.astype(np.float64))) < 2.5
    e = O.math_eval("exp", [0.0, -1e4, 1e4, -2000.0])
    assert e[0] == 1.0 and e[1] == 0.0 and np.isinf(e[2]) and e[3] == 0.0


def test_inverse_trig_accuracy_and_saturation():
    x = np.random.default_rng(2).uniform(-1, 1, 400000).astype(np.float32)
    x64 = x.astype(np.float64)
    assert _ulp_err(O.math_eval("acos", x), np.arccos(x64)) < 3.0
    assert _ulp_err(O.math_eval("asin", x), np.arcsin(x64)) < 4.0
    y = np.random.default_rng(3).uniform(-1, 1, 400000).astype(np.float32)
    assert _ulp_err(O.math_eval("atan2", x, y), np.arctan2(y.astype(np.float64), x64)) < 5.0
    # GLSL leaves acos/asin outside [-1,1] undefined; the parity preset saturates (nrc_math.h header)
    a = O.math_eval("acos", [1.0000001, -1.0000001, 1.0, -1.0])
    assert a[0] == 0.0 and a[2] == 0.0 and abs(a[1] - np.pi) < 1e-6 and abs(a[3] - np.pi) < 1e-6
    assert O.math_eval("atan2", [0.0], [0.0])[0] == 0.0


def test_bit_patterns_are_stable():
    """Known-answer bit patterns: a compiler flag that lets multiplies and adds fuse (or changes rounding)
    would change these, on the host as on the device (tests/test_gpu_math.py checks the same table)."""
    x = np.array([0.5, 1.0, 12.9898, 78.233, 91.2228, 3.0, -2.5, 6.2831855], np.float32)
    got = O.math_eval("sin", x).view(np.uint32)
    ref = np.load(__file__.replace("test_nrc_math.py", "golden/nrc_math_kat.npz"))
    np.testing.assert_array_equal(got, ref["sin"])
    np.testing.assert_array_equal(O.math_eval("cos", x).view(np.uint32), ref["cos"])
    np.testing.assert_array_equal(O.math_eval("exp", -np.abs(x)).view(np.uint32), ref["exp"])
    np.testing.assert_array_equal(O.math_eval("acos", np.clip(x / 100, -1, 1)).view(np.uint32), ref["acos"])
