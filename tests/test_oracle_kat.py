"""Known-answer tests derivable from the reference SOURCE alone: the reference
has no tests or golden vectors, so these pin the oracle to what the shader text fixes."""
import numpy as np

from oracle import oracle as O


def test_hash_known_answers():
    # (x*1 + y*19349663 + z*83492791) mod 2^32 mod T  -- nrc-train.comp:254-260
    kat = [((0, 0, 0), 16384, 0), ((1, 0, 0), 16384, 1), ((0, 1, 0), 16384, 159), ((0, 0, 1), 16384, 16311),
           ((1, 1, 1), 16384, 87), ((16, 16, 16), 16384, 1392), ((511, 511, 511), 16384, 11689),
           ((0, 1, 0), 1 << 22, 2572447), ((0, 0, 1), 1 << 22, 3801015), ((1, 1, 1), 1 << 22, 2179159)]
    for (x, y, z), T, want in kat:
        assert O.hash3(x, y, z, T) == want
        assert (x + y * 19349663 + z * 83492791) % (1 << 32) % T == want


def test_resolution_table():
    # src/MRHE.cpp:111-121, minRes 16, maxRes 512, 16 levels
    want = [16, 20, 25, 32, 40, 50, 64, 80, 101, 128, 161, 203, 256, 322, 406, 512]
    assert list(O.mrhe_resolutions(16, 512, 16)) == want


def test_rng_fixed_point_and_first_draw():
    # rand(vec2(0,0)) = fract(sin(0) * 43758.5453) = 0  (nrc-forward.frag:152-155; frame 0)
    assert O.math_eval("rand", [0.0], [0.0])[0] == 0.0
    a, b = np.float32(0.25), np.float32(0.75)
    d = np.float32(np.float32(a * np.float32(12.9898)) + np.float32(b * np.float32(78.233)))
    want = np.float64(np.sin(np.float64(d))) * 43758.5453
    got = float(O.math_eval("rand", [a], [b])[0])
    # the fract of a ~4e4 float keeps 8 fractional bits: compare at that resolution
    assert abs(got - (want - np.floor(want))) < 2.0 ** -7


def test_parameter_counts():
    assert O.PARAM_COUNT == 5 * 4096 + 192 + 5 * 64 + 3 == 20995
    assert O.BIAS_OFFSET[0] == 20672
