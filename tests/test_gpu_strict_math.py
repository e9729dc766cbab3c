"""-m gpu: the guard-free division / square-root fast paths of csrc/strict.cuh (div_normal, sqrt_normal) against __fdiv_rn /
__fsqrt_rn, exhaustively over the operand ranges their call sites can see (all 2^32 coordinates for the grid lookup)."""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu


def test_fast_paths_equal_the_ieee_routines_exhaustively():
    import neuralradiancecaching_b200 as N
    from neuralradiancecaching_b200 import _lib as L
    ctx = N.Context(0)
    bad = (C.c_uint64 * 4)()
    L.check(L.lib().nrc_selftest_strict_math(ctx.h, bad))
    assert list(bad) == [0, 0, 0, 0], "grid coordinate / sqrt / reciprocal / HG divisions: %s mismatches" % list(bad)
    ctx.destroy()
