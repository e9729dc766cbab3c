"""Checkpoint file format (host logic, CPU) and a GPU round trip."""
import numpy as np
import pytest

from neuralradiancecaching_b200 import checkpoint as K


class _Mlp:
    def __init__(self):
        g = np.random.Generator(np.random.PCG64(1))
        self.p = g.standard_normal(20995).astype(np.float32)
        self.m = g.standard_normal(20995).astype(np.float32)
        self.d = None

    def get_params(self):
        return self.p, np.zeros_like(self.p), self.m

    def set_params(self, p):
        self.p = np.array(p, np.float32)

    def set_state(self, delta=None, momentum=None):
        self.d, self.m = delta, np.array(momentum, np.float32)


class _Tab:
    L, T, F = 4, 64, 2

    def __init__(self):
        self.t = np.random.Generator(np.random.PCG64(2)).standard_normal(self.L * self.T * self.F).astype(np.float32)

    def get_tables(self):
        return self.t, np.zeros_like(self.t)

    def set_tables(self, t, d=None):
        self.t = np.array(t, np.float32)


def test_round_trip_and_rejections(tmp_path):
    a, b = _Mlp(), _Tab()
    path = str(tmp_path / "ck.nrc")
    K.save_checkpoint(path, a, b)
    c, d = _Mlp(), _Tab()
    c.p[:] = 0; c.m[:] = 0; d.t[:] = 0
    K.load_checkpoint(path, c, d)
    np.testing.assert_array_equal(c.p.view(np.uint32), a.p.view(np.uint32))
    np.testing.assert_array_equal(c.m.view(np.uint32), a.m.view(np.uint32))
    np.testing.assert_array_equal(d.t.view(np.uint32), b.t.view(np.uint32))
    assert not c.d.any()
    raw = open(path, "rb").read()
    open(path, "wb").write(raw[:-8])                       # cut short
    with pytest.raises(ValueError):
        K.read_checkpoint(path)
    open(path, "wb").write(b"garbage!" + raw[8:])          # wrong magic
    with pytest.raises(ValueError):
        K.read_checkpoint(path)
    open(path, "wb").write(raw)
    d.T = 128                                              # other table shape
    with pytest.raises(ValueError):
        K.load_checkpoint(path, c, d)


@pytest.mark.gpu
def test_gpu_round_trip_gives_the_same_cache(tmp_path):
    import common as Cm
    import neuralradiancecaching_b200 as N
    ctx = N.Context(0)
    nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=3)
    m = N.MRHE(ctx, 0.01, 0.0, seed=4)
    pos, d = Cm.rand_queries(2000, seed=5)
    tgt = np.random.Generator(np.random.PCG64(6)).random((2000, 3)).astype(np.float32)
    for _ in range(3):                                     # a few steps so that momentum is non-trivial
        N.train_accumulate(nrc, m, pos, d, tgt, 1.0 / 2000)
        N.optimizer_step(nrc, m)
    want = N.infer(nrc, m, pos, d)
    path = str(tmp_path / "ck.nrc")
    K.save_checkpoint(path, nrc, m)
    nrc2 = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.5, seed=9)
    m2 = N.MRHE(ctx, 0.01, 0.0, seed=10)
    K.load_checkpoint(path, nrc2, m2)
    np.testing.assert_array_equal(N.infer(nrc2, m2, pos, d).view(np.uint32), want.view(np.uint32))
    # and training continues identically (momentum restored): one more step on both
    for a, b in ((nrc, m), (nrc2, m2)):
        a.set_state(delta=np.zeros(20995, np.float32))
    N.train_accumulate(nrc, m, pos[:64], d[:64], tgt[:64], 1.0 / 64); N.optimizer_step(nrc, m)
    N.train_accumulate(nrc2, m2, pos[:64], d[:64], tgt[:64], 1.0 / 64); N.optimizer_step(nrc2, m2)
    np.testing.assert_array_equal(nrc.get_params()[2].view(np.uint32), nrc2.get_params()[2].view(np.uint32))
    for o in (nrc, m, nrc2, m2):
        o.Destroy()
    ctx.destroy()
