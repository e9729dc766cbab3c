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


class _AdamMlp(_Mlp):
    adam = True

    def __init__(self):
        super().__init__()
        self.v = np.random.Generator(np.random.PCG64(7)).random(20995).astype(np.float32)
        self.steps = 17

    def get_adam_state(self):
        return self.v, self.steps

    def set_adam_state(self, v, steps):
        self.v, self.steps = np.array(v, np.float32), steps


class _AdamTab(_Tab):
    adam = True

    def __init__(self):
        super().__init__()
        g = np.random.Generator(np.random.PCG64(8))
        self.m1 = g.standard_normal(self.t.size).astype(np.float32)
        self.m2 = g.random(self.t.size).astype(np.float32)
        self.steps = 23

    def get_adam_state(self):
        return self.m1, self.m2, self.steps

    def set_adam_state(self, m1, m2, steps):
        self.m1, self.m2, self.steps = np.array(m1, np.float32), np.array(m2, np.float32), steps


def test_version_2_carries_the_adam_moments(tmp_path):
    path = str(tmp_path / "ck2.nrc")
    a, b = _AdamMlp(), _AdamTab()
    K.save_checkpoint(path, a, b)
    raw = open(path, "rb").read()
    assert raw[8:12] == (2).to_bytes(4, "little")
    c, d = _AdamMlp(), _AdamTab()
    c.v[:] = 0; c.steps = 0; d.m1[:] = 0; d.m2[:] = 0; d.steps = 0
    K.load_checkpoint(path, c, d)
    np.testing.assert_array_equal(c.v.view(np.uint32), a.v.view(np.uint32))
    np.testing.assert_array_equal(d.m1.view(np.uint32), b.m1.view(np.uint32))
    np.testing.assert_array_equal(d.m2.view(np.uint32), b.m2.view(np.uint32))
    assert (c.steps, d.steps) == (17, 23)
    # objects without Adam take the version-1 part of a version-2 file and leave the moments in the returned dict
    e, f = _Mlp(), _Tab()
    got = K.load_checkpoint(path, e, f)
    np.testing.assert_array_equal(e.p.view(np.uint32), a.p.view(np.uint32))
    assert got["adam_mlp"][1] == 17 and got["adam_mrhe"][2] == 23
    # only one of the two with Adam: flags = 2
    K.save_checkpoint(path, _Mlp(), b)
    got = K.read_checkpoint(path)
    assert got["adam_mlp"] is None and got["adam_mrhe"][2] == 23
    # a version-1 writer output is still version 1 and still loads
    K.save_checkpoint(path, _Mlp(), _Tab())
    assert open(path, "rb").read()[8:12] == (1).to_bytes(4, "little")
    assert K.read_checkpoint(path)["adam_mlp"] is None
    # truncated trailer
    K.save_checkpoint(path, a, b)
    raw = open(path, "rb").read()
    open(path, "wb").write(raw[:-4])
    with pytest.raises(ValueError):
        K.read_checkpoint(path)
    open(path, "wb").write(raw + b"\0\0\0\0")
    with pytest.raises(ValueError):
        K.read_checkpoint(path)


@pytest.mark.gpu
def test_gpu_adam_state_round_trip_resumes_training(tmp_path):
    import common as Cm
    import neuralradiancecaching_b200 as N
    ctx = N.Context(0)

    def make(s):
        nrc = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.9, seed=s)
        m = N.MRHE(ctx, 0.01, 0.0, seed=s + 1)
        m.set_fixed_mode(True)
        nrc.set_adam(True)
        m.set_adam(True)
        return nrc, m
    nrc, m = make(3)
    pos, d = Cm.rand_queries(2000, seed=5)
    tgt = np.random.Generator(np.random.PCG64(6)).random((2000, 3)).astype(np.float32)
    for _ in range(3):
        N.train_accumulate(nrc, m, pos, d, tgt, 1.0 / 2000)
        N.optimizer_step(nrc, m)
    path = str(tmp_path / "ck2.nrc")
    K.save_checkpoint(path, nrc, m)
    nrc2, m2 = make(11)
    K.load_checkpoint(path, nrc2, m2)
    assert nrc2.get_adam_state()[1] == 3 and m2.get_adam_state()[2] == 3
    # the restored state is the saved state, bit for bit
    np.testing.assert_array_equal(nrc.get_adam_state()[0].view(np.uint32), nrc2.get_adam_state()[0].view(np.uint32))
    for k in (0, 1):
        np.testing.assert_array_equal(m.get_adam_state()[k].view(np.uint32), m2.get_adam_state()[k].view(np.uint32))
    for a, b in ((nrc, m), (nrc2, m2)):
        N.train_accumulate(a, b, pos[:512], d[:512], tgt[:512], 1.0 / 512)
        N.optimizer_step(a, b)
    # one more step on both: the table gradients are float atomics (summation order differs between launches), so a few ulp
    np.testing.assert_allclose(nrc.get_params()[0], nrc2.get_params()[0], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(m.get_tables()[0], m2.get_tables()[0], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(m.get_adam_state()[1], m2.get_adam_state()[1], rtol=1e-5, atol=1e-12)
    assert nrc2.get_adam_state()[1] == 4 and m2.get_adam_state()[2] == 4
    # the getters refuse when Adam is off
    nrc3 = N.NeuralRadianceCache(ctx, 0.001, 0.0, 0.9, seed=1)
    with pytest.raises(Exception):
        nrc3.get_adam_state()
    for o in (nrc, m, nrc2, m2, nrc3):
        o.Destroy()
    ctx.destroy()


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
