"""Checkpoint of the learnt state (the reference has none -- it retrains from random weights on every start,
src/NeuralRadianceCache.cpp:629-659, src/MRHE.cpp:127-135).

One little-endian file: magic "NRCB200\\0", version, five section sizes, then float32 sections in this order
  MLP params (20 995: W0..W5, b0..b5 -- the C-ABI packing), MLP momentum (20 995), hash tables [L][T][F], table shape, hyper-parameters.
Gradient accumulators are not state (zero between frames).  The C++ facade reads and writes the same layout
(include/nrc_b200.hpp: SaveCheckpoint / LoadCheckpoint)."""
import struct

import numpy as np

MAGIC = b"NRCB200\0"
VERSION = 1


def save_checkpoint(path, nrc, mrhe):
    """nrc: NeuralRadianceCache, mrhe: MRHE (anything with get_params()/get_tables() and L, T, F)."""
    params, _, momentum = nrc.get_params()
    tables, _ = mrhe.get_tables()
    params = np.ascontiguousarray(params, dtype="<f4")
    momentum = np.ascontiguousarray(momentum, dtype="<f4")
    tables = np.ascontiguousarray(tables, dtype="<f4")
    if tables.size != mrhe.L * mrhe.T * mrhe.F:
        raise ValueError("table size does not match L*T*F")
    with open(path, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<5I", VERSION, params.size, tables.size, mrhe.L, mrhe.T))
        f.write(struct.pack("<I", mrhe.F))
        f.write(params.tobytes())
        f.write(momentum.tobytes())
        f.write(tables.tobytes())


def read_checkpoint(path):
    """-> dict(params, momentum, tables, L, T, F); raises ValueError on a file that is not a checkpoint or is cut short."""
    with open(path, "rb") as f:
        head = f.read(len(MAGIC) + 24)
        if len(head) != len(MAGIC) + 24 or head[:len(MAGIC)] != MAGIC:
            raise ValueError("%s: not an nrc-b200 checkpoint" % path)
        version, n_params, n_tables, L, T, F = struct.unpack("<6I", head[len(MAGIC):])
        if version != VERSION:
            raise ValueError("%s: checkpoint version %d, this build reads %d" % (path, version, VERSION))
        if n_tables != L * T * F:
            raise ValueError("%s: inconsistent table shape" % path)
        body = f.read()
    want = 4 * (2 * n_params + n_tables)
    if len(body) != want:
        raise ValueError("%s: %d payload bytes, expected %d" % (path, len(body), want))
    a = np.frombuffer(body, dtype="<f4")
    return dict(params=a[:n_params].copy(), momentum=a[n_params:2 * n_params].copy(), tables=a[2 * n_params:].copy(), L=L, T=T, F=F)


def load_checkpoint(path, nrc, mrhe):
    """Restore params, momentum and tables; the hash-table shape must be the one the MRHE object was created with."""
    c = read_checkpoint(path)
    if (c["L"], c["T"], c["F"]) != (mrhe.L, mrhe.T, mrhe.F):
        raise ValueError("%s: tables are %dx%dx%d, the MRHE object is %dx%dx%d" % (path, c["L"], c["T"], c["F"], mrhe.L, mrhe.T, mrhe.F))
    nrc.set_params(c["params"])
    nrc.set_state(delta=np.zeros_like(c["params"]), momentum=c["momentum"])
    mrhe.set_tables(c["tables"], np.zeros_like(c["tables"]))
    return c
