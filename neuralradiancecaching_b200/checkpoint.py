"""Checkpoint of the learnt state (the reference has none -- it retrains from random weights on every start,
src/NeuralRadianceCache.cpp:629-659, src/MRHE.cpp:127-135).

One little-endian file: magic "NRCB200\\0", uint32 {version, nParams, nTables, L, T, F}, then float32 sections in this order
  MLP params (20 995: W0..W5, b0..b5 -- the C-ABI packing), MLP momentum (20 995), hash tables [L][T][F].
Gradient accumulators are not state (zero between frames).  NOT stored: learning rates / weight decay / beta, the resolution table
and its mode (intended vs as_shipped_std140) and the fixed-preset switch -- they are constructor arguments of the objects the file is
loaded INTO; a file trained under one resolution table or preset loads into objects built with another (the shapes match) and then
decodes differently, so construct the objects as they were when the file was written.  The C++ facade reads and writes the same layout
(include/nrc_b200.hpp: SaveCheckpoint / LoadCheckpoint).

Version 2 = version 1 + a trailer with the Adam state of the opt-in fixed preset, written only when Adam is switched on:
  uint32 flags (bit 0 MLP, bit 1 tables); if bit 0: uint32 steps, second moment (20 995); if bit 1: uint32 steps, first and second
  moments (L*T*F each).  Files without Adam state stay version 1."""
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
    flags = (1 if getattr(nrc, "adam", False) else 0) | (2 if getattr(mrhe, "adam", False) else 0)
    with open(path, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<5I", 2 if flags else VERSION, params.size, tables.size, mrhe.L, mrhe.T))
        f.write(struct.pack("<I", mrhe.F))
        f.write(params.tobytes())
        f.write(momentum.tobytes())
        f.write(tables.tobytes())
        if flags:
            f.write(struct.pack("<I", flags))
            if flags & 1:
                second, steps = nrc.get_adam_state()
                f.write(struct.pack("<I", steps))
                f.write(np.ascontiguousarray(second, dtype="<f4").tobytes())
            if flags & 2:
                m1, m2, steps = mrhe.get_adam_state()
                f.write(struct.pack("<I", steps))
                f.write(np.ascontiguousarray(m1, dtype="<f4").tobytes())
                f.write(np.ascontiguousarray(m2, dtype="<f4").tobytes())


def read_checkpoint(path):
    """-> dict(params, momentum, tables, L, T, F); raises ValueError on a file that is not a checkpoint or is cut short."""
    with open(path, "rb") as f:
        head = f.read(len(MAGIC) + 24)
        if len(head) != len(MAGIC) + 24 or head[:len(MAGIC)] != MAGIC:
            raise ValueError("%s: not an nrc-b200 checkpoint" % path)
        version, n_params, n_tables, L, T, F = struct.unpack("<6I", head[len(MAGIC):])
        if version not in (VERSION, 2):
            raise ValueError("%s: checkpoint version %d, this build reads 1 and 2" % (path, version))
        if n_tables != L * T * F:
            raise ValueError("%s: inconsistent table shape" % path)
        body = f.read()
    want = 4 * (2 * n_params + n_tables)
    out = dict(L=L, T=T, F=F, adam_mlp=None, adam_mrhe=None)
    if version == 2:
        if len(body) < want + 4:
            raise ValueError("%s: version 2 file without its trailer" % path)
        flags, = struct.unpack_from("<I", body, want)
        off = want + 4
        if flags & 1:
            steps, = struct.unpack_from("<I", body, off)
            off += 4
            second = np.frombuffer(body, dtype="<f4", count=n_params, offset=off).copy() if len(body) >= off + 4 * n_params else None
            off += 4 * n_params
            out["adam_mlp"] = (second, steps)
        if flags & 2:
            steps, = struct.unpack_from("<I", body, off)
            off += 4
            ok = len(body) >= off + 8 * n_tables
            m1 = np.frombuffer(body, dtype="<f4", count=n_tables, offset=off).copy() if ok else None
            m2 = np.frombuffer(body, dtype="<f4", count=n_tables, offset=off + 4 * n_tables).copy() if ok else None
            off += 8 * n_tables
            out["adam_mrhe"] = (m1, m2, steps)
        if len(body) != off or (out["adam_mlp"] and out["adam_mlp"][0] is None) or (out["adam_mrhe"] and out["adam_mrhe"][0] is None):
            raise ValueError("%s: %d payload bytes, expected %d" % (path, len(body), off))
    elif len(body) != want:
        raise ValueError("%s: %d payload bytes, expected %d" % (path, len(body), want))
    a = np.frombuffer(body, dtype="<f4", count=2 * n_params + n_tables)
    out.update(params=a[:n_params].copy(), momentum=a[n_params:2 * n_params].copy(), tables=a[2 * n_params:].copy())
    return out


def load_checkpoint(path, nrc, mrhe):
    """Restore params, momentum and tables; the hash-table shape must be the one the MRHE object was created with."""
    c = read_checkpoint(path)
    if (c["L"], c["T"], c["F"]) != (mrhe.L, mrhe.T, mrhe.F):
        raise ValueError("%s: tables are %dx%dx%d, the MRHE object is %dx%dx%d" % (path, c["L"], c["T"], c["F"], mrhe.L, mrhe.T, mrhe.F))
    nrc.set_params(c["params"])
    nrc.set_state(delta=np.zeros_like(c["params"]), momentum=c["momentum"])
    mrhe.set_tables(c["tables"], np.zeros_like(c["tables"]))
    # Adam state is restored into objects that have Adam switched on (set_adam before load_checkpoint); otherwise it is left in `c`
    if c["adam_mlp"] is not None and getattr(nrc, "adam", False):
        nrc.set_adam_state(*c["adam_mlp"])
    if c["adam_mrhe"] is not None and getattr(mrhe, "adam", False):
        mrhe.set_adam_state(*c["adam_mrhe"])
    return c
