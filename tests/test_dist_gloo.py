"""world_size-2 gloo test (CPU) of the data-parallel protocol: interleaved shards of the train grid, all-reduce of the
packed MLP gradient arena (+ loss) and of the delta-table head, identical steps on both ranks == the single-rank result.
The compute stand-in on CPU is the oracle (test infrastructure); on GPUs the same protocol runs in bench.py."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TW, TH = 8, 6


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    import common as Cm
    from neuralradiancecaching_b200 import dist as D
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = Cm.oracle_scene(Cm.scene_cfg())
    sc.set_train_size(TW, TH)
    n = TW * TH
    mine = list(range(rank, n, world))
    assert len(mine) == D.strided_count(rank, n, world)
    recs = np.concatenate([sc.train_trace(TW, TH, i, i + 1)[0] for i in mine])
    sc.train_accumulate(recs[:, 0:3], recs[:, 3:6], recs[:, 6:9])
    _, g_mlp, _ = sc.get_mlp()
    _, g_tab = sc.get_mrhe()
    loss = np.array([sc.mse_loss()], np.float32)
    head = g_tab[:2 * 16384].copy()
    D.allreduce_host([g_mlp, head, loss])
    g_tab[:2 * 16384] = head
    sc.set_mlp_state(delta=g_mlp)
    sc.set_mrhe_delta(g_tab)
    sc.nrc_step(); sc.mrhe_step()
    np.save(os.path.join(out, "p%d.npy" % rank), sc.get_mlp()[0])
    np.save(os.path.join(out, "t%d.npy" % rank), sc.get_mrhe()[0][:2 * 16384])
    np.save(os.path.join(out, "l%d.npy" % rank), loss)
    dist.destroy_process_group()


def test_two_rank_protocol_matches_single_rank(tmp_path):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import common as Cm
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    sc = Cm.oracle_scene(Cm.scene_cfg())
    sc.set_train_size(TW, TH)
    rec, _ = sc.train_trace(TW, TH)
    sc.train_accumulate(rec[:, 0:3], rec[:, 3:6], rec[:, 6:9])
    want_loss = sc.mse_loss()
    sc.nrc_step(); sc.mrhe_step()
    want_p, want_t = sc.get_mlp()[0], sc.get_mrhe()[0][:2 * 16384]
    p0, p1 = np.load(tmp_path / "p0.npy"), np.load(tmp_path / "p1.npy")
    t0, t1 = np.load(tmp_path / "t0.npy"), np.load(tmp_path / "t1.npy")
    np.testing.assert_array_equal(p0.view(np.uint32), p1.view(np.uint32))      # replicas stay bit-identical
    np.testing.assert_array_equal(t0.view(np.uint32), t1.view(np.uint32))
    np.testing.assert_allclose(p0, want_p, rtol=0, atol=1e-7 * max(1.0, float(np.abs(want_p).max())) + 1e-9)
    np.testing.assert_allclose(t0, want_t, rtol=0, atol=1e-6)
    assert abs(float(np.load(tmp_path / "l0.npy")[0]) - want_loss) <= 1e-5 * abs(want_loss) + 1e-12


def test_weak_scaling_frame_shapes():
    from neuralradiancecaching_b200 import dist as D
    assert D.frame_shape(1) == (1920, 1080, 100, 100)
    assert D.frame_shape(4) == (3840, 2160, 100, 400)      # BASELINE config 4
    for n in (2, 8):
        w, h, tw, th = D.frame_shape(n)
        assert abs(w * h / (1920 * 1080) - n) < 0.02 * n and tw * th == 10000 * n
    rows = [D.shard_rows(k, 3, 10) for k in range(3)]
    assert sorted(np.concatenate(rows).tolist()) == list(range(10))


def _gather_worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist
    from neuralradiancecaching_b200 import dist as D
    dist.init_process_group("gloo", rank=rank, world_size=world)
    H, W = 7, 5                                   # odd height: rank 0 has one row more than rank 1
    frame = torch.arange(H * W * 4, dtype=torch.int32).reshape(H, W, 4)
    part = frame[rank::world].contiguous()        # what nrc_read_frame_partition hands a rank
    assert part.shape[0] == D.strided_count(rank, H, world)
    full = D.gather_rows(part, H, rank, world)
    if rank == 0:
        np.save(os.path.join(out, "full.npy"), full.numpy())
    else:
        assert full is None
    dist.destroy_process_group()


def test_frame_assembly_from_interleaved_rows(tmp_path):
    port = 31500 + (os.getpid() % 2000)
    mp.spawn(_gather_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    full = np.load(tmp_path / "full.npy")
    np.testing.assert_array_equal(full, np.arange(7 * 5 * 4, dtype=np.int32).reshape(7, 5, 4))
