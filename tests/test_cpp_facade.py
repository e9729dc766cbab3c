"""The header-only C++ facade (include/nrc_b200.hpp) and examples/headless_main.cpp compile against the C-ABI library;
without a GPU the program fails loudly (no CPU fallback), with one it renders frames."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "neuralradiancecaching_b200")
EXE = os.path.join(ROOT, "examples", "_build", "headless_main")


def _build():
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "headless_main.cpp"), "-o", EXE, "-L" + LIBDIR, "-lnrc_b200",
                           "-Wl,-rpath," + LIBDIR])
    return EXE


def test_facade_compiles_and_fails_loudly_without_gpu():
    import torch
    exe = _build()
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    p = subprocess.run([exe, "1", "32", "32", "/tmp/nrc_x.ppm"], capture_output=True, text=True)
    assert p.returncode == 1 and "no CPU fallback" in p.stderr


@pytest.mark.gpu
def test_headless_main_renders(tmp_path):
    exe = _build()
    out = str(tmp_path / "f.ppm")
    p = subprocess.run([exe, "30", "160", "96", out], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    assert "NRC MSE Loss" in p.stdout
    data = open(out, "rb").read()
    assert data.startswith(b"P6\n160 96\n255\n") and len(data) == len(b"P6\n160 96\n255\n") + 160 * 96 * 3
    assert len(set(data[15:])) > 8   # not a constant image


@pytest.mark.gpu
def test_split_frame_with_data_parallel_facade_matches_single_submit(tmp_path):
    """FrameTrain / DataParallel::ReduceStep / FrameRender with a world of one rank draws the same frames as Render()."""
    exe = _build()
    a, b = str(tmp_path / "a.ppm"), str(tmp_path / "b.ppm")
    p = subprocess.run([exe, "12", "96", "64", a], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    p = subprocess.run([exe, "12", "96", "64", b], capture_output=True, text=True, timeout=300, env=dict(os.environ, NRC_SPLIT_FRAME="1"))
    assert p.returncode == 0 and "timed out" not in p.stderr, p.stderr
    assert open(a, "rb").read() == open(b, "rb").read()


@pytest.mark.gpu
def test_checkpoint_written_by_cpp_is_read_by_python_and_back(tmp_path):
    """headless_main saves a checkpoint after 10 frames; the Python reader parses it; a second run resumes from it."""
    from neuralradiancecaching_b200 import checkpoint as K
    exe = _build()
    ck = str(tmp_path / "cache.nrc")
    env = dict(os.environ, NRC_SAVE_CHECKPOINT=ck)
    p = subprocess.run([exe, "10", "96", "64", str(tmp_path / "a.ppm")], capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0, p.stderr
    c = K.read_checkpoint(ck)
    assert (c["L"], c["T"], c["F"]) == (16, 16384, 2) and c["params"].size == 20995
    assert c["momentum"].any() and c["tables"].any()
    env = dict(os.environ, NRC_LOAD_CHECKPOINT=ck)
    p = subprocess.run([exe, "1", "96", "64", str(tmp_path / "b.ppm")], capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0, p.stderr
