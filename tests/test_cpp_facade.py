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
    p = subprocess.run([exe, "--frames=1", "--width=32", "--height=32", "--out=/tmp/nrc_x.ppm"], capture_output=True, text=True)
    assert p.returncode == 1 and "no CPU fallback" in p.stderr


@pytest.mark.gpu
def test_headless_main_renders(tmp_path):
    exe = _build()
    out = str(tmp_path / "f.ppm")
    pfm = str(tmp_path / "f.pfm")
    p = subprocess.run([exe, "--frames=30", "--width=160", "--height=96", "--out=" + out, "--out-pfm=" + pfm], capture_output=True, text=True,
                       timeout=300)
    assert p.returncode == 0, p.stderr
    assert "NRC MSE Loss" in p.stdout
    raw = open(pfm, "rb").read()
    assert raw.startswith(b"PF\n160 96\n-1.0\n") and len(raw) == len(b"PF\n160 96\n-1.0\n") + 160 * 96 * 12
    data = open(out, "rb").read()
    assert data.startswith(b"P6\n160 96\n255\n") and len(data) == len(b"P6\n160 96\n255\n") + 160 * 96 * 3
    assert len(set(data[15:])) > 8   # not a constant image


@pytest.mark.gpu
def test_split_frame_with_data_parallel_facade_matches_single_submit(tmp_path):
    """FrameTrain / DataParallel::ReduceStep / FrameRender with a world of one rank draws the same frames as Render()."""
    exe = _build()
    a, b = str(tmp_path / "a.ppm"), str(tmp_path / "b.ppm")
    p = subprocess.run([exe, "--frames=12", "--width=96", "--height=64", "--out=" + a], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    p = subprocess.run([exe, "--frames=12", "--width=96", "--height=64", "--out=" + b], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, NRC_SPLIT_FRAME="1"))
    assert p.returncode == 0 and "timed out" not in p.stderr, p.stderr
    assert open(a, "rb").read() == open(b, "rb").read()


@pytest.mark.gpu
def test_checkpoint_written_by_cpp_is_read_by_python_and_back(tmp_path):
    """headless_main saves a checkpoint after 10 frames; the Python reader parses it; a second run resumes from it."""
    from neuralradiancecaching_b200 import checkpoint as K
    exe = _build()
    ck = str(tmp_path / "cache.nrc")
    p = subprocess.run([exe, "--frames=10", "--width=96", "--height=64", "--out=" + str(tmp_path / "a.ppm"), "--save-checkpoint=" + ck],
                       capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    c = K.read_checkpoint(ck)
    assert (c["L"], c["T"], c["F"]) == (16, 16384, 2) and c["params"].size == 20995
    assert c["momentum"].any() and c["tables"].any()
    p = subprocess.run([exe, "--frames=1", "--width=96", "--height=64", "--out=" + str(tmp_path / "b.ppm"), "--load-checkpoint=" + ck],
                       capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr


@pytest.mark.gpu
def test_headless_main_reads_its_data_from_files(tmp_path):
    """--density / --envmap: the C++ readers of include/nrc_host.hpp feed the frame (raw float32 grid, Radiance .hdr), as src/main.cpp:123-139
    does; the fp32 path and the as-shipped resolution table are selectable from the command line."""
    from neuralradiancecaching_b200 import loaders as Ld, scene as S
    exe = _build()
    grid, hdr, out = str(tmp_path / "cloud"), str(tmp_path / "sky.hdr"), str(tmp_path / "f.ppm")
    Ld.WriteFileDensity3D(grid, S.cloud_density())
    open(hdr, "wb").write(Ld.encode_hdr(Ld.float_to_rgbe(S.sky_envmap(256, 128)[..., :3])))
    p = subprocess.run([exe, "--frames=4", "--width=96", "--height=64", "--density=" + grid, "--envmap=" + hdr, "--precision=fp32",
                        "--resolutions=as_shipped_std140", "--stats-every=1", "--out=" + out], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    assert p.stdout.count("NRC MSE Loss") == 4
    data = open(out, "rb").read()
    assert len(set(data[14:])) > 8
    p = subprocess.run([exe, "--frames=1", "--density=" + grid, "--grid=200,200,200"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 1 and "Failed to read" in p.stderr
