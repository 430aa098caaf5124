"""bin/ptb_headless: the reference application's frame loop in C++ over host/integrator.hpp (the reference's own
integrator interface - CreateSampleBuffer, CreateBasicRenderer, ResetBasicRenderer, RunBasicRenderer,
RenderSampleBuffer - on top of the C ABI)."""
import json
import os
import subprocess

import numpy as np
import pytest


@pytest.fixture(scope="module")
def exe():
    from path_tracer_b200 import _build
    return _build.build_headless()


def test_headless_builds_and_fails_loudly_without_a_gpu(exe):
    out = subprocess.run([exe, "--help"], stdout=subprocess.PIPE, text=True)
    assert out.returncode == 0 and "--scene" in out.stdout
    assert subprocess.run([exe, "--tone", "nonsense"], stderr=subprocess.PIPE).returncode == 2
    import torch
    if not torch.cuda.is_available():
        out = subprocess.run([exe, "--scene", "cornell", "--size", "32x32", "--spp", "1"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        assert out.returncode == 1 and "CreateDevice" in out.stderr and out.stdout == ""     # no CPU fallback


@pytest.mark.gpu
@pytest.mark.parametrize("spec,name,params", [("cornell", "cornell", ()), ("terrain:48", "terrain", (48,))])
def test_headless_renders_what_the_python_binding_renders(exe, gpu, scenes, tmp_path, spec, name, params):
    from path_tracer_b200 import _build, integrator as I
    from path_tracer_b200 import scene as S
    sc = scenes(name, *params)                                   # also makes sure the spectrum table file exists
    w, h, spp = 96, 64, 6
    film_path, png_path = str(tmp_path / "film.bin"), str(tmp_path / "frame.png")
    log_path = str(tmp_path / "stderr.txt")
    with open(log_path, "wb") as log:                            # (a file, not a pipe: a misbehaving child must not fill this process)
        out = subprocess.run([exe, "--scene", spec, "--size", "%dx%d" % (w, h), "--spp", str(spp), "--max-frames", "2000", "--tone", "hable",
                              "--brightness", "1.5", "--table", _build.SPECTRUM_TABLE, "--film", film_path, "--out", png_path],
                             stdout=subprocess.PIPE, stderr=log, text=True, timeout=120)
    with open(log_path, "rb") as log:
        assert out.returncode == 0, log.read(4000).decode(errors="replace")
    info = json.loads(out.stdout.strip().splitlines()[-1])
    assert info["paths_finished"] >= spp * w * h and info["rays_traced"] == (info["frames"] + 1) * w * h
    a = np.fromfile(film_path, dtype=np.float32).reshape(h, w, 4)
    assert a[..., 3].sum() == info["paths_finished"]
    # the same calls through the ctypes binding: Reset, Run(2), then Run(1) per frame (application.cpp:104-114)
    film = I.SampleBuffer(gpu, w, h)
    r = I.BasicRenderer(gpu, I.DeviceScene(gpu, sc), film)
    r.reset()
    r.run(2)
    for _ in range(info["frames"] - 1):
        r.run(1)
    b = film.download()
    assert np.array_equal(a[..., 3], b[..., 3]) and np.allclose(a, b, rtol=1e-6, atol=1e-7)
    with open(png_path, "rb") as f:
        assert f.read(8) == b"\x89PNG\r\n\x1a\n"
    lin, srgb = film.resolve(1.5, I.TONE_HABLE, 1.0)
    assert os.path.getsize(png_path) > srgb.size // 2            # stored (uncompressed) deflate blocks


@pytest.mark.gpu
def test_headless_renders_one_image_on_two_devices(exe, gpu, scenes, tmp_path):
    """--devices 0,1: scene packed once and replicated, one host thread and one seed block per GPU, fused peer-memory
    reduce + resolve (host/integrator.hpp device_group over ptb_group_create_local).  The combined film is the sum of what
    each GPU renders alone from its own first frame index."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from path_tracer_b200 import _build, integrator as I
    sc = scenes("cornell")
    w, h, spp, max_frames = 96, 64, 8, 2000
    film_path, png_path = str(tmp_path / "film.bin"), str(tmp_path / "frame.png")
    log_path = str(tmp_path / "stderr.txt")
    with open(log_path, "wb") as log:
        out = subprocess.run([exe, "--scene", "cornell", "--size", "%dx%d" % (w, h), "--spp", str(spp), "--max-frames", str(max_frames), "--devices", "0,1",
                              "--table", _build.SPECTRUM_TABLE, "--film", film_path, "--out", png_path], stdout=subprocess.PIPE, stderr=log, text=True, timeout=300)
    with open(log_path, "rb") as log:
        assert out.returncode == 0, log.read(4000).decode(errors="replace")
    info = json.loads(out.stdout.strip().splitlines()[-1])
    assert info["devices"] == 2 and info["paths_finished"] >= spp * w * h
    a = np.fromfile(film_path, dtype=np.float32).reshape(h, w, 4)
    assert a[..., 3].sum() == info["paths_finished"]
    total = np.zeros_like(a)
    d = I.DeviceScene(gpu, sc)
    for rank in range(2):
        film = I.SampleBuffer(gpu, w, h)
        r = I.BasicRenderer(gpu, d, film)
        r.set_frame_index(I.Group.first_frame(rank, 2, max_frames + 64))
        r.reset()
        r.run(2)
        for _ in range(info["frames_by_device"][rank] - 1):
            r.run(1)
        total = total + film.download()
    # sums agree up to float addition order
    assert np.array_equal(a[..., 3], total[..., 3]) and np.allclose(a, total, rtol=1e-6, atol=1e-7)
    with open(png_path, "rb") as f:
        assert f.read(8) == b"\x89PNG\r\n\x1a\n"
