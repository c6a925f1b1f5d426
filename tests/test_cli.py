"""GPU tier: the C++ drop-in.  build/rt_cli is the reference's main() with OpenCLRaytracer swapped for
CudaRaytracer (IRaytracer interface, reference SceneLoader / ObjectData / PPMExporter compiled unchanged); it is
built in the build container (tools/build_host.sh needs /root/reference) and travels to the GPU box prebuilt.
The PPM it writes must match the oracle's frame quantised by the PPMExporter rule within 1 LSB per channel."""
import os
import subprocess

import numpy as np
import pytest

from helpers import frame_rays
from pyrt import scenes as S
from pyrt import types as T

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "build", "rt_cli")


def _read_p3(path):
    tok = open(path).read().split()
    assert tok[0] == "P3" and tok[3] == "255"
    w, h = int(tok[1]), int(tok[2])
    return w, h, np.array(tok[4:], dtype=np.int64).reshape(h * w, 3)


@pytest.mark.parametrize("name,mode,flags", [("simpleSphere", 2, []), ("roundedCube", 2, ["--fast-ppm"]),
                                             ("multipleSpheres", 1, ["--mode", "shade", "--upload-rays"]),
                                             ("simpleScene", 2, ["--accel", "bvh", "--fast-ppm"])])
def test_cli_writes_the_reference_image(oracle_mod, tmp_path, name, mode, flags):
    scene = os.path.join(ROOT, "baseline", "_ref", name + ".txt")
    if not (os.path.exists(CLI) and os.path.exists(scene)):
        pytest.skip("build/rt_cli or the staged scene files are missing (tools/build_host.sh needs /root/reference)")
    W, H = 256, 144
    out = str(tmp_path / "render.ppm")
    r = subprocess.run([CLI, "--scene", scene, "--width", str(W), "--height", str(H), "--bounces", "30", "--out", out] + flags,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Scene file loaded without any errors." in r.stdout
    w, h, rgb = _read_p3(out)
    assert (w, h) == (W, H)
    objs, lights = S.load_reference_scene(name)
    _, rays = frame_rays(W, H)
    ref, *_ = oracle_mod.port_render(mode, 30, objs, lights, rays)
    d = np.abs(rgb - T.quantize_ppm(ref).astype(np.int64))
    assert d.max() <= 1
    assert (d != 0).mean() < 1e-3


def test_cli_reports_errors(tmp_path):
    if not os.path.exists(CLI):
        pytest.skip("build/rt_cli missing")
    bad = tmp_path / "bad.txt"
    bad.write_text("material m\n===\nprimative cone m\n")
    r = subprocess.run([CLI, "--scene", str(bad), "--out", "none"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 1


def test_cli_multi_gpu_equals_single_gpu(rt, tmp_path):
    """--gpus N: one process, one context and host thread per device, row-blocks of 16 rows round-robin, every device
    copies its rows into the one pinned frame.  Byte-identical PPM to the 1-GPU run (needs >= 2 GPUs in the box)."""
    import ctypes

    n = ctypes.c_int(0)
    assert rt.load().rt_device_count(ctypes.byref(n)) == 0
    scene = os.path.join(ROOT, "baseline", "_ref", "roundedCube.txt")
    if n.value < 2 or not (os.path.exists(CLI) and os.path.exists(scene)):
        pytest.skip("needs two GPUs, build/rt_cli and the staged scene files")
    outs = []
    for gpus in (1, min(n.value, 8)):
        out = str(tmp_path / ("render_%d.ppm" % gpus))
        r = subprocess.run([CLI, "--scene", scene, "--width", "640", "--height", "356", "--bounces", "30", "--accel", "bvh",
                            "--gpus", str(gpus), "--fast-ppm", "--out", out], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        if gpus > 1:
            assert "Frame sharded over %d GPU(s)." % gpus in r.stdout
        outs.append(open(out, "rb").read())
    assert outs[0] == outs[1]


@pytest.mark.parametrize("name", ["roundedCube", "simpleScene"])
def test_fast_ppm_is_byte_identical_to_the_reference_exporter(tmp_path, name):
    """--fast-ppm (quantisation on the device from the device framebuffer, 3 bytes per pixel back, text
    formatted by all host threads) writes the very bytes PPMExporter::ExportP3 (PPMExporter.cpp:7-30, compiled unchanged)
    writes for the same frame -- odd sizes, the unmodified drop-in constructor (pinhole grid recognised) and the uploaded
    ray vector alike."""
    scene = os.path.join(ROOT, "baseline", "_ref", name + ".txt")
    if not (os.path.exists(CLI) and os.path.exists(scene)):
        pytest.skip("build/rt_cli or the staged scene files are missing")
    files = []
    for flags in ([], ["--fast-ppm"], ["--fast-ppm", "--upload-rays"]):
        out = str(tmp_path / ("r%d.ppm" % len(files)))
        r = subprocess.run([CLI, "--scene", scene, "--width", "333", "--height", "187", "--bounces", "30", "--out", out] + flags,
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
        assert ("generated on the device" in r.stdout) == ("--upload-rays" not in flags)
        files.append(open(out, "rb").read())
    assert files[0] == files[1] == files[2]
    assert files[0].startswith(b"P3\n333 187\n255\n")
