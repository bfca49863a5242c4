"""The headless CLI (paths_b200/bin/paths_cli, the stand-in for the reference's SDL2 front end, src/main.rs:39-186): same C ABI as the
Python mirror, so its image must be byte-identical to Renderer.render_rgb24() for the same scene, seed and sample count."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
CLI = ROOT / "paths_b200" / "bin" / "paths_cli"


def run(*args):
    return subprocess.run([str(CLI), *map(str, args)], cwd=ROOT, capture_output=True, text=True, timeout=300)


def read_ppm(path):
    data = Path(path).read_bytes()
    assert data[:2] == b"P6"
    parts, pos = [], 2
    while len(parts) < 3:  # width, height, maxval: whitespace-separated ASCII, then one whitespace byte
        while data[pos:pos + 1].isspace():
            pos += 1
        end = pos
        while not data[end:end + 1].isspace():
            end += 1
        parts.append(int(data[pos:end]))
        pos = end
    w, h, mx = parts
    assert mx == 255
    return np.frombuffer(data[pos + 1:], dtype=np.uint8).reshape(h, w, 3)


def test_cli_usage_and_loud_failure_without_a_device():
    import paths_b200 as pb

    assert CLI.exists(), "paths_cli is built by paths_b200/Makefile (__graft_entry__.build)"
    r = run("--help")
    assert r.returncode == 2 and "usage:" in r.stderr
    r = run("scenes/spheres_on_plane.yml", "--bogus", "1")
    assert r.returncode == 2 and "unknown option" in r.stderr
    r = run("scenes/does_not_exist.yml", "--spp", "1")
    assert r.returncode == 1 and "load" in r.stderr
    if pb.device_count() == 0:  # no CPU path: the renderer refuses to start
        r = run("scenes/spheres_on_plane.yml", "--spp", "1", "--out", "/tmp/never.ppm")
        assert r.returncode == 1 and "no CPU path" in r.stderr


@pytest.mark.gpu
def test_cli_image_equals_the_python_mirror(tmp_path):
    import paths_b200 as pb

    out = tmp_path / "a.ppm"
    r = run("scenes/spheres_on_plane.yml", "--spp", "8", "--width", "160", "--height", "96", "--seed", "7", "--out", out)
    assert r.returncode == 0, r.stderr
    assert "Num rays: %d" % (160 * 96 * 8) in r.stdout  # src/main.rs:111 counts camera samples
    img = read_ppm(out)
    sc = pb.Scene.load("scenes/spheres_on_plane.yml", width=160, height=96)
    rr = pb.Renderer(sc.camera, sc, seed=7)
    rr.render_samples(8)
    ref = rr.render_rgb24()
    rr.shutdown()
    assert np.array_equal(img, np.asarray(ref))


@pytest.mark.gpu
def test_cli_checkpoint_resume_and_stress_scene(tmp_path):
    a, b, ck = tmp_path / "a.ppm", tmp_path / "b.ppm", tmp_path / "s.ck"
    common = ["scenes/bokeh_demo.yml", "--width", "120", "--height", "80", "--seed", "3"]
    assert run(*common, "--spp", "12", "--out", a).returncode == 0
    assert run(*common, "--spp", "5", "--checkpoint", ck, "--out", tmp_path / "half.ppm").returncode == 0
    r = run(*common, "--spp", "7", "--resume", ck, "--out", b)
    assert r.returncode == 0, r.stderr
    assert np.array_equal(read_ppm(a), read_ppm(b))  # 5 + 7 resumed samples are the same 12 samples
    # no scene file: the seeded stress scene, like the reference (src/main.rs:47-50)
    r = run("--stress", "50", "--spp", "2", "--width", "64", "--height", "48", "--out", tmp_path / "s.ppm")
    assert r.returncode == 0 and "generating random stress scene" in r.stdout
    assert read_ppm(tmp_path / "s.ppm").shape == (48, 64, 3)
