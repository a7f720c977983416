"""The C++ host layer (include/ottomarcher.hpp) and the program that plays src/main.rs's role (harness/om_main.cpp):
built with g++ against the C ABI, run as subprocesses.  CPU part: host-side checks and the loud no-device failure;
GPU part: ray/render checks in C++, and om_main's frame against the Python host and the oracle, bit for bit."""
import json
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "harness", "_bin")


@pytest.fixture(scope="module")
def harness(built):
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "harness")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    return BIN


def run(cmd, **kw):
    return subprocess.run(cmd, capture_output=True, text=True, timeout=600, **kw)


def test_cpp_host_checks(harness):
    r = run([os.path.join(harness, "om_test")])
    assert r.returncode == 0, r.stdout + r.stderr
    assert " 0 failed" in r.stdout


def test_om_main_version_and_bad_option(harness):
    r = run([os.path.join(harness, "om_main"), "--version"])
    assert r.returncode == 0 and "sm_100a" in r.stdout
    assert run([os.path.join(harness, "om_main"), "--no-such-option"]).returncode == 2


def test_om_main_fails_loudly_without_a_device(harness):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    r = run([os.path.join(harness, "om_main"), "--width", "32", "--spp", "1", "--quiet"])
    assert r.returncode == 1 and "no CUDA device" in r.stderr


@pytest.mark.gpu
def test_cpp_device_checks(harness):
    r = run([os.path.join(harness, "om_test"), "--gpu"])
    assert r.returncode == 0, r.stdout + r.stderr
    assert " 0 failed" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("fixed", [True, False])
def test_om_main_frame_equals_python_host_and_oracle(harness, api, A, O, tmp_path, fixed):
    from conftest import oracle_scene_from
    W, aspect, spp, seed = 180, 1.5, 12, 3
    H = int(np.float32(W) / np.float32(aspect))
    raw, bmp = str(tmp_path / "px.bin"), str(tmp_path / "frame.bmp")
    cmd = [os.path.join(harness, "om_main"), "--width", str(W), "--aspect", str(aspect), "--spp", str(spp), "--seed", str(seed), "--threads", "2", "--pass", "4",
           "--pixels", raw, "--out", bmp, "--quiet"] + (["--fixed"] if fixed else [])
    r = run(cmd)
    assert r.returncode == 0, r.stdout + r.stderr
    info = json.loads(r.stdout.strip().splitlines()[-1])
    assert info["credited"] == W * H * spp and info["height"] == H
    got = np.fromfile(raw, dtype=A.PIXEL_DTYPE)
    assert got.shape[0] == W * H
    # the same render through the Python host ...
    fz = api.FrozenHittableList.generate("random_scene", seed)
    cam = api.Camera.new((13, 2, 3), (0, 0, 0), (0, 1, 0), 20.0, aspect, 0.1, 10.0)
    p = api.render_params(W, H, spp, seed=seed, flags=0 if fixed else A.RENDER_ADAPTIVE)
    rr = api.Renderer(fz, cam, p)
    for _ in range(spp // 4):
        rr.launch(4)
    want = rr.read_pixels()
    rr.close()
    # ... and through the oracle
    o = oracle_scene_from(O, fz).render(cam, p, nthreads=4)["pixels"]
    for ref in (want, o):
        for f in ("sum", "avg_depth"):
            assert (got[f].view(np.uint32) == np.ascontiguousarray(ref[f]).view(np.uint32)).all(), f
        for f in ("n", "bad_avgs", "bloom", "color"):
            assert (got[f] == ref[f]).all(), f
    assert info["traced"] == int(got["n"].sum())
    with open(bmp, "rb") as f:
        assert f.read(2) == b"BM"
