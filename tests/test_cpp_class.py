"""The C++ drop-in class include/ORBextractor.h (ORB_SLAM3::ORBextractor, both call operators, getters,
mvImagePyramid): compiled with g++ against liborbfe.so.  CPU: it builds and refuses to run without a CUDA device;
GPU: its outputs equal the oracle's."""
import os
import subprocess

import numpy as np
import pytest

from orbslam3lib_b200 import KP_DTYPE, _lib, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "_build", "orbextractor_main")


def _build():
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    src = os.path.join(ROOT, "tests", "cpp", "orbextractor_main.cpp")
    hdr = os.path.join(ROOT, "include", "ORBextractor.h")
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        libdir = os.path.dirname(_lib.SO_PATH)
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"), src, "-o", EXE,
                               "-L", libdir, "-lorbfe", "-Wl,-rpath," + libdir])
    return EXE


def _run(tmp_path, img, nf, nl, lap):
    raw = tmp_path / "in.raw"
    out = tmp_path / "out.bin"
    raw.write_bytes(img.tobytes())
    h, w = img.shape
    r = subprocess.run([_build(), str(raw), str(w), str(h), str(nf), str(nl), str(lap[0]), str(lap[1]), str(out)],
                       capture_output=True, text=True, timeout=300)
    return r, out


def test_cpp_class_builds_and_has_no_fallback(tmp_path):
    img = scenes.make("texture", 160, 120, 1)
    r, out = _run(tmp_path, img, 200, 4, (0, 0))
    if _lib.lib().orbfe_device_count() == 0:
        assert r.returncode == 3 and "no CUDA device" in r.stderr
    else:
        assert r.returncode == 0, r.stderr


def _read_set(buf, pos):
    n, mono = np.frombuffer(buf, np.int32, 2, pos)
    pos += 8
    kps = np.frombuffer(buf, KP_DTYPE, n, pos).copy()
    pos += 28 * n
    desc = np.frombuffer(buf, np.uint8, 32 * n, pos).reshape(n, 32).copy()
    return int(mono), kps, desc, pos + 32 * n


@pytest.mark.gpu
def test_cpp_class_matches_oracle(tmp_path, oracle):
    nf, nl, lap = 500, 6, (100, 300)
    img = scenes.make("texture", 400, 300, 8101)
    r, out = _run(tmp_path, img, nf, nl, lap)
    assert r.returncode == 0, r.stderr
    buf = out.read_bytes()
    ref = oracle.OracleExtractor(nf, 1.2, nl, 20, 7)
    rmono, rk, rd = ref(img, lap)
    mono, kps, desc, pos = _read_set(buf, 0)
    assert mono == rmono
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        np.testing.assert_array_equal(kps[f], rk[f])
    np.testing.assert_array_equal(desc, rd)
    sf = np.frombuffer(buf, np.float32, nl, pos)
    pos += 4 * nl
    np.testing.assert_array_equal(sf, ref.scale)
    for l in range(nl):
        w, h = np.frombuffer(buf, np.int32, 2, pos)
        pos += 8
        np.testing.assert_array_equal(np.frombuffer(buf, np.uint8, w * h, pos).reshape(h, w), ref.level_pixels(l))
        pos += w * h
    fid = np.frombuffer(buf, np.int32, 1, pos)[0]
    pos += 4
    assert fid == 1
    ml, kl, dl, pos = _read_set(buf, pos)
    mr, kr, dr, pos = _read_set(buf, pos)
    assert ml == rmono
    np.testing.assert_array_equal(dl, rd)
    rmono0, rk0, rd0 = ref(img, (0, 0))
    assert mr == rmono0
    np.testing.assert_array_equal(dr, rd0)
    np.testing.assert_array_equal(kr["x"], rk0["x"])
