"""-m gpu: the Lynx wire boundary on the GPU — the four FastRPC skel entry points the reference binds
(orbslam3_open / configure / extractFeatures / bfMatchStereo / close; reference: dsp/src/orbslam_dsp.cpp:866,931,567,
1017-1021, called from LynxHardwareAccelerator.cpp:62,116,146-150) and the LynxHardwareAccelerator class above them
(Python mirror orbslam3lib_b200/wire.py, C++ header include/LynxHardwareAccelerator.h).

Checked against the CPU oracle: side-by-side de-interleave (stride > 2 * width included), integer X / Y, level, the
(int8 sin*64) << 8 | (int8 cos*64) angle word of LHA.cpp:168-201, descriptors, total* / mono*, the equal- and the
different-lapping-area branch, the stereo sub-ranges handed to the matcher, int16 indices / distances, the capacity and
nr > 32767 errors, and the frame-id -> 3-slot matching cache (LHA.cpp:137,207)."""
import os
import subprocess

import numpy as np
import pytest

from orbslam3lib_b200 import KP_DTYPE, LynxHardwareAccelerator, _lib, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _side_by_side(kind, w, h, seed, stride=None):
    left = scenes.make(kind, w, h, seed)
    right = scenes.stereo_right(left, seed)
    stride = stride or 2 * w
    frame = np.full((h, stride), 0xA5, np.uint8)    # bytes past 2 * w must not be read as pixels
    frame[:, :w] = left
    frame[:, w:2 * w] = right
    return frame, left, right


def _angle_word(angle_deg):
    rad = angle_deg.astype(np.float64) * np.pi / 180.0
    c8 = np.rint(np.cos(rad) * 64.0).astype(np.int64)
    s8 = np.rint(np.sin(rad) * 64.0).astype(np.int64)
    return (((s8 & 0xFF) << 8) | (c8 & 0xFF)).astype(np.int32)


def _expect(oracle, left, right, nf, lapL, lapR):
    out = []
    for img, lap in ((left, lapL), (right, lapR)):
        mono, k, d = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)(img, lap)
        out.append((mono, k, d))
    (ml, kl, dl), (mr, kr, dr) = out
    idx, dist = oracle.knn2_hamming(dl[ml:], dr[mr:]) if len(dl) > ml else (np.zeros((0, 2), np.int32),) * 2
    return out, idx, dist


def _check_eye(bufs, total, mono, exp, angle_slack=0):
    emono, ek, ed = exp
    X, Y, A, Lv, D = bufs
    n = len(ek)
    assert (total, mono) == (n, emono)
    np.testing.assert_array_equal(X[:n], ek["x"].astype(np.int32))    # the wire carries integer pixels (LHA.cpp:170-171)
    np.testing.assert_array_equal(Y[:n], ek["y"].astype(np.int32))
    np.testing.assert_array_equal(Lv[:n], ek["octave"])
    np.testing.assert_array_equal(D[:n], ed)
    want = _angle_word(ek["angle"])
    if angle_slack == 0:
        np.testing.assert_array_equal(A[:n], want)
    # the CUDA angle may differ from the oracle's by <= 1e-3 degrees: a word may then differ by one int8 step
    c = ((A[:n] & 0xFF).astype(np.int8).astype(int) - (want & 0xFF).astype(np.int8).astype(int))
    s = (((A[:n] >> 8) & 0xFF).astype(np.int8).astype(int) - ((want >> 8) & 0xFF).astype(np.int8).astype(int))
    assert np.abs(c).max(initial=0) <= 1 and np.abs(s).max(initial=0) <= 1
    assert (A[:n] != want).mean() <= 0.01
    assert not X[n:n + 16].any() and not D[n:n + 4].any(), "entries past total* must stay untouched"


@pytest.mark.gpu
@pytest.mark.parametrize("kind,w,h,seed,nf,stride,lapL,lapR", [
    ("texture", 640, 400, 9101, 1000, None, (0, 0), (0, 0)),                 # the Lynx frame: 1280 x 400, no lapping
    ("texture", 640, 400, 9102, 1500, None, (150, 489), (150, 489)),         # equal lapping areas: one batched call
    ("texture", 752, 480, 9103, 1200, 1600, (290, 751), (0, 461)),           # stride > 2 * width, L != R: two calls
    ("checker", 640, 400, 9104, 1000, 1344, (0, 300), (320, 400)),
])
def test_extract_features_wire(oracle, kind, w, h, seed, nf, stride, lapL, lapR):
    frame, left, right = _side_by_side(kind, w, h, seed, stride)
    lha = LynxHardwareAccelerator(w, h, nf)
    lha.StoreInputBuffer(frame)
    rc, tl, tr, ml, mr = lha.extract_raw(lapL, lapR, slot=1)
    assert rc == _lib.OK
    (el, er), idx, dist = _expect(oracle, left, right, nf, lapL, lapR)
    _check_eye(lha.left, tl, ml, el, angle_slack=1)
    _check_eye(lha.right, tr, mr, er, angle_slack=1)
    ns = tl - ml
    assert ns == len(idx)
    I, D1, D2 = lha.cache[1]
    np.testing.assert_array_equal(I[:ns], idx[:, 0].astype(np.int16))
    np.testing.assert_array_equal(D1[:ns], dist[:, 0].astype(np.int16))
    np.testing.assert_array_equal(D2[:ns], dist[:, 1].astype(np.int16))
    if lapL != (0, 0):
        assert 0 < ns < tl, "the case must have both mono and stereo keypoints"
    lha.close()


@pytest.mark.gpu
def test_extract_orb_and_cache_slots(oracle):
    """ExtractORB x 4 then BFMatchORB(id): ids 2..4 are still cached (3 slots), id 1 was overwritten by id 4."""
    w, h, nf, lap = 640, 400, 800, (120, 520)
    lha = LynxHardwareAccelerator(w, h, nf)
    want = {}
    for f in range(4):
        frame, left, right = _side_by_side("texture", w, h, 9200 + f)
        lha.StoreInputBuffer(frame)
        fid, kl, kr, dl, dr, ml, mr = lha.ExtractORB(lap, lap)
        assert fid == f + 1
        (el, er), idx, dist = _expect(oracle, left, right, nf, lap, lap)
        assert (ml, mr) == (el[0], er[0]) and len(kl) == len(el[1]) and len(kr) == len(er[1])
        np.testing.assert_array_equal(dl, el[2])
        np.testing.assert_array_equal(dr, er[2])
        np.testing.assert_array_equal(kl["x"], np.trunc(el[1]["x"]))
        np.testing.assert_array_equal(kl["octave"], el[1]["octave"])
        # angle after the int8 sin / cos round trip: within one quantisation step (atan(1/64) < 0.9 degrees) of the oracle's
        d = np.abs((kl["angle"] % 360.0) - el[1]["angle"])
        assert np.minimum(d, 360.0 - d).max() < 1.0
        want[fid] = (len(idx), idx[:, 0].astype(np.uint16), dist[:, 0].astype(np.uint16), dist[:, 1].astype(np.uint16))
    for fid in (2, 3, 4):
        n, i, d1, d2 = want[fid]
        gi, g1, g2 = lha.BFMatchORB(fid, n)
        np.testing.assert_array_equal(gi, i)
        np.testing.assert_array_equal(g1, d1)
        np.testing.assert_array_equal(g2, d2)
    n1 = want[1][0]
    gi, g1, g2 = lha.BFMatchORB(1, min(n1, want[4][0]))
    np.testing.assert_array_equal(gi, want[4][1][:len(gi)])    # slot (1 + 3) % 3 == 4 % 3: frame 4 overwrote frame 1
    lha.close()


@pytest.mark.gpu
def test_bf_match_stereo_wire(oracle):
    q, t = scenes.map_descriptors(700, 1300, 9301)
    lha = LynxHardwareAccelerator(64, 64, 100)
    rc, i, d1, d2 = lha.bfMatchStereo(q, t)
    assert rc == _lib.OK
    idx, dist = oracle.knn2_hamming(q, t)
    np.testing.assert_array_equal(i, idx[:, 0].astype(np.int16))
    np.testing.assert_array_equal(d1, dist[:, 0].astype(np.int16))
    np.testing.assert_array_equal(d2, dist[:, 1].astype(np.int16))
    # a single train row: no second neighbour -> distances2 = -1, like knnMatch returning one DMatch
    rc, i, d1, d2 = lha.bfMatchStereo(q[:5], t[:1])
    assert rc == _lib.OK and (i == 0).all() and (d2 == -1).all()
    # empty query is a no-op
    rc, i, d1, d2 = lha.bfMatchStereo(q[:0], t)
    assert rc == _lib.OK
    # output arrays shorter than the query rows: ORBFE_ERR_CAPACITY, nothing written
    rc, i, d1, d2 = lha.bfMatchStereo(q, t, cap=699)
    assert rc == _lib.ERR_CAPACITY and (i == -7).all()
    # more train rows than an int16 index can address (IDL: sequence<short> indices): ORBFE_ERR_ARG
    big = np.zeros((32768, 32), np.uint8)
    rc, i, d1, d2 = lha.bfMatchStereo(q[:4], big)
    assert rc == _lib.ERR_ARG
    rc, i, d1, d2 = lha.bfMatchStereo(q[:4], big[:32767])
    assert rc == _lib.OK
    lha.close()


@pytest.mark.gpu
def test_extract_features_argument_errors():
    w, h = 320, 240
    lha = LynxHardwareAccelerator(w, h, 300)
    frame, _, _ = _side_by_side("texture", w, h, 9401)
    lha.StoreInputBuffer(frame)
    lha.width = 400                                  # stride (640) < 2 * width
    assert lha.extract_raw((0, 0), (0, 0), 0)[0] == _lib.ERR_ARG
    lha.width = w
    lha.inputImage = frame[:h - 1]                   # imageLen too short for `height` rows
    lha.height = h
    assert lha.extract_raw((0, 0), (0, 0), 0)[0] == _lib.ERR_ARG
    lha.close()


EXE = os.path.join(ROOT, "tests", "cpp", "_build", "lynx_main")


def _build():
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    src = os.path.join(ROOT, "tests", "cpp", "lynx_main.cpp")
    hdrs = [os.path.join(ROOT, "include", f) for f in ("LynxHardwareAccelerator.h", "ORBextractor.h", "orbfe_c.h")]
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < max(os.path.getmtime(p) for p in [src] + hdrs):
        libdir = os.path.dirname(_lib.SO_PATH)
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"), src, "-o", EXE,
                               "-L", libdir, "-lorbfe", "-Wl,-rpath," + libdir])
    return EXE


def _run_cpp(tmp_path, frames, w, h, nf, lapL, lapR):
    raw, out = tmp_path / "frames.raw", tmp_path / "out.bin"
    raw.write_bytes(np.ascontiguousarray(frames).tobytes())
    args = [_build(), str(raw), str(w), str(h), str(len(frames)), str(nf)] + [str(v) for v in (*lapL, *lapR)] + [str(out)]
    return subprocess.run(args, capture_output=True, text=True, timeout=300), out


def test_cpp_lynx_class_builds_and_has_no_fallback(tmp_path):
    frame, _, _ = _side_by_side("texture", 160, 120, 1)
    r, out = _run_cpp(tmp_path, frame[None], 160, 120, 200, (0, 0), (0, 0))
    if _lib.lib().orbfe_device_count() == 0:
        assert r.returncode == 3 and "no CUDA device" in r.stderr
    else:
        assert r.returncode == 0, r.stderr


def _read_eye(buf, pos):
    n, count, mono = np.frombuffer(buf, np.int32, 3, pos)
    pos += 12
    kps = np.frombuffer(buf, KP_DTYPE, n, pos).copy()
    pos += 28 * n
    desc = np.frombuffer(buf, np.uint8, 32 * n, pos).reshape(n, 32).copy()
    return int(count), int(mono), kps, desc, pos + 32 * n


@pytest.mark.gpu
def test_cpp_lynx_class_matches_oracle(tmp_path, oracle):
    """include/LynxHardwareAccelerator.h through a g++-built driver: 5 frames, then BFMatchORB for every id."""
    w, h, nf, lapL, lapR, nframes = 640, 400, 1000, (150, 489), (100, 400), 5
    made = [_side_by_side("texture", w, h, 9500 + f) for f in range(nframes)]
    r, out = _run_cpp(tmp_path, np.stack([m[0] for m in made]), w, h, nf, lapL, lapR)
    assert r.returncode == 0, r.stderr
    buf, pos, knn = out.read_bytes(), 0, []
    for f in range(nframes):
        fid = int(np.frombuffer(buf, np.int32, 1, pos)[0])
        pos += 4
        assert fid == f + 1
        (el, er), idx, dist = _expect(oracle, made[f][1], made[f][2], nf, lapL, lapR)
        knn.append((idx, dist))
        for exp in (el, er):
            count, mono, kps, desc, pos = _read_eye(buf, pos)
            assert (count, mono, len(kps)) == (len(exp[1]), exp[0], len(exp[1]))
            np.testing.assert_array_equal(desc, exp[2])
            np.testing.assert_array_equal(kps["x"], np.trunc(exp[1]["x"]))
            np.testing.assert_array_equal(kps["y"], np.trunc(exp[1]["y"]))
            np.testing.assert_array_equal(kps["octave"], exp[1]["octave"])
            d = np.abs((kps["angle"] % 360.0) - exp[1]["angle"])
            assert np.minimum(d, 360.0 - d).max() < 1.0
    for f in range(nframes):
        n = int(np.frombuffer(buf, np.int32, 1, pos)[0])
        pos += 4
        got = np.frombuffer(buf, np.uint16, 3 * n, pos).reshape(3, n)
        pos += 6 * n
        src = f if f >= nframes - 3 else f + 3    # ids 1, 2 were overwritten by ids 4, 5 (slot = id % 3)
        idx, dist = knn[src]
        m = min(n, len(idx))
        assert f < nframes - 3 or n == len(idx)
        np.testing.assert_array_equal(got[0, :m], idx[:m, 0].astype(np.uint16))
        np.testing.assert_array_equal(got[1, :m], dist[:m, 0].astype(np.uint16))
        np.testing.assert_array_equal(got[2, :m], dist[:m, 1].astype(np.uint16))
    assert pos == len(buf)
