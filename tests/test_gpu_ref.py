"""-m gpu: the CUDA extractor, through the C ABI, against the REFERENCE'S OWN extractor (oracle/_ref/libext_old.so =
ORBextractor_old.cc compiled from /root/reference by oracle/ref_ext/make_ref_ext.py; the built library travels to the GPU
box, /root/reference is not read here).  Same bars as test_gpu_extractor.py: keypoints (set, order, every field) and
descriptors bit-exact, angles within 1e-3 degrees, padded pyramid levels bit-exact."""
import numpy as np
import pytest

import ref_ext_lib
from orbslam3lib_b200 import ORBextractor, scenes

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_ext_lib.available(), reason="oracle/_ref/libext_old.so not in the snapshot")]

ANGLE_TOL_DEG = 1e-3

CASES = [
    ("texture", 640, 480, 1000, 1000, 8, (0, 0)),        # BASELINE configs[0]
    ("texture", 752, 480, 1001, 1200, 8, (0, 0)),        # configs[1] image shape
    ("checker", 640, 480, 1002, 1000, 8, (0, 0)),
    ("noise", 752, 480, 1006, 1200, 8, (0, 0)),
    ("flat_blobs", 752, 480, 1004, 1000, 8, (0, 0)),
    ("texture", 640, 480, 1015, 500, 8, (0, 0)),         # configs[2] shape
    ("texture", 640, 400, 1016, 1500, 8, (150, 489)),    # Lynx fisheye frame with a lapping area
    ("texture", 1280, 720, 1007, 2000, 10, (0, 0)),
]


@pytest.mark.parametrize("kind,w,h,seed,nf,nl,lap", CASES)
def test_cuda_extractor_equals_reference_build(kind, w, h, seed, nf, nl, lap):
    img = scenes.make(kind, w, h, seed)
    ref = ref_ext_lib.RefExtractor(nf, 1.2, nl, 20, 7)
    mono_r, kps_r, desc_r = ref(img, lap)
    ex = ORBextractor(nf, 1.2, nl, 20, 7)
    mono_g, kps_g, desc_g = ex(img, lap)
    assert mono_g == mono_r and len(kps_g) == len(kps_r)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        np.testing.assert_array_equal(kps_g[f], kps_r[f], "keypoint field %s" % f)
    d = np.abs(kps_g["angle"] - kps_r["angle"])
    assert np.minimum(d, 360.0 - d).max() <= ANGLE_TOL_DEG
    np.testing.assert_array_equal(desc_g, desc_r, "descriptors")
    for l in range(nl):
        np.testing.assert_array_equal(ex.level_padded(l), ref.level_padded(l), "ComputePyramid level %d" % l)
        kr, kg = ref.level_keypoints(l), ex.level_keypoints(l)
        assert len(kg) == len(kr)
        for f in ("x", "y", "response", "octave", "size"):
            np.testing.assert_array_equal(kg[f], kr[f], "level %d field %s" % (l, f))
