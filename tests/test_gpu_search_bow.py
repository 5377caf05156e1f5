"""-m gpu: ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (ORBmatcher.cc:224-425) on the GPU
(orbfe_search_by_bow_batch) against the CPU oracle: the key-frame feature behind every slot of vpMapPointMatches after the
rotation check, and nmatches.  Index work: bit-exact."""
import numpy as np
import pytest

from orbslam3lib_b200 import ORBmatcher, _lib, scenes
from test_search_rule import bow_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,w,h,nf,stereo,ratio,ori", [
    (8600, 752, 480, 1200, False, 0.7, True),      # Tracking::TrackReferenceKeyFrame: ORBmatcher(0.7, true)
    (8601, 752, 480, 1200, False, 0.75, True),     # relocalisation: ORBmatcher(0.75, true)
    (8602, 640, 400, 1000, True, 0.7, True),       # stereo-camera frame: left and right slots
    (8603, 640, 400, 1000, True, 0.9, False),
    (8604, 320, 240, 400, False, 0.9, True),
])
def test_search_by_bow_matches_oracle(oracle, seed, w, h, nf, stereo, ratio, ori):
    dk, ak, vk, fvk, df, af, nleft, fvf = bow_case(oracle, seed, w, h, nf, stereo)
    rm, rn = oracle.search_by_bow(dk, ak, vk, fvk, df, af, nleft, fvf, ratio, ori)
    nm, m = ORBmatcher(ratio, ori).SearchByBoW(dk, ak, vk, fvk, df, af, fvf, nleft_f=nleft)
    np.testing.assert_array_equal(m, rm)
    assert nm == rn and nm > 30
    if stereo:
        assert (m[nleft:] >= 0).sum() > 10


def test_search_by_bow_contended_node(oracle):
    """every feature in ONE vocabulary node: all key-frame features compete for the frame's features in order."""
    dk, ak, vk, _, df, af, nleft, _ = bow_case(oracle, 8610, 320, 240, 300)
    one = lambda n: (np.array([5], np.int32), np.array([0, n], np.int32), np.arange(n, dtype=np.int32))    # noqa: E731
    rm, rn = oracle.search_by_bow(dk, ak, vk, one(len(dk)), df, af, -1, one(len(df)), 0.9, True)
    nm, m = ORBmatcher(0.9, True).SearchByBoW(dk, ak, vk, one(len(dk)), df, af, one(len(df)))
    np.testing.assert_array_equal(m, rm)
    assert nm == rn and nm > 50


def test_search_by_bow_disjoint_and_empty(oracle):
    dk, ak, vk, fvk, df, af, nleft, fvf = bow_case(oracle, 8620, 320, 240, 300)
    far = (fvf[0] + 1000, fvf[1], fvf[2])    # no common node
    nm, m = ORBmatcher(0.7, True).SearchByBoW(dk, ak, vk, fvk, df, af, far)
    assert nm == 0 and (m == -1).all()
    nm, m = ORBmatcher(0.7, True).SearchByBoW(dk, ak, np.zeros_like(vk), fvk, df, af, fvf)    # no good MapPoint
    assert nm == 0 and (m == -1).all()


def test_search_by_bow_ragged_device_batch(oracle):
    import torch
    cases = [bow_case(oracle, 8630 + k, 320, 240, nf, stereo=(k == 1)) for k, nf in enumerate((500, 300, 400))]
    cap = max(max(len(c[0]), len(c[4])) for c in cases)
    ncap = max(max(len(c[3][0]), len(c[7][0])) for c in cases)
    F = len(cases)
    A = {k: np.zeros(s, d) for k, (s, d) in dict(
        dk=((F, cap, 32), np.uint8), ak=((F, cap), np.float32), vk=((F, cap), np.uint8), kid=((F, ncap), np.int32),
        kst=((F, ncap + 1), np.int32), kix=((F, cap), np.int32), kn=((F,), np.int32), df=((F, cap, 32), np.uint8),
        af=((F, cap), np.float32), nf=((F,), np.int32), nl=((F,), np.int32), fid=((F, ncap), np.int32),
        fst=((F, ncap + 1), np.int32), fix=((F, cap), np.int32), fn=((F,), np.int32)).items()}
    for f, (dk, ak, vk, fvk, df, af, nleft, fvf) in enumerate(cases):
        A["dk"][f, :len(dk)], A["ak"][f, :len(dk)], A["vk"][f, :len(dk)] = dk, ak, vk
        A["kid"][f, :len(fvk[0])], A["kst"][f, :len(fvk[1])], A["kix"][f, :len(fvk[2])], A["kn"][f] = fvk[0], fvk[1], fvk[2], len(fvk[0])
        A["df"][f, :len(df)], A["af"][f, :len(df)], A["nf"][f], A["nl"][f] = df, af, len(df), nleft
        A["fid"][f, :len(fvf[0])], A["fst"][f, :len(fvf[1])], A["fix"][f, :len(fvf[2])], A["fn"][f] = fvf[0], fvf[1], fvf[2], len(fvf[0])
    dev = torch.device("cuda", 0)
    T = {k: torch.from_numpy(v).to(dev) for k, v in A.items()}
    dm = torch.zeros((F, cap), dtype=torch.int32, device=dev)
    dc = torch.zeros(F, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()    # the fills above run on torch's stream, the library on its own
    ORBmatcher(0.7, True).search_by_bow_batch_raw(T["dk"], T["ak"], T["vk"], T["kid"], T["kst"], T["kix"], T["kn"], T["df"], T["af"],
                                                  T["nf"], T["nl"], T["fid"], T["fst"], T["fix"], T["fn"], cap, ncap, F, dm, dc,
                                                  flags=_lib.IN_DEVICE | _lib.OUT_DEVICE)
    m, cnt = dm.cpu().numpy(), dc.cpu().numpy()
    for f, (dk, ak, vk, fvk, df, af, nleft, fvf) in enumerate(cases):
        rm, rn = oracle.search_by_bow(dk, ak, vk, fvk, df, af, nleft, fvf, 0.7, True)
        np.testing.assert_array_equal(m[f, :len(df)], rm)
        assert cnt[f] == rn
