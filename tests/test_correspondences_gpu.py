"""getCorrespondences mirror (correspondences.cpp:148-519): host orchestration around the GPU matcher."""
import numpy as np
import pytest

import oracle.knn as orc

cv2 = pytest.importorskip("cv2")


@pytest.fixture(scope="module")
def m():
    import matchinglib_poselib_b200.matchinglib as mm
    return mm


def _pair(seed=3, shift=(7, 4)):
    rng = np.random.default_rng(seed)
    base = (rng.random((60, 90)) * 255).astype(np.uint8)
    img = cv2.resize(base, (720, 480), interpolation=cv2.INTER_CUBIC)
    img = cv2.GaussianBlur(img, (5, 5), 1.0)
    img2 = np.roll(img, shift, axis=(0, 1))
    return img, img2


def test_error_codes_before_any_device_work(m):
    img, img2 = _pair()
    assert m.getCorrespondences(None, img2)[0] == -1
    assert m.getCorrespondences(img, img2, "MSER", "ORB", "GMBSOF")[0] == -4
    assert m.getCorrespondences(img, img2, "FAST", "KAZE", "LINEAR")[0] == -6
    assert m.getCorrespondences(img, img2, "ORB", "AKAZE", "LINEAR")[0] == -6
    assert m.getCorrespondences(img, img2, "SIFT", "ORB", "LINEAR")[0] == -6
    assert m.getCorrespondences(img, img2, "NOPE", "ORB", "LINEAR")[0] == -2
    assert m.getCorrespondences(img, img2, "ORB", "NOPE", "LINEAR")[0] == -3
    flat = np.full((480, 720), 128, np.uint8)
    assert m.getCorrespondences(flat, flat, "ORB", "ORB", "LINEAR")[0] == -2          # < 10 keypoints
    with pytest.raises(NotImplementedError):
        m.getCorrespondences(img, img2, "ORB", "ORB", "LINEAR", SOFrefine=True)


@pytest.mark.gpu
@pytest.mark.parametrize("feat,ext,matcher", [("ORB", "ORB", "LINEAR"), ("ORB", "ORB", "BRUTEFORCENMS"), ("SIFT", "SIFT", "LINEAR"),
                                              ("AKAZE", "AKAZE", "LINEAR")])
def test_get_correspondences_matches_oracle_chain(m, feat, ext, matcher):
    img, img2 = _pair()
    err, matches, kp1, kp2 = m.getCorrespondences(img, img2, feat, ext, matcher, False, 3000)
    assert err == 0 and len(matches) > 50
    # same descriptors through the oracle's getMatches
    det = m._cv2_detector(feat, 3000); ex = m._cv2_extractor(ext)
    keep = lambda ks: sorted(ks, key=lambda k: -k.response)[:3000] if len(ks) > 3000 else list(ks)     # retainBest
    k1, k2 = keep(det.detect(img)), keep(det.detect(img2))
    k1, d1 = ex.compute(img, k1); k2, d2 = ex.compute(img2, k2)
    oerr, om = orc.get_matches(d1, d2, matcher, True, threads=8)
    assert oerr == 0 and len(om) == len(matches)
    assert np.array_equal(matches["queryIdx"], om["queryIdx"]) and np.array_equal(matches["trainIdx"], om["trainIdx"])
    if matcher == "LINEAR":
        assert np.array_equal(matches["distance"].view(np.int32), om["distance"].view(np.int32))
    # the planted shift (7 rows, 4 columns) is what most matches see
    d = np.array([[kp2[t].pt[0] - kp1[q].pt[0], kp2[t].pt[1] - kp1[q].pt[1]] for q, t in zip(matches["queryIdx"], matches["trainIdx"])])
    good = (np.abs(d[:, 0] - 4) < 1.5) & (np.abs(d[:, 1] - 7) < 1.5)
    assert good.mean() > 0.8


@pytest.mark.gpu
def test_matcher_not_served_maps_to_minus_6(m):
    img, img2 = _pair()
    assert m.getCorrespondences(img, img2, "ORB", "ORB", "GMBSOF", False, 2000)[0] == -6   # -4 + (-2)
