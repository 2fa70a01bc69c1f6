"""Host-side mirror of the reference's matchinglib API for the hot path.

`getMatches` keeps the reference's name, argument meaning, defaults and return codes
(matchinglib_poselib/source/matchinglib/include/matchinglib/matchinglib_matchers.h:61-64,
source/matchers.cpp:115-736) for matcher_name in {"LINEAR", "BRUTEFORCENMS"}; every other matcher
returns -2 (the B200 build has no CPU fallback).  cv::Mat -> numpy array (uint8 or float32, 2-D),
std::vector<cv::KeyPoint> -> anything with len(), std::vector<cv::DMatch> -> list filled in place
with (queryIdx, trainIdx, distance) tuples or a structured numpy array via getMatchesArray.
All arithmetic runs in libmplb200.so on the GPU.
"""
import ctypes

import numpy as np

from . import _lib

CV_8U, CV_32F, CV_64F = 0, 5, 6
DMATCH_DTYPE = np.dtype([("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")])

_NP_TO_CV = {np.dtype(np.uint8): CV_8U, np.dtype(np.float32): CV_32F, np.dtype(np.float64): CV_64F,
             np.dtype(np.int8): 1, np.dtype(np.uint16): 2, np.dtype(np.int16): 3, np.dtype(np.int32): 4}


def pinned_array(shape, dtype):
    """numpy array in the library's huge-page pinned host memory (mpl_host_alloc): descriptor matrices and keypoints
    placed here reach the GPU at the PCIe link rate.  Freed when the array (and every view of it) is gone."""
    import weakref
    lib = _lib.load()
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    ptr = lib.mpl_host_alloc(max(n, 1))
    if not ptr:
        raise MemoryError(_lib.last_error())
    buf = (ctypes.c_char * max(n, 1)).from_address(ptr)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    weakref.finalize(buf, lib.mpl_host_free, ctypes.c_void_p(ptr))
    return arr


def _mat_fields(m):
    """(data pointer keep-alive array, rows, cols, step, cv type) of a 2-D row-major array.
    Like the reference (matchers.cpp:567, 634) the data is read in place with cols as row length;
    arrays whose rows are not unit-stride are made contiguous first."""
    a = np.asarray(m)
    if a.ndim != 2:
        raise ValueError("descriptor matrix must be 2-D")
    if a.dtype not in _NP_TO_CV:
        raise ValueError("unsupported descriptor dtype %s" % a.dtype)
    if a.shape[1] > 0 and a.strides[1] != a.itemsize:
        a = np.ascontiguousarray(a)
    step = a.strides[0] if a.shape[0] > 1 else a.shape[1] * a.itemsize
    if step < a.shape[1] * a.itemsize:
        a = np.ascontiguousarray(a)
        step = a.shape[1] * a.itemsize
    return a, a.shape[0], a.shape[1], step, _NP_TO_CV[a.dtype]


def getMatchesArray(keypoints1, keypoints2, descriptors1, descriptors2, imgSi=None, matcher_name="GMBSOF",
                    VFCrefine=False, ratioTest=True, descriptor_name="", idxPars_NMSLIB="", queryPars_NMSLIB="",
                    nr_threads=0):
    """Returns (err, matches) with matches a numpy array of DMATCH_DTYPE (cv::DMatch layout)."""
    lib = _lib.load()
    a1, r1, c1, s1, t1 = _mat_fields(descriptors1)
    a2, r2, c2, s2, t2 = _mat_fields(descriptors2)
    out = np.empty(max(r1, 1), dtype=DMATCH_DTYPE)        # the library fills the first n_out records
    n_out = ctypes.c_int(0)
    err = lib.mpl_get_matches_api(len(keypoints1), len(keypoints2), _lib.ptr(a1), r1, c1, s1, t1,
                                  _lib.ptr(a2), r2, c2, s2, t2, matcher_name.encode(), int(bool(ratioTest)),
                                  int(nr_threads), _lib.ptr(out), ctypes.byref(n_out))
    if err == -103:
        raise AssertionError(_lib.last_error())          # cv::Exception from CV_Assert in the reference
    final = out[: n_out.value]
    if VFCrefine and err == 0:                            # matchers.cpp:721-733 (the early returns above skip it, as there)
        code, kept = filterWithVFC(keypoints1, keypoints2, final)
        if code == 0 and (len(kept) > 8 or len(final) < 24):
            final = kept
    return err, final


def filterWithVFC(keypL, keypR, matches_in, seed=None):
    """filterWithVFC (matchinglib/source/vfcMatches.cpp:49-86 -> class VFC, SparseVFC) through mpl_vfc_filter.  Returns
    (code, matches_out): code 0 ok, -1 fewer than 5 matches (matches_out empty), -2 fewer than 10 % kept.  seed restarts the rand()
    stream the control points are drawn from (None: it keeps running, like the reference's process-wide rand())."""
    lib = _lib.load()
    k1, k2 = _kp_xy(keypL), _kp_xy(keypR)
    m = np.ascontiguousarray(matches_in, DMATCH_DTYPE)
    if seed is not None:
        lib.mpl_vfc_set_seed(int(seed) & 0xffffffff)
    keep = np.zeros(max(len(m), 1), np.uint8)
    nk, code, it = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0)
    _lib.check(lib.mpl_vfc_filter(_lib.ptr(k1), len(k1), 8, _lib.ptr(k2), len(k2), 8, _lib.ptr(m), len(m), _lib.ptr(keep), ctypes.byref(nk),
                                  ctypes.byref(code), ctypes.byref(it)), "mpl_vfc_filter")
    return code.value, m[keep[: len(m)].astype(bool)]


def getMatches(keypoints1, keypoints2, descriptors1, descriptors2, imgSi, finalMatches, matcher_name="GMBSOF",
               VFCrefine=False, ratioTest=True, descriptor_name="", idxPars_NMSLIB="", queryPars_NMSLIB="",
               nr_threads=0):
    """matchinglib::getMatches.  finalMatches (a list) is cleared and filled; returns the reference's
    int code: 0 ok, -1 wrong input, -2 matcher not supported, -3 matching failed, -4 too few keypoints."""
    err, arr = getMatchesArray(keypoints1, keypoints2, descriptors1, descriptors2, imgSi, matcher_name, VFCrefine,
                               ratioTest, descriptor_name, idxPars_NMSLIB, queryPars_NMSLIB, nr_threads)
    del finalMatches[:]
    if err in (0, -3):
        finalMatches.extend((int(m["queryIdx"]), int(m["trainIdx"]), float(m["distance"])) for m in arr)
    return err


# ---- getCorrespondences (correspondences.cpp:148-519): the orchestrator around getMatches --------------------------
# Keypoint detection and descriptor extraction are OpenCV calls on the host in the reference too (features.cpp:145-244,
# 397-520; SURVEY section 2: outside the hot path); only the matching step runs through libmplb200.  Detector /
# extractor names follow features.cpp:987-1056 for the types the installed cv2 build provides.
def _cv2_detector(name, limit):
    import cv2
    table = {"FAST": lambda: cv2.FastFeatureDetector_create(), "ORB": lambda: cv2.ORB_create(nfeatures=limit),
             "BRISK": lambda: cv2.BRISK_create(), "AKAZE": lambda: cv2.AKAZE_create(), "KAZE": lambda: cv2.KAZE_create(),
             "SIFT": lambda: cv2.SIFT_create(), "MSER": lambda: cv2.MSER_create(),
             "SimpleBlob": lambda: cv2.SimpleBlobDetector_create()}
    return table[name]() if name in table else None


def _cv2_extractor(name):
    import cv2
    table = {"ORB": lambda: cv2.ORB_create(), "BRISK": lambda: cv2.BRISK_create(), "AKAZE": lambda: cv2.AKAZE_create(),
             "KAZE": lambda: cv2.KAZE_create(), "SIFT": lambda: cv2.SIFT_create()}
    return table[name]() if name in table else None


def _get_keypoints(img, featuretype, limit, detector):
    """features.cpp:145-244 without the grid-based response filter of the dynamic branch (retainBest on both branches).
    Returns (err, keypoints)."""
    det = detector if detector is not None else _cv2_detector(featuretype, min(limit, 100000) if featuretype == "ORB" else limit)
    if det is None:
        return -3, []
    kps = list(det.detect(img) if hasattr(det, "detect") else det(img))
    if len(kps) > limit:
        kps = sorted(kps, key=lambda k: -k.response)[:limit]                  # cv::KeyPointsFilter::retainBest
    if len(kps) < 10:
        return -1, kps
    return 0, kps


def getCorrespondences(img1, img2, featuretype="FAST", extractortype="FREAK", matchertype="GMBSOF", dynamicKeypDet=True,
                       limitNrfeatures=8000, VFCrefine=False, GMSrefine=False, ratioTest=True, SOFrefine=False,
                       subPixRefine=0, verbose=0, idxPars_NMSLIB="", queryPars_NMSLIB="", detector=None, extractor=None):
    """matchinglib::getCorrespondences (matchinglib_correspondences.h:61-78, correspondences.cpp:148-519).
    Returns (err, finalMatches, kp1, kp2); err as in the reference: 0 ok, -1 missing input, -2 keypoint error,
    -3 descriptor error, -4 incompatible detector/matcher, -6 incompatible detector/descriptor, and a getMatches failure e
    as -4 + e (:356-359), i.e. -5 wrong input, -6 matcher not served by this build, -7 too few matches, -8 too few keypoints.
    `detector` / `extractor` (objects with detect(img) / compute(img, kps)) replace the cv2 factories when given.
    GMSrefine runs the GMS grid filter (mpl_gms_filter; :378-396) and VFCrefine the vector-field-consensus filter inside
    getMatches (mpl_vfc_filter; matchers.cpp:721-733) on the GPU; SOF and sub-pixel refinement (:415-506) are outside the B200
    hot path."""
    if img1 is None or img2 is None or np.size(img1) == 0 or np.size(img2) == 0:
        return -1, [], [], []
    if SOFrefine or subPixRefine:
        raise NotImplementedError("SOF / sub-pixel refinement are outside the B200 hot path")
    if featuretype in ("MSER", "SimpleBlob") and matchertype == "GMBSOF":
        return -4, [], [], []
    if (featuretype != "KAZE" and extractortype == "KAZE") or (featuretype != "AKAZE" and extractortype == "AKAZE") or \
       (featuretype == "SIFT" and extractortype == "ORB") or (featuretype == "MSD" and extractortype == "SIFT"):
        return -6, [], [], []
    e1, kp1 = _get_keypoints(img1, featuretype, limitNrfeatures, detector)
    e2, kp2 = _get_keypoints(img2, featuretype, limitNrfeatures, detector) if e1 == 0 else (e1, [])
    if e1 or e2:
        return -2, [], kp1, kp2
    ext = extractor if extractor is not None else _cv2_extractor(extractortype)
    if ext is None:
        return -3, [], kp1, kp2
    kp1, d1 = ext.compute(img1, kp1)
    kp2, d2 = ext.compute(img2, kp2)
    if d1 is None or d2 is None or len(kp1) != len(d1) or len(kp2) != len(d2):
        return -3, [], list(kp1 or []), list(kp2 or [])
    err, matches = getMatchesArray(kp1, kp2, d1, d2, None, matchertype, VFCrefine, ratioTest, extractortype,
                                   idxPars_NMSLIB, queryPars_NMSLIB)
    if err != 0:
        return -4 + err, [], list(kp1), list(kp2)
    if GMSrefine:                                                              # correspondences.cpp:378-396
        imgSi = (np.shape(img1)[1], np.shape(img1)[0])                         # cv::Size imgSi = img1.size()
        n_gms, inl = filterMatchesGMS(kp1, imgSi, kp2, imgSi, matches, False, False)
        if n_gms >= 2:                                                         # MIN_FINAL_MATCHES
            matches = matches[inl]
    return 0, matches, list(kp1), list(kp2)


def _kp_xy(keypoints):
    """[n, 2] float32 pixel positions from cv2.KeyPoint objects, (x, y) tuples or an array."""
    if len(keypoints) and hasattr(keypoints[0], "pt"):
        return np.array([k.pt for k in keypoints], np.float32).reshape(-1, 2)
    return np.ascontiguousarray(np.asarray(keypoints, np.float32).reshape(-1, 2))


def filterMatchesGMS(keypoints1, imageSize1, keypoints2, imageSize2, matches, useScale=False, useRotation=False):
    """filterMatchesGMS (matchinglib/source/gms.cpp:54-60 -> thirdparty/gms-1.0 MatchGMS) on the GPU (mpl_gms_filter).
    imageSize = (width, height); matches: DMATCH_DTYPE array.  Returns (n_inliers, inlierMask bool [len(matches)])."""
    lib = _lib.load()
    k1, k2 = _kp_xy(keypoints1), _kp_xy(keypoints2)
    m = np.ascontiguousarray(matches, DMATCH_DTYPE)
    mask = np.zeros(max(len(m), 1), np.uint8)
    n = ctypes.c_int(0)
    _lib.check(lib.mpl_gms_filter(_lib.ptr(k1), len(k1), 8, int(imageSize1[0]), int(imageSize1[1]), _lib.ptr(k2), len(k2), 8,
                                  int(imageSize2[0]), int(imageSize2[1]), _lib.ptr(m), len(m), int(bool(useScale)), int(bool(useRotation)),
                                  _lib.ptr(mask), ctypes.byref(n)), "mpl_gms_filter")
    return n.value, mask[: len(m)].astype(bool)


def knn2_hamming(query, train, k=2, ignore_tail_bytes=0):
    """Raw k-NN (the cvflann / NMSLIB scan): returns (idx[nq,k] int32, dist[nq,k] int32)."""
    lib = _lib.load()
    q, nq, nb, qs, tq = _mat_fields(query)
    t, nt, nb2, ts, tt = _mat_fields(train)
    if tq != CV_8U or tt != CV_8U or (nb != nb2 and nq and nt):
        raise ValueError("knn2_hamming needs uint8 matrices of equal width")
    idx = np.empty((nq, k), np.int32)
    dist = np.empty((nq, k), np.int32)
    _lib.check(lib.mpl_knn2_hamming(_lib.ptr(q), nq, qs, _lib.ptr(t), nt, ts, nb, ignore_tail_bytes, k,
                                    _lib.ptr(idx), _lib.ptr(dist)), "mpl_knn2_hamming")
    return idx, dist


def knn2_l2(query, train, k=2, sqrt_out=False):
    """Raw k-NN for float32 descriptors: (idx[nq,k] int32, dist[nq,k] float32; squared unless sqrt_out)."""
    lib = _lib.load()
    q, nq, d, qs, tq = _mat_fields(query)
    t, nt, d2, ts, tt = _mat_fields(train)
    if tq != CV_32F or tt != CV_32F or (d != d2 and nq and nt):
        raise ValueError("knn2_l2 needs float32 matrices of equal width")
    idx = np.empty((nq, k), np.int32)
    dist = np.empty((nq, k), np.float32)
    _lib.check(lib.mpl_knn2_l2f32(_lib.ptr(q), nq, qs, _lib.ptr(t), nt, ts, d, int(bool(sqrt_out)), k,
                                  _lib.ptr(idx), _lib.ptr(dist)), "mpl_knn2_l2f32")
    return idx, dist
