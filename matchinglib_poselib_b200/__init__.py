"""matchinglib_poselib_b200 -- B200 (sm_100a) implementation of the data-parallel hot path of
josefmaierfl/matchinglib_poselib behind the reference's own API names.

  matchinglib.getMatches            LINEAR / BRUTEFORCENMS brute-force k-NN + Lowe ratio test
  poselib.estimateEssentialMat      RANSAC 5-point essential matrix (+ getPoseTriangPts)

Everything computes in lib/libmplb200.so (hand-written CUDA, C ABI in include/mplb200.h).
"""
from . import _lib, matchinglib, poselib  # noqa: F401
from .matchinglib import filterMatchesGMS, getCorrespondences, getMatches, getMatchesArray, knn2_hamming, knn2_l2, pinned_array  # noqa: F401
from .poselib import estimateEssentialMat, lmeds5pt, ransac5pt  # noqa: F401

__all__ = ["filterMatchesGMS", "getCorrespondences", "pinned_array", "getMatches", "getMatchesArray", "knn2_hamming", "knn2_l2", "matchinglib", "poselib", "estimateEssentialMat", "ransac5pt", "lmeds5pt"]
