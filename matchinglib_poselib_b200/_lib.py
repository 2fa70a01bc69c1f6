"""ctypes binding of libmplb200.so (include/mplb200.h).

The product path FAILS LOUDLY when the CUDA library is missing or no B200 is visible: there is
no CPU fallback and nothing here imports oracle/.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libmplb200.so")

c_int, c_size_t, c_void_p, c_char_p = ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_char_p
c_ll, c_double = ctypes.c_longlong, ctypes.c_double


class MplError(RuntimeError):
    pass


class DMatch(ctypes.Structure):
    """cv::DMatch layout (16 bytes)."""
    _fields_ = [("queryIdx", ctypes.c_int32), ("trainIdx", ctypes.c_int32), ("imgIdx", ctypes.c_int32),
                ("distance", ctypes.c_float)]


class PairIO(ctypes.Structure):
    """mpl_pair_io"""
    _fields_ = [("desc1", c_void_p), ("desc2", c_void_p), ("step1", c_size_t), ("step2", c_size_t), ("pts1", c_void_p), ("pts2", c_void_p),
                ("n1", c_int), ("n2", c_int), ("seed", ctypes.c_ulonglong), ("matches", c_void_p), ("mask", c_void_p), ("samples", c_void_p)]


class PairParams(ctypes.Structure):
    """mpl_pair_params"""
    _fields_ = [("cv_type", c_int), ("cols", c_int), ("matcher_name", c_char_p), ("ratio_test", c_int), ("pts_mode", c_int),
                ("K0", c_double * 9), ("K1", c_double * 9), ("dist0", c_double * 8), ("dist1", c_double * 8),
                ("th", c_double), ("h", c_int), ("conf", c_double), ("refine", c_int)]


class PairResult(ctypes.Structure):
    """mpl_pair_result"""
    _fields_ = [("n_matches", c_int), ("n_corr", c_int), ("n_samples", c_int), ("n_inl", c_int), ("iters_run", c_int), ("status", c_int),
                ("E", c_double * 9)]


# name -> (restype, argtypes).  Every symbol include/mplb200.h declares is listed here; the CPU test
# suite checks that the built library exports each of them.
SIGNATURES = {
    "mpl_init": (c_int, [c_int]),
    "mpl_shutdown": (c_int, []),
    "mpl_last_error": (c_char_p, []),
    "mpl_device_info": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "mpl_kernel_launch_count": (c_ll, []),
    "mpl_host_alloc": (c_void_p, [c_size_t]),
    "mpl_peer_gather_bytes": (c_size_t, [c_int, c_int]),
    "mpl_peer_alloc": (c_int, [c_size_t, c_void_p, c_void_p]),
    "mpl_peer_open": (c_int, [c_void_p, c_void_p]),
    "mpl_peer_close": (c_int, [c_void_p]),
    "mpl_peer_free": (c_int, [c_void_p]),
    "mpl_peer_status": (c_int, [c_void_p, c_void_p, c_void_p]),
    "mpl_allgather_merge_top2_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "mpl_host_free": (c_int, [c_void_p]),
    "mpl_prof_enable": (c_int, [c_int]),
    "mpl_prof_read": (c_int, [c_int, c_void_p, c_void_p]),
    "mpl_knn2_hamming": (c_int, [c_void_p, c_int, c_size_t, c_void_p, c_int, c_size_t, c_int, c_int, c_int, c_void_p, c_void_p]),
    "mpl_knn2_l2f32": (c_int, [c_void_p, c_int, c_size_t, c_void_p, c_int, c_size_t, c_int, c_int, c_int, c_void_p, c_void_p]),
    "mpl_set_l2_path": (c_int, [c_int]),
    "mpl_set_hamming_path": (c_int, [c_int]),
    "mpl_ratio_matches_i32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "mpl_ratio_matches_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "mpl_get_matches_api": (c_int, [c_int, c_int, c_void_p, c_int, c_int, c_size_t, c_int,
                                    c_void_p, c_int, c_int, c_size_t, c_int, c_char_p, c_int, c_int, c_void_p, c_void_p]),
    "mpl_gms_filter": (c_int, [c_void_p, c_int, c_size_t, c_int, c_int, c_void_p, c_int, c_size_t, c_int, c_int, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "mpl_knn_hamming_scratch_bytes": (c_size_t, [c_int, c_int, c_int]),
    "mpl_pack_rows_u8_dev": (c_int, [c_void_p, c_int, c_size_t, c_int, c_void_p, c_int, c_void_p]),
    "mpl_knn_hamming_static_max_bytes": (c_size_t, []),
    "mpl_knn2_hamming_static_dev": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, ctypes.c_longlong, c_void_p, c_void_p]),
    "mpl_knn2_hamming_dev": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_ll, c_void_p, c_size_t, c_void_p, c_void_p]),
    "mpl_merge_top2_dev": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "mpl_knn2_hamming_batch_dev": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "mpl_knn_l2_scratch_bytes": (c_size_t, [c_int, c_int, c_int]),
    "mpl_knn_l2_last_stats": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "mpl_knn_l2_last_stats_host": (c_int, [c_void_p, c_void_p]),
    "mpl_knn2_l2f32_dev": (c_int, [c_void_p, c_int, c_size_t, c_void_p, c_int, c_size_t, c_int, c_int, c_ll, c_void_p, c_size_t, c_void_p, c_void_p]),
    "mpl_set_solver_early_exit": (c_int, [c_int]),
    "mpl_ransac_set_seed": (c_int, [ctypes.c_ulonglong]),
    "mpl_ransac_draw_samples": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "mpl_ransac5pt": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_double, c_int, c_double, c_int,
                              c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "mpl_estimate_essential_api": (c_int, [c_void_p, c_void_p, c_int, c_char_p, c_double, c_int, c_void_p, c_void_p, c_void_p]),
    "mpl_lmeds5pt": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_double, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                             c_void_p, c_void_p]),
    "mpl_arrsac_set_seed": (c_int, [ctypes.c_ulonglong, ctypes.c_ulonglong]),
    "mpl_arrsac5pt": (c_int, [c_void_p, c_void_p, c_int, c_double, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "mpl_vfc_set_seed": (c_int, [ctypes.c_uint]),
    "mpl_vfc_filter": (c_int, [c_void_p, c_int, c_size_t, c_void_p, c_int, c_size_t, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "mpl_usac_params_default": (None, [c_void_p, c_double]),
    "mpl_usac5pt": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "mpl_usac_rand_stream": (c_int, [ctypes.c_uint, c_int, c_void_p]),
    "mpl_match_pose_pair": (c_int, [c_void_p, c_int, c_size_t, c_void_p, c_int, c_size_t, c_int, c_void_p, c_void_p, c_char_p, c_int,
                                    c_double, c_int, c_double, c_int, ctypes.c_ulonglong, c_void_p, c_void_p, c_void_p, c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_void_p]),
    "mpl_match_pose_batch": (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
    "mpl_get_pose_triang_pts_api": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_double, c_int, c_void_p, c_void_p, c_void_p]),
    "mpl_ransac_pack_corr_dev": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "mpl_ransac_solve_dev": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "mpl_ransac_score_dev": (c_int, [c_void_p, c_int, c_void_p, c_int, c_double, c_void_p, c_void_p, c_void_p]),
}

class UsacParams(ctypes.Structure):
    """mpl_usac_params (include/mplb200.h)."""
    _fields_ = [("th", c_double), ("conf", c_double), ("max_hypotheses", c_int), ("prevalidate_sample", c_int),
                ("prosac_beta", c_double), ("prosac_non_rand_conf", c_double), ("prosac_max_samples", c_int), ("prosac_min_stop_len", c_int),
                ("sprt_delta", c_double), ("sprt_epsilon", c_double), ("sprt_tM", c_double), ("sprt_mS", c_double),
                ("lo_inner_sample_size", c_int), ("lo_inner_reps", c_int), ("lo_num_steps", c_int), ("lo_threshold_multiplier", c_double),
                ("estimator", c_int), ("refine", c_int), ("degeneracy", c_int), ("lo_mode", c_int), ("seed", ctypes.c_uint)]


class UsacResult(ctypes.Structure):
    """mpl_usac_result (include/mplb200.h)."""
    _fields_ = [("ok", c_int), ("hyp_count", ctypes.c_uint), ("model_count", ctypes.c_uint), ("rejected_sample_count", ctypes.c_uint),
                ("rejected_model_count", ctypes.c_uint), ("total_points_verified", ctypes.c_uint), ("num_local_optimizations", ctypes.c_uint),
                ("best_inlier_count", ctypes.c_uint), ("sprt_delta", c_double), ("sprt_epsilon", c_double), ("n_batches", c_int), ("n_launches", c_int)]


_lib = None


def load():
    """dlopen libmplb200.so and set the prototypes.  Raises MplError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MplError("libmplb200.so is not built (%s); run `python -c 'import __graft_entry__ as g; g.build()'`. "
                       "There is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)       # AttributeError if the .so is stale
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def last_error():
    return load().mpl_last_error().decode("utf-8", "replace")


def check(rc, what):
    if rc != 0:
        raise MplError("%s failed (%d): %s" % (what, rc, last_error()))


def ptr(a):
    """Host pointer of a numpy array (None -> NULL)."""
    return None if a is None else ctypes.c_void_p(a.ctypes.data)
