"""ctypes loader for libboxmot_b200.so (the C ABI declared in include/boxmot_b200.h).

There is no CPU fallback: if the library is missing or a call fails, this raises."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libboxmot_b200.so")
_lib = None

c_void_p, c_int, c_double = ctypes.c_void_p, ctypes.c_int, ctypes.c_double


class OcsortConfig(ctypes.Structure):
    _fields_ = [(n, c_double) for n in ("det_thresh", "min_conf", "asso_threshold", "inertia",
                                         "q_xy_scaling", "q_s_scaling")] + \
               [(n, c_int) for n in ("max_age", "min_hits", "delta_t", "asso_func", "use_byte",
                                      "img_w", "img_h")]


class DeepocsortConfig(ctypes.Structure):
    _fields_ = [("oc", OcsortConfig)] + \
               [(n, c_double) for n in ("w_association_emb", "alpha_fixed_emb", "aw_param")] + \
               [(n, c_int) for n in ("embedding_off", "aw_off", "emb_dim")]


class BoosttrackConfig(ctypes.Structure):
    _fields_ = [(n, c_double) for n in ("det_thresh", "iou_threshold", "min_box_area", "aspect_ratio_thresh",
                                         "lambda_iou", "lambda_mhd", "lambda_shape", "dlo_boost_coef")] + \
               [(n, c_int) for n in ("max_age", "min_hits", "use_dlo_boost", "use_duo_boost", "s_sim_corr",
                                      "use_rich_s", "use_sb", "use_vt", "with_reid", "emb_dim")]


class BytetrackConfig(ctypes.Structure):
    _fields_ = [(n, c_double) for n in ("min_conf", "track_thresh", "match_thresh")] + \
               [(n, c_int) for n in ("track_buffer", "frame_rate")]


class BotsortConfig(ctypes.Structure):
    _fields_ = [(n, c_double) for n in ("track_high_thresh", "track_low_thresh", "new_track_thresh",
                                         "match_thresh", "proximity_thresh", "appearance_thresh")] + \
               [(n, c_int) for n in ("track_buffer", "frame_rate", "fuse_first_associate", "with_reid",
                                      "emb_dim")]


class StrongsortConfig(ctypes.Structure):
    _fields_ = [(n, c_double) for n in ("min_conf", "max_cos_dist", "max_iou_dist", "mc_lambda", "ema_alpha",
                                         "conf_thresh_high", "conf_thresh_low", "id_preservation_weight")] + \
               [(n, c_int) for n in ("max_age", "n_init", "nn_budget", "born_confirmed", "emb_dim")]


ASSO_FUNCS = {"iou": 0, "giou": 1, "diou": 2, "ciou": 3, "hmiou": 4, "centroid": 5}

_SIGS = {
    "bm_last_error": (ctypes.c_char_p, []),
    "bm_version": (c_int, []),
    "bm_ocsort_create": (c_int, [ctypes.POINTER(OcsortConfig), c_int, c_int, c_int, c_int,
                                 ctypes.POINTER(c_void_p)]),
    "bm_ocsort_destroy": (c_int, [c_void_p]),
    "bm_ocsort_reset": (c_int, [c_void_p, c_void_p]),
    "bm_ocsort_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "bm_ocsort_step_frames": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "bm_ocsort_status": (c_int, [c_void_p, ctypes.POINTER(c_int),
                                 ctypes.POINTER(ctypes.c_ulonglong),
                                 ctypes.POINTER(ctypes.c_ulonglong), c_void_p]),
    "bm_ocsort_counters": (c_int, [c_void_p, ctypes.POINTER(ctypes.c_ulonglong * 4), c_void_p]),
    "bm_ocsort_debug_profile": (c_int, [c_void_p, c_void_p]),
    "bm_ocsort_time_kernels": (c_int, [c_void_p, c_int]),
    "bm_ocsort_kernel_times": (c_int, [c_void_p, ctypes.POINTER(ctypes.c_float * 4)]),
    "bm_ocsort_stream_info": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "bm_ocsort_peek": (c_int, [c_void_p, c_int, ctypes.POINTER(c_int), c_void_p, c_void_p,
                               c_void_p, c_void_p]),
    "bm_deepocsort_create": (c_int, [ctypes.POINTER(DeepocsortConfig), c_int, c_int, c_int,
                                     ctypes.POINTER(c_void_p)]),
    "bm_deepocsort_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p]),
    "bm_deepocsort_kernel_times": (c_int, [c_void_p, ctypes.POINTER(ctypes.c_float * 6)]),
    "bm_boosttrack_create": (c_int, [ctypes.POINTER(BoosttrackConfig), c_int, c_int, c_int,
                                     ctypes.POINTER(c_void_p)]),
    "bm_boosttrack_destroy": (c_int, [c_void_p]),
    "bm_boosttrack_reset": (c_int, [c_void_p, c_void_p]),
    "bm_boosttrack_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p]),
    "bm_boosttrack_counters": (c_int, [c_void_p, ctypes.POINTER(c_int), ctypes.POINTER(ctypes.c_ulonglong * 4),
                                       c_void_p]),
    "bm_boosttrack_time_kernels": (c_int, [c_void_p, c_int]),
    "bm_boosttrack_kernel_times": (c_int, [c_void_p, ctypes.POINTER(ctypes.c_float * 6)]),
    "bm_boosttrack_peek": (c_int, [c_void_p, c_int, ctypes.POINTER(c_int), c_void_p, c_void_p, c_void_p,
                                   c_void_p]),
    "bm_bytetrack_create": (c_int, [ctypes.POINTER(BytetrackConfig), c_int, c_int, c_int,
                                    ctypes.POINTER(c_void_p)]),
    "bm_bytetrack_destroy": (c_int, [c_void_p]),
    "bm_bytetrack_reset": (c_int, [c_void_p, c_void_p]),
    "bm_bytetrack_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "bm_bytetrack_counters": (c_int, [c_void_p, ctypes.POINTER(c_int),
                                      ctypes.POINTER(ctypes.c_ulonglong * 4), c_void_p]),
    "bm_bytetrack_peek": (c_int, [c_void_p, c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_int),
                                  c_void_p, c_void_p, c_void_p, c_void_p]),
    "bm_botsort_create": (c_int, [ctypes.POINTER(BotsortConfig), c_int, c_int, c_int, ctypes.POINTER(c_void_p)]),
    "bm_botsort_destroy": (c_int, [c_void_p]),
    "bm_botsort_reset": (c_int, [c_void_p, c_void_p]),
    "bm_botsort_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "bm_botsort_counters": (c_int, [c_void_p, ctypes.POINTER(c_int), ctypes.POINTER(ctypes.c_ulonglong * 4),
                                    c_void_p]),
    "bm_botsort_peek": (c_int, [c_void_p, c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_int),
                                c_void_p, c_void_p, c_void_p, c_void_p]),
    "bm_lap_solve_batch": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_double,
                                   c_void_p, c_void_p, c_void_p]),
    "bm_lap_rect_batch": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_double,
                                  c_void_p, c_void_p, c_void_p]),
    "bm_iou_matrix": (c_int, [c_int, c_void_p, c_int, c_void_p, c_int, c_double, c_double,
                              c_void_p, c_void_p]),
    "bm_xysr_predict": (c_int, [c_void_p, c_void_p, c_int, c_double, c_double, c_void_p]),
    "bm_xysr_update": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "bm_kf_cv8": (c_int, [c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "bm_kf_cv8_gating": (c_int, [c_int, c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_void_p,
                                 c_void_p]),
    "bm_nn_cosine_distance": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p,
                                      c_void_p]),
    "bm_min_cost_matching_batch": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_double,
                                           c_void_p, c_void_p, c_void_p]),
    "bm_ocsort_set_groups": (c_int, [c_void_p, c_int]),
    "bm_boosttrack_set_groups": (c_int, [c_void_p, c_int]),
    "bm_botsort_set_groups": (c_int, [c_void_p, c_int]),
    "bm_strongsort_gated_metric": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int, c_double,
                                           c_double, c_void_p]),
    "bm_strongsort_enhance_cost": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p]),
    "bm_strongsort_feature_update": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_double, c_void_p,
                                             c_void_p, c_void_p, c_void_p]),
    "bm_strongsort_track": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "bm_strongsort_iou_cost": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "bm_cosine_distance": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "bm_strongsort_create": (c_int, [ctypes.POINTER(StrongsortConfig), c_int, c_int, c_int, c_int,
                                     ctypes.POINTER(c_void_p)]),
    "bm_strongsort_destroy": (c_int, [c_void_p]),
    "bm_linear_sum_assignment_far_batch": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_double,
                                                   c_void_p, c_void_p, c_void_p, c_void_p]),
    "bm_linear_sum_assignment_batch": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                               c_void_p]),
    "bm_strongsort_reset": (c_int, [c_void_p, c_void_p]),
    "bm_strongsort_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "bm_strongsort_counters": (c_int, [c_void_p, ctypes.POINTER(c_int), ctypes.POINTER(ctypes.c_ulonglong * 4),
                                       c_void_p]),
    "bm_bytetrack_select_class": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p]),
    "bm_boosttrack_hold_frame": (c_int, [c_void_p, c_int, c_void_p]),
    "bm_strongsort_set_groups": (c_int, [c_void_p, c_int]),
    "bm_strongsort_usage": (c_int, [c_void_p, ctypes.POINTER(c_int), ctypes.POINTER(c_int), ctypes.POINTER(c_int),
                                    c_void_p]),
    "bm_strongsort_debug_copy": (c_int, [c_void_p, c_int, c_int, c_void_p, ctypes.c_size_t, c_void_p]),
    "bm_strongsort_time_kernels": (c_int, [c_void_p, c_int]),
    "bm_strongsort_kernel_times": (c_int, [c_void_p, ctypes.POINTER(ctypes.c_float * 6)]),
    "bm_strongsort_peek": (c_int, [c_void_p, c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_int), c_void_p, c_void_p,
                                   c_void_p, c_void_p, c_void_p, c_void_p]),
    "bm_pack_mot_rows": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "bm_pack_rows": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "bm_cosine_distance_noalloc": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_void_p, c_int, c_void_p]),
    "bm_xysrb_predict": (c_int, [c_void_p, c_void_p, c_int, c_double, c_double, c_void_p]),
    "bm_xysrb_update": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
}
EXPORTS = tuple(_SIGS)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libboxmot_b200.so is not built (run `python -m boxmot_b200.build`); "
                "there is no CPU fallback for the association path")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(rc, what):
    if rc != 0:
        raise RuntimeError("%s failed (%d): %s" % (what, rc, lib().bm_last_error().decode()))
