"""ctypes binding of include/mmsolver_b200.h (no PyTorch, no numpy required by
the ABI itself; numpy is used for array marshalling).

The CUDA library is the product: importing it fails loudly if it has not been
built, and every compute entry point fails with MMB_ERR_NO_DEVICE when there is
no GPU.  There is no CPU fallback.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmmsolver_b200.so")

MMB_ABI_VERSION = 2
MMB_ATTR_NONE = -(2 ** 31)
MMB_LENS_MAX_COEFFS = 14
MMB_COMM_ID_BYTES = 128

(MMB_OK, MMB_ERR_INVALID_ARGUMENT, MMB_ERR_CUDA, MMB_ERR_NO_DEVICE, MMB_ERR_NOT_ENOUGH_OBJECTS,
 MMB_ERR_MORE_PARAMS_THAN_ERRORS, MMB_ERR_UNSUPPORTED, MMB_ERR_INTERNAL, MMB_ERR_COMM) = (0, -1, -2, -3, -4, -5, -6, -7, -8)

c_i32p = C.POINTER(C.c_int32)
c_f64p = C.POINTER(C.c_double)
c_u8p = C.POINTER(C.c_uint8)


class SceneDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32),
        ("num_frames", C.c_int32), ("frames", c_i32p),
        ("num_static", C.c_int32), ("static_values", c_f64p),
        ("num_anim", C.c_int32), ("anim_values", c_f64p),
        ("num_transforms", C.c_int32), ("tfm_parent", c_i32p), ("tfm_rotate_order", c_i32p),
        ("tfm_attrs", c_i32p),
        ("num_cameras", C.c_int32), ("cam_tfm", c_i32p), ("cam_attrs", c_i32p),
        ("cam_film_fit", c_i32p), ("cam_render_width", c_i32p), ("cam_render_height", c_i32p),
        ("cam_lens", c_i32p),
        ("num_bundles", C.c_int32), ("bnd_tfm", c_i32p),
        ("num_markers", C.c_int32), ("mkr_cam", c_i32p), ("mkr_bnd", c_i32p),
        ("mkr_x", c_f64p), ("mkr_y", c_f64p), ("mkr_weight", c_f64p), ("mkr_enable", c_u8p),
        ("num_lenses", C.c_int32), ("lens_model", c_i32p), ("lens_attrs", c_i32p),
        ("lens_camera_params", c_f64p),
        ("num_attrs", C.c_int32), ("attr_ref", c_i32p), ("attr_min", c_f64p), ("attr_max", c_f64p),
        ("attr_offset", c_f64p), ("attr_scale", c_f64p),
        ("num_stiffness", C.c_int32), ("stiff_attr", c_i32p), ("stiff_weight", c_i32p), ("stiff_variance", c_i32p),
        ("stiff_value", c_i32p),
        ("num_smoothness", C.c_int32), ("smooth_attr", c_i32p), ("smooth_weight", c_i32p), ("smooth_variance", c_i32p),
        ("smooth_value", c_i32p),
        ("current_frame", C.c_int32),
    ]


class Options(C.Structure):
    _fields_ = [
        ("iter_max", C.c_int32), ("tau", C.c_double), ("eps1", C.c_double), ("eps2", C.c_double),
        ("eps3", C.c_double), ("delta", C.c_double), ("auto_diff_type", C.c_int32),
        ("auto_param_scale", C.c_int32), ("robust_loss_type", C.c_int32),
        ("robust_loss_scale", C.c_double), ("scene_graph_mode", C.c_int32), ("solver_type", C.c_int32),
        ("time_eval_mode", C.c_int32), ("accept_only_better", C.c_int32), ("image_width", C.c_double),
        ("frame_solve_mode", C.c_int32), ("remove_unused_markers", C.c_int32),
        ("remove_unused_attributes", C.c_int32), ("lens_index_literal", C.c_int32), ("device", C.c_int32),
    ]


MMB_FMT_ERROR_PER_MARKER_PER_FRAME = 1
MMB_FMT_ERROR_PER_FRAME = 2
MMB_FMT_SOLVE_VALUES = 4
MMB_FMT_INPUT_STATS = 8


class Result(C.Structure):
    _fields_ = [
        ("success", C.c_int32), ("reason_num", C.c_int32), ("reason_str", C.c_char_p),
        ("error_final", C.c_double), ("error_avg", C.c_double), ("error_min", C.c_double),
        ("error_max", C.c_double), ("iteration_num", C.c_int32), ("iteration_function_num", C.c_int32),
        ("iteration_jacobian_num", C.c_int32), ("iteration_attempt_num", C.c_int32),
        ("user_interrupted", C.c_int32), ("error_is_better", C.c_int32),
        ("timer_solve_s", C.c_double), ("timer_jacobian_s", C.c_double), ("timer_function_s", C.c_double),
        ("timer_linear_s", C.c_double),
    ]


# every symbol include/mmsolver_b200.h declares: (name, restype, argtypes)
SYMBOLS = [
    ("mmb_options_default", None, [C.POINTER(Options)]),
    ("mmb_problem_create", C.c_int32, [C.POINTER(SceneDesc), C.POINTER(Options), C.POINTER(C.c_void_p)]),
    ("mmb_problem_destroy", None, [C.c_void_p]),
    ("mmb_problem_sizes", C.c_int32, [C.c_void_p, c_i32p, c_i32p, c_i32p]),
    ("mmb_get_index_maps", C.c_int32, [C.c_void_p, c_i32p, c_i32p]),
    ("mmb_affects", C.c_int32, [C.POINTER(SceneDesc), c_u8p, c_u8p, c_u8p]),
    ("mmb_problem_used_objects", C.c_int32, [C.c_void_p, c_i32p, c_i32p, c_i32p, c_i32p]),
    ("mmb_get_marker_weights", C.c_int32, [C.c_void_p, c_f64p]),
    ("mmb_get_initial_parameters", C.c_int32, [C.c_void_p, c_f64p]),
    ("mmb_eval_residuals", C.c_int32, [C.c_void_p, c_f64p, c_f64p, c_f64p]),
    ("mmb_eval_jacobian_dense", C.c_int32, [C.c_void_p, c_f64p, c_f64p, C.c_int32]),
    ("mmb_eval_jacobian_blocks", C.c_int32, [C.c_void_p, c_f64p, c_f64p, c_i32p, c_i32p]),
    ("mmb_solve", C.c_int32, [C.c_void_p, C.POINTER(Result), c_f64p, c_f64p, c_f64p]),
    ("mmb_solve_per_frame", C.c_int32, [C.c_void_p, C.POINTER(Result), C.POINTER(Result), c_f64p, c_f64p, c_f64p]),
    ("mmb_shard_frames", C.c_int32, [C.c_int32, C.c_int32, C.c_int32, c_i32p, c_i32p]),
    ("mmb_shard_param_map", C.c_int32, [C.POINTER(SceneDesc), C.c_int32, C.c_int32, c_i32p, c_i32p]),
    ("mmb_problem_create_shard_local", C.c_int32, [C.POINTER(SceneDesc), C.POINTER(Options), C.c_int32, C.c_int32,
                                                   C.POINTER(C.c_void_p)]),
    ("mmb_problem_create_shard", C.c_int32, [C.POINTER(SceneDesc), C.POINTER(Options), C.c_int32, C.c_int32,
                                             C.POINTER(C.c_void_p)]),
    ("mmb_comm_unique_id", C.c_int32, [c_u8p, C.c_int32]),
    ("mmb_problem_attach_comm", C.c_int32, [C.c_void_p, c_u8p, C.c_int32, C.c_int32]),
    ("mmb_comm_create", C.c_int32, [c_u8p, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    ("mmb_comm_destroy", None, [C.c_void_p]),
    ("mmb_problem_use_comm", C.c_int32, [C.c_void_p, C.c_void_p]),
    ("mmb_aggregate_results", C.c_int32, [C.POINTER(Result), C.c_int32, C.POINTER(Result)]),
    ("mmb_get_last_residuals", C.c_int32, [C.c_void_p, c_f64p]),
    ("mmb_format_results", C.c_int32, [C.POINTER(Result), C.c_int32, C.c_int32, c_i32p, c_i32p, C.c_int32,
                                       C.POINTER(C.c_char_p), C.c_int32, c_f64p, c_f64p, c_f64p, C.c_int32,
                                       C.c_char_p, C.c_int64, C.POINTER(C.c_int64)]),
    ("mmb_request_cancel", None, [C.c_void_p]),
    ("mmb_bench_eval", C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, c_f64p, c_f64p, c_i32p]),
    ("mmb_lens_apply", C.c_int32, [C.c_int32, C.c_int32, c_f64p, c_f64p, C.c_int32, C.c_int64, c_f64p, c_f64p]),
    ("mmb_lens_apply_layered", C.c_int32, [C.c_int32, C.c_int32, c_i32p, c_f64p, c_f64p, C.c_int32, C.c_int64, c_f64p,
                                           c_f64p]),
    ("mmb_lens_apply_layered_no_alloc", C.c_int32, [C.c_int32, C.c_int32, c_i32p, c_f64p, c_f64p, C.c_int32, C.c_int64,
                                                    c_f64p, c_f64p]),
    ("mmb_bench_lens", C.c_int32, [C.c_int32, C.c_int32, c_f64p, c_f64p, C.c_int32, C.c_int64, C.c_int32, c_f64p]),
    ("mmb_debug_chol_trace", C.c_int32, [C.POINTER(C.c_uint64), C.c_int32]),
    ("mmb_bench_stream_floor", C.c_int32, [C.c_int32, C.c_int64, C.c_int64, C.c_int32, c_f64p]),
    ("mmb_bench_fp64_peaks", C.c_int32, [C.c_int32, c_f64p, c_f64p]),
    ("mmb_launch_count", C.c_int64, [C.c_void_p]),
    ("mmb_solve_stats", C.c_int32, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    ("mmb_release_cached_memory", C.c_int32, [C.c_int32]),
    ("mmb_status_string", C.c_char_p, [C.c_int32]),
    ("mmb_last_error", C.c_char_p, []),
    ("mmb_version", C.c_char_p, []),
]

_lib = None


def bind(lib, symbols=SYMBOLS):
    for name, res, args in symbols:
        fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    return lib


def load():
    """Load libmmsolver_b200.so; raise (never fall back) when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "%s is not built. Run `python -m mayamatchmovesolver_b200.build` "
                "(needs nvcc). There is no CPU fallback." % LIB_PATH)
        _lib = bind(C.CDLL(LIB_PATH))
    return _lib


class MmbError(RuntimeError):
    def __init__(self, status, where):
        lib = load()
        msg = lib.mmb_status_string(status).decode()
        detail = lib.mmb_last_error().decode()
        super().__init__("%s: %s (%d)%s" % (where, msg, status, (": " + detail) if detail else ""))
        self.status = status


def check(status, where):
    if status != MMB_OK:
        raise MmbError(status, where)
