"""Loader for the in-tree CUDA library (elaina_b200/libwost.so).

There is NO CPU fallback: if the library is missing or cannot be loaded, importing the product API fails loudly.
"""
import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libwost.so")
CSRC = os.path.join(HERE, "csrc")

c_void_p, c_int, c_float, c_u32, c_u64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_uint32, ctypes.c_uint64
c_char_p = ctypes.c_char_p


class ProblemDesc(ctypes.Structure):
    _fields_ = [
        ("dim", c_int),
        ("dirichlet", c_void_p),
        ("neumann", c_void_p),
        ("color_dirichlet", c_void_p),
        ("n_color_dirichlet", c_int),
        ("color_neumann", c_void_p),
        ("n_color_neumann", c_int),
        ("dirichlet_intensity", c_float),
        ("neumann_intensity", c_float),
        ("source_intensity", c_float),
        ("source", c_void_p),
        ("source_dims", c_int * 3),
        ("source_origin", c_float * 3),
        ("source_voxel", c_float),
        ("grid_scale", c_float),
        ("grid_pos", c_float * 3),
        ("grid_up", c_float * 3),
        ("grid_right", c_float * 3),
        ("mask", c_void_p),
        ("mask_w", c_int),
        ("mask_h", c_int),
        ("aabb_min", c_float * 3),
        ("aabb_max", c_float * 3),
    ]


class Settings(ctypes.Structure):
    _fields_ = [
        ("frame_w", c_int),
        ("frame_h", c_int),
        ("max_walking_depth", c_int),
        ("samples_per_pixel", c_int),
        ("epsilon_shell", c_float),
        ("debug_pixel", c_int),
        ("save_spp_metrics_duration", c_int),
        ("save_spp_metrics_until", c_int),
        ("save_time_metrics_duration", c_int),
        ("train_spp_count", c_int),
        ("max_guided_depth_training", c_int),
        ("max_guided_depth_guiding", c_int),
        ("uniform_fraction_training", c_float),
        ("uniform_fraction_guiding", c_float),
        ("seed", c_u32),
        ("walk_mode", c_int),
        ("spp_batch", c_int),
        ("pixel_begin", c_int),
        ("pixel_end", c_int),
        ("pixel_stride_rows", c_int),
        ("num_shards", c_int),
        ("shard", c_int),
        ("mlp_impl", c_int),
        ("max_train_depth", c_int),
        ("train_batch_size", c_int),
        ("train_min_batch", c_int),
        ("train_batches_per_frame", c_int),
    ]


# every symbol include/wost.h declares: (name, restype, argtypes)
_PP = ctypes.POINTER
SYMBOLS = [
    ("wost_last_error", c_char_p, []),
    ("wost_version", c_char_p, []),
    ("wost_set_device", c_int, [c_int]),
    ("wost_device_count", c_int, []),
    ("wost_scene_create", c_int, [c_int, c_void_p, c_int, c_void_p, c_int, _PP(c_void_p)]),
    ("wost_scene_destroy", None, [c_void_p]),
    ("wost_scene_info", c_int, [c_void_p, _PP(c_int), _PP(c_int), _PP(c_int), _PP(c_int)]),
    ("wost_scene_dump_bvh", c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    ("wost_scene_dump_sorted", c_int, [c_void_p, c_void_p, c_void_p]),
    ("wost_scene_dump_traversal_layout", c_int, [c_void_p, c_void_p, _PP(c_int), c_void_p]),
    ("wost_lbvh_nearest", c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    ("wost_lbvh_nearest_static", c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    ("wost_lbvh_silhouette", c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p]),
    ("wost_lbvh_ray", c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    ("wost_lbvh_ray_any", c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    ("wost_lbvh_sample_sphere", c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    ("wost_lbvh_nearest_dev", c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    ("wost_lbvh_silhouette_dev", c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    ("wost_lbvh_ray_dev", c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    ("wost_problem_create", c_int, [_PP(ProblemDesc), _PP(c_void_p)]),
    ("wost_problem_destroy", None, [c_void_p]),
    ("wost_integrator_create", c_int, [c_void_p, _PP(Settings), c_int, _PP(c_void_p)]),
    ("wost_integrator_destroy", None, [c_void_p]),
    ("wost_solve", c_int, [c_void_p, _PP(c_u64), _PP(c_u64), _PP(c_float)]),
    ("wost_render_channel", c_int, [c_void_p, c_int]),
    ("wost_film_download", c_int, [c_void_p, c_int, c_void_p]),
    ("wost_film_stderr", c_int, [c_void_p, c_void_p]),
    ("wost_traversal_stats", c_int, [c_void_p, c_void_p, c_void_p, c_int]),
    ("wost_debug_counters", c_int, [c_void_p, c_void_p]),
    ("wost_scene_traversal_stats", c_int, [c_void_p, c_void_p, c_void_p, c_int]),
    ("wost_film_device_ptr", c_int, [c_void_p, c_int, _PP(c_void_p)]),
    ("wost_last_launch_count", c_int, [c_void_p, _PP(c_u64)]),
    ("wost_l2_window_bytes", c_int, [c_void_p, _PP(c_u64)]),
    ("wost_last_step_kernel_time", c_int, [c_void_p, _PP(c_u64), _PP(c_float)]),
    ("wost_guided_last_samples", c_int, [c_void_p, c_void_p, c_u64, _PP(c_u64)]),
    ("wost_run_config", c_int, [c_char_p]),
    ("wost_comm_unique_id", c_int, [c_void_p]),
    ("wost_comm_create", c_int, [c_int, c_int, c_void_p, _PP(c_void_p)]),
    ("wost_comm_create_from_env", c_int, [_PP(c_void_p)]),
    ("wost_comm_destroy", None, [c_void_p]),
    ("wost_comm_info", c_int, [c_void_p, _PP(c_int), _PP(c_int), _PP(c_int)]),
    ("wost_comm_shard_settings", c_int, [c_void_p, _PP(Settings), c_int]),
    ("wost_integrator_set_comm", c_int, [c_void_p, c_void_p]),
    ("wost_film_gather", c_int, [c_void_p, c_void_p, c_int, c_int]),
    ("wost_comm_allreduce_f64", c_int, [c_void_p, c_void_p, c_int, c_int]),
    ("wost_reset_network", c_int, [c_void_p, c_char_p]),
    ("wost_reset_training", c_int, [c_void_p]),
    ("wost_query_network", c_int, [c_void_p, c_void_p, c_void_p]),
    ("wost_guided_network_info", c_int, [c_void_p, _PP(c_int), _PP(c_int), _PP(c_int)]),
    ("wost_guided_get_params", c_int, [c_void_p, c_int, c_void_p]),
    ("wost_guided_set_params", c_int, [c_void_p, c_void_p]),
    ("wost_guided_infer", c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    ("wost_guided_train_batch", c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, _PP(c_float), c_void_p]),
    ("wost_vmm_eval", c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_u32, c_void_p, c_void_p, c_void_p]),
    ("wost_set_allreduce", c_int, [c_void_p, c_void_p, c_void_p]),
    ("wost_set_frame_callback", c_int, [c_void_p, c_void_p, c_void_p]),
    ("wost_trim_allocation_cache", c_int, []),
    ("wost_nvdb_read_vec3f", c_int, [ctypes.c_char_p, c_void_p, c_void_p, c_void_p, c_void_p, c_u64]),
    ("wost_guided_last_profile", c_int, [c_void_p, c_void_p]),
    ("wost_guided_last_stats", c_int, [c_void_p, _PP(c_u64), _PP(c_u64), _PP(c_float)]),
]
ALLREDUCE_FN = ctypes.CFUNCTYPE(c_int, c_void_p, c_void_p, c_u64, c_void_p)
FRAME_FN = ctypes.CFUNCTYPE(None, c_void_p, c_int, c_int, c_u64, ctypes.POINTER(ctypes.c_float), c_int, c_int)

_lib = None


def build(verbose=False):
    """Compile every CUDA translation unit for sm_100a into elaina_b200/libwost.so (in-tree)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", CSRC, "-j", "4"], stdout=out)
    return LIB_PATH


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "elaina_b200/libwost.so is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the walker)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, res, args in SYMBOLS:
        fn = getattr(lib, name)  # raises AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class WostError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("wost status %d: %s" % (status, message))
        self.status = status


def check(status):
    if status != 0:
        msg = load().wost_last_error()
        raise WostError(status, msg.decode() if msg else "")
