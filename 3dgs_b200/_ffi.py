"""ctypes binding of libgsx.so (the C ABI declared in include/gsx.h).

There is no CPU fallback: if the shared library is missing this raises, and every entry point needs a CUDA
device.  PyTorch is used only to own device memory and streams.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgsx.so")

GSX_OK = 0
GSX_E_CAPACITY = 6
GSX_SH_DSPH_CORRECT = 1
GSX_GRAD_OVERWRITE = 2
GSX_GRAD_NO_SH = 4

# every symbol include/gsx.h declares (tests check the library exports all of them)
SYMBOLS = [
    "gsx_version", "gsx_last_error",
    "gsx_compute_camera_space_points", "gsx_project_to_screen", "gsx_cull_gaussians", "gsx_compute_sigma",
    "gsx_compute_conic", "gsx_precompute_spherical_harmonics", "gsx_sorted_list_size",
    "gsx_sorted_list_workspace_bytes", "gsx_get_sorted_gaussian_list", "gsx_render_image",
    "gsx_render_image_backward", "gsx_precompute_spherical_harmonics_backward", "gsx_compute_conic_backward",
    "gsx_compute_projection_jacobian_backward", "gsx_compute_sigma_backward", "gsx_project_to_screen_backward",
    "gsx_compute_camera_space_points_backward", "gsx_sort_workspace_bytes", "gsx_sort_pairs_u64", "gsx_sort_pairs_u32", "gsx_tile_ranges",
    "gsx_ctx_create", "gsx_ctx_destroy", "gsx_forward", "gsx_backward", "gsx_ctx_stats", "gsx_ctx_reserve", "gsx_ctx_overflow",
    "gsx_ctx_state", "gsx_render_view_host", "gsx_ctx_enable_timing", "gsx_ctx_timing", "gsx_ctx_count_work", "gsx_ctx_count_hits", "gsx_measure_fp32_tflops",
    "gsx_batch_reserve", "gsx_batch_forward_params", "gsx_batch_forward_view", "gsx_batch_backward_view",
    "gsx_batch_backward_params", "gsx_batch_render_host", "gsx_batch_graph_capture", "gsx_graph_launch", "gsx_graph_destroy", "gsx_batch_extract_view_rgb_grad",
    "gsx_sh_grads_from_view_rgb", "gsx_ctx_build_sorted_keys", "gsx_batch_extract_view_rgb_grad_peers",
    "gsx_allreduce_multimem_f32", "gsx_push_view_rgb_peers", "gsx_adam_step", "gsx_adam_step_fused",
    "gsx_fused_loss_workspace_bytes", "gsx_fused_loss", "gsx_fused_loss_static",
    "gsx_compute_morton_codes", "gsx_morton_order_workspace_bytes", "gsx_morton_order", "gsx_gather_rows", "gsx_gather_rows_multi",
    "gsx_dp_unique_id", "gsx_dp_init", "gsx_dp_destroy", "gsx_dp_grads", "gsx_dp_step", "gsx_dp_render_host", "gsx_dp_render_host_async", "gsx_dp_allreduce",
    "gsx_dp_enable_timing", "gsx_dp_timing",
    "gsx_clone_gaussians", "gsx_split_gaussians", "gsx_densify_workspace_bytes", "gsx_densify_plan", "gsx_densify_apply",
]


class GsxView(C.Structure):
    _fields_ = [("view", C.c_float * 16), ("proj", C.c_float * 16), ("campos", C.c_float * 3),
                ("focal_x", C.c_float), ("focal_y", C.c_float), ("tan_fovx", C.c_float), ("tan_fovy", C.c_float),
                ("width", C.c_int), ("height", C.c_int), ("near_thresh", C.c_float), ("cull_padding", C.c_int),
                ("mh_dist", C.c_float), ("background", C.c_float), ("l_max", C.c_int), ("flags", C.c_int)]


class GsxParams(C.Structure):
    _fields_ = [("xyz", C.c_void_p), ("scale", C.c_void_p), ("quaternion", C.c_void_p), ("opacity", C.c_void_p),
                ("band0", C.c_void_p), ("sh", C.c_void_p), ("sh_stride", C.c_int)]


class GsxGrads(C.Structure):
    _fields_ = [("xyz", C.c_void_p), ("scale", C.c_void_p), ("quaternion", C.c_void_p), ("opacity", C.c_void_p),
                ("band0", C.c_void_p), ("sh", C.c_void_p), ("uv", C.c_void_p)]


class GsxStats(C.Structure):
    _fields_ = [("N", C.c_int), ("P", C.c_int), ("K", C.c_longlong), ("capacity", C.c_longlong)]


class GsxState(C.Structure):
    _fields_ = [("image", C.c_void_p), ("weight_per_pixel", C.c_void_p), ("splats_per_pixel", C.c_void_p),
                ("rec0", C.c_void_p), ("rec1", C.c_void_p), ("rec2", C.c_void_p), ("radius", C.c_void_p),
                ("depth", C.c_void_p), ("mask", C.c_void_p), ("tiles_touched", C.c_void_p),
                ("sorted_gaussians", C.c_void_p), ("sorted_keys", C.c_void_p), ("ranges", C.c_void_p),
                ("num_instances", C.c_void_p), ("packed_grads", C.c_void_p)]


class GsxAdamGroup(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("width", C.c_int), ("lr", C.c_float)]


GSX_ADAM_MAX_GROUPS = 8


class GsxAdamEpilogue(C.Structure):
    _fields_ = [("params", GsxGrads), ("exp_avg", C.c_void_p * 6), ("exp_avg_sq", C.c_void_p * 6), ("lr", C.c_float * 6),
                ("b1", C.c_float), ("b2", C.c_float), ("eps", C.c_float), ("bias1", C.c_float), ("bias2", C.c_float)]


GSX_DP_ID_BYTES = 128


class GsxDensityConfig(C.Structure):
    _fields_ = [("uv_grad_threshold", C.c_float), ("delete_opacity_threshold", C.c_float), ("scene_extent", C.c_float),
                ("split_scale_factor", C.c_float), ("max_gaussians", C.c_int)]


class GsxDensityCounts(C.Structure):
    _fields_ = [("n_in", C.c_int), ("n_prune", C.c_int), ("n_clone", C.c_int), ("n_split", C.c_int), ("n_out", C.c_int)]


class GsxGaussianSet(C.Structure):
    _fields_ = [("xyz", C.c_void_p), ("scale", C.c_void_p), ("quaternion", C.c_void_p), ("opacity", C.c_void_p),
                ("band0", C.c_void_p), ("sh", C.c_void_p), ("sh_stride", C.c_int), ("sh_width", C.c_int),
                ("exp_avg", C.c_void_p * 6), ("exp_avg_sq", C.c_void_p * 6)]


class GsxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"gsx status {code}: {msg}")
        self.code = code


_lib = None


def lib():
    """Loads libgsx.so; raises if it has not been built (python __graft_entry__.py build / make -C 3dgs_b200/csrc)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `make -C 3dgs_b200/csrc` "
                              "(there is no CPU fallback for the splatting path)")
        try:
            import torch  # noqa: F401  (makes libcudart.so.12 resolvable before dlopen)
        except Exception:
            pass
        L = C.CDLL(LIB_PATH)
        L.gsx_last_error.restype = C.c_char_p
        L.gsx_sorted_list_workspace_bytes.restype = C.c_size_t
        L.gsx_sort_workspace_bytes.restype = C.c_size_t
        L.gsx_sort_workspace_bytes.argtypes = [C.c_longlong]
        L.gsx_morton_order_workspace_bytes.restype = C.c_size_t
        L.gsx_morton_order_workspace_bytes.argtypes = [C.c_int]
        L.gsx_fused_loss_workspace_bytes.restype = C.c_size_t
        L.gsx_fused_loss_workspace_bytes.argtypes = [C.c_int, C.c_int]
        L.gsx_sorted_list_workspace_bytes.argtypes = [C.c_int, C.c_longlong]
        L.gsx_densify_workspace_bytes.restype = C.c_size_t
        L.gsx_densify_workspace_bytes.argtypes = [C.c_int]
        _lib = L
    return _lib


def check(rc):
    if rc != GSX_OK:
        raise GsxError(rc, lib().gsx_last_error().decode())


def f32(x):
    return C.c_float(float(x))
