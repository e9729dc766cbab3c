"""ctypes binding of libnrc_b200.so -- the stub a Python host needs for include/nrc_b200.h.

There is no fallback: if the shared library is missing this raises, and every call that returns a non-zero
nrc_status raises NrcError with nrc_last_error()."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libnrc_b200.so")

PARAM_COUNT = 20995
WEIGHT_COUNT = 20672
FRAME_RGBA32F, FRAME_BGRA8 = 0, 1
RES_INTENDED, RES_AS_SHIPPED_STD140 = 0, 1
MLP_FP32, MLP_BF16_TC = 0, 1


class NrcError(RuntimeError):
    pass


class VolumeParams(C.Structure):   # nrc_volume_params
    _fields_ = [("random", C.c_float * 4), ("useNN", C.c_uint32), ("showNonNN", C.c_uint32), ("densityFactor", C.c_float),
                ("g", C.c_float), ("noNnSpp", C.c_int32), ("withNnSpp", C.c_int32)]


class DirLightData(C.Structure):   # nrc_dir_light
    _fields_ = [("color", C.c_float * 3), ("zenith", C.c_float), ("dir", C.c_float * 3), ("azimuth", C.c_float),
                ("strength", C.c_float)]


class PointLightData(C.Structure):   # nrc_point_light
    _fields_ = [("pos", C.c_float * 3), ("strength", C.c_float), ("color", C.c_float * 3)]


class CameraData(C.Structure):   # nrc_camera
    _fields_ = [("projView", C.c_float * 16), ("invProjView", C.c_float * 16), ("pos", C.c_float * 3)]


# every symbol include/nrc_b200.h declares (tests check the library exports all of them)
SYMBOLS = """nrc_last_error nrc_version nrc_ctx_create nrc_ctx_destroy nrc_sync nrc_ctx_stream nrc_ctx_sm_count
nrc_ctx_set_precision nrc_ctx_launch_count nrc_mlp_create nrc_mlp_destroy nrc_mlp_set_config nrc_mlp_set_params
nrc_mlp_set_state nrc_mlp_get_params nrc_mlp_reset_stats nrc_mlp_get_stats nrc_mlp_grad_ptr nrc_mlp_grad_count
nrc_mrhe_resolutions nrc_mrhe_create nrc_mrhe_destroy nrc_mrhe_set_config nrc_mrhe_table_bytes nrc_mrhe_set_tables
nrc_mrhe_get_tables nrc_mrhe_grad_ptr nrc_mrhe_grad_extent nrc_encode nrc_encode_dir nrc_infer nrc_train_accumulate
nrc_optimizer_step nrc_volume_create nrc_volume_destroy nrc_volume_set_params nrc_volume_sample nrc_envmap_create
nrc_envmap_destroy nrc_envmap_set_strengths nrc_envmap_set_importance_sampling nrc_hdr4f_to_cdf nrc_renderer_create nrc_renderer_destroy
nrc_renderer_set_camera nrc_renderer_set_dir_light nrc_renderer_set_point_light nrc_renderer_set_partition nrc_resize
nrc_render nrc_frame_train nrc_frame_render nrc_frame_bytes nrc_frame_ptr nrc_read_frame nrc_read_render_probes
nrc_read_train_probes nrc_frame_stats nrc_frame_totals nrc_renderer_set_partition_strided nrc_read_frame_partition
nrc_frame_partition_bytes nrc_kernel_count nrc_kernel_name nrc_selftest_strict_math
nrc_ctx_profile_enable nrc_ctx_profile_read nrc_ctx_profile_timeline nrc_dp_handle_bytes nrc_dp_create nrc_dp_export
nrc_renderer_set_pool_capacity nrc_renderer_get_pool_capacity nrc_renderer_set_train_env_light nrc_read_frame_partition_async nrc_readback_wait nrc_mlp_get_stats_async nrc_dp_connect nrc_dp_reduce_step nrc_dp_reduce_step_phase nrc_dp_status nrc_dp_destroy
nrc_ctx_set_infer_mode nrc_feature_tile_bytes nrc_encode_tiles nrc_mlp_forward_tiles
nrc_mrhe_set_fixed_mode nrc_mlp_set_adam nrc_mrhe_set_adam
nrc_mlp_get_adam_state nrc_mlp_set_adam_state nrc_mrhe_get_adam_state nrc_mrhe_set_adam_state
nrc_train_accumulate_static nrc_train_accumulate_static_max_samples""".split()

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise NrcError("libnrc_b200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                           "there is no CPU fallback")
        L = C.CDLL(SO_PATH)
        L.nrc_last_error.restype = C.c_char_p
        L.nrc_version.restype = C.c_char_p
        L.nrc_kernel_name.restype = C.c_char_p
        L.nrc_ctx_stream.restype = C.c_void_p
        L.nrc_ctx_launch_count.restype = C.c_uint64
        for n in ("nrc_mlp_grad_ptr", "nrc_mrhe_grad_ptr", "nrc_frame_ptr"):
            getattr(L, n).restype = C.c_void_p
        for n in ("nrc_mlp_grad_count", "nrc_mrhe_table_bytes", "nrc_mrhe_grad_extent", "nrc_frame_bytes", "nrc_frame_partition_bytes",
                  "nrc_dp_handle_bytes", "nrc_feature_tile_bytes", "nrc_train_accumulate_static_max_samples"):
            getattr(L, n).restype = C.c_size_t
        _lib = L
    return _lib


def check(status):
    if status != 0:
        raise NrcError("nrc_status %d: %s" % (status, lib().nrc_last_error().decode()))
