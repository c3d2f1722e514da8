"""ctypes binding of the C ABI declared in ``include/egr_b200.h`` (no torch types cross this boundary)."""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

_FP = C.c_void_p  # every device pointer travels as a plain address


def _ptr_struct(name, fields):
    return type(name, (C.Structure,), {"_fields_": [(f, _FP) for f in fields]})


CAMERA_FIELDS = ["origin", "vertical_fov_radians", "rotation_c2w", "rotation_w2c", "znear", "zfar"]
CONFIG_FIELDS = [
    "exp_power", "alpha_threshold", "transmittance_threshold", "accumulate_samples", "jitter_primary_rays",
    "num_bounces", "global_scale_factor", "loss_weight_diffuse", "loss_weight_specular", "loss_weight_depth",
    "loss_weight_normal", "loss_weight_f0", "loss_weight_roughness", "eps_forward_normalization", "eps_scale_grad",
    "eps_ray_surface_offset", "eps_min_roughness", "reflection_invalid_normal_threshold",
    "backfacing_invalid_normal_threshold", "backfacing_max_dist",
]
FRAMEBUFFER_FIELDS = [
    "output_rgb", "output_depth", "output_normal", "output_f0", "output_roughness", "output_transmittance",
    "output_total_transmittance", "output_ray_origin", "output_ray_direction", "output_final", "output_denoised",
    "accumulated_rgb", "accumulated_transmittance", "accumulated_total_transmittance", "accumulated_depth",
    "accumulated_normal", "accumulated_f0", "accumulated_roughness", "accumulated_sample_count",
    "target_diffuse", "target_specular", "target_depth", "target_normal", "target_f0", "target_roughness",
]
GAUSSIAN_FIELDS = ["rgb", "normal", "f0", "roughness", "opacity", "scale", "mean", "rotation"]
METADATA_FIELDS = ["grads_enabled", "total_num_calls", "random_seeds"]
STATS_FIELDS = ["num_accumulated_per_pixel", "num_traversed_per_pixel", "num_hits_per_pixel",
                "num_composited_per_pixel", "num_steps_per_pixel", "hitlog_status"]

EgrCamera = _ptr_struct("EgrCamera", CAMERA_FIELDS)
class EgrEdit(C.Structure):  # include/egr_b200.h (row f2)
    _fields_ = [("translate", C.c_float * 3), ("center", C.c_float * 3), ("scale", C.c_float),
                ("rotation", C.c_float * 9), ("quaternion", C.c_float * 4), ("kill", C.c_int)]


class EgrAdam(C.Structure):  # include/egr_b200.h (row f1)
    _fields_ = [("lr", C.c_double * 8), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double),
                ("scale_decay", C.c_double), ("step", C.c_uint32), ("clamp", C.c_int), ("zero_grads", C.c_int),
                ("zero_total_weight", C.c_int)]


GRAD_TARGET_FIELDS = ["mean", "opacity", "scale", "roughness", "rotation", "rgb", "normal", "f0"]
EgrGradTargets = _ptr_struct("EgrGradTargets", GRAD_TARGET_FIELDS)
EgrConfig = _ptr_struct("EgrConfig", CONFIG_FIELDS)
EgrFramebuffer = _ptr_struct("EgrFramebuffer", FRAMEBUFFER_FIELDS)
EgrMetadata = _ptr_struct("EgrMetadata", METADATA_FIELDS)
EgrStats = _ptr_struct("EgrStats", STATS_FIELDS)


class EgrGaussians(C.Structure):
    _fields_ = [("count", C.c_uint32)] + [(f, _FP) for f in GAUSSIAN_FIELDS] + [("grad_records", _FP)]


class EgrParams(C.Structure):
    _fields_ = [
        ("image_width", C.c_uint32), ("image_height", C.c_uint32), ("num_steps", C.c_uint32), ("tile_rows", C.c_uint32),
        ("camera", EgrCamera), ("config", EgrConfig), ("framebuffer", EgrFramebuffer), ("gaussians", EgrGaussians),
        ("metadata", EgrMetadata), ("stats", EgrStats),
        ("bvh_storage", _FP), ("bvh_storage_bytes", C.c_size_t),
        ("pixel_state", _FP), ("pixel_state_bytes", C.c_size_t),
        ("hitlog", _FP), ("hitlog_bytes", C.c_size_t),
        ("used_ray_origin", _FP), ("used_ray_direction", _FP), ("step_executed", _FP),
        ("override_ray_origin", _FP), ("override_ray_direction", _FP), ("override_step_executed", _FP),
        ("early_stop_transmittance", C.c_float), ("reserved1", C.c_uint32),
    ]


EGR_ABI_VERSION = 2
EGR_MAX_STEPS = 8
EGR_GRAD_STRIDE = 24
EGR_RECORD_FLOATS = 16
EGR_NOALLOC_MAX_GAUSSIANS = 65536
EGR_COMPACT_MAX_ARRAYS = 40
EGR_PHASE_BEGIN = 0x1
EGR_PHASE_FINISH = 0x10000
EGR_PHASE_ALL = 0xFFFFFFFF


def EGR_PHASE_STEP(s: int) -> int:
    return 0x2 << s

# every symbol include/egr_b200.h declares: (restype, argtypes)
SYMBOLS = {
    "egr_abi_version": (C.c_int, []),
    "egr_last_error": (C.c_char_p, []),
    "egr_max_alpha": (C.c_double, []),
    "egr_roughness_downweight_grad": (C.c_int, []),
    "egr_roughness_downweight_grad_power": (C.c_double, []),
    "egr_buffer_size": (C.c_int, []),
    "egr_max_iterations": (C.c_int, []),
    "egr_bvh_storage_bytes": (C.c_size_t, [C.c_uint32]),
    "egr_bvh_scratch_bytes": (C.c_size_t, [C.c_uint32]),
    "egr_bvh_build": (C.c_int, [C.POINTER(EgrGaussians), C.POINTER(EgrConfig), _FP, C.c_size_t, _FP, C.c_size_t, _FP]),
    "egr_bvh_build_noalloc": (C.c_int, [C.POINTER(EgrGaussians), C.POINTER(EgrConfig), _FP, C.c_size_t, _FP]),
    "egr_bvh_refit": (C.c_int, [C.POINTER(EgrGaussians), C.POINTER(EgrConfig), _FP, C.c_size_t, _FP]),
    "egr_pixel_state_bytes": (C.c_size_t, [C.c_uint32, C.c_uint32, C.c_uint32]),
    "egr_hitlog_bytes_for_hits": (C.c_size_t, [C.c_uint64]),
    "egr_camera_set_pose": (C.c_int, [C.POINTER(EgrCamera), C.POINTER(C.c_float), C.POINTER(C.c_float), _FP]),
    "egr_gradients_accumulate": (C.c_int, [_FP, C.c_uint32, C.POINTER(EgrGradTargets), _FP]),
    "egr_params_export": (C.c_int, [C.POINTER(EgrGradTargets), C.POINTER(EgrGradTargets), C.c_uint32, _FP]),
    "egr_knn3_scratch_bytes": (C.c_size_t, [C.c_uint32]),
    "egr_knn3_mean_dist2": (C.c_int, [_FP, C.c_uint32, _FP, _FP, C.c_size_t, _FP]),
    "egr_prune_keep_mask": (C.c_int, [_FP, C.c_uint32, C.c_float, C.c_float, _FP, _FP]),
    "egr_prune_near_cameras": (C.c_int, [_FP, C.c_uint32, _FP, _FP, C.c_uint32, _FP, _FP]),
    "egr_compact_scratch_bytes": (C.c_size_t, [C.c_uint32]),
    "egr_compact_rows": (C.c_int, [_FP, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                   C.POINTER(C.c_uint32), _FP, C.c_size_t, _FP, _FP]),
    "egr_duplicate_rows": (C.c_int, [_FP, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                     C.POINTER(C.c_uint32), C.c_int, C.POINTER(C.c_float), _FP, C.c_size_t, _FP, _FP]),
    "egr_select_box": (C.c_int, [_FP, C.c_uint32, C.POINTER(C.c_float), C.POINTER(C.c_float), _FP, _FP]),
    "egr_select_cylinder": (C.c_int, [_FP, C.c_uint32, C.POINTER(C.c_float), C.POINTER(C.c_float), _FP, _FP]),
    "egr_edit_apply": (C.c_int, [_FP, C.c_uint32, _FP, _FP, _FP, _FP, C.POINTER(EgrEdit), _FP]),
    "egr_adam_step": (C.c_int, [_FP, C.c_uint32, C.POINTER(EgrGradTargets), C.POINTER(EgrGradTargets),
                                C.POINTER(EgrGradTargets), C.POINTER(EgrAdam), _FP]),
    "egr_raytrace": (C.c_int, [C.POINTER(EgrParams), C.c_int, _FP]),
    "egr_raytrace_phases": (C.c_int, [C.POINTER(EgrParams), C.c_int, C.c_uint32, _FP]),
    "egr_reset_accumulators": (C.c_int, [C.POINTER(EgrParams), _FP]),
    "egr_denoise_passthrough": (C.c_int, [C.POINTER(EgrParams), _FP]),
    "egr_denoise": (C.c_int, [C.POINTER(EgrParams), _FP]),
    "egr_hitlog_export": (C.c_int, [C.POINTER(EgrParams), C.c_int, C.c_int, _FP, _FP, _FP, _FP]),
    "egr_bvh_export": (C.c_int, [_FP, C.c_size_t, C.c_uint32, _FP, _FP, _FP, _FP, _FP]),
    "egr_bvh_export_wide": (C.c_int, [_FP, C.c_size_t, C.c_uint32, _FP, _FP, _FP, _FP]),
    "egr_bvh_wide_capacity": (C.c_uint32, [C.c_uint32]),
}

_lib = None


def lib_path() -> str:
    return _build.LIB_PATH


def load(build_if_missing: bool = True):
    """Load ``libegr_b200.so``.  There is no fallback: a missing/unbuildable extension raises."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    if not os.path.exists(path) or (build_if_missing and _build.needs_build()):
        if not build_if_missing:
            raise RuntimeError(f"CUDA extension missing: {path} (run __graft_entry__.build())")
        try:
            _build.build()
        except Exception as e:  # no CPU fallback exists, by design
            if not os.path.exists(path):
                raise RuntimeError(f"CUDA extension missing and cannot be built: {e}") from e
    lib = C.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        try:
            fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        except AttributeError:
            if os.environ.get("EGR_B200_LIB"):  # an experiment's variant built from older sources
                continue
            raise
        fn.restype = res
        fn.argtypes = args
    if lib.egr_abi_version() != EGR_ABI_VERSION:
        raise RuntimeError(f"ABI mismatch: library {lib.egr_abi_version()} != binding {EGR_ABI_VERSION}")
    _lib = lib
    return lib


def check(rc: int) -> None:
    """Status code -> RuntimeError (the reference surfaces C++ exceptions as RuntimeError)."""
    if rc != 0:
        msg = load().egr_last_error()
        raise RuntimeError(f"egr_b200 error {rc}: {msg.decode(errors='replace') if msg else ''}")
