"""ORACLE — TEST INFRASTRUCTURE ONLY (never imported by the product package).

ctypes binding of ``oracle/egr_oracle.cpp``, the CPU restatement of the reference's differentiable
Gaussian ray tracer (see the header of that file for the reference file:line map).  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may
import this module, and only as the checker.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libegr_oracle.so")
_lib = None

CONFIG_DEFAULTS = dict(  # core/config.h:32-51
    exp_power=3.0,
    alpha_threshold=0.005,
    transmittance_threshold=0.01,
    global_scale_factor=1.0,
    loss_weight_diffuse=1.0,
    loss_weight_specular=1.0,
    loss_weight_depth=1.0,
    loss_weight_normal=1.0,
    loss_weight_f0=1.0,
    loss_weight_roughness=1.0,
    eps_forward_normalization=1e-12,
    eps_scale_grad=1e-12,
    eps_ray_surface_offset=0.01,
    eps_min_roughness=0.01,
    reflection_invalid_normal_threshold=0.7,
    backfacing_invalid_normal_threshold=0.9,
    backfacing_max_dist=0.1,
    num_bounces=2,
    jitter_primary_rays=1,
    accumulate_samples=0,
    max_bounces=2,  # flags.h:5
)
_FLOAT_KEYS = [k for k, v in CONFIG_DEFAULTS.items() if isinstance(v, float)]
_INT_KEYS = [k for k, v in CONFIG_DEFAULTS.items() if isinstance(v, int)]


class OrcConfig(C.Structure):
    _fields_ = [(k, C.c_float) for k in _FLOAT_KEYS] + [(k, C.c_int32) for k in _INT_KEYS]


class OrcCamera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("rot_c2w", C.c_float * 9), ("fovy", C.c_float), ("znear", C.c_float),
                ("zfar", C.c_float)]


_FP = C.POINTER(C.c_float)
_IP = C.POINTER(C.c_int32)
_UP = C.POINTER(C.c_uint32)
SCENE_KEYS = ["mean", "scale", "rotation", "opacity", "rgb", "normal", "f0", "roughness"]
SCENE_CH = dict(mean=3, scale=3, rotation=4, opacity=1, rgb=3, normal=3, f0=3, roughness=1)


class OrcScene(C.Structure):
    _fields_ = [("n", C.c_int32)] + [(k, _FP) for k in SCENE_KEYS]


TARGET_KEYS = ["diffuse", "specular", "depth", "normal", "f0", "roughness"]
TARGET_CH = dict(diffuse=3, specular=3, depth=1, normal=3, f0=3, roughness=1)


class OrcTargets(C.Structure):
    _fields_ = [(k, _FP) for k in TARGET_KEYS]


OUT_PLANES = dict(rgb=3, depth=1, normal=3, f0=3, roughness=1, transmittance=1, total_transmittance=1, ray_origin=3,
                  ray_direction=3)
ACC_PLANES = dict(acc_rgb=3, acc_transmittance=1, acc_total_transmittance=1, acc_depth=1, acc_normal=3, acc_f0=3,
                  acc_roughness=1)


class OrcOutputs(C.Structure):
    _fields_ = ([(k, _FP) for k in OUT_PLANES] + [("final_", _FP)] + [(k, _FP) for k in ACC_PLANES] +
                [("acc_count", _IP), ("random_seeds", _UP), ("num_accumulated", _IP), ("num_hits_per_pixel", _IP),
                 ("num_composited_per_pixel", _IP), ("num_steps_per_pixel", _IP), ("used_ray_origin", _FP),
                 ("used_ray_direction", _FP), ("step_executed", _IP)])


GRAD_KEYS = ["dL_drgb", "dL_dnormal", "dL_df0", "dL_droughness", "dL_dopacity", "dL_dscale", "dL_dmean",
             "dL_drotation", "total_weight"]
GRAD_CH = dict(dL_drgb=3, dL_dnormal=3, dL_df0=3, dL_droughness=1, dL_dopacity=1, dL_dscale=3, dL_dmean=3,
               dL_drotation=4, total_weight=1)


class OrcGrads(C.Structure):
    _fields_ = [(k, _FP) for k in GRAD_KEYS]


class OrcHitLists(C.Structure):
    _fields_ = [("cap", C.c_int32), ("counts", _IP), ("ids", _IP), ("ts", _FP)]


class OrcOverride(C.Structure):
    _fields_ = [("ray_origin", _FP), ("ray_direction", _FP), ("step_executed", _IP), ("records", _FP),
                ("leaf_rank", _UP)]


def build(force: bool = False) -> str:
    """Compile the oracle with g++ (no GPU, no reference needed)."""
    src = os.path.join(_HERE, "egr_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_num_threads.restype = C.c_int
        for _n in ("orc_det_exp", "orc_det_log", "orc_det_sin", "orc_det_cos"):
            getattr(_lib, _n).restype = C.c_float
            getattr(_lib, _n).argtypes = [C.c_float]
        _lib.orc_tea4.restype = C.c_uint32
        _lib.orc_tea4.argtypes = [C.c_uint32, C.c_uint32]
        _lib.orc_rnd.restype = C.c_float
        _lib.orc_rnd.argtypes = [C.POINTER(C.c_uint32)]
        _lib.orc_prepare.argtypes = [C.POINTER(OrcScene), C.POINTER(OrcConfig), _FP]
        _lib.orc_render.argtypes = [
            C.POINTER(OrcScene), C.POINTER(OrcConfig), C.POINTER(OrcCamera), C.c_int, C.c_int, C.c_uint32, C.c_int,
            C.POINTER(OrcTargets), C.POINTER(OrcOutputs), C.POINTER(OrcGrads), C.POINTER(OrcHitLists),
            C.POINTER(OrcOverride), C.c_int, C.c_int]
    return _lib


def knn3_mean_dist2(points) -> np.ndarray:
    """Oracle of ``distCUDA2`` (simple_knn.cu:141-199): brute force, O(N^2); keep N at a few ten thousand."""
    p = _f32(points).reshape(-1, 3)
    out = np.zeros(p.shape[0], np.float32)
    f = lib().orc_knn3_mean_dist2
    f.argtypes = [_FP, C.c_int, _FP]
    f.restype = None
    f(_fp(p), int(p.shape[0]), _fp(out))
    return out


def num_threads() -> int:
    return int(lib().orc_num_threads())


def _f32(a, shape=None):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float32))
    if shape is not None:
        a = a.reshape(shape)
    return a


def _fp(a):
    return a.ctypes.data_as(_FP)


def make_config(**kw) -> OrcConfig:
    d = dict(CONFIG_DEFAULTS)
    for k, v in kw.items():
        if k not in d:
            raise KeyError(k)
        d[k] = v
    c = OrcConfig()
    for k in _FLOAT_KEYS:
        setattr(c, k, float(d[k]))
    for k in _INT_KEYS:
        setattr(c, k, int(d[k]))
    return c


def make_camera(origin, rot_c2w, fovy, znear=0.01, zfar=999.9) -> OrcCamera:
    cam = OrcCamera()
    cam.origin[:] = [float(v) for v in np.asarray(origin).reshape(3)]
    cam.rot_c2w[:] = [float(v) for v in np.asarray(rot_c2w, dtype=np.float32).reshape(9)]
    cam.fovy, cam.znear, cam.zfar = float(fovy), float(znear), float(zfar)
    return cam


def _scene_struct(scene):
    keep = {}
    s = OrcScene()
    n = int(np.asarray(scene["mean"]).shape[0])
    s.n = n
    for k in SCENE_KEYS:
        keep[k] = _f32(scene[k], (n, SCENE_CH[k]))
        setattr(s, k, _fp(keep[k]))
    return s, keep


def prepare(scene, cfg: OrcConfig) -> np.ndarray:
    """Packed per-Gaussian geometry records [N,16] (same layout as the CUDA geometry record)."""
    s, keep = _scene_struct(scene)
    rec = np.zeros((s.n, 16), np.float32)
    lib().orc_prepare(C.byref(s), C.byref(cfg), _fp(rec))
    return rec


def render(scene, cfg: OrcConfig, cam: OrcCamera, width: int, height: int, total_num_calls: int = 1,
           grads: bool = False, targets=None, accum=None, hit_cap: int = 0, override=None, pixel_range=None):
    """One ``raytrace()`` call of the reference, on the CPU.  Returns a dict of numpy arrays.

    ``accum``: dict with the 7 ``acc_*`` planes [S,P,C] + ``acc_count`` (int) updated in place when
    ``cfg.accumulate_samples``.  ``override``: dict(ray_origin, ray_direction, step_executed, records).
    """
    L = lib()
    P = width * height
    S = cfg.max_bounces + 1
    s, keep = _scene_struct(scene)
    n = s.n
    out = OrcOutputs()
    res = {}
    for k, ch in OUT_PLANES.items():
        res[k] = np.zeros((S, height, width, ch), np.float32)
        if k in ("transmittance", "total_transmittance"):
            res[k][:] = 1.0
        setattr(out, k, _fp(res[k]))
    res["final"] = np.zeros((1, height, width, 3), np.float32)
    out.final_ = _fp(res["final"])
    if accum is not None:
        cnt = np.array([int(accum["acc_count"])], np.int32)
        for k in ACC_PLANES:
            setattr(out, k, _fp(accum[k]))
        out.acc_count = cnt.ctypes.data_as(_IP)
    elif cfg.accumulate_samples:
        raise ValueError("accumulate_samples needs accum buffers")
    res["random_seeds"] = np.zeros((height, width, 1), np.uint32)
    out.random_seeds = res["random_seeds"].ctypes.data_as(_UP)
    for k in ("num_accumulated", "num_hits_per_pixel", "num_composited_per_pixel", "num_steps_per_pixel"):
        res[k] = np.zeros((height, width), np.int32)
        setattr(out, k, res[k].ctypes.data_as(_IP))
    res["used_ray_origin"] = np.zeros((S, height, width, 3), np.float32)
    res["used_ray_direction"] = np.zeros((S, height, width, 3), np.float32)
    res["step_executed"] = np.zeros((S, height, width), np.int32)
    out.used_ray_origin = _fp(res["used_ray_origin"])
    out.used_ray_direction = _fp(res["used_ray_direction"])
    out.step_executed = res["step_executed"].ctypes.data_as(_IP)

    tg = OrcTargets()
    keep_t = {}
    if grads:
        if targets is None:
            targets = {}
        for k in TARGET_KEYS:
            keep_t[k] = _f32(targets.get(k, np.zeros((height, width, TARGET_CH[k]))), (height, width, TARGET_CH[k]))
            setattr(tg, k, _fp(keep_t[k]))
    gr = OrcGrads()
    for k in GRAD_KEYS:
        res[k] = np.zeros((n, GRAD_CH[k]), np.float32)
        setattr(gr, k, _fp(res[k]))
    hl = OrcHitLists()
    hl.cap = int(hit_cap)
    if hit_cap > 0:
        res["hit_counts"] = np.zeros((S, height, width), np.int32)
        res["hit_ids"] = np.full((S, height, width, hit_cap), -1, np.int32)
        res["hit_ts"] = np.zeros((S, height, width, hit_cap), np.float32)
        hl.counts = res["hit_counts"].ctypes.data_as(_IP)
        hl.ids = res["hit_ids"].ctypes.data_as(_IP)
        hl.ts = _fp(res["hit_ts"])
    ov = OrcOverride()
    keep_o = {}
    ovp = None
    if override is not None:
        if override.get("ray_origin") is not None:
            keep_o["o"] = _f32(override["ray_origin"], (S, P, 3))
            keep_o["d"] = _f32(override["ray_direction"], (S, P, 3))
            ov.ray_origin, ov.ray_direction = _fp(keep_o["o"]), _fp(keep_o["d"])
        if override.get("step_executed") is not None:
            keep_o["e"] = np.ascontiguousarray(np.asarray(override["step_executed"], np.int32).reshape(S, P))
            ov.step_executed = keep_o["e"].ctypes.data_as(_IP)
        if override.get("records") is not None:
            keep_o["r"] = _f32(override["records"], (n, 16))
            ov.records = _fp(keep_o["r"])
        if override.get("leaf_rank") is not None:
            keep_o["k"] = np.ascontiguousarray(np.asarray(override["leaf_rank"], np.uint32).reshape(n))
            ov.leaf_rank = keep_o["k"].ctypes.data_as(_UP)
        ovp = C.byref(ov)
    pb, pe = (0, P) if pixel_range is None else pixel_range
    L.orc_render(C.byref(s), C.byref(cfg), C.byref(cam), width, height, C.c_uint32(total_num_calls), int(bool(grads)),
                 C.byref(tg), C.byref(out), C.byref(gr), C.byref(hl), ovp, int(pb), int(pe))
    return res


# ---- oracle 2: the reference's own shader sources compiled for the host (oracle/ref_shim) -------------------
_REF_PATH = os.path.join(_HERE, "_ref", "libref_tracer.so")
_ref = None


def ref_available(try_build: bool = True) -> bool:
    """True when oracle/_ref/libref_tracer.so exists (built here from /root/reference, or shipped prebuilt)."""
    if not os.path.exists(_REF_PATH) and try_build and os.path.isdir("/root/reference/editable_gauss_refl/cuda/csrc"):
        try:
            subprocess.check_call(["make", "-s", "-C", _HERE, "ref"])
        except Exception:
            return False
    return os.path.exists(_REF_PATH)


def ref_lib():
    global _ref
    if _ref is None:
        if not ref_available():
            raise RuntimeError("oracle/_ref/libref_tracer.so is absent (reference checkout not available)")
        _ref = C.CDLL(_REF_PATH)
        _ref.ref_max_bounces.restype = C.c_int
        _ref.ref_num_threads.restype = C.c_int
        _ref.ref_render.restype = C.c_int
        _ref.ref_render.argtypes = [
            C.POINTER(OrcScene), C.POINTER(OrcConfig), C.POINTER(OrcCamera), C.c_int, C.c_int, C.c_uint32, C.c_int,
            C.POINTER(OrcTargets), C.POINTER(OrcOutputs), C.POINTER(OrcGrads), C.c_int, C.c_int]
    return _ref


def ref_render(scene, cfg: OrcConfig, cam: OrcCamera, width: int, height: int, total_num_calls: int = 1,
               grads: bool = False, targets=None, accum=None, pixel_range=None):
    """Same call as :func:`render`, executed by the reference's literal shader code on the host."""
    L = ref_lib()
    P = width * height
    S = cfg.max_bounces + 1
    s, keep = _scene_struct(scene)
    n = s.n
    out = OrcOutputs()
    res = {}
    for k, ch in OUT_PLANES.items():
        res[k] = np.zeros((S, height, width, ch), np.float32)
        setattr(out, k, _fp(res[k]))
    res["final"] = np.zeros((1, height, width, 3), np.float32)
    out.final_ = _fp(res["final"])
    cnt = np.array([int(accum["acc_count"]) if accum is not None else 0], np.int32)
    out.acc_count = cnt.ctypes.data_as(_IP)
    if accum is not None:
        for k in ACC_PLANES:
            setattr(out, k, _fp(accum[k]))
    elif cfg.accumulate_samples:
        raise ValueError("accumulate_samples needs accum buffers")
    res["random_seeds"] = np.zeros((height, width, 1), np.uint32)
    out.random_seeds = res["random_seeds"].ctypes.data_as(_UP)
    for k in ("num_accumulated", "num_hits_per_pixel"):
        res[k] = np.zeros((height, width), np.int32)
        setattr(out, k, res[k].ctypes.data_as(_IP))
    tg = OrcTargets()
    keep_t = {}
    if grads:
        targets = targets or {}
        for k in TARGET_KEYS:
            keep_t[k] = _f32(targets.get(k, np.zeros((height, width, TARGET_CH[k]))), (height, width, TARGET_CH[k]))
            setattr(tg, k, _fp(keep_t[k]))
    gr = OrcGrads()
    for k in GRAD_KEYS:
        res[k] = np.zeros((n, GRAD_CH[k]), np.float32)
        setattr(gr, k, _fp(res[k]))
    pb, pe = (0, P) if pixel_range is None else pixel_range
    rc = L.ref_render(C.byref(s), C.byref(cfg), C.byref(cam), width, height, C.c_uint32(total_num_calls),
                      int(bool(grads)), C.byref(tg), C.byref(out), C.byref(gr), int(pb), int(pe))
    if rc != 0:
        raise RuntimeError(f"ref_render failed ({rc}): max_bounces must equal the reference's compile-time value")
    return res


# ---- row f3: the reference's own simple-knn compiled in place for sm_100 (oracle/ref_shim/ref_knn.cu) -------
_REF_KNN_PATH = os.path.join(_HERE, "_ref", "libref_knn.so")


def ref_knn_available(try_build: bool = True) -> bool:
    """True when oracle/_ref/libref_knn.so exists (built here with nvcc from /root/reference, or shipped prebuilt).
    It is CUDA code: usable on the GPU box only."""
    if not os.path.exists(_REF_KNN_PATH) and try_build and os.path.isfile(
            "/root/reference/submodules/simple-knn/simple_knn.cu"):
        try:
            subprocess.check_call(["make", "-s", "-C", _HERE, "refknn"])
        except Exception:
            return False
    return os.path.exists(_REF_KNN_PATH)


def ref_knn3_mean_dist2(points_dev_ptr: int, n: int, out_dev_ptr: int) -> None:
    """Run the reference's SimpleKNN::knn on device pointers (synchronises the device)."""
    L = C.CDLL(_REF_KNN_PATH)
    L.ref_knn3_mean_dist2.restype = C.c_int
    L.ref_knn3_mean_dist2.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
    rc = L.ref_knn3_mean_dist2(int(n), points_dev_ptr, out_dev_ptr)
    if rc:
        raise RuntimeError(f"reference simple-knn failed with CUDA error {rc}")
