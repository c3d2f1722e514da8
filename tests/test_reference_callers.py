"""Boundary conformance with the reference's UNMODIFIED callers (SURVEY 8 row a21 / 8b caller usage matrix).

CPU tier; skipped when the reference checkout is absent (it does not exist on the GPU box).  Two parts:

1. *Surface walk*: the reference's caller sources are parsed (``ast``) and every method called on ``cuda_module`` and
   every attribute reached through a ``get_*()`` holder (directly or through a local variable bound to one) is
   collected; each must exist on ``egr_b200.Raytracer`` and its holders, as a tensor of the shape / dtype SURVEY
   Appendix B lists.
2. *Unmodified facade*: the reference's own ``editable_gauss_refl/renderer/{gaussian_raytracer,gaussian_renderer}.py``
   are imported as they are -- with ``editable_gauss_refl.make_raytracer`` re-pointed at this package (the seam of
   INTEGRATION.md) and stand-ins for the model / camera modules that need plyfile, kornia, simple_knn -- and driven
   through construction, parameter export, a grad-mode and a no-grad ``render()``, gradient import, ``zero_grad`` and
   ``rebuild_bvh``.  The holders are the package's real holder classes on the CPU; the five methods that launch
   device work are replaced by recorders (there is no CPU tracer), so what is checked is the call protocol: which
   holder tensors the reference mutates in place, in which layout, and which device methods it calls in which order.
"""
import ast
import importlib
import os
import sys
import types
from unittest import mock

import numpy as np
import pytest
import torch

import egr_b200
from egr_b200 import scenes
from egr_b200.raytracer import Raytracer

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "editable_gauss_refl")),
                                reason="reference checkout not available")

GETTERS = {"get_camera": "camera_data", "get_config": "config_data", "get_framebuffer": "framebuffer_data",
           "get_gaussians": "gaussian_data", "get_metadata": "meta_data", "get_stats": "stats_data",
           "get_ppll_forward_data": "ppll_forward_data", "get_ppll_backward_data": "ppll_backward_data"}
CALLERS = ["editable_gauss_refl/renderer/gaussian_raytracer.py", "editable_gauss_refl/renderer/gaussian_renderer.py",
           "train.py", "render.py", "measure_fps.py", "tools/render_novel_views.py", "gaussian_viewer.py"]


class SurfaceOnly(Raytracer):
    """The package's holders on the CPU; device work is recorded instead of launched."""

    def __init__(self, W, H, N, fwd=180_000_000, bwd=120_000_000):
        self._make_holders(torch.device("cpu"), W, H, N, fwd, bwd, None, False)
        self.calls = []
        self._bvh_count = N

    def raytrace(self):
        self.calls.append(("raytrace", torch.is_grad_enabled()))

    def update_bvh(self):
        self.calls.append(("update_bvh",))

    def rebuild_bvh(self):
        self.calls.append(("rebuild_bvh", self.gaussian_data.count))

    def denoise(self):
        self.calls.append(("denoise",))

    def reset_accumulators(self):
        self.calls.append(("reset_accumulators",))

    def resize(self, n):
        self.calls.append(("resize", int(n)))
        self.gaussian_data.resize(int(n))


def _collect_surface(path):
    """(methods called on cuda_module, {getter: attributes reached through it})."""
    tree = ast.parse(open(path).read())
    methods, attrs = set(), {}

    def getter_of(node, env):
        # <expr>.get_x()  or a name bound to one
        if isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute) and node.func.attr in GETTERS:
            return node.func.attr
        if isinstance(node, ast.Name) and node.id in env:
            return env[node.id]
        return None

    for fn in [n for n in ast.walk(tree) if isinstance(n, (ast.FunctionDef, ast.Module))]:
        env = {}
        for node in ast.walk(fn):
            if isinstance(node, ast.Assign) and len(node.targets) == 1 and isinstance(node.targets[0], ast.Name):
                g = getter_of(node.value, {})
                if g:
                    env[node.targets[0].id] = g
        for node in ast.walk(fn):
            if isinstance(node, ast.Attribute):
                v = node.value
                if isinstance(v, ast.Attribute) and v.attr == "cuda_module":
                    methods.add(node.attr)
                if isinstance(v, ast.Name) and v.id == "cuda_module":
                    methods.add(node.attr)
                g = getter_of(v, env)
                if g:
                    attrs.setdefault(g, set()).add(node.attr)
    return methods, attrs


def test_every_member_the_reference_callers_touch_exists():
    W, H, N, S = 12, 8, 5, 3
    rt = SurfaceOnly(W, H, N)
    methods, attrs = set(), {}
    for rel in CALLERS:
        p = os.path.join(REF, rel)
        if not os.path.exists(p):
            continue
        m, a = _collect_surface(p)
        methods |= m
        for k, v in a.items():
            attrs.setdefault(k, set()).update(v)
    # the walk really found the protocol (guards against a silently empty parse)
    assert {"raytrace", "update_bvh", "rebuild_bvh", "resize", "denoise", "reset_accumulators", "get_config",
            "get_framebuffer", "get_gaussians", "get_camera"} <= methods, methods
    assert {"num_bounces", "accumulate_samples"} <= attrs["get_config"]
    assert "total_weight" in attrs["get_gaussians"] and "output_denoised" in attrs["get_framebuffer"]
    for m in methods:
        assert callable(getattr(Raytracer, m, None)), f"Raytracer.{m} missing"
    for getter, names in attrs.items():
        holder = getattr(rt, getter)()
        assert holder is getattr(rt, GETTERS[getter])
        for name in names:
            assert hasattr(holder, name), f"{getter}().{name} missing"
    # shapes / dtypes of SURVEY Appendix B for everything the callers reach
    f32, i32, b8 = torch.float32, torch.int32, torch.bool
    want = {
        "get_camera": {"vertical_fov_radians": ((1,), f32), "znear": ((1,), f32), "zfar": ((1,), f32)},
        "get_config": dict([(k, ((1,), f32)) for k in (
            "exp_power", "alpha_threshold", "transmittance_threshold", "global_scale_factor", "loss_weight_diffuse",
            "loss_weight_specular", "loss_weight_depth", "loss_weight_normal", "loss_weight_f0", "loss_weight_roughness",
            "eps_forward_normalization", "eps_scale_grad", "eps_ray_surface_offset", "eps_min_roughness",
            "reflection_invalid_normal_threshold", "backfacing_invalid_normal_threshold", "backfacing_max_dist")] +
            [("accumulate_samples", ((1,), b8)), ("jitter_primary_rays", ((1,), b8)), ("num_bounces", ((1,), i32))]),
        "get_framebuffer": dict(
            [("output_" + k, ((S, H, W, 3), f32)) for k in ("rgb", "normal", "f0", "ray_origin", "ray_direction")] +
            [("output_" + k, ((S, H, W, 1), f32)) for k in ("depth", "roughness", "transmittance", "total_transmittance")] +
            [("output_final", ((1, H, W, 3), f32)), ("output_denoised", ((1, H, W, 3), f32))] +
            [("accumulated_" + k, ((S, H, W, 3), f32)) for k in ("rgb", "normal", "f0")] +
            [("accumulated_" + k, ((S, H, W, 1), f32)) for k in ("transmittance", "total_transmittance", "depth", "roughness")] +
            [("accumulated_sample_count", ((1,), i32))] +
            [("target_" + k, ((H, W, 3), f32)) for k in ("diffuse", "specular", "normal", "f0")] +
            [("target_" + k, ((H, W, 1), f32)) for k in ("depth", "roughness")]),
        "get_gaussians": dict(
            [(k, ((N, 3), f32)) for k in ("rgb", "normal", "f0", "scale", "mean")] +
            [(k, ((N, 1), f32)) for k in ("roughness", "opacity", "total_weight")] + [("rotation", ((N, 4), f32))] +
            [("dL_d" + k, ((N, 3), f32)) for k in ("rgb", "normal", "f0", "scale", "mean")] +
            [("dL_d" + k, ((N, 1), f32)) for k in ("roughness", "opacity")] + [("dL_drotation", ((N, 4), f32))]),
        "get_metadata": {"grads_enabled": ((1,), b8), "total_num_calls": ((1,), i32), "random_seeds": ((H, W, 1), i32)},
        "get_stats": {"num_accumulated_per_pixel": ((H, W), i32), "num_traversed_per_pixel": ((H, W), i32)},
    }
    for getter, fields in want.items():
        holder = getattr(rt, getter)()
        for name, (shape, dtype) in fields.items():
            t = getattr(holder, name)
            assert isinstance(t, torch.Tensor) and tuple(t.shape) == shape and t.dtype == dtype, (getter, name, t.shape, t.dtype)
    for name in set().union(*attrs.values()):  # every reached attribute is covered by the table above
        assert any(name in f for f in want.values()) or name in ("set_pose",), name
    # constants and introspection of raytracer.cpp:145-204
    assert Raytracer.MAX_ALPHA() == pytest.approx(0.9999) and Raytracer.ROUGHNESS_DOWNWEIGHT_GRAD() is True
    assert len(rt.describe_output_buffers()) == 11 and len(rt.describe_gaussian_attributes()) == 8


@pytest.fixture()
def reference_renderer():
    """The reference's renderer modules, unmodified, with the single seam re-pointed at this package."""
    saved = {k: v for k, v in sys.modules.items() if k == "editable_gauss_refl" or k.startswith("editable_gauss_refl.")}
    for k in saved:
        del sys.modules[k]
    sys.path.insert(0, REF)
    made = []
    try:
        pkg = importlib.import_module("editable_gauss_refl")  # defines make_raytracer only (torch.classes loader)

        def make_raytracer(W, H, N, fwd=180_000_000, bwd=120_000_000):
            made.append(SurfaceOnly(W, H, N, fwd, bwd))
            return made[-1]

        pkg.make_raytracer = make_raytracer  # the one line a maintainer re-points
        # the model / camera modules are only named in type annotations here; the real ones need plyfile, kornia,
        # simple_knn, cv2 ... (absent from this image)
        scene_pkg = types.ModuleType("editable_gauss_refl.scene")
        scene_pkg.__path__ = []
        gm = types.ModuleType("editable_gauss_refl.scene.gaussian_model")
        gm.GaussianModel = type("GaussianModel", (), {})
        cm = types.ModuleType("editable_gauss_refl.scene.cameras")
        cm.Camera = type("Camera", (), {})
        sys.modules.update({"editable_gauss_refl.scene": scene_pkg, "editable_gauss_refl.scene.gaussian_model": gm,
                            "editable_gauss_refl.scene.cameras": cm})
        rnd = importlib.import_module("editable_gauss_refl.renderer")
        assert rnd.__file__.startswith(REF)
        yield rnd, made
    finally:
        sys.path.remove(REF)
        for k in [k for k in sys.modules if k == "editable_gauss_refl" or k.startswith("editable_gauss_refl.")]:
            del sys.modules[k]
        sys.modules.update(saved)


def test_unmodified_reference_facade_drives_the_binding(reference_renderer):
    rnd, made = reference_renderer
    W, H, N = 16, 12, 40
    sc = scenes.random_scene(N, seed=5)
    pc = scenes.SyntheticModel(sc, device="cpu")
    cam = scenes.random_camera(W, H)
    tg = scenes.random_targets(W, H, seed=2)
    view = scenes.SyntheticCamera(cam, tg, device="cpu")
    rtr = rnd.GaussianRaytracer(pc, W, H)  # gaussian_raytracer.py:11-30
    rt = made[0]
    assert rtr.cuda_module is rt and rt.calls == [("rebuild_bvh", N)]
    cfg, g = rt.get_config(), rt.get_gaussians()
    assert float(cfg.loss_weight_diffuse) == 5.0 and float(cfg.loss_weight_specular) == 3.0
    assert float(cfg.alpha_threshold) == pytest.approx(0.005) and float(cfg.exp_power) == 3.0
    for k, attr in (("mean", "_xyz"), ("scale", "_scaling"), ("rotation", "_rotation"), ("opacity", "_opacity"),
                    ("rgb", "_diffuse"), ("normal", "_normal"), ("roughness", "_roughness"), ("f0", "_f0")):
        assert torch.equal(getattr(g, k), getattr(pc, attr)), k  # _export_param_values, :41-51

    # __call__ / render(): the reference moves host values to "cuda"; on this CPU tier that is the identity
    real_tensor = torch.tensor

    def tensor_no_device(*a, **kw):
        kw.pop("device", None)
        return real_tensor(*a, **kw)

    rt.calls.clear()
    with mock.patch.object(torch.Tensor, "cuda", lambda self, *a, **k: self), mock.patch.object(torch, "tensor", tensor_no_device):
        g.grad_records.fill_(0.25)  # what a backward would have left in the holders
        with torch.enable_grad():
            pkg = rnd.render(view, rtr, denoise=False)  # gaussian_renderer.py:21-92 -> gaussian_raytracer.py:75-151
        assert rt.calls == [("update_bvh",), ("raytrace", True)]
        cam_h, fb = rt.get_camera(), rt.get_framebuffer()
        assert float(cam_h.znear) == pytest.approx(0.01) and float(cam_h.zfar) == pytest.approx(999.9)
        assert float(cam_h.vertical_fov_radians) == pytest.approx(cam["fovy"])
        assert np.allclose(cam_h.rotation_c2w.numpy(), cam["c2w"], atol=1e-6)  # -R with the first column flipped back
        assert torch.equal(cam_h.rotation_w2c, cam_h.rotation_c2w.t()) and torch.equal(cam_h.origin, view.camera_center)
        for k in ("diffuse", "specular", "depth", "normal", "f0", "roughness"):
            assert np.array_equal(getattr(fb, "target_" + k).numpy(), tg[k]), k  # CHW -> HWC copy
        # _import_param_gradients (:53-62): model .grad += holder .grad, which is a view of the flat record
        for attr in ("_xyz", "_opacity", "_scaling", "_rotation", "_diffuse", "_normal", "_roughness", "_f0"):
            assert float(getattr(pc, attr).grad.min()) == 0.25 and float(getattr(pc, attr).grad.max()) == 0.25, attr
        assert pkg.rgb.shape == (3, 3, H, W) and pkg.final.shape == (1, 3, H, W) and pkg.depth.shape == (3, 1, H, W)
        assert pkg.target_diffuse is view.diffuse_image
        rtr.zero_grad()  # :64-73
        assert float(g.grad_records[:, [c for c in range(24) if c not in (15, 19, 23)]].abs().max()) == 0.0
        assert float(g.total_weight.min()) == 0.25  # total_weight is not a gradient: train.py:243 zeroes it itself
        rt.calls.clear()
        with torch.no_grad():
            rnd.render(view, rtr, targets_available=False, denoise=True)
        assert rt.calls == [("raytrace", False), ("denoise",)]  # no refit outside grad mode (:139-140)
        assert float(fb.target_diffuse.abs().max()) == 0.0  # absent targets are zeroed
        rt.calls.clear()
        with torch.no_grad():
            rnd.render(view, rtr, targets_available=False, force_update_bvh=True)
        assert rt.calls == [("update_bvh",), ("raytrace", False)]
    # GaussianRaytracer.rebuild_bvh (:32-38) after the model changed size: resize -> export -> rebuild
    keep = torch.arange(N) % 3 != 0
    for attr in ("_xyz", "_opacity", "_scaling", "_rotation", "_diffuse", "_normal", "_roughness", "_f0"):
        setattr(pc, attr, getattr(pc, attr)[keep].contiguous())
    rt.calls.clear()
    rtr.rebuild_bvh()
    n2 = int(keep.sum())
    assert rt.calls == [("resize", n2), ("rebuild_bvh", n2)]
    assert g.mean.shape == (n2, 3) and torch.equal(g.mean, pc._xyz) and g.mean.grad.shape == (n2, 3)
    assert g.mean.grad.data_ptr() == g.dL_dmean.data_ptr()
    # train.py:202-209, 240-243: the scalars and total_weight the training loop touches directly
    assert isinstance(cfg.num_bounces.item(), int)
    cfg.num_bounces.fill_(0)
    cfg.num_bounces.copy_(1)
    mask = (g.total_weight / 125 < 0.1).squeeze(1)
    assert mask.shape == (n2,) and mask.dtype == torch.bool
    g.total_weight.zero_()
    # render.py:195-209 / tools/render_novel_views.py:59-72
    cfg.accumulate_samples.copy_(True)
    rt.reset_accumulators()
    assert rt.calls[-1] == ("reset_accumulators",)
