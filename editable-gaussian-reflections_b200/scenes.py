"""Synthetic, seeded scenes of the shapes named in BASELINE.json + duck-typed model / camera.

Everything here is host-side numpy (no dataset exists on the box).  ``random_scene`` is config C1
("random-50k"), ``kitchen_scene`` config C2 ("kitchen-1M": dense-init-shaped points on the surfaces of a closed
4x4x3 room with six objects, isotropic scale = lattice spacing, opacity 0.1, roughness 0.1, f0 0.04 -- the
values ``GaussianModel`` initialises with, reference ``scene/gaussian_model.py:197-212``, ``config.py:42-50``).
"""
from __future__ import annotations

import math
from types import SimpleNamespace

import numpy as np

PARAM_KEYS = ["mean", "scale", "rotation", "opacity", "rgb", "normal", "f0", "roughness"]


def look_at(eye, target, up=(0.0, 0.0, 1.0)):
    """c2w rotation (columns = camera x, y, z axes; the camera looks along -z: d = c2w @ (x, y, -1))."""
    eye, target, up = (np.asarray(v, np.float64) for v in (eye, target, up))
    f = target - eye
    f /= np.linalg.norm(f)
    x = np.cross(f, up)
    x /= np.linalg.norm(x)
    y = np.cross(x, f)
    return np.stack([x, y, -f], axis=1).astype(np.float32)


def c2w_to_reference_R(c2w):
    """Inverse of ``R_c2w_blender = -R; R_c2w_blender[:,0] *= -1`` (gaussian_raytracer.py:95-97)."""
    R = -np.asarray(c2w, np.float32).copy()
    R[:, 0] = -R[:, 0]
    return R


def random_scene(n=50_000, seed=0):
    """C1: uniformly random anisotropic Gaussians in [-1,1]^3 with coherent normals so that bounces fire."""
    g = np.random.default_rng(seed)
    mean = g.uniform(-1, 1, (n, 3)).astype(np.float32)
    scale = np.log(g.uniform(0.01, 0.05, (n, 3))).astype(np.float32)
    rotation = g.normal(size=(n, 4)).astype(np.float32)
    opacity = g.uniform(-2, 4, (n, 1)).astype(np.float32)
    rgb = g.uniform(0, 1, (n, 3)).astype(np.float32)
    nrm = mean / np.maximum(np.linalg.norm(mean, axis=1, keepdims=True), 1e-6)
    normal = (nrm * g.uniform(0.8, 1.0, (n, 1))).astype(np.float32)
    f0 = g.uniform(0.04, 1, (n, 3)).astype(np.float32)
    roughness = g.uniform(0, 0.5, (n, 1)).astype(np.float32)
    return dict(mean=mean, scale=scale, rotation=rotation, opacity=opacity, rgb=rgb, normal=normal, f0=f0,
                roughness=roughness)


def random_camera(width=256, height=256, distance=4.0, fovy=0.6911):
    """C1 camera: at (0,0,distance) looking at the origin."""
    c2w = look_at((0.0, 0.0, distance), (0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0))
    return dict(origin=np.array([0, 0, distance], np.float32), c2w=c2w, fovy=fovy, width=width, height=height)


def random_targets(width, height, seed=1):
    g = np.random.default_rng(seed)
    t = dict(diffuse=g.uniform(0, 1, (height, width, 3)), specular=g.uniform(0, 1, (height, width, 3)),
             depth=g.uniform(2, 6, (height, width, 1)), normal=g.uniform(-1, 1, (height, width, 3)),
             f0=g.uniform(0, 1, (height, width, 3)), roughness=g.uniform(0, 1, (height, width, 1)))
    return {k: v.astype(np.float32) for k, v in t.items()}


# ---- kitchen-shaped dense-init scene ---------------------------------------------------------------------------
def _rect(origin, eu, ev, nrm, h):
    """Lattice points of spacing h on the rectangle origin + a*eu + b*ev."""
    lu, lv = np.linalg.norm(eu), np.linalg.norm(ev)
    nu, nv = max(int(round(lu / h)), 1), max(int(round(lv / h)), 1)
    a = (np.arange(nu) + 0.5) / nu
    b = (np.arange(nv) + 0.5) / nv
    A, Bm = np.meshgrid(a, b, indexing="ij")
    pts = origin[None] + A.reshape(-1, 1) * eu[None] + Bm.reshape(-1, 1) * ev[None]
    return pts, np.broadcast_to(nrm, pts.shape)


def _box(center, size, h, inward=False):
    c, s = np.asarray(center, np.float64), np.asarray(size, np.float64) / 2
    pts, nrms = [], []
    for ax in range(3):
        u, v = (ax + 1) % 3, (ax + 2) % 3
        for sgn in (-1.0, 1.0):
            o = c.copy()
            o[ax] += sgn * s[ax]
            o[u] -= s[u]
            o[v] -= s[v]
            eu, ev = np.zeros(3), np.zeros(3)
            eu[u], ev[v] = 2 * s[u], 2 * s[v]
            n = np.zeros(3)
            n[ax] = -sgn if inward else sgn
            p, q = _rect(o, eu, ev, n, h)
            pts.append(p), nrms.append(q)
    return np.concatenate(pts), np.concatenate(nrms)


def _sphere(center, radius, h):
    n = max(int(round(4 * math.pi * radius * radius / (h * h))), 8)
    i = np.arange(n) + 0.5
    phi = np.arccos(1 - 2 * i / n)
    th = math.pi * (1 + 5 ** 0.5) * i
    d = np.stack([np.cos(th) * np.sin(phi), np.sin(th) * np.sin(phi), np.cos(phi)], 1)
    return np.asarray(center)[None] + radius * d, d


def kitchen_scene(n=1_000_000, seed=0):
    """C2: closed 4x4x3 room (z up, floor z=0) + 3 spheres + 3 boxes; exactly ``n`` Gaussians."""
    g = np.random.default_rng(seed)
    objs = [("sphere", (0.9, 0.8, 0.45), 0.45), ("sphere", (-1.0, -0.6, 0.35), 0.35), ("sphere", (0.2, -1.2, 0.6), 0.6),
            ("box", (-1.1, 1.0, 0.4), (0.9, 0.7, 0.8)), ("box", (1.2, -0.9, 0.5), (0.6, 0.6, 1.0)),
            ("box", (0.0, 0.3, 0.15), (1.2, 0.8, 0.3))]
    area = 2 * 16 + 4 * 12
    for kind, c, s in objs:
        area += 4 * math.pi * s * s if kind == "sphere" else 2 * (s[0] * s[1] + s[1] * s[2] + s[0] * s[2])
    h = math.sqrt(area / n) * 0.995  # slightly denser, trimmed to n below
    parts = [_box((0, 0, 1.5), (4, 4, 3), h, inward=True)]
    for kind, c, s in objs:
        parts.append(_sphere(c, s, h) if kind == "sphere" else _box(c, s, h))
    pts = np.concatenate([p for p, _ in parts])
    nrm = np.concatenate([q for _, q in parts])
    m = pts.shape[0]
    if m >= n:
        keep = np.sort(g.choice(m, n, replace=False))
        pts, nrm = pts[keep], nrm[keep]
    else:
        extra = g.choice(m, n - m, replace=True)
        pts = np.concatenate([pts, pts[extra] + g.normal(scale=h * 0.25, size=(n - m, 3))])
        nrm = np.concatenate([nrm, nrm[extra]])
    mean = pts.astype(np.float32)
    scale = np.full((n, 3), math.log(h), np.float32)  # sqrt(mean 3-NN dist^2) of a square lattice = h
    rotation = np.zeros((n, 4), np.float32)
    rotation[:, 0] = 1.0
    opacity = np.full((n, 1), math.log(0.1 / 0.9), np.float32)
    base = 0.5 + 0.5 * np.sin(mean * np.array([2.1, 1.7, 2.9], np.float32)[None])  # smooth albedo pattern
    rgb = (0.15 + 0.7 * base).astype(np.float32)
    normal = nrm.astype(np.float32)
    f0 = np.full((n, 3), 0.04, np.float32)
    roughness = np.full((n, 1), 0.1, np.float32)
    return dict(mean=mean, scale=scale, rotation=rotation, opacity=opacity, rgb=rgb, normal=normal, f0=f0,
                roughness=roughness, spacing=h)


def catacaustics_scene(n=3_000_000, seed=0):
    """C4: the kitchen geometry with real-scene-shaped Gaussians: mixed anisotropy (per-axis log-scale spread,
    random orientation), opacities spread over the logit range, varied roughness / f0."""
    sc = kitchen_scene(n, seed=seed)
    g = np.random.default_rng(seed + 17)
    sc["scale"] = (sc["scale"] + g.normal(0.0, 0.6, (n, 3))).clip(-9.0, math.log(5.0)).astype(np.float32)  # clamp_max 5
    q = g.normal(size=(n, 4))
    sc["rotation"] = (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(np.float32)
    sc["opacity"] = g.uniform(-3.0, 4.0, (n, 1)).astype(np.float32)
    sc["roughness"] = g.uniform(0.02, 0.6, (n, 1)).astype(np.float32)
    sc["f0"] = g.uniform(0.02, 0.5, (n, 3)).astype(np.float32)
    return sc


def kitchen_cameras(num=100, width=800, height=800, fovy=0.6911):
    """Inward-looking orbit inside the room (radius 1.6 m, height 1.5 m, looking slightly down at the centre)."""
    cams = []
    for i in range(num):
        a = 2 * math.pi * i / num
        eye = np.array([1.6 * math.cos(a), 1.6 * math.sin(a), 1.5], np.float32)
        tgt = np.array([-0.3 * math.cos(a), -0.3 * math.sin(a), 0.6], np.float32)
        cams.append(dict(origin=eye, c2w=look_at(eye, tgt), fovy=fovy, width=width, height=height))
    return cams


# ---- duck-typed model / camera used by the renderer facade --------------------------------------------------------
DEFAULT_CFG = dict(loss_weight_diffuse=5.0, loss_weight_specular=3.0, loss_weight_depth=2.5, loss_weight_normal=2.5,
                   loss_weight_f0=1.0, loss_weight_roughness=1.0, transmittance_threshold=0.01, alpha_threshold=0.005,
                   exp_power=3)  # reference config.py:37-58


class SyntheticModel:
    """The slice of ``GaussianModel`` (scene/gaussian_model.py:53-177) the renderer touches; raw parameters with
    ``.grad`` buffers, activations as in the reference (exp scaling, identity elsewhere: the tracer activates)."""

    def __init__(self, scene: dict, device="cuda", cfg: dict | None = None):
        import torch

        self.cfg = SimpleNamespace(**{**DEFAULT_CFG, **(cfg or {})})
        names = dict(mean="_xyz", scale="_scaling", rotation="_rotation", opacity="_opacity", rgb="_diffuse",
                     normal="_normal", roughness="_roughness", f0="_f0")
        for k, attr in names.items():
            t = torch.from_numpy(np.ascontiguousarray(scene[k])).to(device)
            t.grad = torch.zeros_like(t)
            setattr(self, attr, t)

    get_scaling = property(lambda s: __import__("torch").exp(s._scaling))
    _get_scaling = property(lambda s: s._scaling)
    _get_rotation = property(lambda s: s._rotation)
    get_xyz = property(lambda s: s._xyz)
    get_diffuse = property(lambda s: s._diffuse)
    get_normal = property(lambda s: s._normal)
    get_roughness = property(lambda s: s._roughness)
    get_f0 = property(lambda s: s._f0)


class SyntheticCamera:
    """The slice of ``Camera`` (scene/cameras.py:22-152) the renderer touches."""

    def __init__(self, cam: dict, targets: dict | None = None, device="cuda", pin=False):
        import torch

        self.R = c2w_to_reference_R(cam["c2w"])
        self.FoVy = float(cam["fovy"])
        self.camera_center = torch.from_numpy(np.asarray(cam["origin"], np.float32)).to(device)
        self.image_width, self.image_height = cam["width"], cam["height"]
        self.original_image = None
        for k, attr in (("diffuse", "diffuse_image"), ("specular", "specular_image"), ("depth", "depth_image"),
                        ("normal", "normal_image"), ("f0", "f0_image"), ("roughness", "roughness_image")):
            v = None
            if targets is not None:
                v = torch.from_numpy(np.ascontiguousarray(np.moveaxis(targets[k], -1, 0)))  # CHW like the datasets
                v = v.pin_memory() if pin else v.to(device)
            setattr(self, attr, v)
