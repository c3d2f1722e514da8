"""B200-native differentiable Gaussian ray tracer: drop-in for the hot path of ``editable_gauss_refl``.

``make_raytracer`` is the single seam of the reference (``editable_gauss_refl/__init__.py:15-27``); it returns
an object with the surface of ``torch.classes.raytracer.Raytracer``.  The directory name is not a Python
identifier; import it as ``egr_b200`` (the shim module at the repository root) or through ``importlib``.
"""
from . import _capi, build, scenes  # noqa: F401
from .raytracer import DEFAULT_MAX_BOUNCES, Raytracer  # noqa: F401
from .renderer import GaussianRaytracer, render  # noqa: F401
from . import distributed  # noqa: F401
from .knn import distCUDA2  # noqa: F401
from . import prune  # noqa: F401
from . import edit  # noqa: F401
from . import optim  # noqa: F401
from . import trainer  # noqa: F401


def make_raytracer(image_width: int, image_height: int, num_gaussians: int, ppll_forward_size: int = 180_000_000,
                   ppll_backward_size: int = 120_000_000, **kw):
    """Same signature and defaults as the reference's ``make_raytracer``."""
    return Raytracer(image_width, image_height, num_gaussians, ppll_forward_size, ppll_backward_size, **kw)


__all__ = ["make_raytracer", "Raytracer", "GaussianRaytracer", "render", "scenes", "distributed", "build", "_capi",
           "DEFAULT_MAX_BOUNCES", "distCUDA2", "prune", "edit", "optim", "trainer"]
