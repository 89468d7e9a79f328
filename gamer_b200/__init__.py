"""gamer_b200 — B200-native implementation of GAMER's per-pixel galaxy volume integrator.

Only the hot path (ray march, component evaluation, simplex-noise octaves, dust absorption / RGB
emission, post-process) lives here, behind the C ABI of include/gamer_b200.h.  The Python layer
mirrors the reference's caller-facing classes (Rasterizer, RenderingParams, GamerCamera,
GalaxyInstance, Galaxy/.gax) and is plumbing, not the product.
"""
from .gax import ComponentParams, Galaxy, GalaxyParams           # noqa: F401
from .model import GalaxyInstance, GamerCamera, RenderingParams, Scene, Spectra, build_scene  # noqa: F401


def __getattr__(name):          # Rasterizer/Context need the built library; import lazily
    if name in ("Rasterizer", "Context"):
        from . import rasterizer
        return getattr(rasterizer, name)
    raise AttributeError(name)
