"""Shared scene builders for tests / smoke / bench: the reference's fixtures and canonical views."""
import math
import os

import numpy as np

from gamer_b200.gax import ComponentParams, Galaxy, GalaxyParams
from gamer_b200.model import GalaxyInstance, GamerCamera, RenderingParams, build_scene

GAX_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gax")
FIXTURES = ["Spiral", "Sombrero", "SB0", "RedBar", "TonsOfArms", "Irregular", "Irregular2", "VortexCloud", "WheelGalaxy"]

RAY_STEP = float(np.float32(0.025))      # CLI values pass through QString::toFloat (consolerenderer.cpp:55-61)

# name -> (camera, target, up, fov)
VIEWS = {
    "cli":     ((0.5, 0.0, 0.0), (0, 0, 0), (0, 1, 0), 90.0),          # singleGalaxy.sh:1 — edge-on, from inside
    "oblique": ((0.5196, 0.5196, 0.5196), (0, 0, 0), (0, 1, 0), 60.0),  # SURVEY §8d C2b — exterior
    "faceon":  ((0.0, 0.9, 0.001), (0, 0, 0), (0, 0, 1), 60.0),
    "edge":    ((0.9, 0.02, 0.0), (0, 0, 0), (0, 1, 0), 60.0),
    "outside": ((1.6, 0.7, -1.1), (0, 0, 0), (0, 1, 0), 50.0),
}


def load_galaxy(name: str) -> Galaxy:
    return Galaxy.Load(os.path.join(GAX_DIR, name + ".gax"))


def rendering_params(size, view="cli", ray_step=RAY_STEP) -> RenderingParams:
    cam, target, up, fov = VIEWS[view] if isinstance(view, str) else view
    rp = RenderingParams(m_size=size, m_rayStep=ray_step, m_camera=GamerCamera(cam, target, up, fov))
    rp.spectra().PopulateSpectra()
    return rp


def single_galaxy_scene(name_or_galaxy, size, view="cli", ray_step=RAY_STEP, preview=False,
                        position=(0, 0, 0), orientation=(0, 1, 0), iscale=1.0):
    g = load_galaxy(name_or_galaxy) if isinstance(name_or_galaxy, str) else name_or_galaxy
    rp = rendering_params(size, view, ray_step)
    gi = GalaxyInstance(g, g.galaxyParams.name, position, orientation, iscale, 0.0)
    return build_scene(rp, [gi], preview=preview)


def multi_galaxy_scene(size, view="outside", ray_step=RAY_STEP):
    """Three galaxies at different positions/orientations: exercises sorting, rotationTo, axis-angle twirl."""
    rp = rendering_params(size, view, ray_step)
    gis = [
        GalaxyInstance(load_galaxy("Spiral"), "a", (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 1.0, 0.0),
        GalaxyInstance(load_galaxy("Sombrero"), "b", (-0.9, 0.3, 0.6), (0.3, 0.8, -0.5), 0.7, 0.0),
        GalaxyInstance(load_galaxy("RedBar"), "c", (0.7, -0.2, -0.5), (-0.6, 0.5, 0.62), 1.3, 0.0),
    ]
    return build_scene(rp, gis)


def synthetic_galaxy(kind="allclasses") -> Galaxy:
    """Classes / branches no shipped fixture covers: dust positive, stars with
    noiseOffset != 0, stars small (no-op), ellipsoidal axis, negative persistence, 1 and 3 arms."""
    g = Galaxy()
    p = g.galaxyParams
    p.name = kind
    p.windingB, p.windingN, p.noArms = 0.8, 3.0, 3.0
    p.arm1, p.arm2, p.arm3, p.arm4 = 0.0, 2.1, 4.2, 5.0
    if kind == "ellipsoid":
        p.axis = [1.0, 0.6, 0.8]
    C = ComponentParams
    g.componentParams = [
        C("bulge", 40.0, "Yellow", 1.0, 0.02, 6.0, 1.0, 0.0, 0.1, 1.0, 0.0, 1.0, 1.0, 0.1, "b"),
        C("disk", 2500.0, "Blue", 0.4, 0.03, 0.45, 1.0, 0.3, 0.15, 8.0, -0.05, 0.7, -0.5, 0.1, "d"),
        C("dust positive", 120.0, "Purple", 0.3, 0.025, 0.4, 1.0, 0.0, 0.1, 6.0, 1.1, 1.4, 1.0, 0.15, "dp"),
        C("dust", 200.0, "Cyan", 0.2, 0.02, 0.45, 1.0, 1.0, 0.2, 4.0, 0.1, 1.3, -0.8, 0.0, "du"),
        C("stars", 8.0, "Red", 0.2, 0.04, 0.35, 1.0, 0.0, 0.3, 40.0, 0.35, 9.0, 1.0, 0.0, "s"),
        C("stars small", 5.0, "White", 0.2, 0.04, 0.35, 1.0, 0.0, 0.3, 40.0, 0.0, 2.0, 1.0, 0.0, "ss"),
        C("dust2", 90.0, "Blue", 0.0, 0.02, 0.4, 1.0, 0.0, 0.1, 5.0, 1.05, 1.5, 1.0, 0.3, "d2 no arm"),
        C("disk", 300.0, "White", 0.5, 0.02, 0.3, 0.0, 0.0, 0.1, 3.0, 0.0, 1.0, 1.0, 0.0, "inactive"),
    ]
    return g
