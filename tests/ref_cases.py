"""The scenes rendered by the reference's own compiled code for tests/golden/ref_path_frames.npz
(tests/golden/make_fixtures.py) and re-rendered by the oracle / the CUDA path in the tests.
A case is (RenderingParams, [GalaxyInstance], preview); sizes are small because the reference build used for
the fixtures is single-threaded."""
import copy
import math

from gamer_b200.model import GalaxyInstance, build_scene

import scenes

SIZE = 40


def _one(name, view, size=SIZE, preview=False, position=(0, 0, 0), orientation=(0, 1, 0), iscale=1.0, galaxy=None):
    g = galaxy if galaxy is not None else scenes.load_galaxy(name)
    rp = scenes.rendering_params(size, view)
    return rp, [GalaxyInstance(g, name, position, orientation, iscale, 0.0)], preview


def _synthetic(kind):
    g = scenes.synthetic_galaxy(kind)
    # "stars small" draws from libc rand() per sample (galaxycomponents.cpp:159-170): not reproducible, dropped
    g.componentParams = [c for c in g.componentParams if c.className != "stars small"]
    return g


def _multi(size, n, view="outside"):
    rp = scenes.rendering_params(size, view)
    gis = []
    for i in range(n):
        a = 2.0 * math.pi * i / n
        pos = (1.1 * math.cos(a), 0.35 * math.sin(3 * a), 1.1 * math.sin(a))
        ori = (math.sin(a + 0.3), 0.4 + 0.6 * math.cos(2 * a) ** 2, math.cos(1.7 * a))
        gis.append(GalaxyInstance(scenes.load_galaxy(scenes.FIXTURES[i % len(scenes.FIXTURES)]), "g%d" % i, pos, ori,
                                  0.6 + 0.1 * (i % 5), 0.0))
    return rp, gis, False


def cases():
    out = {}
    views = ["cli", "oblique", "edge", "faceon", "outside"]
    for k, name in enumerate(scenes.FIXTURES):              # every shipped galaxy, two views each
        for v in (views[k % 5], views[(k + 2) % 5]):
            out["%s/%s" % (name, v)] = _one(name, v)
    out["Spiral/cli/preview"] = _one("Spiral", "cli", preview=True)          # RenderDirect: minRayStep 0.01
    out["Spiral/tilted"] = _one("Spiral", "outside", position=(0.2, -0.1, 0.3), orientation=(0.3, 0.8, -0.5), iscale=0.7)
    out["Sombrero/flipped"] = _one("Sombrero", "oblique", orientation=(0.0, -1.0, 0.0))   # rotationTo's 180 degree branch
    out["allclasses/edge"] = _one("allclasses", "edge", galaxy=_synthetic("allclasses"))
    out["allclasses/oblique"] = _one("allclasses", "oblique", galaxy=_synthetic("allclasses"))
    out["ellipsoid/outside"] = _one("ellipsoid", "outside", galaxy=_synthetic("ellipsoid"))
    rp = scenes.rendering_params(48, "outside")
    out["three/outside"] = (rp, [
        GalaxyInstance(scenes.load_galaxy("Spiral"), "a", (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 1.0, 0.0),
        GalaxyInstance(scenes.load_galaxy("Sombrero"), "b", (-0.9, 0.3, 0.6), (0.3, 0.8, -0.5), 0.7, 0.0),
        GalaxyInstance(scenes.load_galaxy("RedBar"), "c", (0.7, -0.2, -0.5), (-0.6, 0.5, 0.62), 1.3, 0.0)], False)
    out["eleven/outside"] = _multi(32, 11)                   # more galaxies than one kernel pass holds
    return out


def negarm_cases():
    """Negative arm exponents (tests/golden/ref_negarm_frames.npz): three arms (fmax drops the NaN arm) and one (the NaN
    reaches the contribution test), each from two views."""
    out = {}
    for kind in ("negarm3", "negarm1"):
        for v in ("oblique", "faceon"):
            out["%s/%s" % (kind, v)] = _one(kind, v, galaxy=_synthetic(kind))
    return out


# AssembleImage settings checked on top of some of the frames: (exposure, gamma, saturation)
TONEMAPS = [(1.0, 1.0, 1.0), (0.5, 0.8, 1.3), (2.5, 1.6, 0.4), (0.1, 0.45, 2.0)]
TONEMAP_CASES = ["Spiral/cli", "three/outside", "allclasses/edge"]


def scene_of(case, no_radial_gate=False):
    rp, gis, preview = case
    return build_scene(rp, [copy.copy(g) for g in gis], preview=preview, no_radial_gate=no_radial_gate)
