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
                        position=(0, 0, 0), orientation=(0, 1, 0), iscale=1.0, no_radial_gate=False):
    g = load_galaxy(name_or_galaxy) if isinstance(name_or_galaxy, str) else name_or_galaxy
    rp = rendering_params(size, view, ray_step)
    gi = GalaxyInstance(g, g.galaxyParams.name, position, orientation, iscale, 0.0)
    return build_scene(rp, [gi], preview=preview, no_radial_gate=no_radial_gate)


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
    if kind in ("negarm3", "negarm1"):
        # a NEGATIVE arm exponent with a strong winding and a large delta: the wrapped phase difference passes
        # 3.14159265, where getArm's base goes negative and pow gives NaN (galaxycomponent.h:140-146): with three arms
        # fmax drops the NaN arm, with one arm the NaN fails the `val * weight > 0.0005` test
        p.windingB, p.windingN, p.noArms = 0.45, 5.0, (3.0 if kind == "negarm3" else 1.0)
        g.componentParams = [
            C("bulge", 30.0, "Yellow", 1.0, 0.02, 6.0, 1.0, 0.0, 0.1, 1.0, 0.0, 1.0, 1.0, 0.1, "b"),
            C("disk", 40.0, "Blue", -0.05, 0.03, 0.45, 1.0, 2.5, 0.15, 8.0, 0.0, 0.7, 1.0, 0.1, "d neg arm"),
            C("dust2", 60.0, "Red", -0.02, 0.025, 0.4, 1.0, 1.2, 0.1, 5.0, 1.05, 1.5, 1.0, 0.2, "d2 neg arm"),
            C("stars", 4.0, "White", 0.3, 0.04, 0.35, 1.0, 0.0, 0.3, 40.0, 0.0, 9.0, 1.0, 0.0, "s"),
        ]
        return g
    if kind == "sharing":
        # twirl sharing (DComp::tgroup): three components with winding 0.12 and two without arms (winding factor 0),
        # interleaved with a fourth factor so that the kept point is replaced between members of a group
        g.componentParams = [
            C("bulge", 30.0, "Yellow", 1.0, 0.02, 6.0, 1.0, 0.0, 0.1, 1.0, 0.0, 1.0, 1.0, 0.1, "b"),
            C("disk", 2500.0, "Blue", 0.4, 0.03, 0.45, 1.0, 0.3, 0.12, 8.0, -0.05, 0.7, 1.0, 0.1, "d1"),
            C("dust", 200.0, "Cyan", 0.2, 0.02, 0.45, 1.0, 1.0, 0.3, 4.0, 0.1, 1.3, 1.0, 0.0, "du"),
            C("dust2", 90.0, "Blue", 0.3, 0.025, 0.4, 1.0, 0.0, 0.12, 5.0, 1.05, 1.5, 1.0, 0.2, "d2"),
            C("dust positive", 120.0, "Purple", 0.0, 0.025, 0.4, 1.0, 0.0, 0.12, 6.0, 1.1, 1.4, 1.0, 0.15, "dp no arm"),
            C("disk", 900.0, "White", 0.1, 0.02, 0.35, 1.0, 0.0, 0.12, 2.0, 0.0, 1.0, 1.0, 0.0, "d3"),
            C("dust2", 60.0, "Red", 0.0, 0.02, 0.4, 1.0, 0.0, 0.3, 3.0, 1.0, 1.2, 1.0, 0.3, "d2 no arm"),
        ]
        return g
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


# ---- BASELINE configs 4 and 5 as concrete inputs -----------------------------------------

def orbit_camera(k, n_frames=360):
    """C4: cam_k = 0.9 (cos t cos 0.35, sin 0.35, sin t cos 0.35), t = 2 pi k / n, target origin, fov 60."""
    t = 2.0 * math.pi * k / n_frames
    return ((0.9 * math.cos(t) * math.cos(0.35), 0.9 * math.sin(0.35), 0.9 * math.sin(t) * math.cos(0.35)),
            (0, 0, 0), (0, 1, 0), 60.0)


def hd_frame_scene(k, width=1920, height=1080, galaxy="Spiral", n_frames=360):
    """C4: the reference only renders squares, so 1920x1080 is the centred crop (rows 420..1499) of the
    full_size = 1920 square image: every pixel is the reference's pixel at size 1920."""
    full = single_galaxy_scene(galaxy, width, orbit_camera(k, n_frames))
    return full.tile(0, (width - height) // 2, width, height)


def random_galaxy(rng) -> Galaxy:
    """C5: one galaxy of the seeded sweep (ranges from SURVEY.md §8d, row C5)."""
    def logu(a, b):
        return float(math.exp(rng.uniform(math.log(a), math.log(b))))
    g = Galaxy()
    p = g.galaxyParams
    p.name = "random"
    p.windingB = float(rng.uniform(0.4, 3.0))
    p.windingN = float(rng.choice([2, 4, 5]))
    p.noArms = float(rng.choice([1, 2, 3, 4]))
    p.arm1, p.arm2, p.arm3, p.arm4 = 0.0, 3.14, 1.57, 4.71
    spectra = ["Red", "Yellow", "Blue", "White", "Cyan", "Purple"]
    n = int(rng.integers(3, 9))
    classes = ["bulge", "disk"] + [str(rng.choice(["bulge", "disk", "dust", "dust2", "stars"])) for _ in range(n - 2)]
    comps = []
    for cls in classes:
        c = ComponentParams(className=cls, name=cls, spectrum=str(rng.choice(spectra)))
        c.strength = {"bulge": logu(5, 5000), "disk": logu(5, 5000), "dust": logu(10, 500), "dust2": logu(10, 500),
                      "stars": logu(0.5, 100)}[cls]
        c.arm = float(rng.uniform(0.01, 1.0))
        c.z0 = float(rng.uniform(0.01, 0.07))
        c.r0 = float(rng.uniform(3, 25)) if cls == "bulge" else float(rng.uniform(0.15, 0.6))
        c.inner = float(rng.uniform(0, 0.6))
        c.active = 1.0
        c.delta = float(rng.uniform(0, 2.5))
        c.winding = float(rng.uniform(0, 0.5))
        c.scale = float(rng.uniform(5, 150)) if cls == "stars" else float(rng.uniform(1, 20))
        c.noiseOffset = float(rng.uniform(1, 1.5)) if cls == "dust2" else 0.0
        c.noiseTilt = {"disk": float(rng.uniform(0.01, 1.0)), "dust2": float(rng.uniform(1, 1.5)),
                       "stars": float(rng.uniform(8, 20))}.get(cls, 1.0)
        c.ks = 1.0 if rng.uniform() < 0.5 else float(rng.choice([-0.5, -0.8]))
        comps.append(c)
    g.componentParams = comps
    return g


def random_galaxy_scene(index, size=1024, seed=20261104):
    """C5: galaxy `index` of the sweep and its camera (uniform on the |cam| = 0.9 sphere, fov 60).
    Every galaxy has its own generator stream, so any rank can build any galaxy."""
    rng = np.random.default_rng([seed, index])
    g = random_galaxy(rng)
    v = rng.normal(size=3)
    v = 0.9 * v / np.linalg.norm(v)
    up = (0, 1, 0) if abs(v[1]) < 0.85 else (1, 0, 0)
    return single_galaxy_scene(g, size, ((float(v[0]), float(v[1]), float(v[2])), (0, 0, 0), up, 60.0))


def many_galaxy_scene(size, n=11, view="outside", ray_step=RAY_STEP):
    """More galaxies than one kernel pass holds (8): a ring of fixtures with varied orientations, so the
    host has to split the sorted list into far -> near passes."""
    rp = rendering_params(size, view, ray_step)
    gis = []
    for i in range(n):
        a = 2.0 * math.pi * i / n
        pos = (1.1 * math.cos(a), 0.35 * math.sin(3 * a), 1.1 * math.sin(a))
        ori = (math.sin(a + 0.3), 0.4 + 0.6 * math.cos(2 * a) ** 2, math.cos(1.7 * a))
        gis.append(GalaxyInstance(load_galaxy(FIXTURES[i % len(FIXTURES)]), f"g{i}", pos, ori, 0.6 + 0.1 * (i % 5), 0.0))
    return build_scene(rp, gis)
