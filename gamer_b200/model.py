"""Host-side mirror of the reference's caller-facing model for the render path.

Same names and meaning as the reference so that a caller (or a test) reads like the reference:
  Spectra / ComponentSpectrum   source/galaxy/spectrum.h:9-72
  GamerCamera                   source/galaxy/gamercamera.h:7-45, gamercamera.cpp:185-217
  RenderingParams               source/galaxy/renderingparams.h:19-64
  GalaxyInstance                source/galaxy/galaxyinstance.h:8-45
  Rasterizer                    source/galaxy/rasterizer.h:20-127 (the seam; see rasterizer.py)
`build_scene` is Rasterizer::Prepare (rasterizer.cpp:190-213) + Galaxy::SetupComponents/SetupSpectra
(galaxy.cpp:75-109) flattened into the three POD arrays of include/gamer_b200.h.
"""
import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

from . import _lib
from .gax import Galaxy


def _f32(v):
    return float(np.float32(v))


class Spectra:
    """spectrum.h:40-72. PopulateSpectra() installs the six named RGB triples."""

    def __init__(self):
        self.m_spectra = []           # list of (name, (r,g,b))
        self.m_default = ("default", (1.0, 1.0, 1.0))

    def PopulateSpectra(self):
        self.m_spectra = [("Red", (1, 0.6, 0.4)), ("Yellow", (1, 0.9, 0.45)), ("Blue", (0.4, 0.6, 1.0)),
                          ("White", (1.0, 1.0, 1.0)), ("Cyan", (0.3, 0.7, 1.0)), ("Purple", (1.0, 0.3, 0.8))]

    def FindSpectrum(self, name: str):
        for n, rgb in self.m_spectra:
            if n.lower() == name.lower():
                return rgb
        return None


# Host math of the parameter model (camera matrix, galaxy rotation): the product library's by default. bench.py's
# reference arm installs the CPU oracle's equivalents instead, so that arm never loads libgamer_b200.so at all.
_host_math = None


def set_host_math(camera_inv_vp=None, rotation_to=None):
    """camera_inv_vp(cam, target, up, fov) -> 16 floats [row*4+col]; rotation_to(orient) -> (w, x, y, z).
    None restores the product library (gamer_b200_camera_inv_vp / gamer_b200_rotation_to_orientation)."""
    global _host_math
    _host_math = (camera_inv_vp, rotation_to) if camera_inv_vp and rotation_to else None


@dataclass
class GamerCamera:
    m_camera: Sequence[float] = (3.0, 3.0, 3.0)
    m_target: Sequence[float] = (0.0, 0.0, 0.0)
    m_up: Sequence[float] = (0.0, 1.0, 0.0)
    m_perspective: float = 70.0

    def setupViewmatrix(self) -> np.ndarray:
        """inverse(projection * view) as 16 float32 [row*4+col] (gamercamera.cpp:185-208)."""
        if _host_math:
            return np.array(_host_math[0](self.m_camera, self.m_target, self.m_up, float(self.m_perspective)), dtype=np.float32)
        lib = _lib.load()
        v3 = C.c_float * 3
        out = (C.c_float * 16)()
        rc = lib.gamer_b200_camera_inv_vp(v3(*self.m_camera), v3(*self.m_target), v3(*self.m_up),
                                          float(self.m_perspective), out)
        _lib.check(None, rc)
        return np.array(out[:], dtype=np.float32)


@dataclass
class RenderingParams:
    m_size: int = 128
    m_exposure: float = 1.0
    m_gamma: float = 1.0
    m_saturation: float = 1.0
    m_rayStep: float = 0.001
    m_noStars: int = 0
    m_starSize: float = 1.0            # renderingparams.h:29-30
    m_starSizeSpread: float = 1.0
    m_starStrength: float = 1.0
    m_previewSize: int = 64            # the rest of renderingparams.h:19-39 (file round trips; not used by the path)
    m_detailLevel: float = 0.01
    m_noiseDetail: float = 1.0
    m_nside: int = 32
    m_renderType: str = ""
    m_galaxyDirectory: str = "galaxies/"
    m_sceneDirectory: str = "scenes/"
    m_imageDirectory: str = "images/"
    m_currentGalaxy: str = ""
    m_sceneMode: str = "galaxy"
    m_camera: GamerCamera = field(default_factory=GamerCamera)
    m_spectra: Spectra = field(default_factory=Spectra)

    def camera(self):
        return self.m_camera

    def spectra(self):
        return self.m_spectra

    def size(self):
        return self.m_size

    def rayStep(self):
        return self.m_rayStep


@dataclass
class GalaxyInstance:
    m_galaxy: Galaxy
    m_galaxyName: str = ""
    m_position: Sequence[float] = (0.0, 0.0, 0.0)
    m_orientation: Sequence[float] = (0.0, 1.0, 0.0)
    m_intensityScale: float = 1.0
    m_redshift: float = 0.0

    def rotmat(self):
        """setupQuaternions (galaxyinstance.cpp:69-71)."""
        if _host_math:
            return [float(v) for v in _host_math[1](self.m_orientation)]
        lib = _lib.load()
        q = (C.c_float * 4)()
        _lib.check(None, lib.gamer_b200_rotation_to_orientation((C.c_float * 3)(*self.m_orientation), q))
        return [q[0], q[1], q[2], q[3]]


@dataclass
class Scene:
    frame: _lib.GamerFrame
    galaxies: C.Array
    components: C.Array

    def tile(self, x0, y0, width, height, band_rows=0, band_stride=0) -> "Scene":
        f = _lib.GamerFrame.from_buffer_copy(self.frame)
        f.x0, f.y0, f.width, f.height, f.band_rows, f.band_stride = x0, y0, width, height, band_rows, band_stride
        return Scene(f, self.galaxies, self.components)


def build_scene(rp: RenderingParams, instances: List[GalaxyInstance], preview: bool = False,
                inv_vp: Optional[np.ndarray] = None, no_radial_gate: bool = False) -> Scene:
    """Rasterizer::Prepare: sort far->near by |position - camera| (stable, like std::sort on the
    reference's small lists is in practice), build components with resolved spectra, camera matrix."""
    cam = np.array(rp.camera().m_camera, dtype=np.float32)

    def dist(gi):
        d = np.array(gi.m_position, dtype=np.float32) - cam
        return float(np.float32(np.sqrt(np.float64(d[0]) ** 2 + np.float64(d[1]) ** 2 + np.float64(d[2]) ** 2)))

    ordered = sorted(instances, key=dist, reverse=True)
    comps, gals = [], []
    for gi in ordered:
        gp = gi.m_galaxy.galaxyParams
        g = _lib.GamerGalaxy()
        g.pos[:] = [_f32(v) for v in gi.m_position]
        g.orient[:] = [_f32(v) for v in gi.m_orientation]
        g.rotmat[:] = gi.rotmat()
        g.axis[:] = [_f32(v) for v in gp.axis]
        g.intensity_scale = gi.m_intensityScale
        g.winding_b, g.winding_n, g.no_arms = gp.windingB, gp.windingN, gp.noArms
        g.arm[:] = [gp.arm1, gp.arm2, gp.arm3, gp.arm4]
        g.comp_begin = len(comps)
        for cp in gi.m_galaxy.componentParams:
            cid = _lib.CLASS_IDS.get(cp.className)
            if cid is None:            # Galaxy::instantiate prints an error and skips (galaxy.cpp:38-41)
                continue
            c = _lib.GamerComponent()
            c.class_id = cid
            c.active = cp.active
            c.strength, c.arm, c.z0, c.r0, c.inner = cp.strength, cp.arm, cp.z0, cp.r0, cp.inner
            c.delta, c.winding, c.scale = cp.delta, cp.winding, cp.scale
            c.noise_offset, c.noise_tilt, c.ks = cp.noiseOffset, cp.noiseTilt, cp.ks
            rgb = rp.spectra().FindSpectrum(cp.spectrum) or rp.spectra().m_default[1]   # galaxy.cpp:75-85
            c.spec[:] = [_f32(v) for v in rgb]
            comps.append(c)
        g.comp_count = len(comps) - g.comp_begin
        gals.append(g)
    f = _lib.GamerFrame()
    n = int(rp.size())
    f.width = f.height = f.full_size = n
    f.x0 = f.y0 = f.band_rows = f.band_stride = 0
    m = rp.camera().setupViewmatrix() if inv_vp is None else np.asarray(inv_vp, dtype=np.float32)
    f.inv_vp[:] = [float(v) for v in m]
    f.cam[:] = [float(v) for v in cam]
    f.ray_step = float(rp.rayStep())
    f.min_ray_step = 0.01 if preview else 0.001     # rasterizer.cpp:437-441
    f.flags = _lib.FRAME_NO_RADIAL_GATE if no_radial_gate else 0   # gamer_b200.h: the test g++ -O1+ deletes
    f.n_galaxies, f.n_components = len(gals), len(comps)
    ga = (_lib.GamerGalaxy * max(1, len(gals)))(*gals)
    ca = (_lib.GamerComponent * max(1, len(comps)))(*comps)
    return Scene(f, ga, ca)


def split_passes(scene: Scene) -> List[Scene]:
    """Scenes beyond the per-pass limits of the kernel (8 galaxies / 64 active components) as a list of
    far -> near chunks with GAMER_FRAME_CONTINUE / GAMER_FRAME_PARTIAL set, to be rendered in order into the
    same context (Context.render_passes). One chunk (no flags) when the scene fits."""
    f = scene.frame
    chunks, cur, cur_active = [], [], 0
    for gi in range(f.n_galaxies):
        g = scene.galaxies[gi]
        active = sum(1 for k in range(g.comp_begin, g.comp_begin + g.comp_count) if scene.components[k].active == 1.0)
        if active > _lib.MAX_COMPONENTS:
            raise ValueError(f"galaxy {gi} has {active} active components (limit {_lib.MAX_COMPONENTS})")
        if cur and (len(cur) == _lib.MAX_GALAXIES or cur_active + active > _lib.MAX_COMPONENTS):
            chunks.append(cur)
            cur, cur_active = [], 0
        cur.append(gi)
        cur_active += active
    if cur or not chunks:
        chunks.append(cur)
    if len(chunks) == 1:
        return [scene]
    out = []
    for ci, idxs in enumerate(chunks):
        comps, gals = [], []
        for gi in idxs:
            g = _lib.GamerGalaxy.from_buffer_copy(scene.galaxies[gi])
            begin = len(comps)
            for k in range(g.comp_begin, g.comp_begin + g.comp_count):
                comps.append(_lib.GamerComponent.from_buffer_copy(scene.components[k]))
            g.comp_begin = begin
            gals.append(g)
        fr = _lib.GamerFrame.from_buffer_copy(f)
        fr.n_galaxies, fr.n_components = len(gals), len(comps)
        fr.flags = ((f.flags & _lib.FRAME_NO_RADIAL_GATE) | (_lib.FRAME_CONTINUE if ci > 0 else 0)
                    | (_lib.FRAME_PARTIAL if ci < len(chunks) - 1 else 0))
        out.append(Scene(fr, (_lib.GamerGalaxy * max(1, len(gals)))(*gals), (_lib.GamerComponent * max(1, len(comps)))(*comps)))
    return out
