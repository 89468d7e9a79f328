"""Qt-free reader/writer of GAMER's RenderingParams files (`RenderParams.dat`).

QDataStream layout (source/galaxy/renderingparams.h:41-64): GamerCamera (gamercamera.h:31-39: camera,
target, up as QVector3D = 3 x f64 each, perspective f64), size i32, previewSize i32, exposure, gamma,
saturation, detailLevel, noiseDetail f64, noStars i32, starSize, starSizeSpread, starStrength, rayStep f64,
galaxyDirectory, sceneDirectory, currentGalaxy, sceneMode, imageDirectory QString, Spectra (spectrum.h:74-93:
u32 count + (QString name, QVector3D rgb) each), then nside i32 and renderType QString — which older files
(the shipped publish/data/RenderParams.dat) do not have: the reader keeps the defaults for them.
"""
from .gax import _Reader, _Writer
from .model import GamerCamera, RenderingParams, Spectra


def _i32(r):
    v = r.u32()
    return v - (1 << 32) if v & 0x80000000 else v


def from_bytes(data: bytes) -> RenderingParams:
    r = _Reader(data)
    rp = RenderingParams()
    rp.m_camera = GamerCamera(tuple(r.vec3()), tuple(r.vec3()), tuple(r.vec3()), r.f64())
    rp.m_size, rp.m_previewSize = _i32(r), _i32(r)
    rp.m_exposure, rp.m_gamma, rp.m_saturation, rp.m_detailLevel, rp.m_noiseDetail = (r.f64() for _ in range(5))
    rp.m_noStars = _i32(r)
    rp.m_starSize, rp.m_starSizeSpread, rp.m_starStrength, rp.m_rayStep = (r.f64() for _ in range(4))
    rp.m_galaxyDirectory, rp.m_sceneDirectory, rp.m_currentGalaxy, rp.m_sceneMode, rp.m_imageDirectory = (
        r.qstring() for _ in range(5))
    sp = Spectra()
    for _ in range(r.u32()):
        name = r.qstring()
        sp.m_spectra.append((name, tuple(r.vec3())))
    rp.m_spectra = sp
    if r.o < len(data):                     # newer files
        rp.m_nside = _i32(r)
        rp.m_renderType = r.qstring()
    if r.o != len(data):
        raise ValueError(f"RenderingParams: {len(data) - r.o} trailing bytes")
    return rp


def to_bytes(rp: RenderingParams) -> bytes:
    w = _Writer()
    c = rp.m_camera
    w.vec3(c.m_camera); w.vec3(c.m_target); w.vec3(c.m_up); w.f64(float(c.m_perspective))
    w.u32(rp.m_size & 0xFFFFFFFF); w.u32(rp.m_previewSize & 0xFFFFFFFF)
    for v in (rp.m_exposure, rp.m_gamma, rp.m_saturation, rp.m_detailLevel, rp.m_noiseDetail):
        w.f64(float(v))
    w.u32(rp.m_noStars & 0xFFFFFFFF)
    for v in (rp.m_starSize, rp.m_starSizeSpread, rp.m_starStrength, rp.m_rayStep):
        w.f64(float(v))
    for s in (rp.m_galaxyDirectory, rp.m_sceneDirectory, rp.m_currentGalaxy, rp.m_sceneMode, rp.m_imageDirectory):
        w.qstring(s)
    w.u32(len(rp.m_spectra.m_spectra))
    for name, rgb in rp.m_spectra.m_spectra:
        w.qstring(name)
        w.vec3(rgb)
    w.u32(rp.m_nside & 0xFFFFFFFF)
    w.qstring(rp.m_renderType)
    return b"".join(w.parts)


def Load(path: str) -> RenderingParams:
    with open(path, "rb") as f:
        return from_bytes(f.read())


def Save(rp: RenderingParams, path: str) -> None:
    with open(path, "wb") as f:
        f.write(to_bytes(rp))
