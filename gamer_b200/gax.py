"""Qt-free reader/writer of GAMER's `.gax` galaxy files and the parameter model they hold.

Mirrors, by name, the reference's parameter classes:
  ComponentParams  source/galaxy/componentparams.h:7-44
  GalaxyParams     source/galaxy/galaxyparams.h:10-43
  Galaxy           source/galaxy/galaxy.h:39-68 (stream operators), galaxy.cpp:173-206 (Save/Load)
A `.gax` file is a raw big-endian QDataStream dump: QString = u32 byte length (0xFFFFFFFF = null)
+ UTF-16BE; double = IEEE f64; QVector3D = three floats written as f64 (QDataStream's default
DoublePrecision); QVector<T> = u32 count + items.
"""
import struct
from dataclasses import dataclass, field
from typing import List


class _Reader:
    def __init__(self, data: bytes):
        self.d, self.o = data, 0

    def u32(self):
        v = struct.unpack_from(">I", self.d, self.o)[0]
        self.o += 4
        return v

    def f64(self):
        v = struct.unpack_from(">d", self.d, self.o)[0]
        self.o += 8
        return v

    def qstring(self):
        n = self.u32()
        if n == 0xFFFFFFFF:
            return ""
        s = self.d[self.o:self.o + n].decode("utf-16-be")
        self.o += n
        return s

    def vec3(self):
        return [self.f64(), self.f64(), self.f64()]


class _Writer:
    def __init__(self):
        self.parts = []

    def u32(self, v):
        self.parts.append(struct.pack(">I", v))

    def f64(self, v):
        self.parts.append(struct.pack(">d", v))

    def qstring(self, s):
        b = s.encode("utf-16-be")
        self.u32(len(b))
        self.parts.append(b)

    def vec3(self, v):
        for x in v:
            self.f64(float(x))


@dataclass
class ComponentParams:
    className: str = "bulge"
    strength: float = 1.0
    spectrum: str = "White"
    arm: float = 0.5
    z0: float = 0.02
    r0: float = 0.5
    active: float = 1.0
    delta: float = 0.0
    winding: float = 0.1
    scale: float = 1.0
    noiseOffset: float = 0.0
    noiseTilt: float = 1.0
    ks: float = 1.0
    inner: float = 0.0
    name: str = "New component"

    _ORDER = ("arm", "z0", "r0", "active", "delta", "winding", "scale", "noiseOffset", "noiseTilt", "ks", "inner")


@dataclass
class GalaxyParams:
    name: str = "NewGalaxy"
    axis: List[float] = field(default_factory=lambda: [1.0, 1.0, 1.0])
    bulgeDust: float = 0.025
    bulgeAxis: List[float] = field(default_factory=lambda: [1.0, 1.0, 1.0])
    windingB: float = 0.5
    windingN: float = 4.0
    noArms: float = 2.0
    arm1: float = 0.0
    arm2: float = 3.141592653589793
    arm3: float = 2 * 3.141592653589793
    arm4: float = 3 * 3.141592653589793
    innerTwirl: float = 0.0
    warpAmplitude: float = 0.0
    warpScale: float = 0.0


@dataclass
class Galaxy:
    displayName: str = ""
    galaxyParams: GalaxyParams = field(default_factory=GalaxyParams)
    componentParams: List[ComponentParams] = field(default_factory=list)

    @staticmethod
    def Load(path: str) -> "Galaxy":
        with open(path, "rb") as f:
            return Galaxy.from_bytes(f.read())

    @staticmethod
    def from_bytes(data: bytes) -> "Galaxy":
        r = _Reader(data)
        g = Galaxy()
        g.displayName = r.qstring()
        p = g.galaxyParams
        p.name = r.qstring()
        p.axis = r.vec3()
        p.bulgeDust = r.f64()
        p.bulgeAxis = r.vec3()
        for k in ("windingB", "windingN", "noArms", "arm1", "arm2", "arm3", "arm4", "innerTwirl", "warpAmplitude", "warpScale"):
            setattr(p, k, r.f64())
        n = r.u32()
        for _ in range(n):
            c = ComponentParams()
            c.className = r.qstring()
            c.strength = r.f64()
            c.spectrum = r.qstring()
            for k in ComponentParams._ORDER:
                setattr(c, k, r.f64())
            c.name = r.qstring()
            g.componentParams.append(c)
        if r.o != len(data):
            raise ValueError(f".gax: {len(data) - r.o} trailing bytes")
        return g

    def to_bytes(self) -> bytes:
        w = _Writer()
        w.qstring(self.displayName)
        p = self.galaxyParams
        w.qstring(p.name)
        w.vec3(p.axis)
        w.f64(p.bulgeDust)
        w.vec3(p.bulgeAxis)
        for k in ("windingB", "windingN", "noArms", "arm1", "arm2", "arm3", "arm4", "innerTwirl", "warpAmplitude", "warpScale"):
            w.f64(getattr(p, k))
        w.u32(len(self.componentParams))
        for c in self.componentParams:
            w.qstring(c.className)
            w.f64(c.strength)
            w.qstring(c.spectrum)
            for k in ComponentParams._ORDER:
                w.f64(getattr(c, k))
            w.qstring(c.name)
        return b"".join(w.parts)

    def Save(self, path: str) -> None:
        with open(path, "wb") as f:
            f.write(self.to_bytes())
