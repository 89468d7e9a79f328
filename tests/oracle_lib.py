"""ctypes door onto the CPU oracle (oracle/gamer_oracle.cpp). TEST INFRASTRUCTURE: only tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs import this module."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ODIR = os.path.join(ROOT, "oracle")
STRICT = os.path.join(ODIR, "_build", "libgamer_oracle.so")
FAST = os.path.join(ODIR, "_build", "native", "libgamer_oracle_fast.so")
REF_NOISE = os.path.join(ODIR, "_ref", "libgamer_ref_noise.so")


class Oracle:
    def __init__(self, path):
        self.path = path
        lib = self.lib = C.CDLL(path)
        p = C.c_void_p
        lib.go_render.argtypes = [p, p, p, p, p, C.c_int]
        lib.go_tonemap.argtypes = [p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, p, p, p]
        lib.go_raw3d.restype = C.c_double
        lib.go_raw3d.argtypes = [C.c_double] * 3
        lib.go_simplex.argtypes = [p, C.c_int, C.c_double, p, p]
        lib.go_fbm.argtypes = [p, C.c_int, C.c_int, C.c_double, C.c_double, p]
        lib.go_ridged.argtypes = [p, C.c_int, C.c_double, C.c_int, C.c_double, C.c_double, C.c_double, p]
        lib.go_camera_inv_vp.argtypes = [p, p, p, C.c_double, p]
        lib.go_rotation_to_orientation.argtypes = [p, p]
        lib.go_coord2ray.argtypes = [p, C.c_int, C.c_int, p]
        lib.go_stamp_stars.argtypes = [C.c_int, p, C.c_int, p]

    def max_threads(self):
        return self.lib.go_max_threads()

    def render(self, scene, threads=0):
        """scene: gamer_b200.model.Scene. Returns (rgb[h,w,3] float32, counters dict)."""
        f = scene.frame
        out = np.empty((f.height, f.width, 3), dtype=np.float32)
        cnt = (C.c_uint64 * 8)()
        rc = self.lib.go_render(C.addressof(f), C.addressof(scene.galaxies), C.addressof(scene.components),
                                out.ctypes.data, C.addressof(cnt), threads)
        assert rc == 0
        return out, {"rays": cnt[0], "samples": cnt[1], "gates": cnt[2], "comp_evals": cnt[3], "octaves": cnt[4]}

    def tonemap(self, rgb, exposure=1.0, gamma=1.0, saturation=1.0, stars=None):
        rgb = np.ascontiguousarray(rgb, dtype=np.float32)
        h, w = rgb.shape[:2]
        a = np.empty((h, w), dtype=np.uint32)
        s = np.empty((h, w), dtype=np.uint32)
        sp = None
        if stars is not None:
            stars = np.ascontiguousarray(stars, dtype=np.float32)
            sp = stars.ctypes.data
        assert self.lib.go_tonemap(rgb.ctypes.data, w, h, exposure, gamma, saturation, sp, a.ctypes.data, s.ctypes.data) == 0
        return a, s

    def stamp_stars(self, size, stars):
        """stars: ctypes array of gamer_b200._lib.GamerStar. Returns the star layer [size,size,3] float32."""
        out = np.empty((size, size, 3), dtype=np.float32)
        assert self.lib.go_stamp_stars(size, C.addressof(stars) if len(stars) else None, len(stars), out.ctypes.data) == 0
        return out

    def simplex(self, xyz, frequency=1.0):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        n = xyz.shape[0]
        lat = np.empty((n, 10), dtype=np.int32)
        val = np.empty(n, dtype=np.float64)
        assert self.lib.go_simplex(xyz.ctypes.data, n, frequency, lat.ctypes.data, val.ctypes.data) == 0
        return lat, val

    def fbm(self, xyz, octaves, persistence, scale):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        val = np.empty(xyz.shape[0], dtype=np.float64)
        assert self.lib.go_fbm(xyz.ctypes.data, xyz.shape[0], octaves, persistence, scale, val.ctypes.data) == 0
        return val

    def ridged(self, xyz, frequency, octaves, lacunarity, offset, gain):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        val = np.empty(xyz.shape[0], dtype=np.float64)
        assert self.lib.go_ridged(xyz.ctypes.data, xyz.shape[0], frequency, octaves, lacunarity, offset, gain, val.ctypes.data) == 0
        return val

    def camera_inv_vp(self, cam, target, up, fov):
        v = lambda a: (C.c_float * 3)(*a)
        out = (C.c_float * 16)()
        self.lib.go_camera_inv_vp(v(cam), v(target), v(up), fov, out)
        return np.array(out[:], dtype=np.float32)

    def rotation_to(self, orient):
        out = (C.c_float * 4)()
        self.lib.go_rotation_to_orientation((C.c_float * 3)(*orient), out)
        return np.array(out[:], dtype=np.float32)


def build(target="all", force=False):
    subprocess.run(["make", "-C", ODIR] + (["-B"] if force else []) + [target], check=True, stdout=subprocess.DEVNULL)


_fast_built = False


def load(fast=False):
    """strict: the parity oracle (-O2, IEEE). fast: the timed CPU baseline with the reference's flags
    (-Ofast -march=native): always rebuilt once per process ON THE MACHINE THAT TIMES IT — a library that
    travelled from another box has the wrong -march=native and possibly an older source."""
    global _fast_built
    if fast:
        if not _fast_built:
            build("native", force=True)
            _fast_built = True
        return Oracle(FAST)
    if not os.path.exists(STRICT):
        build("all")
    try:
        return Oracle(STRICT)
    except AttributeError:          # stale build from an older source: rebuild once
        build("all", force=True)
        return Oracle(STRICT)
