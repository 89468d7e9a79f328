"""ctypes door onto oracle/_ref/libgamer_ref_path*.so — the reference's OWN render path compiled in place
(oracle/Makefile `make ref`, wrapper oracle/ref_path_wrap.cpp). TEST INFRASTRUCTURE: used by
tests/golden/make_fixtures.py to produce the committed golden frames and by tests/test_ref_path.py where the
libraries exist (they are built only where /root/reference exists; the golden fixtures travel instead).

Two builds (see oracle/Makefile): "src" keeps the `intensity > 0.001` test of galaxycomponent.cpp:62 as
written, "o2" is the plain g++ -O2 build in which that test is compiled away (GAMER_FRAME_NO_RADIAL_GATE)."""
import ctypes as C
import os
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBS = {"src": os.path.join(ROOT, "oracle", "_ref", "libgamer_ref_path.so"),
        "o2": os.path.join(ROOT, "oracle", "_ref", "libgamer_ref_path_o2.so"),
        # the timing build: -Ofast -fopenmp -march=x86-64-v3 (bench.py --impl reference / cpu_baseline)
        "fast": os.path.join(ROOT, "oracle", "_ref", "libgamer_ref_path_fast.so")}
_loaded = {}


def available(kind="src"):
    if not os.path.exists(LIBS[kind]):
        return False
    if kind == "fast":          # built for x86-64-v3: needs AVX2/FMA/BMI2 on the machine it runs on
        try:
            flags = next(l for l in open("/proc/cpuinfo") if l.startswith("flags")).split()
        except (OSError, StopIteration):
            return False
        return all(f in flags for f in ("avx2", "fma", "bmi2", "movbe", "f16c"))
    return True


def _lib(kind):
    if kind not in _loaded:
        L = C.CDLL(LIBS[kind])
        p = C.c_void_p
        L.refpath_create.restype = p
        L.refpath_create.argtypes = [C.c_int, C.c_double, p, p, p, C.c_double]
        L.refpath_add_galaxy.argtypes = [p, C.c_char_p, p, p, C.c_double, C.c_double]
        L.refpath_render.argtypes = [p, C.c_int, p]
        L.refpath_render_rows.argtypes = [p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, p]
        L.refpath_pixel.argtypes = [p, C.c_int, p]
        L.refpath_ray.argtypes = [p, C.c_int, p]
        L.refpath_trace.argtypes = [p, C.c_int, C.c_int, C.c_int, p]
        L.refpath_assemble.argtypes = [p, C.c_double, C.c_double, C.c_double, p, p]
        L.refpath_destroy.argtypes = [p]
        L.refpath_stamp_stars.argtypes = [C.c_int, p, C.c_int, p]
        _loaded[kind] = L
    return _loaded[kind]


def stamp_stars(size, stars, kind="src"):
    """Buffer2D::RenderGaussian (util/buffer2d.cpp:224-243) of the reference itself for every star of a GamerStar
    ctypes array, on a zeroed size x size buffer -> float32 [size, size, 3]."""
    out = np.zeros((size, size, 3), np.float32)
    assert _lib(kind).refpath_stamp_stars(int(size), C.addressof(stars) if len(stars) else None, len(stars), out.ctypes.data) == 0
    return out


def _f3(v):
    return np.ascontiguousarray(v, dtype=np.float32)


class RefRasterizer:
    """The reference's Rasterizer fed from this repo's model objects (gamer_b200.model.RenderingParams,
    GalaxyInstance). Galaxies reach it as .gax files written by gamer_b200.gax (so the writer is exercised
    against the reference's own reader, galaxy.cpp:191-206)."""

    def __init__(self, rp, instances, kind="src"):
        self.L = _lib(kind)
        cam = rp.camera()
        c, t, u = _f3(cam.m_camera), _f3(cam.m_target), _f3(cam.m_up)
        self.size = int(rp.size())
        self.h = self.L.refpath_create(self.size, float(rp.rayStep()), c.ctypes.data, t.ctypes.data, u.ctypes.data,
                                       float(cam.m_perspective))
        self._tmp = tempfile.TemporaryDirectory()
        for k, gi in enumerate(instances):
            path = os.path.join(self._tmp.name, "g%d.gax" % k)
            gi.m_galaxy.Save(path)
            p, o = _f3(gi.m_position), _f3(gi.m_orientation)
            rc = self.L.refpath_add_galaxy(self.h, path.encode(), p.ctypes.data, o.ctypes.data,
                                           float(gi.m_intensityScale), float(gi.m_redshift))
            assert rc == 0, "reference could not load " + path

    def render(self, preview=False):
        """Rasterizer::Render (or RenderDirect when preview) -> linear float RGB [size, size, 3]."""
        out = np.zeros((self.size, self.size, 3), np.float32)
        assert self.L.refpath_render(self.h, int(bool(preview)), out.ctypes.data) == 0
        return out

    def render_rows(self, row0, stride, n_rows, threads=0, prepare=True):
        """Rows row0, row0+stride, ... through (Prepare +) setupCamera + renderPixel, OpenMP over pixels in the
        "fast" build. Returns (rgb [n_rows, size, 3], threads used)."""
        out = np.zeros((n_rows, self.size, 3), np.float32)
        used = self.L.refpath_render_rows(self.h, int(row0), int(stride), int(n_rows), int(threads), int(bool(prepare)),
                                          out.ctypes.data)
        return out, used

    def ray(self, idx):
        d = np.zeros(3, np.float32)
        self.L.refpath_ray(self.h, int(idx), d.ctypes.data)
        return d

    def assemble(self, exposure, gamma, saturation):
        """Rasterizer::AssembleImage -> (image, shadow image) as ARGB32 [size, size] uint32."""
        a = np.zeros((self.size, self.size), np.uint32)
        s = np.zeros((self.size, self.size), np.uint32)
        assert self.L.refpath_assemble(self.h, float(exposure), float(gamma), float(saturation), a.ctypes.data, s.ctypes.data) == 0
        return a, s

    def trace(self, idx, gal=0, max_samples=8192):
        out = np.zeros((max_samples, 7), np.float32)
        n = self.L.refpath_trace(self.h, int(idx), int(gal), max_samples, out.ctypes.data)
        return out[:min(n, max_samples)], n

    def close(self):
        if self.h:
            self.L.refpath_destroy(self.h)
            self.h = None
            self._tmp.cleanup()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
