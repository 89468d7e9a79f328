"""Regenerates tests/golden/ from the reference checkout (run in the build container only).

  gax/*.gax        the nine galaxy fixtures of publish/data/galaxies/ (binary parameter files, 0.6-1.3 KB
                   each; data, not source) — the fixture set both reference harnesses sweep
                   (tools/galaxy_repro.py:1053-1066)
  galaxies.json    the same files decoded by gamer_b200.gax (human-readable)
  simplex_ref.npz  raw_noise_3d / octave_noise_3d values produced by the reference's OWN
                   source/noise/simplexnoise.cpp compiled in place (oracle/_ref, `make -C oracle ref`)
  noise_class_ref.npz  Noise::octave_noise_3d and Noise::getRidgedMf of the reference's OWN noise.cpp / simplex.cpp
                   (class Simplex(1,1,1,1), as Rasterizer constructs it), compiled in place against oracle/qtshim,
                   with the argument patterns of galaxycomponents.cpp (disk/dust/stars fBm, dust2 ridged)
  ref_star_layers.npz  star layers stamped by the reference's OWN Buffer2D::RenderGaussian (util/buffer2d.cpp:224-243) from
                   star lists of gamer_b200_star_list (`python make_fixtures.py stars` regenerates this file alone)
  ref_path_frames.npz  whole frames rendered by the reference's OWN Rasterizer (Prepare, RenderPixels, renderPixel,
                   getIntensity, the GalaxyComponent classes, GamerCamera, Buffer2D::ToColorBuffer ...) compiled in
                   place against oracle/qtshim (oracle/ref_path_wrap.cpp), for the scenes of tests/ref_cases.py:
                   "<case>|src" as the source reads, "<case>|o2" as g++ -O2 compiles it (see oracle/Makefile),
                   plus "<case>|tm<k>|image/shadow" ARGB32 images of AssembleImage
"""
import ctypes as C
import glob
import json
import os
import shutil
import sys
from dataclasses import asdict

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from gamer_b200.gax import Galaxy  # noqa: E402

REF = os.environ.get("GAMER_REFERENCE", "/root/reference")


def main():
    out = {}
    for f in sorted(glob.glob(os.path.join(REF, "publish/data/galaxies/*.gax"))):
        shutil.copy(f, os.path.join(HERE, "gax", os.path.basename(f)))
        out[os.path.basename(f)] = asdict(Galaxy.Load(f))
    with open(os.path.join(HERE, "galaxies.json"), "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)

    ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libgamer_ref_noise.so"))
    ref.ref_raw_noise_3d.restype = C.c_double
    ref.ref_raw_noise_3d.argtypes = [C.c_double] * 3
    ref.ref_octave_noise_3d.restype = C.c_double
    ref.ref_octave_noise_3d.argtypes = [C.c_double] * 6
    rng = np.random.default_rng(20261104)
    pts = np.concatenate([
        rng.uniform(-4, 4, (600, 3)), rng.uniform(-300, 300, (300, 3)), rng.uniform(-80000, 80000, (100, 3)),
        np.array([[0, 0, 0], [1, 2, 3], [-1, -2, -3], [0.5, 0.5, 0.5], [-2, 0, 0], [3, -3, 0], [1e-9, -1e-9, 0],
                  [255, 256, 257], [-255.5, 0.25, 1024.75]], dtype=np.float64)])
    pts = pts.astype(np.float32).astype(np.float64)          # float-representable, like every caller's input
    raw = np.array([ref.ref_raw_noise_3d(*p) for p in pts])
    cases = [(10, 1.0, 1.0), (9, -0.8, 0.5), (10, 1.0, 15.0), (4, -2.0, 0.2), (10, -0.5, 0.5)]
    octv = np.array([[ref.ref_octave_noise_3d(o, p, s, *q) for q in pts[:300]] for (o, p, s) in cases])
    np.savez_compressed(os.path.join(HERE, "simplex_ref.npz"), pts=pts, raw=raw, oct_cases=np.array(cases), oct=octv)
    # the class-level entry points the components call (galaxycomponent.h:93-98, galaxycomponents.cpp:45,68,92-94,141)
    cls = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libgamer_ref_noise_class.so"))
    p = C.c_void_p
    cls.ref_class_octave_noise_3d.argtypes = [p, C.c_int, C.c_double, C.c_double, C.c_double, p]
    cls.ref_class_ridged.argtypes = [p, C.c_int, C.c_double, C.c_int, C.c_double, C.c_double, C.c_double, p]
    cls.ref_class_raw_3d.argtypes = [p, C.c_int, p]
    pf = np.ascontiguousarray(np.concatenate([rng.uniform(-1.2, 1.2, (700, 3)), rng.uniform(-30, 30, (200, 3)),
                                              np.array([[0, 0, 0], [0.5, 0.25, -0.125], [-1, 0, 2], [3, 3, 3]])]), dtype=np.float32)
    n = len(pf)
    fbm_cases = [(10, 1.0, float(np.float32(10) * np.float32(0.1))), (10, 1.0, float(np.float32(1) * np.float32(0.1))),
                 (9, -0.8, float(np.float32(4) * np.float32(0.1))), (10, -0.5, float(np.float32(5) * np.float32(0.1))),
                 (10, 1.0, float(np.float32(0.01) * np.float32(6) * 100)), (4, -2.0, 2.0 * float(np.float32(0.1)))]
    fbm = np.empty((len(fbm_cases), n))
    for k, (o, pe, sc) in enumerate(fbm_cases):
        cls.ref_class_octave_noise_3d(pf.ctypes.data, n, o, pe, sc, fbm[k].ctypes.data)
    # dust2: getRidgedMf(r * scale, ks, 9, 2.5, noiseOffset, noiseTilt); the point already carries the scale
    rid_cases = [(1.0, 9, 2.5, 1.0, 1.5, 5.0), (1.0, 9, 2.5, 1.2, 1.5, 4.0), (1.0, 9, 2.5, 1.05, 1.5, 6.0), (-0.5, 9, 2.5, 1.1, 1.2, 8.0)]
    rid = np.empty((len(rid_cases), n))
    for k, (fr, o, lac, off, gain, scale) in enumerate(rid_cases):
        q = np.ascontiguousarray(pf * np.float32(scale))
        cls.ref_class_ridged(q.ctypes.data, n, fr, o, lac, off, gain, rid[k].ctypes.data)
    raw2 = np.empty(len(pts))
    pd = np.ascontiguousarray(pts)
    cls.ref_class_raw_3d(pd.ctypes.data, len(pts), raw2.ctypes.data)
    assert np.array_equal(raw2, raw)            # Simplex::raw_3d is raw_noise_3d
    np.savez_compressed(os.path.join(HERE, "noise_class_ref.npz"), pts=pf, fbm_cases=np.array(fbm_cases), fbm=fbm,
                        rid_cases=np.array(rid_cases), rid=rid)
    print("wrote", len(out), "galaxies,", len(pts), "noise points,", n, "class-level noise points")
    ref_path_frames()


def ref_path_frames():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ref_cases
    import ref_path_lib
    frames = {}
    for name, case in ref_cases.cases().items():
        rp, gis, preview = case
        for kind in ("src", "o2"):
            with ref_path_lib.RefRasterizer(rp, gis, kind) as r:
                frames[name + "|" + kind] = r.render(preview)
                if kind == "src" and name in ref_cases.TONEMAP_CASES:
                    for k, (e, g, s) in enumerate(ref_cases.TONEMAPS):
                        img, shadow = r.assemble(e, g, s)
                        frames["%s|tm%d|image" % (name, k)] = img
                        frames["%s|tm%d|shadow" % (name, k)] = shadow
        print(name, float(frames[name + "|src"].max()), float(np.abs(frames[name + "|src"] - frames[name + "|o2"]).max()))
    np.savez_compressed(os.path.join(HERE, "ref_path_frames.npz"), **frames)
    print("wrote", len(frames), "reference frames/images")


STAR_CASES = {   # name: (size, no_stars, base_size, size_spread, strength, seed)
    # stamp width = star size * image size / 245 (buffer2d.cpp:165): sizes chosen so stamps are 4..20 pixels wide,
    # overlap heavily (the per-channel max decides) and many hang over the image edges
    "dense96": (96, 400, 20.0, 8.0, 1.0, 7),
    "big_stamps64": (64, 60, 60.0, 20.0, 2.5, 11),
    "edges33": (33, 120, 40.0, 20.0, 0.7, 3),
}


def star_layers():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ref_path_lib
    from gamer_b200 import rasterizer
    out = {}
    for name, (size, n, base, spread, strength, seed) in STAR_CASES.items():
        lst = rasterizer.star_list(size, n, base, spread, strength, seed)
        out[name + "|xyw"] = np.array([[s.x, s.y, s.w] for s in lst], dtype=np.int32)
        out[name + "|cs"] = np.array([list(s.cs) for s in lst], dtype=np.float32)
        out[name + "|layer"] = ref_path_lib.stamp_stars(size, lst)
        print(name, size, n, float(out[name + "|layer"].max()), int((out[name + "|layer"] > 0).sum()))
    np.savez_compressed(os.path.join(HERE, "ref_star_layers.npz"), **out)


def negarm_frames():
    """tests/golden/ref_negarm_frames.npz: the negative-arm synthetic galaxies rendered by the reference's own code
    (a file of its own so that ref_path_frames.npz need not be rewritten)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ref_cases
    import ref_path_lib
    frames = {}
    for name, case in ref_cases.negarm_cases().items():
        rp, gis, preview = case
        for kind in ("src", "o2"):
            with ref_path_lib.RefRasterizer(rp, gis, kind) as r:
                frames[name + "|" + kind] = r.render(preview)
        print(name, float(frames[name + "|src"].max()), int(np.isnan(frames[name + "|src"]).sum()))
    np.savez_compressed(os.path.join(HERE, "ref_negarm_frames.npz"), **frames)


if __name__ == "__main__":
    if sys.argv[1:] == ["negarm"]:
        negarm_frames()
    elif sys.argv[1:] == ["stars"]:
        star_layers()
    else:
        main()
        star_layers()
        negarm_frames()
