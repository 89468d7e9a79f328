"""Regenerates tests/golden/ from the reference checkout (run in the build container only).

  gax/*.gax        the nine galaxy fixtures of publish/data/galaxies/ (binary parameter files, 0.6-1.3 KB
                   each; data, not source) — the fixture set both reference harnesses sweep
                   (tools/galaxy_repro.py:1053-1066)
  galaxies.json    the same files decoded by gamer_b200.gax (human-readable)
  simplex_ref.npz  raw_noise_3d / octave_noise_3d values produced by the reference's OWN
                   source/noise/simplexnoise.cpp compiled in place (oracle/_ref, `make -C oracle ref`)
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
    print("wrote", len(out), "galaxies,", len(pts), "noise points")


if __name__ == "__main__":
    main()
