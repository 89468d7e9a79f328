"""Prints parity diagnostics (no asserts) — run on the GPU box."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_lib, scenes
from gamer_b200.rasterizer import Context
import test_gpu_parity as T

o = oracle_lib.load(); ctx = Context([0])
rnd, adv = T.noise_points()
for freq in (1.0, 0.7, 150.0):
    lg, vg = ctx.probe_simplex(rnd, freq); lc, vc = o.simplex(rnd, freq)
    bad = np.where(~np.all(lg == lc, axis=1))[0]
    print("simplex freq", freq, "mismatch rows", len(bad), "max val err", np.abs(vg - vc).max())
    for b in bad[:5]:
        print("   pt", rnd[b], "gpu", lg[b], "cpu", lc[b], vg[b], vc[b])
lg, vg = ctx.probe_simplex(adv, 1.0); lc, vc = o.simplex(adv, 1.0)
bad = np.where(~np.all(lg[:, [0,1,2,5,8]] == lc[:, [0,1,2,5,8]], axis=1))[0]
print("adv cell mismatches", len(bad), "max val err", np.abs(vg - vc).max())
for b in bad[:8]:
    print("   pt", adv[b], "gpu", lg[b], "cpu", lc[b], vg[b], vc[b])
for oc, pe, sc in [(10, 1.0, 1.0), (9, -0.8, 0.5), (10, 1.0, 150.0), (4, -2.0, 0.2), (10, -0.5, 0.7)]:
    pts = np.random.default_rng(3).uniform(-1, 1, (100000, 3)).astype(np.float32)
    g = ctx.probe_fbm(pts, oc, pe, sc); c = o.fbm(pts, oc, pe, sc)
    e = np.abs(g - c); print("fbm", oc, pe, sc, "max err", e.max(), "n>4e-6", (e > 4e-6).sum())
for fr, off, gain, sc in [(1.0, 1.0, 1.5, 5.0), (1.0, 1.2, 1.5, 4.0), (1.0, 1.05, 1.5, 20.0), (2.0, 0.9, 2.0, 8.0)]:
    pts = (np.random.default_rng(4).uniform(-1, 1, (100000, 3)) * sc).astype(np.float32)
    g = ctx.probe_ridged(pts, fr, 9, 2.5, off, gain); c = o.ridged(pts, fr, 9, 2.5, off, gain)
    e = np.abs(g - c); print("ridged", fr, off, gain, sc, "max err", e.max(), "n>2e-5", (e > 2e-5).sum(), "p99.9", np.quantile(e, 0.999))

def scene_diag(sc, what):
    ig, ic, cg, cc = T.render_both(ctx, o, sc)
    err = np.abs(ig.astype(np.float64) - ic); bound = T.RTOL * np.abs(ic) + T.ATOL
    r = err / bound; idx = np.unravel_index(np.argmax(r), r.shape)
    print(f"{what:28s} samples {cg['samples']}/{cc['samples']} evals {cg['comp_evals']}/{cc['comp_evals']} oct {cg['octaves']}/{cc['octaves']} "
          f"max err/bound {r.max():.3g} at {idx} gpu {ig[idx]:.6g} cpu {ic[idx]:.6g} n>1 {(r > 1).sum()}")
for name in scenes.FIXTURES:
    scene_diag(scenes.single_galaxy_scene(name, 64, "cli"), name + "/cli")
for name in ["Spiral", "Sombrero", "RedBar", "TonsOfArms", "Irregular", "WheelGalaxy"]:
    for view in ["oblique", "faceon", "edge", "outside"]:
        scene_diag(scenes.single_galaxy_scene(name, 96, view), f"{name}/{view}")
for kind in ["allclasses", "ellipsoid"]:
    for view in ["oblique", "cli"]:
        scene_diag(scenes.single_galaxy_scene(scenes.synthetic_galaxy(kind), 80, view), f"{kind}/{view}")
scene_diag(scenes.multi_galaxy_scene(96), "multi")
