"""The reference harness's own pass criterion on its own workload: every shipped galaxy at 64x64 from the singleGalaxy.sh
view (tools/galaxy_repro.py:1024-1032), 8-bit RGB values of the CUDA path against the CPU oracle's, which the harness
requires to be IDENTICAL (compare_images, :988-995). Prints one line per fixture.  python tests/tools/eight_bit_report.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_lib, scenes
from gamer_b200.rasterizer import Context

ctx, oracle = Context([0]), oracle_lib.load()
print("fixture        view   size  8-bit equal  max diff  max err/bound")
tot = eq = 0
for name in scenes.FIXTURES:
    for view, size in (("cli", 64), ("oblique", 256)):
        sc = scenes.single_galaxy_scene(name, size, view)
        ctx.upload(sc)
        img = ctx.render()
        _, shadow = ctx.tonemap(1.0, 1.0, 1.0)
        ref, _ = oracle.render(sc)
        _, rshadow = oracle.tonemap(ref)
        g = shadow.view(np.uint8).reshape(-1, 4)[:, :3].astype(np.int16)
        c = rshadow.view(np.uint8).reshape(-1, 4)[:, :3].astype(np.int16)
        err = np.abs(img.astype(np.float64) - ref) / (1e-3 * np.abs(ref) + 1e-5)
        tot += g.size; eq += int((g == c).sum())
        print(f"{name:14s} {view:7s} {size:4d}  {float((g == c).mean()):.6f}    {int(np.abs(g - c).max())}        {float(err.max()):.4f}")
print(f"all: {eq} of {tot} 8-bit values identical ({eq / tot:.6f})")
ctx.close()
