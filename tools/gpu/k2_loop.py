"""K2 alone, for ncu: render one 2048^2 frame, then tonemap_device N times (python tools/gpu/k2_loop.py [size] [n])."""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "./tests")
import scenes
from gamer_b200.rasterizer import Context
size = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
n = int(sys.argv[2]) if len(sys.argv) > 2 else 8
ctx = Context([0])
ctx.upload(scenes.single_galaxy_scene("Spiral", size, "oblique")); ctx.render_device()
for _ in range(n):
    ctx.tonemap_device(1.0, 1.0, 1.0)
print("k2 ms (events, warm L2):", ctx.kernel_ms()[1])
ctx.close()
