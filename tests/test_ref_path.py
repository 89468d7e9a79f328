"""The oracle against frames rendered by the reference's OWN render path compiled in place
(tests/golden/ref_path_frames.npz, made by tests/golden/make_fixtures.py from oracle/_ref/libgamer_ref_path*.so).
Bit-for-bit: every float of every frame, both readings of galaxycomponent.cpp:62 (gamer_b200.h,
GAMER_FRAME_NO_RADIAL_GATE), and every ARGB32 pixel of AssembleImage.  The -m gpu tests hold the CUDA path
to the same frames within the linear-intensity tolerance."""
import os

import numpy as np
import pytest

import ref_cases
import ref_path_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_path_frames.npz")
CASES = ref_cases.cases()
KINDS = {"src": False, "o2": True}       # fixture suffix -> no_radial_gate


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


def test_golden_covers_every_case(golden):
    for name in CASES:
        for kind in KINDS:
            assert name + "|" + kind in golden.files
    # the two readings of the return-less path do differ on the shipped galaxies
    assert any(not np.array_equal(golden[n + "|src"], golden[n + "|o2"]) for n in CASES)


@pytest.mark.parametrize("kind", sorted(KINDS))
@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_equals_reference_frame(golden, oracle, name, kind):
    sc = ref_cases.scene_of(CASES[name], no_radial_gate=KINDS[kind])
    rgb, cnt = oracle.render(sc)
    ref = golden[name + "|" + kind]
    assert rgb.shape == ref.shape
    assert ref.max() > 1.0                                   # a real picture, not an empty frame
    assert np.array_equal(rgb.view(np.uint32), ref.view(np.uint32)), \
        "oracle differs from the reference's own code: max |d| = %g" % float(np.abs(rgb - ref).max())


@pytest.mark.parametrize("name", ref_cases.TONEMAP_CASES)
def test_oracle_tonemap_equals_reference_assemble(golden, oracle, name):
    rgb = golden[name + "|src"]
    for k, (e, g, s) in enumerate(ref_cases.TONEMAPS):
        img, shadow = oracle.tonemap(rgb, e, g, s)
        assert np.array_equal(np.asarray(img).reshape(rgb.shape[:2]), golden["%s|tm%d|image" % (name, k)])
        assert np.array_equal(np.asarray(shadow).reshape(rgb.shape[:2]), golden["%s|tm%d|shadow" % (name, k)])


@pytest.mark.skipif(not (ref_path_lib.available("src") and ref_path_lib.available("o2")),
                    reason="oracle/_ref/libgamer_ref_path*.so are built only where the reference checkout exists")
def test_fixture_is_what_the_reference_library_renders_now(golden):
    """The committed frames are reproducible from the compiled reference (where it is present)."""
    for name in ("Spiral/cli", "three/outside"):
        rp, gis, preview = CASES[name]
        for kind in KINDS:
            with ref_path_lib.RefRasterizer(rp, gis, kind) as r:
                assert np.array_equal(r.render(preview), golden[name + "|" + kind])


@pytest.mark.skipif(not ref_path_lib.available("src"), reason="needs oracle/_ref/libgamer_ref_path.so")
def test_timing_door_renders_the_same_rows(golden):
    """bench.py's CPU arm goes through refpath_render_rows (Prepare + setupCamera + renderPixel per pixel):
    it has to produce the rows Rasterizer::Render produced."""
    rp, gis, preview = CASES["three/outside"]
    with ref_path_lib.RefRasterizer(rp, gis, "src") as r:
        rows, used = r.render_rows(1, 5, 9)
    assert np.array_equal(rows, golden["three/outside|src"][1:1 + 5 * 9:5])


@pytest.mark.skipif(not ref_path_lib.available("src"), reason="needs oracle/_ref/libgamer_ref_path.so")
def test_rays_equal_reference_camera(oracle):
    """GamerCamera::setupViewmatrix + coord2ray (gamercamera.cpp:185-217) through the reference's own code."""
    import ctypes as C
    for name in ("Spiral/cli", "Spiral/tilted", "eleven/outside"):
        rp, gis, preview = CASES[name]
        sc = ref_cases.scene_of(CASES[name])
        n = rp.size()
        with ref_path_lib.RefRasterizer(rp, gis) as r:
            r.render(preview)                               # Prepare() sets the camera matrices up
            for idx in list(range(0, n * n, 37)) + [n * n - 1]:
                d = np.zeros(3, np.float32)
                oracle.lib.go_coord2ray(C.addressof(sc.frame), idx % n, idx // n, d.ctypes.data)
                assert np.array_equal(d, r.ray(idx))


# ---- the CUDA path against the same frames -------------------------------------------------------------------
REL, ABS = 1e-3, 1e-5      # |gpu - ref| <= REL * |ref| + ABS on linear intensity (BASELINE.json north_star)


@pytest.mark.gpu
@pytest.mark.parametrize("kind", sorted(KINDS))
def test_gpu_equals_reference_frames(golden, gpu_ctx, kind):
    for name in sorted(CASES):
        sc = ref_cases.scene_of(CASES[name], no_radial_gate=KINDS[kind])
        rgb = gpu_ctx.render_passes(sc)
        ref = golden[name + "|" + kind]
        err = np.abs(rgb.reshape(ref.shape) - ref) - (REL * np.abs(ref) + ABS)
        assert err.max() <= 0, "%s|%s: exceeds tolerance by %g" % (name, kind, float(err.max()))


@pytest.mark.gpu
def test_gpu_tonemap_equals_reference_assemble(golden, gpu_ctx):
    """K2 on the reference's own frames: an empty scene renders zeros, the frame enters as the added layer
    (0 + v == v exactly), so every ARGB32 pixel has to equal Buffer2D::ToColorBuffer's."""
    import scenes
    from gamer_b200.model import build_scene
    for name in ref_cases.TONEMAP_CASES:
        rgb = golden[name + "|src"]
        gpu_ctx.upload(build_scene(scenes.rendering_params(rgb.shape[0], "cli"), []))
        assert not gpu_ctx.render().any()
        for k, (e, g, s) in enumerate(ref_cases.TONEMAPS):
            img, shadow = gpu_ctx.tonemap(e, g, s, rgb)
            assert np.array_equal(img, golden["%s|tm%d|image" % (name, k)])
            assert np.array_equal(shadow, golden["%s|tm%d|shadow" % (name, k)])


# ---- the star layer's stamping, pinned by the reference's own Buffer2D::RenderGaussian ------------------------------
def _golden_star_cases():
    g = np.load(os.path.join(ROOT, "tests", "golden", "ref_star_layers.npz"))
    from gamer_b200._lib import GamerStar
    for name in sorted({k.split("|")[0] for k in g.files}):
        xyw, cs, layer = g[name + "|xyw"], g[name + "|cs"], g[name + "|layer"]
        lst = (GamerStar * len(xyw))()
        for s, (x, y, w), c in zip(lst, xyw, cs):
            s.x, s.y, s.w = int(x), int(y), int(w)
            s.cs[:] = [float(v) for v in c]
        yield name, lst, layer


def test_oracle_star_stamp_equals_reference_rendergaussian(oracle):
    """tests/golden/ref_star_layers.npz holds layers stamped by the reference's OWN Buffer2D::RenderGaussian
    (util/buffer2d.cpp:224-243, compiled in place: oracle/ref_path_wrap.cpp refpath_stamp_stars) from three star lists
    with overlapping and border-crossing stamps. The oracle's restatement go_stamp_stars reproduces every float."""
    n = 0
    for name, lst, layer in _golden_star_cases():
        assert (layer > 0).mean() > 0.2, name
        assert np.array_equal(oracle.stamp_stars(layer.shape[0], lst), layer), name
        n += 1
    assert n == 3


def test_reference_star_stamp_reproduces_golden():
    """Where the reference library is present (the build container), it reproduces the committed layers."""
    if not ref_path_lib.available("src"):
        pytest.skip("oracle/_ref not built here")
    for name, lst, layer in _golden_star_cases():
        assert np.array_equal(ref_path_lib.stamp_stars(layer.shape[0], lst), layer), name


@pytest.mark.gpu
def test_gpu_star_layer_equals_reference_rendergaussian(gpu_ctx):
    """K4 against the reference's own stamping: within one float ulp (the device's double exp and glibc's agree to an
    ulp of double; the product with the colour is then rounded once), including tiles of the frame."""
    import scenes
    for name, lst, layer in _golden_star_cases():
        size = layer.shape[0]
        full = scenes.single_galaxy_scene("Spiral", size, "oblique")
        gpu_ctx.upload(full)
        got = gpu_ctx.render_stars(lst, want_layer=True)
        ulp = np.spacing(np.abs(layer).astype(np.float32)).astype(np.float64)
        assert np.all(np.abs(got.astype(np.float64) - layer) <= ulp), name
        assert (got == layer).mean() > 0.98, name                  # and nearly all of it bit for bit
        h = size // 2
        gpu_ctx.upload(full.tile(3, 5, size - 7, h))
        part = gpu_ctx.render_stars(lst, want_layer=True)
        assert np.all(np.abs(part.astype(np.float64) - layer[5:5 + h, 3:size - 4]) <= ulp[5:5 + h, 3:size - 4]), name


# ---- negative arm exponents: pow of a negative base is NaN in the reference, and fmax / the contribution test quench it ----
NEGARM = ref_cases.negarm_cases()


@pytest.fixture(scope="module")
def negarm_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_negarm_frames.npz"))


@pytest.mark.parametrize("kind", sorted(KINDS))
@pytest.mark.parametrize("name", sorted(NEGARM))
def test_oracle_equals_reference_frame_negative_arm(negarm_golden, oracle, name, kind):
    """getArm with a negative exponent and a wrapped phase difference beyond 3.14159265 (galaxycomponent.h:120-146):
    frames of the reference's own code, three arms (fmax drops the NaN arm) and one arm (the NaN fails the
    `val * weight > 0.0005` test). The oracle reproduces them bit for bit."""
    sc = ref_cases.scene_of(NEGARM[name], no_radial_gate=KINDS[kind])
    rgb, _ = oracle.render(sc)
    ref = negarm_golden[name + "|" + kind]
    assert ref.max() > 1.0 and np.isfinite(ref).all()
    assert np.array_equal(rgb.view(np.uint32), ref.view(np.uint32)), float(np.abs(rgb - ref).max())


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(NEGARM))
def test_gpu_equals_reference_frame_negative_arm(negarm_golden, gpu_ctx, oracle, name):
    """The CUDA path on the same scenes: finite, within the 1e-3 bar of the reference's frame, sample counts equal the
    oracle's (pow(0, negative) = +inf used to reach the pixel)."""
    sc = ref_cases.scene_of(NEGARM[name])
    gpu_ctx.upload(sc)
    cnt = gpu_ctx.render_counted()
    img = gpu_ctx.render()
    ref = negarm_golden[name + "|src"]
    _, rcnt = oracle.render(sc)
    assert np.isfinite(img).all()
    assert cnt["samples"] == rcnt["samples"] and cnt["gates"] == rcnt["gates"]
    err = np.abs(img.astype(np.float64) - ref) / (1e-3 * np.abs(ref.astype(np.float64)) + 1e-5)
    assert err.max() <= 1.0, (name, float(err.max()))
