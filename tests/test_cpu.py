"""CPU-only tests: the oracle against the reference-produced golden vectors, the host logic, and
that the C-ABI library loads and exports every symbol include/gamer_b200.h declares."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle_lib
import scenes
from gamer_b200 import _lib
from gamer_b200.gax import Galaxy

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---- oracle pinned by the reference's own simplexnoise.cpp ------------------------------------------
def test_oracle_simplex_equals_reference_golden(oracle):
    g = np.load(os.path.join(ROOT, "tests", "golden", "simplex_ref.npz"))
    raw = np.array([oracle.lib.go_raw3d(*p) for p in g["pts"]])
    assert np.array_equal(raw, g["raw"])                      # bit for bit
    for (oc, pe, sc), ref in zip(g["oct_cases"], g["oct"]):
        v = oracle.fbm(g["pts"][:300].astype(np.float32), int(oc), float(pe), float(sc))
        assert np.array_equal(v, ref)


@pytest.mark.skipif(not os.path.exists(oracle_lib.REF_NOISE), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_simplex_equals_reference_library(oracle):
    ref = C.CDLL(oracle_lib.REF_NOISE)
    ref.ref_raw_noise_3d_batch.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    rng = np.random.default_rng(11)
    pts = np.concatenate([rng.uniform(-50, 50, (200000, 3)), rng.integers(-5, 5, (2000, 3)).astype(np.float64)])
    out = np.empty(len(pts))
    ref.ref_raw_noise_3d_batch(pts.ctypes.data, len(pts), out.ctypes.data)
    mine = np.array([oracle.lib.go_raw3d(*p) for p in pts[:20000]])
    assert np.array_equal(mine, out[:20000])


def test_simplex_known_properties(oracle):
    assert oracle.lib.go_raw3d(0.0, 0.0, 0.0) == 0.0
    pts = np.random.default_rng(5).uniform(-100, 100, (20000, 3)).astype(np.float32)
    _, v = oracle.simplex(pts)
    assert np.all(np.abs(v) < 1.0)
    lat, _ = oracle.simplex(np.array([[0, 0, 0], [-2, 0, 0], [1, 2, 3]], dtype=np.float32))
    assert lat[0, :3].tolist() == [-1, -1, -1]                # fastfloor quirk at exact non-positive integers
    assert lat[2, :3].tolist() == [3, 4, 5]


# ---- .gax reader ------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", scenes.FIXTURES)
def test_gax_roundtrip(name):
    path = os.path.join(scenes.GAX_DIR, name + ".gax")
    g = Galaxy.Load(path)
    assert g.to_bytes() == open(path, "rb").read()
    assert g.galaxyParams.name == name
    assert 3 <= len(g.componentParams) <= 8


def test_gax_contents_spiral():
    g = scenes.load_galaxy("Spiral")
    c = g.componentParams
    assert [x.className for x in c] == ["bulge", "disk", "dust2", "bulge", "stars", "disk"]
    assert c[1].strength == 5000 and c[1].spectrum == "Blue" and c[3].active == 0
    assert (g.galaxyParams.windingB, g.galaxyParams.windingN, g.galaxyParams.noArms) == (0.5, 4.0, 2.0)


# ---- C ABI ------------------------------------------------------------------------------------------
def test_abi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "gamer_b200.h")).read()
    declared = set(re.findall(r"\b(gamer_b200_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    lib = _lib.load()
    for s in declared:
        assert hasattr(lib, s), s
    assert lib.gamer_b200_abi_version() == 1


def test_struct_layout_matches_header():
    assert C.sizeof(_lib.GamerComponent) == 120
    assert C.sizeof(_lib.GamerGalaxy) == 128
    assert C.sizeof(_lib.GamerFrame) == 144


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    h = C.c_void_p()
    rc = lib.gamer_b200_create(C.byref(h), None, 0)
    assert rc == _lib.E_CUDA and not h.value
    assert b"CUDA" in lib.gamer_b200_last_error(None) or b"cuda" in lib.gamer_b200_last_error(None)


def test_missing_library_fails_loudly(tmp_path):
    """no extension, no render path: the binding raises instead of falling back to anything"""
    import subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import sys; sys.path.insert(0, %r)\n"
            "from gamer_b200.rasterizer import Context\n"
            "try:\n    Context([0])\nexcept ImportError as e:\n    print('ImportError', 'no CPU fallback' in str(e))\n" % root)
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120,
                       env=dict(os.environ, GAMER_B200_LIB=str(tmp_path / "nowhere.so")))
    assert p.stdout.strip() == "ImportError True", p.stdout + p.stderr


# ---- host math vs oracle's independent restatement -----------------------------------------------------
@pytest.mark.parametrize("view", list(scenes.VIEWS))
def test_camera_matrix(oracle, view):
    cam, target, up, fov = scenes.VIEWS[view]
    rp = scenes.rendering_params(64, view)
    m = rp.camera().setupViewmatrix()
    ref = oracle.camera_inv_vp(cam, target, up, fov)
    assert np.allclose(m, ref, rtol=2e-6, atol=2e-6)


def test_rotation_to(oracle):
    from gamer_b200.model import GalaxyInstance
    for o in [(0, 1, 0), (0.3, 0.8, -0.5), (-0.6, 0.5, 0.62), (0, -1, 0), (1, 0, 0)]:
        q = np.array(GalaxyInstance(Galaxy(), "", (0, 0, 0), o).rotmat(), dtype=np.float32)
        assert np.allclose(q, oracle.rotation_to(o), atol=1e-6)
    assert GalaxyInstance(Galaxy(), "", (0, 0, 0), (0, 1, 0)).rotmat() == [1.0, 0.0, 0.0, 0.0]


def test_scene_sorted_far_to_near():
    sc = scenes.multi_galaxy_scene(32)
    cam = np.array(sc.frame.cam[:])
    d = [np.linalg.norm(np.array(sc.galaxies[i].pos[:]) - cam) for i in range(sc.frame.n_galaxies)]
    assert d == sorted(d, reverse=True)
    assert sc.frame.n_components == sum(sc.galaxies[i].comp_count for i in range(3))


# ---- oracle behaviour the reference documents ----------------------------------------------------------
def test_oracle_render_basics(oracle):
    sc = scenes.single_galaxy_scene("Spiral", 24, "cli")
    img, cnt = oracle.render(sc)
    assert img.shape == (24, 24, 3) and np.all(img >= 0) and img.max() > 1
    assert cnt["rays"] == 24 * 24 and cnt["gates"] == 5 * cnt["samples"]      # 5 active components
    img1, _ = oracle.render(sc, threads=1)
    assert np.array_equal(img, img1)                                          # thread-count independent
    # tiles and interleaved bands index the same pixels
    t = sc.tile(5, 7, 11, 9)
    part, _ = oracle.render(t)
    assert np.array_equal(part, img[7:16, 5:16])
    b = sc.tile(0, 2, 24, 6, band_rows=2, band_stride=8)
    part, _ = oracle.render(b)
    assert np.array_equal(part, img[[2, 3, 10, 11, 18, 19]])


def test_oracle_tonemap(oracle):
    rgb = np.array([[[0.0, 1.0, 2.55], [30.0, 0.05, 12.345]]], dtype=np.float32)
    a, s = oracle.tonemap(rgb)
    assert s[0, 0] == 0xFF000000 | (0 << 16) | (10 << 8) | 25
    assert s[0, 1] == 0xFF000000 | (255 << 16) | (0 << 8) | 123
    assert a[0, 1] == 0xFF000000 | (123 << 16) | (0 << 8) | 255


def test_star_list_is_deterministic_and_follows_renderstars():
    """gamer_b200_star_list: buffer2d.cpp:140-160's expressions over a counter-based generator."""
    from gamer_b200.rasterizer import star_list
    a = star_list(512, 4000, 0.51, 1.5, 1.0, seed=3)
    b = star_list(512, 4000, 0.51, 1.5, 1.0, seed=3)
    c = star_list(512, 4000, 0.51, 1.5, 1.0, seed=4)
    assert bytes(a) == bytes(b) and bytes(a) != bytes(c)
    assert bytes(star_list(512, 10, 0.51, 1.5, 1.0, seed=3)) == bytes(a)[:10 * 24]      # star i does not depend on n
    x = np.array([s.x for s in a]); y = np.array([s.y for s in a]); w = np.array([s.w for s in a])
    cs = np.array([list(s.cs) for s in a])
    assert x.min() >= 0 and x.max() < 512 and y.min() >= 0 and y.max() < 512
    assert abs(x.mean() - 255.5) < 10 and abs(y.mean() - 255.5) < 10
    # int baseSize = 0, int sizeSpread = 1: sz = max(N(0,1), 0); w = int(int(sz*512)/245.0) is 0 below sz = 245/512
    import math
    p0 = 0.5 * (1 + math.erf((245 / 512) / math.sqrt(2)))
    assert abs((w == 0).mean() - p0) < 0.03 and w.min() == 0
    assert (cs >= 0).all() and (cs[:, 1] <= cs[:, 0] + 1e-6).all()        # cy = min(u + 0.6f, cx)
    big = star_list(2048, 2000, 3.0, 1.0, 2.0, seed=1)                      # baseSize 3: sz >= 1
    assert min(s.w for s in big) >= int(int(1.0 * 2048) / 245.0)


def test_oracle_star_stamp(oracle):
    """RenderGaussian: peak = cs at the centre pixel, per-channel max where stamps overlap, clipped at the border."""
    from gamer_b200._lib import GamerStar
    stars = (GamerStar * 3)()
    stars[0].x, stars[0].y, stars[0].w = 10, 12, 8
    stars[0].cs[:] = [1.0, 0.5, 0.25]
    stars[1].x, stars[1].y, stars[1].w = 12, 12, 8
    stars[1].cs[:] = [0.5, 2.0, 0.25]
    stars[2].x, stars[2].y, stars[2].w = 0, 31, 6              # clipped at two borders
    stars[2].cs[:] = [3.0, 3.0, 3.0]
    layer = oracle.stamp_stars(32, stars)
    g = np.float32(np.exp(-(2 / 8) ** 2 / np.float64(np.float32(0.01))))      # two pixels from a centre, w = 8
    assert np.array_equal(layer[12, 10], np.float32([1.0, 0.5, 0.25]))          # max(own centre, neighbour's tail)
    assert np.array_equal(layer[12, 12], np.float32([0.5, 2.0, 0.25]))
    assert layer[31, 0, 0] == 3.0
    assert layer[12, 14, 0] == np.float32(0.5) * g        # star 0 stops at x < 4: only star 1 reaches column 14
    assert layer[12 + 4].sum() == 0 and layer[:, 20:].sum() == 0
    assert oracle.stamp_stars(16, (GamerStar * 0)()).sum() == 0


def test_fits_export_roundtrip(tmp_path):
    """FitsIO::Savedouble's content (one BITPIX -64 plane per channel, rows bottom-up) as valid FITS."""
    from gamer_b200 import fits
    rgb = np.random.default_rng(3).uniform(0, 40, (37, 37, 3)).astype(np.float32)
    names = fits.save_rgb(str(tmp_path / "img"), rgb)
    assert [os.path.basename(n) for n in names] == ["img_red.fits", "img_green.fits", "img_blue.fits"]
    for no, n in enumerate(names):
        raw = open(n, "rb").read()
        assert len(raw) % 2880 == 0 and raw[:30] == b"SIMPLE  =                    T"
        assert np.array_equal(fits.read_plane(n), rgb[:, :, no].astype(np.float64))
        # first stored value is the bottom-left pixel, big-endian double
        first = np.frombuffer(raw, dtype=">f8", count=1, offset=2880)[0]
        assert first == np.float64(rgb[-1, 0, no])


# ---- oracle pinned by the reference's own Noise / Simplex CLASSES (noise.cpp, simplex.cpp in place) --------
def test_oracle_fbm_and_ridged_equal_reference_class_golden(oracle):
    """tests/golden/noise_class_ref.npz was produced by Noise::octave_noise_3d (noise.cpp:162-180) and
    Noise::getRidgedMf (noise.cpp:81-128) of the reference itself, compiled in place against oracle/qtshim with
    the argument patterns of galaxycomponents.cpp. The restated fBm and ridged multifractal must match bit for bit."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "noise_class_ref.npz"))
    pts = g["pts"]
    for (oc, pe, sc), ref in zip(g["fbm_cases"], g["fbm"]):
        assert np.array_equal(oracle.fbm(pts, int(oc), float(pe), float(sc)), ref), (oc, pe, sc)
    for (fr, oc, lac, off, gain, scale), ref in zip(g["rid_cases"], g["rid"]):
        q = np.ascontiguousarray(pts * np.float32(scale))
        # a negative frequency (ks < 0) makes pow(frequency, -0.05) NaN in the reference too: compared as equal NaNs
        assert np.array_equal(oracle.ridged(q, float(fr), int(oc), float(lac), float(off), float(gain)), ref, equal_nan=True), (fr, off, gain, scale)


def test_dmath_host_build_matches_libm(tmp_path):
    """csrc/dmath.cuh (the call-free FP64 exp / atan / sincos and the fast atan2f of K1) compiled for the host and
    compared with libm over the argument ranges the kernel uses: <= 1 ulp in double, identical once narrowed to float
    (which is all the reference consumes: the winding angle and the twirl quaternion). atan2f_glibc / atanf_glibc
    (the theta of a negative arm exponent) must equal glibc's own float functions bit for bit (glibc < 2.41: the
    fdlibm versions; skipped with a newer libm, whose atan2f is a different, correctly rounded algorithm)."""
    import subprocess
    src = tmp_path / "dm.cpp"
    hdr = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gamer_b200", "csrc", "dmath.cuh")
    src.write_text(r'''
#include "%s"
#include <cstdio>
#include <cstdlib>
#include <cmath>
static double ulps(double a, double b) { if (a == b) return 0; double u = std::fabs(std::nextafter(b, INFINITY) - b); return std::fabs(a - b) / u; }
int main() {
    srand(1);
    double me = 0, ma = 0, ms = 0; float mf = 0; long mism = 0, mg = 0;
    for (long i = 0; i < 1000000; i++) {
        double u = rand() / (double)RAND_MAX;
        double x = -12.0 * u, e = dmath::exp_nc(x), er = exp(x);
        me = fmax(me, ulps(e, er)); mism += ((float)e != (float)er);
        double y = 6.0 * u * u * (i & 1 ? 1 : -1), a = dmath::atan_nc(y), ar = atan(y);
        ma = fmax(ma, ulps(a, ar)); mism += ((float)a != (float)ar);
        double an = (float)(600.0 * (u - 0.5)), s, c; dmath::sincos_nc(an, s, c);
        ms = fmax(ms, fmax(std::fabs(s - sin(an)), std::fabs(c - cos(an))));
        mism += ((float)s != (float)sin(an)) + ((float)c != (float)cos(an));
        float fy = (float)(2 * (rand() / (double)RAND_MAX) - 1), fx = (float)(2 * (rand() / (double)RAND_MAX) - 1);
        mf = fmaxf(mf, fabsf(dmath::atan2f_fast(fy, fx) - (float)atan2((double)fy, (double)fx)));
        if (i & 2) fy *= 1e-3f;
        if (i & 4) fx *= 1e-4f;
        float ag = dmath::atan2f_glibc(fy, fx), al = atan2f(fy, fx), tg = dmath::atanf_glibc(6.0f * fy), tl = atanf(6.0f * fy);
        mg += (memcmp(&ag, &al, 4) != 0) + (memcmp(&tg, &tl, 4) != 0);    // the libm's float atan2f / atanf, bit for bit
    }
    printf("%%.4f %%.4f %%.3g %%.3g %%ld %%ld\n", me, ma, ms, (double)mf, mism, mg);
}
''' % hdr)
    exe = tmp_path / "dm"
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-o", str(exe), str(src)], check=True)
    me, ma, ms, mf, mism, mg = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    assert float(me) <= 1.0 and float(ma) <= 1.0, (me, ma)
    assert float(ms) <= 3e-16 and float(mf) <= 5e-7, (ms, mf)
    assert int(mism) == 0
    import platform
    lib, ver = platform.libc_ver()
    if lib == "glibc" and tuple(int(v) for v in ver.split(".")[:2]) < (2, 41):
        assert int(mg) == 0


def test_scene_builder_twirl_groups_and_exact_thresholds(tmp_path):
    """csrc/scene_build.cpp compiled for the host: components with the same winding factor form a twirl group
    (DComp::tgroup = first member, -1 when alone), and the float thresholds of the step clamp are exact: the
    reference's double expression clamp(len * rayStep, min, 0.01) (rasterizer.cpp:449) flips exactly there."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "sb.cpp"
    src.write_text(r'''
#include "%s/gamer_b200/csrc/scene_build.cpp"
#include <cmath>
static GamerComponent comp(int cls, double arm, double winding, double active = 1.0) {
    GamerComponent c; memset(&c, 0, sizeof c);
    c.class_id = cls; c.active = active; c.strength = 100; c.arm = arm; c.z0 = 0.02; c.r0 = 0.4; c.inner = 0.1; c.delta = 0;
    c.winding = winding; c.scale = 5; c.noise_offset = 0; c.noise_tilt = 1; c.ks = 1; c.spec[0] = c.spec[1] = c.spec[2] = 1;
    return c;
}
int main() {
    GamerComponent cs[9] = {
        comp(GAMER_CLASS_BULGE, 1.0, 0.12), comp(GAMER_CLASS_DISK, 0.4, 0.12), comp(GAMER_CLASS_DUST, 0.2, 0.3),
        comp(GAMER_CLASS_DISK, 0.3, 0.5, 0.0) /* inactive */, comp(GAMER_CLASS_DUST2, 0.3, 0.12),
        comp(GAMER_CLASS_DUST_POSITIVE, 0.0, 0.12) /* no arm: factor 0 */, comp(GAMER_CLASS_STARS, 0.2, 0.12),
        comp(GAMER_CLASS_DISK, 0.1, 0.12), comp(GAMER_CLASS_DUST2, 0.0, 0.3) /* no arm: factor 0 */ };
    GamerGalaxy g; memset(&g, 0, sizeof g);
    g.orient[1] = 1; g.rotmat[0] = 1; g.axis[0] = g.axis[1] = g.axis[2] = 1; g.intensity_scale = 1;
    g.winding_b = 0.5; g.winding_n = 2; g.no_arms = 2; g.arm[1] = 3.14; g.comp_begin = 0; g.comp_count = 9;
    GamerFrame f; memset(&f, 0, sizeof f);
    f.width = f.height = f.full_size = 64; f.cam[0] = 0.5f; f.ray_step = (double)0.025f; f.min_ray_step = 0.001;
    f.n_galaxies = 1; f.n_components = 9;
    for (int i = 0; i < 4; i++) f.inv_vp[5 * i] = 1;
    static DScene sc; char err[256] = "";
    if (gk_build_scene(&f, &g, cs, &sc, err, sizeof err) != GAMER_OK) { printf("ERR %%s\n", err); return 1; }
    const DGalaxy &d = sc.gal[0];
    printf("%%d %%d", d.comp_count, d.any_tshare);
    for (int k = 0; k < d.comp_count; k++) printf(" %%d:%%d", sc.comp[d.comp_begin + k].cls, sc.comp[d.comp_begin + k].tgroup);
    printf("\n");
    // the step clamp as the reference evaluates it, on both sides of each threshold
    auto step = [&](float L) { double v = (double)L * f.ray_step; v = v < 0.01 ? v : 0.01; v = f.min_ray_step < v ? v : f.min_ray_step; return v; };
    const float hi = sc.len_hi_thr, lo = sc.len_lo_thr;
    printf("%%d %%d %%d %%d\n", step(hi) < 0.01, step(std::nextafter(hi, 2.0f)) == 0.01, step(lo) == f.min_ray_step,
           step(std::nextafter(lo, 2.0f)) > f.min_ray_step);
    printf("%%d\n", sc.est2_lo <= lo * lo && sc.est2_hi >= hi * hi);
    // the slab gate |height / z0| > 2 (galaxycomponent.h:61-64) and the radial gate intensity > 0.001
    // (galaxycomponent.h:72-81, galaxycomponent.cpp:58-65) of the first disk, on both sides of their thresholds
    const DComp &q = sc.comp[d.comp_begin + 1];
    auto slab = [&](float a) { return !(std::fabs((double)a / cs[1].z0) > 2.0); };
    auto radial = [&](float plen) { float radius = plen / g.axis[0]; double r = std::exp(-(double)radius / (cs[1].r0 * 0.5f));
                                    double in = r - 0.01f; in = in < 1 ? in : 1; in = 0 < in ? in : 0; if (in > 0.1) in = 0.1; return in > 0.001; };
    printf("%%d %%d %%d %%d %%d\n", slab(q.h_gate), !slab(std::nextafter(q.h_gate, 9.0f)), radial(q.plen_thr),
           !radial(std::nextafter(q.plen_thr, 9.0f)), q.plen_lo < q.plen_thr && q.plen_thr < q.plen_hi);
    // validation: operands the kernels' slow-path-free division / sqrt cannot take, and scenes whose march cannot
    // make progress, are refused with GAMER_E_ARG instead of producing NaN or a kernel that never ends
    int bad = 0, tried = 0;
    auto refused = [&](GamerGalaxy gg, GamerComponent c1, GamerFrame ff) {
        GamerComponent cc[9]; memcpy(cc, cs, sizeof cc); cc[1] = c1;
        tried++; bad += gk_build_scene(&ff, &gg, cc, &sc, err, sizeof err) == GAMER_E_ARG;
    };
    { GamerGalaxy x = g; x.axis[0] = 0; refused(x, cs[1], f); }
    { GamerGalaxy x = g; x.axis[1] = NAN; refused(x, cs[1], f); }
    { GamerGalaxy x = g; x.winding_b = 0; refused(x, cs[1], f); }
    { GamerGalaxy x = g; x.pos[2] = 1e6f; refused(x, cs[1], f); }
    { GamerComponent c = cs[1]; c.z0 = 0; refused(g, c, f); }
    { GamerComponent c = cs[1]; c.scale = 1e6; refused(g, c, f); }
    { GamerComponent c = cs[1]; c.r0 = INFINITY; refused(g, c, f); }
    { GamerFrame x = f; x.min_ray_step = 1e-7; refused(g, cs[1], x); }
    printf("%%d %%d %%d\n", bad, tried, gk_build_scene(&f, &g, cs, &sc, err, sizeof err) == GAMER_OK);
}
''' % root)
    exe = tmp_path / "sb"
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-I", os.path.join(root, "gamer_b200", "csrc"), "-o", str(exe),
                    str(src), os.path.join(root, "gamer_b200", "csrc", "host_math.cpp")], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines()
    # active components in file order: bulge, disk, dust, dust2, dust positive, stars, disk, dust2
    # groups: {disk, dust2, disk} share 0.12 -> first member (index 1); the arm-less pair shares factor 0 -> index 4
    assert out[0] == "8 1 0:-1 1:1 2:-1 3:1 4:4 5:-1 1:1 3:4", out[0]
    assert out[1] == "1 1 1 1", out[1]
    assert out[2] == "1", out[2]
    assert out[3] == "1 1 1 1 1", out[3]
    assert out[4] == "8 8 1", out[4]


def test_qtf_twirl_y_equals_qt_sequence_and_oracle_shim(tmp_path):
    """csrc/qtf.cuh compiled for the host: (1) twirl_y — the short sequence K1 uses when the galaxy's orientation is
    exactly (0,1,0) — equals the full Qt 5.15 sequence twirl_qt (galaxycomponent.h:86-90 via QQuaternion::fromAxisAndAngle,
    normalized, two operator*) bit for bit on 4 M random points/angles and never leaves its two known quaternion lengths;
    (2) twirl_qt equals the oracle's Qt stand-in (oracle/qt5_math.h, what the reference's own sources were compiled
    against) on every one of them: the half-angle sine / cosine are glibc's sinf / cosf restated (dmath::sincosf_glibc),
    not the correctly rounded values, which differ from glibc's on 2.6 % of angles; (3) the three double constants that say when float(sqrt(len)) is 1 or 1 - 2^-24 are re-derived."""
    import subprocess
    src = tmp_path / "tw.cpp"
    src.write_text(r'''
#include "%s/gamer_b200/csrc/qtf.cuh"
#include "%s/oracle/qt5_math.h"
#include <cstdio>
#include <cstdlib>
#include <cmath>
static double lowest(bool (*pred)(double), double lo, double hi) {   // smallest double in (lo, hi] with pred true
    uint64_t a = dmath::to_bits(lo), b = dmath::to_bits(hi);
    while (b - a > 1) { uint64_t m = a + (b - a) / 2; if (pred(dmath::from_bits(m))) b = m; else a = m; }
    return dmath::from_bits(b);
}
int main() {
    srand(7);
    long bad = 0, notok = 0, shim_bad = 0, sc_diff = 0, n = 4000000;
    const float ax[3] = {0.f, 1.f, 0.f};
    for (long i = 0; i < n; i++) {
        qtf::f3 p = {(float)(2.0 * rand() / RAND_MAX - 1), (float)(0.2 * rand() / RAND_MAX - 0.1), (float)(2.0 * rand() / RAND_MAX - 1)};
        double tw = 1.3 * rand() / RAND_MAX * ((i & 1) ? 1 : -0.3);
        if (i %% 1000 == 0) tw = 0.0;
        bool ok;
        qtf::f3 a = qtf::twirl_qt(ax, p, tw), b = qtf::twirl_y(p, tw, &ok);
        notok += !ok;
        bad += (a.x != b.x || a.y != b.y || a.z != b.z);
        // the oracle's path: QQuaternion::fromAxisAndAngle(orientation, float(tw * 180)) * p with the host's sinf / cosf
        const float deg = (float)(tw * 180.0);
        q5::V3 r = q5::rotate(q5::from_axis_angle(q5::V3(0.f, 1.f, 0.f), deg), q5::V3(p.x, p.y, p.z));
        float s, c; qtf::half_angle_sc(tw, s, c);
        const float ha = q5::deg2rad_f(deg / 2.0f);
        const bool same_sc = (s == std::sin(ha)) && (c == std::cos(ha));
        sc_diff += !same_sc;
        if (same_sc) shim_bad += (r.x != a.x || r.y != a.y || r.z != a.z);
    }
    printf("%%ld %%ld %%ld %%ld %%ld\n", bad, notok, shim_bad, sc_diff, n);
    auto l_of = [](double d) { return (float)std::sqrt(d); };
    (void)l_of;
    const double one_lo = lowest([](double d) { return (float)std::sqrt(d) >= 1.0f; }, 0.9999, 1.0);
    const double one_hi = lowest([](double d) { return (float)std::sqrt(d) > 1.0f; }, 1.0, 1.001);
    const double below_lo = lowest([](double d) { return (float)std::sqrt(d) >= std::nextafter(1.0f, 0.0f); }, 0.9999, 1.0);
    printf("%%d %%d %%d\n", one_lo == QT_LEN_ONE_LO, one_hi == QT_LEN_ONE_HI, below_lo == QT_LEN_BELOW_LO);
    // x / (1 - 2^-24) is the next float away from zero for every normal float
    long div_bad = 0; const float l = std::nextafter(1.0f, 0.0f);
    for (long i = 0; i < 4000000; i++) {
        float x = (float)(2.0 * rand() / RAND_MAX - 1) * std::ldexp(1.0f, -(rand() %% 40));
        if (x == 0.f || std::fabs(x) < 1.2e-38f) continue;
        div_bad += (x / l != std::nextafter(x, x > 0 ? 2.0f : -2.0f));
    }
    printf("%%ld\n", div_bad);
}
''' % (ROOT, ROOT))
    exe = tmp_path / "tw"
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-o", str(exe), str(src)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines()
    bad, notok, shim_bad, sc_diff, n = (int(v) for v in out[0].split())
    assert bad == 0 and notok == 0, out[0]
    assert shim_bad == 0 and sc_diff == 0, out[0]
    assert out[1] == "1 1 1", out[1]
    assert out[2] == "0", out[2]


def test_sincosf_glibc_equals_host_libm(tmp_path):
    """dmath::sincosf_glibc (csrc/dmath.cuh) — the restatement of glibc's sinf / cosf that K1's twirl uses — against the
    host's libm on every 16th float of [0, 16) with both signs, plus every float of [0.01, 0.02) and [1, 1.25): bit
    equal (the whole of [0, 16), 2.2 G values, was checked once the same way: one difference, the sign of sin(-0))."""
    import subprocess
    src = tmp_path / "sc.cpp"
    src.write_text(r'''
#include "%s/gamer_b200/csrc/dmath.cuh"
#include <cstdio>
#include <cstring>
static inline uint32_t asu(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float asf(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static long bad = 0, n = 0;
static void range(float lo, float hi, uint32_t stride) {
    for (uint32_t b = asu(lo); b < asu(hi); b += stride)
        for (int sg = 0; sg < 2; sg++) {
            const float y = asf(b | (sg ? 0x80000000u : 0));
            if (y == 0.0f) continue;
            float s, c; volatile float vy = y;
            const bool ok = dmath::sincosf_glibc(y, s, c);
            bad += !ok || asu(sinf(vy)) != asu(s) || asu(cosf(vy)) != asu(c); n++;
        }
}
int main() {
    range(0.0f, 16.0f, 16); range(0.01f, 0.02f, 1); range(1.0f, 1.25f, 1); range(16.0f, 119.0f, 64);
    float s, c;
    printf("%%ld %%ld %%d\n", bad, n, (int)dmath::sincosf_glibc(121.0f, s, c));
}
''' % ROOT)
    exe = tmp_path / "sc"
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-o", str(exe), str(src)], check=True)
    bad, n, big = (int(v) for v in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split())
    assert bad == 0 and n > 100_000_000 and big == 0, (bad, n, big)


def test_winding_table_is_certified_against_libm(tmp_path):
    """csrc/scene_build.cpp, gk_build_winding_table: getWinding (galaxycomponent.h:156-165) as a piecewise degree-9
    table in s = 1/(radius + 0.05). Evaluated the way the kernel evaluates it, on 1 M random float radii per parameter
    set, it stays within 2^-46 of the reference's double expression with libm — the bound the kernel's guard (2^-44 on
    the float twirl angle) relies on. Also for an ellipsoidal galaxy (radius up to axis.z / axis.x)."""
    import subprocess
    src = tmp_path / "wt.cpp"
    src.write_text(r'''
#include "%s/gamer_b200/csrc/scene_build.cpp"
#include <cstdlib>
int main() {
    GamerGalaxy g; memset(&g, 0, sizeof g);
    double worst = 0; int tables = 0, bytes = 0;
    for (double ax : {1.0, 0.4}) for (double wb : {0.4, 0.54, 1.0, 3.0}) for (double N : {2.0, 4.0, 5.0}) {
        g.axis[0] = (float)ax; g.axis[1] = 1; g.axis[2] = 1; g.winding_b = wb; g.winding_n = N;
        DGalaxy d; memset(&d, 0, sizeof d); std::vector<double> c;
        gk_build_winding_table(g, d, c);
        if (d.wt_n <= 0) { printf("no table for ax %%g wb %%g N %%g\n", ax, wb, N); return 1; }
        tables++; bytes = (int)(c.size() * 8);
        const double rmax = 1.02 / ax; srand(3);
        for (int i = 0; i < 1000000; i++) {
            const float radius = (float)(rmax * rand() / RAND_MAX * (i & 1 ? 1.0 : (double)rand() / RAND_MAX));
            const double a = gk_winding_from_table(d, c.data(), radius), b = winding_direct(g, radius);
            worst = fmax(worst, fabs(a - b) / fabs(b));
        }
    }
    printf("%%d %%.3g %%d %%d\n", tables, worst, worst <= 0x1p-46, bytes);
}
''' % ROOT)
    exe = tmp_path / "wt"
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-I", os.path.join(ROOT, "gamer_b200", "csrc"), "-o", str(exe),
                    str(src), os.path.join(ROOT, "gamer_b200", "csrc", "host_math.cpp")], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    assert out[0] == "24" and out[2] == "1", out
    assert int(out[3]) <= 96 * 10 * 8


def test_tonemap_fp32_decision_never_disagrees_with_the_exact_chain(tmp_path):
    """csrc/tone.cuh compiled for the host: K2 decides a byte in FP32 only where its error bound leaves no integer of
    1..255 within reach of 10 v, and goes through the reference's double chain otherwise. On random pixels, on pixels
    constructed to land within +-4 ulp of every byte boundary (grey and coloured, nine saturations), and on
    zeros / subnormals / overflow / infinities / NaN: a decided byte always equals the exact chain's, the combined
    function always does, and nearly every random pixel is decided."""
    import subprocess
    hdr = os.path.join(ROOT, "gamer_b200", "csrc", "tone.cuh")
    src = tmp_path / "tn.cpp"
    src.write_text(r'''
#include "@HDR@"
#include <cstdio>
#include <random>
int main() {
    std::mt19937_64 rng(7);
    std::uniform_real_distribution<double> U(0, 1);
    const double sats[] = {1.0, 1.467, 0.5, 0.0, -1.0, 3.7, 1e-3, 0.9999999, 25.0};
    long n = 0, decided = 0, bad = 0, undec_adv = 0, nadv = 0;
    auto check = [&](float x, float y, float z, double S, bool adv) {
        uint32_t ea, es, fa, fs;
        { const uint32_t p = tone::tone_exact(x, y, z, S); ea = 0xFF000000u | p; es = 0xFF000000u | ((p & 255u) << 16) | (p & 0xFF00u) | (p >> 16); }
        const tone::ToneSat k = tone::tone_sat(S);
        uint32_t r, g, b;
        const bool ok = tone::tone_decide(x, y, z, k, r, g, b);
        if (adv) { nadv++; undec_adv += !ok; } else { n++; decided += ok; }
        if (ok) { tone::pack_px(r, g, b, fa, fs); if (fa != ea || fs != es) { if (bad < 10) printf("BAD %a %a %a S %g: %08x %08x\n", x, y, z, S, fa, ea); bad++; } }
        uint32_t ma, ms; tone::tone_mapped(x, y, z, S, k, ma, ms);
        if (ma != ea || ms != es) bad++;
    };
    for (long i = 0; i < 2000000; i++) {
        const double S = sats[i % 9];
        double sc = (i & 1) ? 30.0 : pow(10.0, -6 + 8 * U(rng));
        check((float)(sc * U(rng)), (float)(sc * U(rng)), (float)(sc * U(rng)), S, false);
    }
    for (long i = 0; i < 2000000; i++) {
        const double S = sats[i % 9];
        const int m = 1 + (int)(rng() % 256);          // 1..256
        float x, y, z;
        if (i % 3 == 0) { x = y = z = (float)(m / 10.0); }
        else {
            y = (float)(26.0 * U(rng)); z = (float)(26.0 * U(rng));
            const double den = (1 - S) / 3 + S;
            if (fabs(den) < 1e-3) continue;
            x = (float)((m / 10.0 - (1 - S) * ((double)y + z) / 3) / den);
        }
        int ul = (int)(rng() % 9) - 4;
        for (int q = 0; q < (ul < 0 ? -ul : ul); q++) x = nextafterf(x, ul < 0 ? -INFINITY : INFINITY);
        if (i % 3 == 0) { y = z = x; if (i % 2) y = nextafterf(y, INFINITY); }
        check(x, y, z, S, true);
        check(y, x, z, S, true);
        check(z, y, x, S, true);
    }
    const float sp[] = {0.0f, -0.0f, -1.0f, 1e-40f, 255.0f, 25.5f, 25.6f, 1e30f, 3e38f, INFINITY, -INFINITY, NAN, 1e-45f};
    for (float a : sp) for (float b : sp) for (float c : sp) for (double S : sats) check(a, b, c, S, true);
    printf("random: %ld of %ld decided in FP32 (%.6f); boundary cases %ld, undecided %ld; wrong %ld\n", decided, n, decided / (double)n, nadv, undec_adv, bad);
    return bad != 0;
}
'''.replace("@HDR@", hdr))
    exe = tmp_path / "tn"
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-o", str(exe), str(src)], check=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout
    frac = float(r.stdout.split("(")[1].split(")")[0])
    assert frac > 0.999, r.stdout
