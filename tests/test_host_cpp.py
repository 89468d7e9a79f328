"""The Qt-free C++ host side (gamer_b200/host): .gax layer, PNG writer, and the `galaxy` CLI that is
argument-compatible with the reference's consolerenderer (source/consolerenderer.cpp:16-101)."""
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

import scenes
from gamer_b200.gax import Galaxy

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "gamer_b200", "lib", "gamer_b200_cli")
GAX = os.path.join(ROOT, "tests", "golden", "gax")


def _run(*args, timeout=300, env=None):
    return subprocess.run([CLI, *map(str, args)], capture_output=True, text=True, timeout=timeout,
                          env=dict(os.environ, **env) if env else None)


def _read_png_rgba(path):
    d = open(path, "rb").read()
    assert d[:8] == b"\x89PNG\r\n\x1a\n"
    o, idat, w, h = 8, b"", 0, 0
    while o < len(d):
        n, typ = struct.unpack(">I4s", d[o:o + 8])
        body = d[o + 8:o + 8 + n]
        assert zlib.crc32(typ + body) == struct.unpack(">I", d[o + 8 + n:o + 12 + n])[0]
        if typ == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
            assert (depth, ctype) == (8, 6)
        elif typ == b"IDAT":
            idat += body
        o += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(h, 1 + 4 * w)
    assert not raw[:, 0].any()          # filter type None on every row
    return raw[:, 1:].reshape(h, w, 4)


def test_cli_exists():
    assert os.access(CLI, os.X_OK), "build it: make -C gamer_b200/host"


@pytest.mark.parametrize("name", scenes.FIXTURES)
def test_gax_roundtrip_is_byte_exact(name, tmp_path):
    src = os.path.join(GAX, name + ".gax")
    out = tmp_path / "copy.gax"
    assert _run("gaxcopy", src, out).returncode == 0
    assert open(src, "rb").read() == open(out, "rb").read()


@pytest.mark.parametrize("name", scenes.FIXTURES)
def test_gax_dump_matches_python_reader(name):
    p = _run("gaxdump", os.path.join(GAX, name + ".gax"))
    assert p.returncode == 0
    lines = p.stdout.splitlines()
    g = Galaxy.Load(os.path.join(GAX, name + ".gax"))
    assert lines[1] == "name=" + g.galaxyParams.name
    n = int(lines[6].split("=")[1])
    assert n == len(g.componentParams)
    for line, c in zip(lines[7:], g.componentParams):
        cls, strength, spectrum, nums, cname = line.split("|")
        assert (cls, spectrum, cname) == (c.className, c.spectrum, c.name)
        assert float(strength) == c.strength
        want = [c.arm, c.z0, c.r0, c.active, c.delta, c.winding, c.scale, c.noiseOffset, c.noiseTilt, c.ks, c.inner]
        assert [float(v) for v in nums.split()] == want


def test_gax_truncated_file_is_an_error(tmp_path):
    d = open(os.path.join(GAX, "Spiral.gax"), "rb").read()
    bad = tmp_path / "bad.gax"
    bad.write_bytes(d[:-5])
    p = _run("gaxdump", bad)
    assert p.returncode == 1 and "ERROR" in p.stdout


def test_png_writer(tmp_path):
    out = tmp_path / "t.png"
    assert _run("pngtest", out, 300, 250).returncode == 0
    img = _read_png_rgba(out)
    y, x = np.mgrid[0:250, 0:300]
    assert np.array_equal(img[..., 0], (x * 7 + y) & 255)
    assert np.array_equal(img[..., 1], (y * 5) & 255)
    assert np.array_equal(img[..., 2], (x ^ y) & 255)
    assert (img[..., 3] == 255).all()


def test_fits_writer_matches_python_writer(tmp_path):
    """C++ fits_write.hpp against gamer_b200/fits.py byte for byte, and through the reader."""
    from gamer_b200 import fits
    w, h = 37, 23                                            # ragged: data not a multiple of the 2880-byte block
    base = tmp_path / "p"
    assert _run("fitstest", base, w, h).returncode == 0
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    rgb = np.stack([(x * np.float32(0.37) - y * np.float32(1.25) + np.float32(c * 1000.0)) + np.float32(1e-3) * (x * y)
                    for c in range(3)], axis=-1).astype(np.float32)
    for no, suffix in enumerate(("_red.fits", "_green.fits", "_blue.fits")):
        got = open(str(base) + suffix, "rb").read()
        assert len(got) % 2880 == 0
        plane = fits.read_plane(str(base) + suffix)
        assert plane.shape == (h, w)
        # the pattern is evaluated in float by both sides; compare through the reader, then the bytes of the Python writer
        assert np.array_equal(got, fits.encode_plane(plane.astype(np.float32)))
        assert np.allclose(plane, rgb[..., no], rtol=1e-6, atol=1e-3)


def test_cli_usage_errors(tmp_path):
    assert _run().returncode == 1
    p = _run("galaxy", "omp", 1, 2, 3)                      # consolerenderer.cpp:31-37: exactly 19 words
    assert p.returncode == 1 and "Incorrect usage" in p.stdout
    args = ["galaxy", "cuda", 0.5, 0, 0, 0, 0, 0, 0, 1, 0, 90, 1, 1, 1.0, 0.025, os.path.join(GAX, "Spiral.gax"), 64, tmp_path / "o"]
    p = _run(*args)
    assert p.returncode == 1 and "Cannot recognize" in p.stdout
    args[1], args[16] = "omp", tmp_path / "missing.gax"
    p = _run(*args)
    assert p.returncode == 1 and "Cannot open" in p.stdout


def test_rendering_params_file_cpp_equals_python():
    """RenderParams.dat (the reference's shipped file, older layout without nside/renderType)."""
    from gamer_b200 import renderparams
    path = os.path.join(ROOT, "tests", "golden", "RenderParams.dat")
    p = _run("rpdump", path)
    assert p.returncode == 0
    rp = renderparams.Load(path)
    l1, l2 = p.stdout.splitlines()
    cam = [float(v) for v in l1.split("target=")[0].replace("camera=", "").split()]
    assert np.allclose(cam, rp.camera().m_camera, rtol=1e-7)
    kv = dict(t.split("=") for t in l2.split())
    assert int(kv["size"]) == rp.m_size == 384 and int(kv["noStars"]) == rp.m_noStars == 25000
    assert float(kv["starSize"]) == rp.m_starSize and float(kv["spread"]) == rp.m_starSizeSpread == 1.5
    assert int(kv["spectra"]) == len(rp.m_spectra.m_spectra) == 7 and kv["galaxy"] == rp.m_currentGalaxy == "Spiral"
    assert np.isclose(float(kv["gamma"]), rp.m_gamma) and np.isclose(float(kv["rayStep"]), rp.m_rayStep)
    # SURVEY.md Appendix A's decoding of the same file
    assert np.allclose(rp.camera().m_up, (-0.527, 0.342, -0.778), atol=1e-3) and rp.camera().m_perspective == 90
    assert np.allclose((rp.m_exposure, rp.m_gamma, rp.m_saturation, rp.m_rayStep), (0.767, 0.867, 1.467, 0.02), atol=1e-3)
    assert rp.m_spectra.FindSpectrum("green") == (np.float32(0.7), 1.0, np.float32(0.7))
    # write -> read is the identity; the writer always emits the current layout (with nside/renderType)
    b = renderparams.to_bytes(rp)
    assert renderparams.to_bytes(renderparams.from_bytes(b)) == b and len(b) == 520 + 4 + 4


@pytest.mark.gpu
def test_cli_skybox_equals_python_pipeline(tmp_path):
    """`skybox omp RenderParams.dat Spiral.gax 48`: six PNGs, each equal to the ctypes pipeline's face."""
    from gamer_b200 import renderparams
    from gamer_b200.model import GalaxyInstance
    from gamer_b200.rasterizer import Rasterizer, RenderSkybox
    rpfile = os.path.join(ROOT, "tests", "golden", "RenderParams.dat")
    p = subprocess.run([CLI, "skybox", "omp", rpfile, os.path.join(GAX, "Spiral.gax"), "48"], capture_output=True, text=True,
                       timeout=300, cwd=tmp_path)
    assert p.returncode == 0, p.stdout + p.stderr
    rp = renderparams.Load(rpfile)
    rp.m_size = 48
    r = Rasterizer()
    r.AddGalaxy(GalaxyInstance(Galaxy.Load(os.path.join(GAX, "Spiral.gax"))))
    faces = RenderSkybox(r, rp)
    assert sorted(faces) == ["SkyboxX+", "SkyboxX-", "SkyboxY+", "SkyboxY-", "SkyboxZ+", "SkyboxZ-"]
    lit = 0
    for name, sh in faces.items():
        png = _read_png_rgba(str(tmp_path / (name + ".png")))
        want = np.stack([(sh >> 16) & 255, (sh >> 8) & 255, sh & 255, sh >> 24], axis=-1).astype(np.uint8)
        assert np.array_equal(png, want), name
        lit += int(want[..., :3].max() > 20)
    assert lit >= 2          # the camera sits off-centre: some faces look away from the galaxy


@pytest.mark.gpu
@pytest.mark.parametrize("name,size,gamma", [("Spiral", 128, 1.0), ("Sombrero", 96, 0.8)])
def test_cli_image_equals_python_pipeline(name, size, gamma, tmp_path):
    """singleGalaxy.sh's command line -> PNG; pixels equal the ctypes pipeline's shadow image."""
    from gamer_b200.model import GalaxyInstance, RenderingParams
    from gamer_b200.rasterizer import Rasterizer
    out = tmp_path / "img"
    p = _run("galaxy", "omp", 0.5, 0, 0, 0, 0, 0, 0, 1, 0, 90, 1.2, gamma, 1.1, 0.025, os.path.join(GAX, name + ".gax"), size, out,
             env={"GAMER_B200_FITS": "1"})
    assert p.returncode == 0, p.stdout + p.stderr
    png = _read_png_rgba(str(out) + ".png")
    rp = RenderingParams(m_size=size, m_exposure=float(np.float32(1.2)), m_gamma=float(np.float32(gamma)),
                         m_saturation=float(np.float32(1.1)), m_rayStep=float(np.float32(0.025)))
    rp.camera().m_camera, rp.camera().m_target, rp.camera().m_up, rp.camera().m_perspective = (0.5, 0, 0), (0, 0, 0), (0, 1, 0), 90.0
    rp.spectra().PopulateSpectra()
    r = Rasterizer()
    r.setRenderingParams(rp)
    r.AddGalaxy(GalaxyInstance(Galaxy.Load(os.path.join(GAX, name + ".gax"))))
    r.Render()
    while r.getState() != Rasterizer.State.done:
        pass
    r.AssembleImage()
    sh = r.getImageShadowBuffer()
    want = np.stack([(sh >> 16) & 255, (sh >> 8) & 255, sh & 255, sh >> 24], axis=-1).astype(np.uint8)
    assert np.array_equal(png, want)
    assert want[..., :3].max() > 50        # a real image, not black
    # GAMER_B200_FITS=1: the linear planes, bit for bit the render buffer widened to double
    from gamer_b200 import fits
    rb = np.asarray(r.getRenderBuffer()).reshape(size, size, 3)
    for no, suffix in enumerate(("_red.fits", "_green.fits", "_blue.fits")):
        assert np.array_equal(fits.read_plane(str(out) + suffix), rb[..., no].astype(np.float64)), suffix
