"""Rasterizer — the reference's seam (source/galaxy/rasterizer.h:20-127) over the C ABI.

Method names, argument meaning and call order follow the reference class so a caller written for
`Rasterizer` (ConsoleRenderer::RenderSingleGalaxy, consolerenderer.cpp:29-101) ports line by line:
setRenderingParams / AddGalaxy / Render (or RenderDirect) / getState / getPercentDone / Abort /
AssembleImage / getRenderBuffer / getBuffer / getImageShadowBuffer.  RenderPixels() is one blocking
FFI call into the CUDA kernels; there is no CPU fallback.
"""
import ctypes as C
import threading
from typing import List, Optional

import numpy as np

from . import _lib
from .gax import Galaxy
from .model import GalaxyInstance, RenderingParams, Scene, build_scene, split_passes


class Context:
    """Owns one gamer_ctx (device memory + streams). One per Rasterizer."""

    def __init__(self, device_ids: Optional[List[int]] = None):
        self.lib = _lib.load()
        self.h = C.c_void_p()
        ids = device_ids or [0]
        arr = (C.c_int32 * len(ids))(*ids)
        _lib.check(None, self.lib.gamer_b200_create(C.byref(self.h), arr, len(ids)))
        self.shape = None

    def close(self):
        if self.h:
            self.lib.gamer_b200_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, scene: Scene):
        _lib.check(self.h, self.lib.gamer_b200_upload(self.h, C.byref(scene.frame), scene.galaxies, scene.components))
        self.shape = (scene.frame.height, scene.frame.width)

    def render(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        h, w = self.shape
        if out is None:
            out = np.empty((h, w, 3), dtype=np.float32)
        assert out.dtype == np.float32 and out.size == h * w * 3 and out.flags.c_contiguous
        _lib.check(self.h, self.lib.gamer_b200_render(self.h, out.ctypes.data))
        return out

    def render_passes(self, scene: Scene, out: Optional[np.ndarray] = None, counted=False):
        """Render a scene of any number of galaxies: far -> near chunks accumulate in the device framebuffer
        (GAMER_FRAME_CONTINUE / _PARTIAL), the last pass applies the final scale and copies the frame out."""
        passes = split_passes(scene)
        total = {}
        for p in passes[:-1]:
            self.upload(p)
            if counted:
                for k, v in self.render_counted().items():
                    total[k] = total.get(k, 0) + v
            else:
                self.render_device()
        self.upload(passes[-1])
        if counted:
            for k, v in self.render_counted().items():
                total[k] = total.get(k, 0) + v
            return total
        return self.render(out)

    def render_ptr(self, host_ptr: int):
        _lib.check(self.h, self.lib.gamer_b200_render(self.h, host_ptr))

    def render_device(self):
        _lib.check(self.h, self.lib.gamer_b200_render_device(self.h))

    def render_counted(self) -> dict:
        _lib.check(self.h, self.lib.gamer_b200_render_counted(self.h))
        c = (C.c_uint64 * _lib.CNT_N)()
        _lib.check(self.h, self.lib.gamer_b200_counters(self.h, c))
        return {"rays": c[0], "samples": c[1], "gates": c[2], "comp_evals": c[3], "octaves": c[4]}

    def tonemap(self, exposure=1.0, gamma=1.0, saturation=1.0, stars: Optional[np.ndarray] = None):
        h, w = self.shape
        argb = np.empty((h, w), dtype=np.uint32)
        shadow = np.empty((h, w), dtype=np.uint32)
        sp = None
        if stars is not None:
            stars = np.ascontiguousarray(stars, dtype=np.float32)
            assert stars.size == h * w * 3
            sp = stars.ctypes.data
        _lib.check(self.h, self.lib.gamer_b200_tonemap(self.h, exposure, gamma, saturation, sp,
                                                       argb.ctypes.data, shadow.ctypes.data))
        return argb, shadow

    def render_stars(self, stars, want_layer=False):
        """K4: stamp a GamerStar array (see star_list) into the device star layer of the uploaded frame."""
        n = len(stars)
        out = np.empty(self.shape + (3,), dtype=np.float32) if want_layer else None
        _lib.check(self.h, self.lib.gamer_b200_render_stars(self.h, stars, n, out.ctypes.data if want_layer else None))
        return out

    def tonemap_layer(self, exposure=1.0, gamma=1.0, saturation=1.0):
        """AssembleImage with the device star layer (GAMER_STARS_LAYER)."""
        h, w = self.shape
        argb = np.empty((h, w), dtype=np.uint32)
        shadow = np.empty((h, w), dtype=np.uint32)
        _lib.check(self.h, self.lib.gamer_b200_tonemap(self.h, exposure, gamma, saturation, _lib.STARS_LAYER,
                                                       argb.ctypes.data, shadow.ctypes.data))
        return argb, shadow

    def tonemap_ptr(self, exposure, gamma, saturation, argb_ptr, shadow_ptr):
        _lib.check(self.h, self.lib.gamer_b200_tonemap(self.h, exposure, gamma, saturation, None, argb_ptr, shadow_ptr))

    def tonemap_device(self, exposure=1.0, gamma=1.0, saturation=1.0):
        _lib.check(self.h, self.lib.gamer_b200_tonemap_device(self.h, exposure, gamma, saturation))

    def kernel_ms(self):
        a, b = C.c_float(), C.c_float()
        _lib.check(self.h, self.lib.gamer_b200_last_kernel_ms(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def launches(self) -> int:
        n = C.c_uint64()
        _lib.check(self.h, self.lib.gamer_b200_launch_count(self.h, C.byref(n)))
        return n.value

    def stream(self, which=0) -> int:
        s = C.c_void_p()
        _lib.check(self.h, self.lib.gamer_b200_stream(self.h, which, C.byref(s)))
        return s.value

    def device_buffers(self, which=0):
        p = [C.c_void_p(), C.c_void_p(), C.c_void_p()]
        rows = C.c_int32()
        _lib.check(self.h, self.lib.gamer_b200_device_buffers(self.h, which, C.byref(p[0]), C.byref(p[1]), C.byref(p[2]),
                                                              C.byref(rows)))
        return p[0].value, p[1].value, p[2].value, rows.value

    def copy_bands(self, which_buf: int, host_full_frame) -> None:
        """gamer_b200_copy_bands: this context's tile into its place in a full_size x full_size host image
        (numpy array or raw address; which_buf 0 float RGB, 1 ARGB32, 2 shadow)."""
        ptr = host_full_frame if isinstance(host_full_frame, int) else host_full_frame.ctypes.data
        _lib.check(self.h, self.lib.gamer_b200_copy_bands(self.h, which_buf, ptr))

    def device_tensors(self, which=0):
        """Zero-copy torch views of the device framebuffers of the last render: (rgb [rows,W,3] float32,
        argb [rows,W] int32, shadow [rows,W] int32). For callers that keep results on the GPU (NCCL gather)."""
        import torch

        class _Cai:
            def __init__(self, ptr, shape, typestr):
                self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 2}

        rgb, argb, shadow, rows = self.device_buffers(which)
        w = self.shape[1]
        dev = torch.device("cuda", torch.cuda.current_device())
        return (torch.as_tensor(_Cai(rgb, (rows, w, 3), "<f4"), device=dev),
                torch.as_tensor(_Cai(argb, (rows, w), "<i4"), device=dev),
                torch.as_tensor(_Cai(shadow, (rows, w), "<i4"), device=dev))

    def progress(self) -> float:
        f = C.c_double()
        _lib.check(self.h, self.lib.gamer_b200_progress(self.h, C.byref(f)))
        return f.value

    def abort(self):
        _lib.check(self.h, self.lib.gamer_b200_abort(self.h))

    # probes of the device noise code (parity tests)
    def probe_simplex(self, xyz: np.ndarray, frequency=1.0):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        n = xyz.shape[0]
        lat = np.empty((n, 10), dtype=np.int32)
        val = np.empty(n, dtype=np.float32)
        _lib.check(self.h, self.lib.gamer_b200_probe_simplex(self.h, xyz.ctypes.data, n, frequency, lat.ctypes.data, val.ctypes.data))
        return lat, val

    def probe_fbm(self, xyz, octaves, persistence, scale):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        val = np.empty(xyz.shape[0], dtype=np.float32)
        _lib.check(self.h, self.lib.gamer_b200_probe_fbm(self.h, xyz.ctypes.data, xyz.shape[0], octaves, persistence, scale, val.ctypes.data))
        return val

    def probe_ridged(self, xyz, frequency, octaves, lacunarity, offset, gain):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        val = np.empty(xyz.shape[0], dtype=np.float32)
        _lib.check(self.h, self.lib.gamer_b200_probe_ridged(self.h, xyz.ctypes.data, xyz.shape[0], frequency, octaves, lacunarity,
                                                            offset, gain, val.ctypes.data))
        return val


def star_list(size, no_stars, base_size, size_spread, strength, seed=0):
    """gamer_b200_star_list: the deterministic stand-in for the reference's rand()-driven star parameters."""
    lib = _lib.load()
    arr = (_lib.GamerStar * max(1, int(no_stars)))()
    _lib.check(None, lib.gamer_b200_star_list(int(size), int(no_stars), float(base_size), float(size_spread), float(strength),
                                              int(seed), arr))
    return (_lib.GamerStar * int(no_stars)).from_buffer(arr) if no_stars > 0 else (_lib.GamerStar * 0)()


class Rasterizer:
    """rasterizer.h:20-127."""

    class State:
        idle, rendering, done = 0, 1, 2

    def __init__(self, device_ids: Optional[List[int]] = None):
        self.m_galaxies: List[GalaxyInstance] = []
        self.m_renderingParams: Optional[RenderingParams] = None
        self.m_state = Rasterizer.State.idle
        self.m_isPreview = False
        self.m_renderBuffer: Optional[np.ndarray] = None      # Buffer2D: float RGB [y][x][3]
        self.m_starsBuffer: Optional[np.ndarray] = None
        self.m_imageBuffer: Optional[np.ndarray] = None       # QImage ARGB32
        self.m_imageShadowBuffer: Optional[np.ndarray] = None
        self._ctx = Context(device_ids)
        self._thread: Optional[threading.Thread] = None
        self._error: Optional[BaseException] = None

    # -- inputs --------------------------------------------------------------------------------
    def setRenderingParams(self, rp: RenderingParams):
        self.m_renderingParams = rp

    def AddGalaxy(self, gi_or_file, position=(0, 0, 0), orientation=(0, 1, 0), iscale=1.0, redshift=0.0, name=""):
        if isinstance(gi_or_file, GalaxyInstance):
            self.m_galaxies.append(gi_or_file)
        else:                                              # rasterizer.cpp:229-236
            g = Galaxy.Load(gi_or_file)
            self.m_galaxies.append(GalaxyInstance(g, name, position, orientation, iscale, redshift))

    def Clear(self):
        self.m_galaxies = []

    def CopyFrom(self, other: "Rasterizer"):
        self.m_galaxies = list(other.m_galaxies)

    def setNewSize(self, s: int):
        self.m_renderingParams.m_size = int(s)

    # -- render ---------------------------------------------------------------------------------
    def Prepare(self) -> Scene:
        return build_scene(self.m_renderingParams, self.m_galaxies, preview=self.m_isPreview)

    def RenderPixels(self):
        scene = self.Prepare()
        try:
            self.m_renderBuffer = self._ctx.render_passes(scene)
        except _lib.GamerError as e:
            if e.code != _lib.E_ABORTED:
                raise
            self.m_renderBuffer = np.zeros(self._ctx.shape + (3,), dtype=np.float32)

    def run(self):
        self.m_state = Rasterizer.State.rendering
        try:
            self.RenderPixels()
        except BaseException as e:          # surfaced by getState()
            self._error = e
        self.m_state = Rasterizer.State.done

    def Render(self):
        """Asynchronous, like QThread::start (rasterizer.cpp:350-362)."""
        self.m_isPreview = False
        self._error = None
        self.m_state = Rasterizer.State.rendering
        self._thread = threading.Thread(target=self.run, daemon=True)
        self._thread.start()

    def RenderDirect(self):
        """Synchronous preview (rasterizer.cpp:365-370)."""
        self.m_isPreview = True
        self.run()
        if self._error:
            raise self._error

    def getState(self):
        if self.m_state == Rasterizer.State.done and self._error:
            err, self._error = self._error, None
            raise err
        return self.m_state

    def getPercentDone(self) -> float:
        return self._ctx.progress()

    def Abort(self):
        self._ctx.abort()

    def RenderStars(self, stars: Optional[np.ndarray] = None, seed: int = 0):
        """Rasterizer::RenderStars (rasterizer.cpp:215-225). Without an argument the layer is generated on the
        device from RenderingParams' noStars / starSize / starSizeSpread / starStrength (K4; the star
        parameters come from the documented counter-based generator, the reference's are rand()-driven and
        not reproducible). A caller may instead hand in its own float layer [N,N,3]."""
        self.m_starsBuffer = stars
        self._star_seed = seed
        self._device_stars = stars is None

    def AssembleImage(self):
        rp = self.m_renderingParams
        if not self.m_isPreview and getattr(self, "_device_stars", False) and rp.m_noStars > 0:
            n = self._ctx.shape[0] if self._ctx.shape else rp.m_size
            lst = star_list(rp.m_size, rp.m_noStars, rp.m_starSize, rp.m_starSizeSpread, rp.m_starStrength, self._star_seed)
            self._ctx.render_stars(lst)
            self.m_imageBuffer, self.m_imageShadowBuffer = self._ctx.tonemap_layer(rp.m_exposure, rp.m_gamma, rp.m_saturation)
            return
        stars = None if self.m_isPreview else self.m_starsBuffer
        self.m_imageBuffer, self.m_imageShadowBuffer = self._ctx.tonemap(rp.m_exposure, rp.m_gamma, rp.m_saturation, stars)

    # -- outputs --------------------------------------------------------------------------------
    def getRenderBuffer(self):
        return self.m_renderBuffer

    def getBuffer(self):
        return self.m_imageBuffer

    def getImageShadowBuffer(self):
        return self.m_imageShadowBuffer


SKYBOX_FACES = [("Z-", (0, 0, -1), (0, 1, 0)), ("Z+", (0, 0, 1), (0, 1, 0)), ("Y-", (0, 1, 0), (0, 0, -1)),
                ("Y+", (0, -1, 0), (0, 0, 1)), ("X-", (1, 0, 0), (0, 1, 0)), ("X+", (-1, 0, 0), (0, 1, 0))]


def RenderSkybox(rasterizer: "Rasterizer", rp: RenderingParams):
    """RenderQueue::RenderSkybox (galaxy/renderqueue.cpp:129-173) driven like ConsoleRenderer::RenderSkybox
    (consolerenderer.cpp:103-172): six fov-90 views from rp's camera position; returns {"SkyboxZ-": image, ...}
    with the ARGB32 shadow image the reference saves as <name>.png. The star layer (PrepareBuffer renders one,
    rasterizer.cpp:174) uses face index as seed."""
    import copy
    out = {}
    cam = tuple(rp.camera().m_camera)
    for i, (name, plane, up) in enumerate(SKYBOX_FACES):
        face = copy.deepcopy(rp)
        c = face.camera()
        c.m_camera, c.m_target, c.m_up, c.m_perspective = cam, tuple(a + b for a, b in zip(cam, plane)), up, 90.0
        rasterizer.setRenderingParams(face)
        rasterizer.RenderStars(seed=i)
        rasterizer.Render()
        while rasterizer.getState() != Rasterizer.State.done:
            pass
        rasterizer.AssembleImage()
        out["Skybox" + name] = rasterizer.getImageShadowBuffer()
    return out
