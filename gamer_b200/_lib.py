"""ctypes binding of libgamer_b200.so (include/gamer_b200.h).

The library is the product: if it is missing or does not load, importing the compute entry points
fails loudly — there is no Python/CPU fallback for the hot path.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GAMER_B200_LIB") or os.path.join(_HERE, "lib", "libgamer_b200.so")   # env: A/B builds

CNT_RAYS, CNT_SAMPLES, CNT_GATES, CNT_COMP_EVALS, CNT_OCTAVES, CNT_N = 0, 1, 2, 3, 4, 8
E_ARG, E_CUDA, E_ABORTED, E_STATE = -1, -2, -3, -4
FRAME_CONTINUE, FRAME_PARTIAL, FRAME_NO_RADIAL_GATE = 1, 2, 4
MAX_GALAXIES, MAX_COMPONENTS = 8, 64

CLASS_IDS = {          # factory strings of source/galaxy/galaxy.cpp:15-42
    "bulge": 0, "disk": 1, "dust": 2, "dust2": 3, "dust positive": 4, "stars": 5, "stars small": 6,
}


class GamerComponent(C.Structure):
    _fields_ = [("class_id", C.c_int32), ("_pad0", C.c_int32), ("active", C.c_double),
                ("strength", C.c_double), ("arm", C.c_double), ("z0", C.c_double), ("r0", C.c_double),
                ("inner", C.c_double), ("delta", C.c_double), ("winding", C.c_double), ("scale", C.c_double),
                ("noise_offset", C.c_double), ("noise_tilt", C.c_double), ("ks", C.c_double),
                ("spec", C.c_float * 3), ("_pad1", C.c_float)]


class GamerGalaxy(C.Structure):
    _fields_ = [("pos", C.c_float * 3), ("orient", C.c_float * 3), ("rotmat", C.c_float * 4),
                ("axis", C.c_float * 3), ("_pad0", C.c_float),
                ("intensity_scale", C.c_double), ("winding_b", C.c_double), ("winding_n", C.c_double),
                ("no_arms", C.c_double), ("arm", C.c_double * 4),
                ("comp_begin", C.c_int32), ("comp_count", C.c_int32)]


class GamerFrame(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("full_size", C.c_int32),
                ("x0", C.c_int32), ("y0", C.c_int32), ("band_rows", C.c_int32), ("band_stride", C.c_int32),
                ("flags", C.c_int32),
                ("inv_vp", C.c_float * 16), ("cam", C.c_float * 3), ("_pad1", C.c_float),
                ("ray_step", C.c_double), ("min_ray_step", C.c_double),
                ("n_galaxies", C.c_int32), ("n_components", C.c_int32), ("_pad2", C.c_int32 * 2)]


class GamerStar(C.Structure):
    _fields_ = [("x", C.c_int32), ("y", C.c_int32), ("w", C.c_int32), ("cs", C.c_float * 3)]


STARS_LAYER = C.c_void_p(1)          # GAMER_STARS_LAYER


# every symbol include/gamer_b200.h declares (tests check the library exports all of them)
SYMBOLS = [
    "gamer_b200_abi_version", "gamer_b200_device_count", "gamer_b200_create", "gamer_b200_destroy",
    "gamer_b200_last_error", "gamer_b200_upload", "gamer_b200_render", "gamer_b200_render_device",
    "gamer_b200_render_counted", "gamer_b200_counters", "gamer_b200_tonemap", "gamer_b200_tonemap_device",
    "gamer_b200_progress", "gamer_b200_abort", "gamer_b200_device_buffers", "gamer_b200_last_kernel_ms",
    "gamer_b200_launch_count", "gamer_b200_stream", "gamer_b200_camera_inv_vp", "gamer_b200_rotation_to_orientation",
    "gamer_b200_host_alloc", "gamer_b200_host_free",
    "gamer_b200_probe_simplex", "gamer_b200_probe_fbm", "gamer_b200_probe_ridged",
    "gamer_b200_star_list", "gamer_b200_render_stars",
    "gamer_b200_fbm_static_max_points", "gamer_b200_fbm_static", "gamer_b200_copy_bands",
]

_lib = None


class GamerError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"gamer_b200 error {code}: {msg}")
        self.code = code


def load():
    """Load the C-ABI library; raises if it has not been built (python __graft_entry__.py build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `make -C gamer_b200/csrc` "
                          "(there is no CPU fallback for the render path)")
    lib = C.CDLL(LIB_PATH)
    p = C.c_void_p
    lib.gamer_b200_abi_version.restype = C.c_int
    lib.gamer_b200_last_error.restype = C.c_char_p
    lib.gamer_b200_last_error.argtypes = [p]
    lib.gamer_b200_device_count.argtypes = [C.POINTER(C.c_int32)]
    lib.gamer_b200_create.argtypes = [C.POINTER(p), C.POINTER(C.c_int32), C.c_int32]
    lib.gamer_b200_destroy.argtypes = [p]
    lib.gamer_b200_destroy.restype = None
    lib.gamer_b200_upload.argtypes = [p, C.POINTER(GamerFrame), C.POINTER(GamerGalaxy), C.POINTER(GamerComponent)]
    lib.gamer_b200_render.argtypes = [p, p]
    lib.gamer_b200_render_device.argtypes = [p]
    lib.gamer_b200_render_counted.argtypes = [p]
    lib.gamer_b200_counters.argtypes = [p, C.POINTER(C.c_uint64)]
    lib.gamer_b200_tonemap.argtypes = [p, C.c_double, C.c_double, C.c_double, p, p, p]
    lib.gamer_b200_tonemap_device.argtypes = [p, C.c_double, C.c_double, C.c_double]
    lib.gamer_b200_progress.argtypes = [p, C.POINTER(C.c_double)]
    lib.gamer_b200_abort.argtypes = [p]
    lib.gamer_b200_device_buffers.argtypes = [p, C.c_int32, C.POINTER(p), C.POINTER(p), C.POINTER(p), C.POINTER(C.c_int32)]
    lib.gamer_b200_last_kernel_ms.argtypes = [p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.gamer_b200_stream.argtypes = [p, C.c_int32, C.POINTER(p)]
    lib.gamer_b200_launch_count.argtypes = [p, C.POINTER(C.c_uint64)]
    lib.gamer_b200_camera_inv_vp.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_double,
                                             C.POINTER(C.c_float)]
    lib.gamer_b200_rotation_to_orientation.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.gamer_b200_host_alloc.argtypes = [C.POINTER(p), C.c_uint64]
    lib.gamer_b200_host_free.argtypes = [p]
    lib.gamer_b200_probe_simplex.argtypes = [p, p, C.c_int32, C.c_double, p, p]
    lib.gamer_b200_probe_fbm.argtypes = [p, p, C.c_int32, C.c_int32, C.c_double, C.c_double, p]
    lib.gamer_b200_probe_ridged.argtypes = [p, p, C.c_int32, C.c_double, C.c_int32, C.c_double, C.c_double, C.c_double, p]
    lib.gamer_b200_star_list.argtypes = [C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_uint64,
                                         C.POINTER(GamerStar)]
    lib.gamer_b200_render_stars.argtypes = [p, C.POINTER(GamerStar), C.c_int32, p]
    if hasattr(lib, "gamer_b200_copy_bands"):
        lib.gamer_b200_copy_bands.argtypes = [p, C.c_int32, p]
    if hasattr(lib, "gamer_b200_fbm_static"):       # absent only from older A/B builds selected through GAMER_B200_LIB
        lib.gamer_b200_fbm_static_max_points.restype = C.c_int
        lib.gamer_b200_fbm_static.argtypes = [p, C.c_int32, C.c_int32, C.c_double, C.c_double, p]
    _lib = lib
    return lib


def check(ctx, rc):
    if rc != 0:
        msg = load().gamer_b200_last_error(ctx)
        raise GamerError(rc, msg.decode() if msg else "")
