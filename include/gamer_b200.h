/*
 * gamer_b200.h — C ABI of the B200-native galaxy volume integrator.
 *
 * This is the drop-in boundary for ONE path of MichaelMoroz/gamer: the per-pixel
 * ray march + component evaluation + simplex-noise octaves + dust absorption /
 * RGB emission (K1) and the exposure/gamma/saturation pass (K2).  Everything a
 * caller of the reference's `Rasterizer` sets or reads for that path crosses
 * this header as plain pointers and sizes; no C++/torch/Qt types, no
 * exceptions, no exit(), no logging.
 *
 * Reference seam replaced (paths relative to the reference checkout):
 *   class Rasterizer            source/galaxy/rasterizer.h:20-127
 *   Rasterizer::RenderPixels    source/galaxy/rasterizer.cpp:279-313   -> gamer_b200_render*
 *   Rasterizer::AssembleImage   source/galaxy/rasterizer.cpp:315-327   -> gamer_b200_tonemap*
 *   Rasterizer::Abort/getPercentDone  rasterizer.h:91-118              -> gamer_b200_abort / _progress
 *   the parameter model (ComponentParams / GalaxyParams / GalaxyInstance /
 *   RenderingParams / GamerCamera)                                     -> the three POD structs below
 *
 * Precision of the structs follows the reference's own storage: what it keeps
 * in `double` (componentparams.h:7-23, galaxyparams.h:12-29, renderingparams.h:19-37)
 * is `double` here, what it keeps in a float Qt value class (QVector3D,
 * QQuaternion, QMatrix4x4) is `float` here, so "identical parameters" means
 * identical bits on both sides of a parity test.
 *
 * Every function returns 0 on success or a negative GAMER_E_* code; the text of
 * the last error of a context is available from gamer_b200_last_error().
 * There is NO CPU fallback: without a CUDA device every compute entry point
 * fails with GAMER_E_CUDA.
 */
#ifndef GAMER_B200_H
#define GAMER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GAMER_B200_ABI_VERSION 1

/* error classes */
#define GAMER_OK          0
#define GAMER_E_ARG      -1   /* bad argument / scene out of supported range  */
#define GAMER_E_CUDA     -2   /* CUDA runtime error (text in last_error)      */
#define GAMER_E_ABORTED  -3   /* gamer_b200_abort() was called during render  */
#define GAMER_E_STATE    -4   /* call order (e.g. render before upload)       */

/* component classes — the factory strings of source/galaxy/galaxy.cpp:15-42 */
#define GAMER_CLASS_BULGE          0   /* "bulge"          galaxycomponents.cpp:5-39    */
#define GAMER_CLASS_DISK           1   /* "disk"           galaxycomponents.cpp:41-59   */
#define GAMER_CLASS_DUST           2   /* "dust"           galaxycomponents.cpp:64-84   */
#define GAMER_CLASS_DUST2          3   /* "dust2"          galaxycomponents.cpp:85-114  */
#define GAMER_CLASS_DUST_POSITIVE  4   /* "dust positive"  galaxycomponents.cpp:116-136 */
#define GAMER_CLASS_STARS          5   /* "stars"          galaxycomponents.cpp:139-155 */
#define GAMER_CLASS_STARS_SMALL    6   /* "stars small"    galaxycomponents.cpp:159-170 — libc rand() per sample,
                                          nondeterministic; evaluated as a no-op exactly as the reference's own
                                          parity harnesses do (tools/galaxy_repro.py:734-737)                  */

#define GAMER_MAX_GALAXIES    8
#define GAMER_MAX_COMPONENTS  64      /* over all galaxies of one frame */

/* One GalaxyComponent: ComponentParams (componentparams.h:7-23) + its resolved spectrum (spectrum.h:50-59). */
typedef struct GamerComponent {
    int32_t class_id;           /* GAMER_CLASS_*                                              */
    int32_t _pad0;
    double  active;             /* evaluated iff == 1.0  (rasterizer.cpp:458)                 */
    double  strength, arm, z0, r0, inner, delta, winding, scale,
            noise_offset, noise_tilt, ks;
    float   spec[3];            /* RGB of the named spectrum; white if the name is unknown    */
    float   _pad1;
} GamerComponent;

/* One GalaxyInstance (galaxyinstance.h:8-23) + its GalaxyParams (galaxyparams.h:12-29). */
typedef struct GamerGalaxy {
    float   pos[3];             /* GalaxyInstance::position                                   */
    float   orient[3];          /* GalaxyInstance::orientation (as given; not re-normalised)  */
    float   rotmat[4];          /* QQuaternion rotationTo((0,1,0), orient) as w,x,y,z         */
    float   axis[3];            /* GalaxyParams::axis (ellipsoid semi-axes)                   */
    float   _pad0;
    double  intensity_scale;
    double  winding_b, winding_n, no_arms;
    double  arm[4];             /* arm1..arm4 (radians)                                       */
    int32_t comp_begin;         /* range into the component array, file order                 */
    int32_t comp_count;
} GamerGalaxy;

/* Scenes with more than GAMER_MAX_GALAXIES galaxies (or GAMER_MAX_COMPONENTS components) are rendered in
 * several passes over far -> near chunks of the sorted galaxy list: the reference accumulates rp.I across
 * galaxies in one float triple per pixel (rasterizer.cpp:383-409); the passes keep that triple in the device
 * framebuffer between them, unscaled, so the result is bit-identical to a single pass. */
#define GAMER_FRAME_CONTINUE 1   /* start every pixel from the unscaled I the previous pass left          */
#define GAMER_FRAME_PARTIAL  2   /* leave the unscaled I (skip the final * 0.01 / rayStep): more to come  */
/* GalaxyComponent::calculateIntensity (galaxycomponent.cpp:45-88) has no return statement on the path where
 * `intensity > 0.001` (:62) fails. As written (and as the reference's tools/galaxy_repro.py:609 restates it)
 * such a sample contributes nothing: that is the default here. g++ and clang at -O1 and above — the reference's
 * own build uses -Ofast, gamer.pro:13 — treat the return-less path as unreachable and delete the test, so in
 * those binaries every sample with intensity > 0 goes on to the `val * weight > 0.0005` test (:79). Set this
 * flag to reproduce such a build. Both behaviours are pinned against the reference's own code compiled in
 * place (oracle/Makefile, tests/test_ref_path.py). */
#define GAMER_FRAME_NO_RADIAL_GATE 4

/* One frame: camera + size + step (RenderingParams, GamerCamera). Galaxies must already be sorted
 * far -> near as Rasterizer::Prepare does (rasterizer.cpp:190-201). */
typedef struct GamerFrame {
    int32_t width, height;      /* tile extent in pixels; the reference has width==height==full_size  */
    int32_t full_size;          /* the `size` used by coord2ray (gamercamera.cpp:210-212)             */
    int32_t x0, y0;             /* tile origin inside the full_size x full_size image                 */
    int32_t band_rows;          /* 0: rows y0..y0+height-1. >0: interleaved bands — local row L is image */
    int32_t band_stride;        /* row y0 + (L / band_rows) * band_stride + (L % band_rows)            */
    int32_t flags;              /* GAMER_FRAME_*: multi-pass rendering of scenes beyond the per-pass limits */
    float   inv_vp[16];         /* inverse(projection*view), element [row*4+col] = QMatrix4x4(row,col)*/
    float   cam[3];             /* GamerCamera::camera                                                */
    float   _pad1;
    double  ray_step;           /* RenderingParams::rayStep                                           */
    double  min_ray_step;       /* 0.001; 0.01 in preview (rasterizer.cpp:437-441)                    */
    int32_t n_galaxies;
    int32_t n_components;
    int32_t _pad2[2];
} GamerFrame;

/* counters filled by gamer_b200_counters (same definitions as the oracle's) */
#define GAMER_CNT_RAYS       0   /* (pixel, galaxy) pairs that reach getIntensity              */
#define GAMER_CNT_SAMPLES    1   /* iterations of the march loop (rasterizer.cpp:447-475)      */
#define GAMER_CNT_GATES      2   /* active-component visits (getRadius+getHeightModulation)    */
#define GAMER_CNT_COMP_EVALS 3   /* componentIntensity calls that do work                      */
#define GAMER_CNT_OCTAVES    4   /* raw simplex evaluations                                    */
#define GAMER_CNT_N          8

typedef struct gamer_ctx gamer_ctx;

/* lifecycle ------------------------------------------------------------------------------- */
int  gamer_b200_abi_version(void);
int  gamer_b200_device_count(int32_t *count);
/* One context per Rasterizer. device_ids/n_devices: the CUDA devices this context renders on
 * (row bands are interleaved across them and gathered into the caller's buffer). */
int  gamer_b200_create (gamer_ctx **out, const int32_t *device_ids, int32_t n_devices);
void gamer_b200_destroy(gamer_ctx *ctx);
const char *gamer_b200_last_error(gamer_ctx *ctx);     /* ctx may be NULL: last create() error */

/* scene ------------------------------------------------------------------------------------ */
/* Copies the scene; the caller may free its arrays on return. Replaces setRenderingParams/
 * AddGalaxy/Prepare (rasterizer.cpp:48-51,190-213,229-240). */
int  gamer_b200_upload (gamer_ctx *ctx, const GamerFrame *frame,
                        const GamerGalaxy *galaxies, const GamerComponent *components);

/* K1: ray march ---------------------------------------------------------------------------- */
/* Blocking. Renders the uploaded frame and writes width*height*3 floats, row-major
 * (x + y*width), RGB interleaved — the layout of Buffer2D (buffer2d.cpp:26-31) — into HOST memory. */
int  gamer_b200_render (gamer_ctx *ctx, float *rgb_linear_out);
/* Same, but leaves the result in the context's device framebuffer (no D2H). */
int  gamer_b200_render_device(gamer_ctx *ctx);
/* Like render_device but also accumulates the GAMER_CNT_* counters (slower instrumented kernel). */
int  gamer_b200_render_counted(gamer_ctx *ctx);
int  gamer_b200_counters(gamer_ctx *ctx, uint64_t out[GAMER_CNT_N]);

/* K2: post-process ------------------------------------------------------------------------- */
/* Buffer2D::PostProcess + ToColorBuffer (buffer2d.cpp:65-85,106-126) on the last rendered frame.
 * stars_rgb_or_null: optional host float[width*height*3] added first (Buffer2D::Add, :204-215).
 * argb_out gets QImage `image` (0xFF, c.red=z, c.green=y, c.blue=x), argb_shadow_out gets the
 * R/B-swapped `shadowImage` the CLI saves as PNG (consolerenderer.cpp:98). Either may be NULL. */
int  gamer_b200_tonemap(gamer_ctx *ctx, double exposure, double gamma, double saturation,
                        const float *stars_rgb_or_null,
                        uint32_t *argb_out, uint32_t *argb_shadow_out);
int  gamer_b200_tonemap_device(gamer_ctx *ctx, double exposure, double gamma, double saturation);

/* ---- background star layer ---------------------------------------------------
 * Rasterizer::RenderStars (rasterizer.cpp:215-225) -> Buffer2D::RenderStars (util/buffer2d.cpp:140-173)
 * -> Buffer2D::RenderGaussian (:224-243): every star is a w x w Gaussian stamp exp(-(dx^2+dy^2)/0.01f)
 * times an RGB strength, combined into the layer with a per-channel MAX, so the layer does not depend
 * on the order of the stars. AssembleImage adds the layer to the frame (rasterizer.cpp:320-321).
 *
 * The reference draws the star parameters from libc rand() and a process-global std::mt19937
 * (util/random.h:49-53) whose state depends on everything drawn before: that sequence is not
 * reproducible, so parity of the star LIST can only be statistical. gamer_b200_star_list defines it
 * instead by a counter-based generator (splitmix64 of (seed, star index, draw index); Gaussians by
 * Box-Muller), with the reference's expressions for position, colour, size and strength — including
 * RenderStars' `int baseSize, int sizeSpread` parameters, which truncate the doubles it is called with.
 * The STAMPING (gamer_b200_render_stars) follows the reference exactly and is checked against the oracle. */
typedef struct {
    int32_t x, y;               /* centre pixel, in the full_size x full_size image              */
    int32_t w;                  /* stamp width `asz` (pixels); 0 draws nothing                   */
    float   cs[3];              /* ss * (cx, cy, cz) as the float QVector3D RenderGaussian gets  */
} GamerStar;

/* Host: fill out[no_stars] for an image of `size` pixels. Deterministic in all arguments. */
int  gamer_b200_star_list(int32_t size, int32_t no_stars, double base_size, double size_spread,
                          double strength, uint64_t seed, GamerStar *out);
/* K4: stamp the stars into the context's device star layer for the uploaded frame (tile and band
 * addressing honoured); optionally copy the layer to the host (width*height*3 floats). */
int  gamer_b200_render_stars(gamer_ctx *ctx, const GamerStar *stars, int32_t n_stars,
                             float *stars_rgb_out_or_null);
/* Pass this as `stars_rgb_or_null` of gamer_b200_tonemap to add the device layer made by
 * gamer_b200_render_stars instead of a host buffer. */
#define GAMER_STARS_LAYER ((const float *)(uintptr_t)1)

/* control (callable from any thread; rasterizer.h:91-118) ----------------------------------- */
int  gamer_b200_progress(gamer_ctx *ctx, double *fraction);
/* Rasterizer::Abort. Like m_abort (set at any time, reset when RenderPixelsOMP ends, rasterizer.cpp:296-297,311) the
 * request is sticky: it stops the render in flight or, made while none is, cancels the next one, and that render
 * consumes it. A render returns GAMER_E_ABORTED only if it was actually cut short (tiles left undrawn are 0); an abort
 * that arrives after the last tile leaves a complete frame, which is delivered with GAMER_OK. */
int  gamer_b200_abort   (gamer_ctx *ctx);

/* device-side access for callers that keep data on the GPU (multi-GPU gather, benchmarks) ---- */
/* Device pointers of device `which` (0..n_devices-1): linear float RGB framebuffer and the two
 * ARGB images; rows_out = number of rows that device holds (its bands, in band order). */
int  gamer_b200_device_buffers(gamer_ctx *ctx, int32_t which, void **rgb_dev, void **argb_dev,
                               void **argb_shadow_dev, int32_t *rows_out);
/* The final gather of the tile partition, host side: copies this (single-device) context's tile — its rows at
 * x0 / y0 / band_rows / band_stride — to their place in a full_size x full_size HOST image. which_buf: 0 = linear float
 * RGB (12 B/pixel), 1 = ARGB32, 2 = shadow ARGB32. With one process per GPU, every process calls it on the SAME host
 * frame (shared memory): no device collective and no serial copy by one rank. Reference analogue: RasterThread's
 * threads writing their chunk of the shuffled render list into the one m_renderBuffer (rasterthread.cpp:6-25,36-51). */
int  gamer_b200_copy_bands(gamer_ctx *ctx, int32_t which_buf, void *host_full_frame);
/* The CUDA stream (cudaStream_t) device `which` of this context launches on, so a caller can order its
 * own work (events, copies, collectives) with the kernels. */
int  gamer_b200_stream(gamer_ctx *ctx, int32_t which, void **stream_out);
/* CUDA-event time of the last K1 / K2 launch on device 0 (milliseconds). */
int  gamer_b200_last_kernel_ms(gamer_ctx *ctx, float *k1_ms, float *k2_ms);
/* number of kernel launches issued by this context since creation */
int  gamer_b200_launch_count(gamer_ctx *ctx, uint64_t *launches);

/* host-side helpers of the parameter model (pure host code, no device needed) ---------------- */
/* GamerCamera::setupViewmatrix (gamercamera.cpp:185-208) with m_rotMatrix = identity:
 * inverse(perspective(fov,1,1,100) * lookAt(target, camera, up)) — eye/centre swapped as in the
 * reference — as 16 floats [row*4+col], ready for GamerFrame.inv_vp. */
int  gamer_b200_camera_inv_vp(const float cam[3], const float target[3], const float up[3],
                              double fov_deg, float inv_vp_out[16]);
/* GalaxyInstance::setupQuaternions (galaxyinstance.cpp:69-71): rotationTo((0,1,0), orient) as w,x,y,z */
int  gamer_b200_rotation_to_orientation(const float orient[3], float wxyz_out[4]);
/* Pinned host memory for fast host<->device copies of framebuffers (optional; any host pointer works). */
int  gamer_b200_host_alloc(void **ptr, uint64_t bytes);
int  gamer_b200_host_free (void *ptr);

/* function-level probes of the device noise code (parity tests; same device functions as K1) -- */
/* raw simplex (simplexnoise.cpp:243-333) at n points given as float coordinates widened to
 * double exactly like the reference's callers do. lattice_out: 10 int32 per point
 * (i,j,k, i1+2*j1+4*k1, i2+2*j2+4*k2, gi0,gi1,gi2,gi3, 0). value_out: noise value. */
int  gamer_b200_probe_simplex(gamer_ctx *ctx, const float *xyz, int32_t n, double frequency,
                              int32_t *lattice_out, float *value_out);
/* Noise::octave_noise_3d (noise.cpp:162-180) and Noise::getRidgedMf (:81-128) */
int  gamer_b200_probe_fbm   (gamer_ctx *ctx, const float *xyz, int32_t n, int32_t octaves,
                             double persistence, double scale, float *value_out);
int  gamer_b200_probe_ridged(gamer_ctx *ctx, const float *xyz, int32_t n, double frequency,
                             int32_t octaves, double lacunarity, double offset, double gain,
                             float *value_out);

/* allocation-free entry point ------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The same evaluation as gamer_b200_probe_fbm (Noise::octave_noise_3d, noise.cpp:162-180, at n points), but
 * without a context and without any cudaMalloc / cudaFree / allocating library call: the points and the
 * values are staged in a static device array inside the library (4 MB, held by every GPU that loads the
 * library until the process exits). Runs on the calling thread's CURRENT CUDA device, on the legacy default
 * stream, and blocks until value_out (host, n floats) is written; xyz is host memory, 3 floats per point.
 *   limit   n <= gamer_b200_fbm_static_max_points() (262144); a larger n returns GAMER_E_ARG with a message
 *           naming the limit (gamer_b200_last_error(NULL)) — split the call.
 *   sharing every call uses the same array, which keeps the previous call's contents (nothing in it needs
 *           to start at zero). Calls are serialised inside one process; calls of DIFFERENT processes or
 *           any other work that overlaps on the GPU and touches the array would race. */
int  gamer_b200_fbm_static_max_points(void);
int  gamer_b200_fbm_static(const float *xyz, int32_t n, int32_t octaves, double persistence, double scale,
                           float *value_out);

#ifdef __cplusplus
}
#endif
#endif /* GAMER_B200_H */
