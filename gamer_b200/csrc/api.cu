// api.cu — the extern "C" boundary (include/gamer_b200.h): contexts, scene upload, K1/K2 launches,
// multi-device band split + gather, progress/abort, probes.  No exceptions, exit() or logging cross
// this file; every CUDA error is turned into GAMER_E_CUDA with its text kept in the context.
#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/gamer_b200.h"
#include "kernels.h"
#include "scene_build.h"

#define MD_BAND_ROWS 32      // rows per interleaved band when one context spans several devices

namespace {

struct Dev {
    int device = 0;
    int sm_count = 0;
    cudaStream_t st = nullptr, ctrl = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};   // K1 begin/end, K2 begin/end
    float *rgb = nullptr, *stars = nullptr;
    GamerStar *star_list = nullptr;            // device copy of the last star list (K4 input)
    size_t cap_star_list = 0;
    bool layer_valid = false;                  // d.stars holds a K4 layer for the uploaded frame
    uint32_t *argb = nullptr, *shadow = nullptr;
    size_t cap_px = 0, cap_stars = 0;
    unsigned long long *counters = nullptr;
    int *abort_flag = nullptr;
    unsigned int *progress = nullptr;
    unsigned int *progress_host = nullptr;    // pinned
    double *wtab = nullptr;                    // the frame's getWinding tables (scene_build.cpp), DEV_MAX_GALAXIES of them
    unsigned int n_blocks = 0;
    int rows = 0;                              // local rows this device renders
    DScene scene;
};

std::string g_create_error;
std::mutex g_create_mutex;

}  // namespace

struct gamer_ctx {
    std::vector<Dev> devs;
    GamerFrame frame;
    bool uploaded = false, rendered = false, counted = false;
    bool skipped = false;                      // the last render was not launched at all (abort requested before it)
    std::string err;
    std::atomic<int> abort_req{0};
    std::atomic<int> rendering{0};
    uint64_t launches = 0;
    float k1_ms = 0.f, k2_ms = 0.f;
    uint64_t counters[GAMER_CNT_N] = {0, 0, 0, 0, 0, 0, 0, 0};
    std::mutex mu;                             // render/tonemap/upload are not re-entrant per context
};

namespace {

int set_err(gamer_ctx *c, int code, const std::string &m) {
    if (c) c->err = m;
    else { std::lock_guard<std::mutex> l(g_create_mutex); g_create_error = m; }
    return code;
}
int cuda_fail(gamer_ctx *c, cudaError_t e, const char *what) {
    return set_err(c, GAMER_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CK(call)                                                   \
    do {                                                           \
        cudaError_t _e = (call);                                   \
        if (_e != cudaSuccess) return cuda_fail(ctx, _e, #call);   \
    } while (0)

void free_dev(Dev &d) {
    cudaSetDevice(d.device);
    if (d.rgb) cudaFree(d.rgb);
    if (d.stars) cudaFree(d.stars);
    if (d.star_list) cudaFree(d.star_list);
    if (d.argb) cudaFree(d.argb);
    if (d.shadow) cudaFree(d.shadow);
    if (d.counters) cudaFree(d.counters);
    if (d.abort_flag) cudaFree(d.abort_flag);
    if (d.progress) cudaFree(d.progress);
    if (d.wtab) cudaFree(d.wtab);
    if (d.progress_host) cudaFreeHost(d.progress_host);
    for (auto &e : d.ev) if (e) cudaEventDestroy(e);
    if (d.st) cudaStreamDestroy(d.st);
    if (d.ctrl) cudaStreamDestroy(d.ctrl);
    d = Dev();
}

// rows of a `height`-row frame that device `which` of `n` owns (interleaved MD_BAND_ROWS bands)
int rows_of_device(int height, int which, int n) {
    if (n == 1) return height;
    int rows = 0;
    for (int b = which; b * MD_BAND_ROWS < height; b += n) {
        int r0 = b * MD_BAND_ROWS, r1 = r0 + MD_BAND_ROWS;
        rows += (r1 < height ? r1 : height) - r0;
    }
    return rows;
}

int ensure_capacity(gamer_ctx *ctx, Dev &d, size_t px) {
    if (px <= d.cap_px) return GAMER_OK;
    CK(cudaSetDevice(d.device));
    if (d.rgb) cudaFree(d.rgb);
    if (d.argb) cudaFree(d.argb);
    if (d.shadow) cudaFree(d.shadow);
    d.rgb = nullptr; d.argb = d.shadow = nullptr; d.cap_px = 0;
    CK(cudaMalloc(&d.rgb, px * 3 * sizeof(float)));
    CK(cudaMalloc(&d.argb, px * sizeof(uint32_t)));
    CK(cudaMalloc(&d.shadow, px * sizeof(uint32_t)));
    d.cap_px = px;
    return GAMER_OK;
}

// launch K1 on every device of the context (no host sync)
int launch_render(gamer_ctx *ctx, bool counted) {
    for (Dev &d : ctx->devs) {
        if (d.rows <= 0) continue;
        CK(cudaSetDevice(d.device));
        CK(cudaMemsetAsync(d.progress, 0, 2 * sizeof(unsigned int), d.st));
        if (counted) CK(cudaMemsetAsync(d.counters, 0, GAMER_CNT_N * sizeof(unsigned long long), d.st));
        if (!(d.scene.flags & GAMER_FRAME_CONTINUE))
            CK(cudaMemsetAsync(d.rgb, 0, (size_t)d.rows * d.scene.width * 3 * sizeof(float), d.st));
        CK(cudaEventRecord(d.ev[0], d.st));
        CK(gk::launch_k1(d.scene, d.rgb, d.counters, d.abort_flag, d.progress, counted, d.sm_count, d.st, &d.n_blocks));
        CK(cudaEventRecord(d.ev[1], d.st));
        ctx->launches += 1;
    }
    return GAMER_OK;
}

int sync_all(gamer_ctx *ctx) {
    for (Dev &d : ctx->devs) {
        CK(cudaSetDevice(d.device));
        CK(cudaStreamSynchronize(d.st));
    }
    return GAMER_OK;
}

int finish_render(gamer_ctx *ctx, bool counted) {
    int rc = sync_all(ctx);
    if (rc) return rc;
    Dev &d0 = ctx->devs[0];
    CK(cudaSetDevice(d0.device));
    if (d0.rows > 0 && !ctx->skipped) CK(cudaEventElapsedTime(&ctx->k1_ms, d0.ev[0], d0.ev[1]));
    if (counted && !ctx->skipped) {
        for (auto &v : ctx->counters) v = 0;
        for (Dev &d : ctx->devs) {
            if (d.rows <= 0) continue;
            unsigned long long h[GAMER_CNT_N];
            CK(cudaSetDevice(d.device));
            CK(cudaMemcpy(h, d.counters, sizeof(h), cudaMemcpyDeviceToHost));
            for (int q = 0; q < GAMER_CNT_N; q++) ctx->counters[q] += h[q];
        }
    }
    ctx->counted = counted;
    ctx->rendered = true;
    // Rasterizer::Abort sets m_abort at any time and RenderPixelsOMP resets it when it ends (rasterizer.cpp:296-297,311):
    // a request cancels the render in flight or, made while idle, the next one, and is consumed by that render. Whether
    // THIS render was cut short is read from the device (tiles finished < tiles), not from the request: an abort that
    // lands after the last tile leaves a complete frame, which is delivered.
    if (ctx->abort_req.exchange(0) || ctx->skipped) {
        bool cut = ctx->skipped;
        for (Dev &d : ctx->devs) {
            if (d.rows <= 0) continue;
            CK(cudaSetDevice(d.device));
            if (!ctx->skipped) {
                unsigned int done = 0;
                CK(cudaMemcpy(&done, d.progress, sizeof(done), cudaMemcpyDeviceToHost));
                if (done < d.n_blocks) cut = true;
            }
            CK(cudaMemset(d.abort_flag, 0, sizeof(int)));      // only here: never in the render stream of a live render
        }
        ctx->skipped = false;
        if (cut) return set_err(ctx, GAMER_E_ABORTED, "render aborted");
    }
    return GAMER_OK;
}

// copy every device's bands into a full-frame host array with `elem` bytes per pixel
int gather_to_host(gamer_ctx *ctx, char *host, size_t elem, int which_buf) {
    const int W = ctx->frame.width, H = ctx->frame.height, n = (int)ctx->devs.size();
    for (int di = 0; di < n; di++) {
        Dev &d = ctx->devs[di];
        if (d.rows <= 0) continue;
        CK(cudaSetDevice(d.device));
        const char *src = which_buf == 0 ? (const char *)d.rgb : which_buf == 1 ? (const char *)d.argb
                        : which_buf == 2 ? (const char *)d.shadow : (const char *)d.stars;
        if (n == 1) {
            CK(cudaMemcpyAsync(host, src, (size_t)W * H * elem, cudaMemcpyDeviceToHost, d.st));
            continue;
        }
        int local = 0;
        for (int b = di; b * MD_BAND_ROWS < H; b += n) {
            int r0 = b * MD_BAND_ROWS, r1 = r0 + MD_BAND_ROWS;
            if (r1 > H) r1 = H;
            CK(cudaMemcpyAsync(host + (size_t)r0 * W * elem, src + (size_t)local * W * elem, (size_t)(r1 - r0) * W * elem,
                               cudaMemcpyDeviceToHost, d.st));
            local += r1 - r0;
        }
    }
    return sync_all(ctx);
}

}  // namespace

extern "C" {

int gamer_b200_abi_version(void) { return GAMER_B200_ABI_VERSION; }

int gamer_b200_device_count(int32_t *count) {
    if (!count) return GAMER_E_ARG;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { *count = 0; return cuda_fail(nullptr, e, "cudaGetDeviceCount"); }
    *count = n;
    return GAMER_OK;
}

const char *gamer_b200_last_error(gamer_ctx *ctx) {
    if (ctx) return ctx->err.c_str();
    return g_create_error.c_str();
}

void gamer_b200_destroy(gamer_ctx *ctx) {
    if (!ctx) return;
    for (Dev &d : ctx->devs) free_dev(d);
    delete ctx;
}

int gamer_b200_create(gamer_ctx **out, const int32_t *device_ids, int32_t n_devices) {
    if (!out) return GAMER_E_ARG;
    *out = nullptr;
    gamer_ctx *ctx = nullptr;          // CK() reports into g_create_error while ctx is null
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (ndev <= 0) return set_err(nullptr, GAMER_E_CUDA, "no CUDA device: this library has no CPU fallback");
    int32_t dflt = 0;
    if (!device_ids || n_devices <= 0) { device_ids = &dflt; n_devices = 1; }
    for (int i = 0; i < n_devices; i++)
        if (device_ids[i] < 0 || device_ids[i] >= ndev) return set_err(nullptr, GAMER_E_ARG, "device id out of range");
    gamer_ctx *c = new (std::nothrow) gamer_ctx();
    if (!c) return set_err(nullptr, GAMER_E_ARG, "out of host memory");
    c->devs.resize(n_devices);
    for (int i = 0; i < n_devices; i++) {
        Dev &d = c->devs[i];
        d.device = device_ids[i];
        cudaError_t e = cudaSetDevice(d.device);
        cudaDeviceProp prop;
        if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, d.device);
        if (e == cudaSuccess && prop.major != 10) {
            gamer_b200_destroy(c);
            return set_err(nullptr, GAMER_E_CUDA, "device is not sm_100 (B200): kernels are built for sm_100a only");
        }
        if (e == cudaSuccess) e = gk::prepare_device();
        if (e == cudaSuccess) { d.sm_count = prop.multiProcessorCount; e = cudaStreamCreateWithFlags(&d.st, cudaStreamNonBlocking); }
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d.ctrl, cudaStreamNonBlocking);
        for (int k = 0; k < 4 && e == cudaSuccess; k++) e = cudaEventCreate(&d.ev[k]);
        if (e == cudaSuccess) e = cudaMalloc(&d.counters, GAMER_CNT_N * sizeof(unsigned long long));
        if (e == cudaSuccess) e = cudaMalloc(&d.abort_flag, sizeof(int));
        if (e == cudaSuccess) e = cudaMalloc(&d.progress, 2 * sizeof(unsigned int));
        if (e == cudaSuccess) e = cudaMalloc(&d.wtab, (size_t)DEV_MAX_GALAXIES * WT_DOUBLES_PER_GALAXY * sizeof(double));
        if (e == cudaSuccess) e = cudaHostAlloc(&d.progress_host, sizeof(unsigned int), cudaHostAllocDefault);
        if (e == cudaSuccess) e = cudaMemset(d.abort_flag, 0, sizeof(int));
        if (e == cudaSuccess) e = cudaMemset(d.progress, 0, 2 * sizeof(unsigned int));
        if (e == cudaErrorNotSupported) {
            gamer_b200_destroy(c);
            return set_err(nullptr, GAMER_E_CUDA, "the dynamic shared-memory window does not start at the offset the noise tables are "
                                                  "addressed with (a debugger, sanitizer or driver that reserves a different amount)");
        }
        if (e != cudaSuccess) {
            gamer_b200_destroy(c);
            return cuda_fail(nullptr, e, "context creation");
        }
    }
    *out = c;
    return GAMER_OK;
}

int gamer_b200_upload(gamer_ctx *ctx, const GamerFrame *frame, const GamerGalaxy *galaxies, const GamerComponent *components) {
    if (!ctx || !frame) return GAMER_E_ARG;
    std::lock_guard<std::mutex> lock(ctx->mu);
    const int n = (int)ctx->devs.size();
    if (n > 1 && frame->band_rows != 0) return set_err(ctx, GAMER_E_ARG, "a multi-device context splits bands itself: band_rows must be 0");
    if (frame->flags & GAMER_FRAME_CONTINUE) {     // a later pass of a multi-pass frame: same geometry, previous pass left raw I
        const GamerFrame &p = ctx->frame;
        if (!ctx->rendered || !(p.flags & GAMER_FRAME_PARTIAL) || p.width != frame->width || p.height != frame->height ||
            p.full_size != frame->full_size || p.x0 != frame->x0 || p.y0 != frame->y0 || p.band_rows != frame->band_rows ||
            p.band_stride != frame->band_stride)
            return set_err(ctx, GAMER_E_STATE, "GAMER_FRAME_CONTINUE needs a rendered GAMER_FRAME_PARTIAL pass of the same frame geometry");
    }
    char msg[256] = {0};
    for (int di = 0; di < n; di++) {
        Dev &d = ctx->devs[di];
        GamerFrame f = *frame;
        d.rows = rows_of_device(frame->height, di, n);
        if (n > 1) {
            f.y0 = frame->y0 + di * MD_BAND_ROWS;
            f.band_rows = MD_BAND_ROWS;
            f.band_stride = MD_BAND_ROWS * n;
            f.height = d.rows > 0 ? d.rows : 1;
        }
        std::vector<double> wtab_host;
        int rc = gk_build_scene(&f, galaxies, components, &d.scene, msg, sizeof(msg), &wtab_host);
        if (rc) return set_err(ctx, rc, msg);
        d.scene.wtab = d.wtab;
        if (!wtab_host.empty()) {      // pageable source: staged before the call returns; ordered before the next kernel
            CK(cudaSetDevice(d.device));
            CK(cudaMemcpyAsync(d.wtab, wtab_host.data(), wtab_host.size() * sizeof(double), cudaMemcpyHostToDevice, d.st));
        }
        if (d.rows > 0) {
            rc = ensure_capacity(ctx, d, (size_t)d.rows * frame->width);
            if (rc) return rc;
        }
    }
    ctx->frame = *frame;
    for (Dev &d : ctx->devs) d.layer_valid = false;      // the star layer belongs to one uploaded frame
    ctx->uploaded = true;
    ctx->rendered = false;
    return GAMER_OK;
}

static int render_common(gamer_ctx *ctx, bool counted, float *host_out) {
    if (!ctx) return GAMER_E_ARG;
    std::lock_guard<std::mutex> lock(ctx->mu);
    if (!ctx->uploaded) return set_err(ctx, GAMER_E_STATE, "render before upload");
    ctx->rendering.store(1);           // before the request is looked at: an abort() that misses the look sees rendering == 1
    int rc = GAMER_OK;
    if (ctx->abort_req.load()) {       // requested before this render began: nothing is launched (every pixel "skipped")
        ctx->skipped = true;
        for (Dev &d : ctx->devs) {
            if (d.rows <= 0 || (d.scene.flags & GAMER_FRAME_CONTINUE)) continue;
            if (cudaSetDevice(d.device) == cudaSuccess)
                cudaMemsetAsync(d.rgb, 0, (size_t)d.rows * d.scene.width * 3 * sizeof(float), d.st);
        }
    } else rc = launch_render(ctx, counted);
    if (!rc) rc = finish_render(ctx, counted);
    ctx->rendering.store(0);
    if (rc) return rc;
    if (host_out) return gather_to_host(ctx, (char *)host_out, 3 * sizeof(float), 0);
    return GAMER_OK;
}

int gamer_b200_render(gamer_ctx *ctx, float *rgb_linear_out) {
    if (!rgb_linear_out) return GAMER_E_ARG;
    return render_common(ctx, false, rgb_linear_out);
}
int gamer_b200_render_device(gamer_ctx *ctx) { return render_common(ctx, false, nullptr); }
int gamer_b200_render_counted(gamer_ctx *ctx) { return render_common(ctx, true, nullptr); }

int gamer_b200_counters(gamer_ctx *ctx, uint64_t out[GAMER_CNT_N]) {
    if (!ctx || !out) return GAMER_E_ARG;
    if (!ctx->rendered || !ctx->counted) return set_err(ctx, GAMER_E_STATE, "counters need a preceding render_counted");
    for (int q = 0; q < GAMER_CNT_N; q++) out[q] = ctx->counters[q];
    return GAMER_OK;
}

static int tonemap_common(gamer_ctx *ctx, double exposure, double gamma, double saturation, const float *stars_host,
                          uint32_t *argb_out, uint32_t *shadow_out) {
    if (!ctx) return GAMER_E_ARG;
    std::lock_guard<std::mutex> lock(ctx->mu);
    if (!ctx->rendered) return set_err(ctx, GAMER_E_STATE, "tonemap before render");
    if (!(exposure != 0.0)) return set_err(ctx, GAMER_E_ARG, "exposure must be non-zero");
    const int W = ctx->frame.width, H = ctx->frame.height, n = (int)ctx->devs.size();
    for (int di = 0; di < n; di++) {
        Dev &d = ctx->devs[di];
        if (d.rows <= 0) continue;
        CK(cudaSetDevice(d.device));
        size_t px = (size_t)d.rows * W;
        const float *stars_dev = nullptr;
        if (stars_host == GAMER_STARS_LAYER) {
            if (!d.layer_valid) return set_err(ctx, GAMER_E_STATE, "GAMER_STARS_LAYER without a preceding render_stars for this frame");
            stars_dev = d.stars;
        } else if (stars_host) {
            d.layer_valid = false;
            if (d.cap_stars < px) {
                if (d.stars) cudaFree(d.stars);
                d.stars = nullptr; d.cap_stars = 0;
                CK(cudaMalloc(&d.stars, px * 3 * sizeof(float)));
                d.cap_stars = px;
            }
            if (n == 1) CK(cudaMemcpyAsync(d.stars, stars_host, px * 3 * sizeof(float), cudaMemcpyHostToDevice, d.st));
            else {
                int local = 0;
                for (int b = di; b * MD_BAND_ROWS < H; b += n) {
                    int r0 = b * MD_BAND_ROWS, r1 = r0 + MD_BAND_ROWS;
                    if (r1 > H) r1 = H;
                    CK(cudaMemcpyAsync(d.stars + (size_t)local * W * 3, stars_host + (size_t)r0 * W * 3,
                                       (size_t)(r1 - r0) * W * 3 * sizeof(float), cudaMemcpyHostToDevice, d.st));
                    local += r1 - r0;
                }
            }
            stars_dev = d.stars;
        }
        CK(cudaEventRecord(d.ev[2], d.st));
        int k2_launches = 1;          // 2 when the pixel count leaves a remainder modulo 4 (the scalar kernel takes it)
        CK(gk::launch_k2(d.rgb, stars_dev, d.argb, d.shadow, px, exposure, gamma, saturation, d.sm_count, d.st, &k2_launches));
        CK(cudaEventRecord(d.ev[3], d.st));
        ctx->launches += k2_launches;
    }
    int rc = sync_all(ctx);
    if (rc) return rc;
    Dev &d0 = ctx->devs[0];
    CK(cudaSetDevice(d0.device));
    if (d0.rows > 0) CK(cudaEventElapsedTime(&ctx->k2_ms, d0.ev[2], d0.ev[3]));
    if (argb_out) { rc = gather_to_host(ctx, (char *)argb_out, sizeof(uint32_t), 1); if (rc) return rc; }
    if (shadow_out) { rc = gather_to_host(ctx, (char *)shadow_out, sizeof(uint32_t), 2); if (rc) return rc; }
    return GAMER_OK;
}

int gamer_b200_tonemap(gamer_ctx *ctx, double exposure, double gamma, double saturation, const float *stars_rgb_or_null,
                       uint32_t *argb_out, uint32_t *argb_shadow_out) {
    return tonemap_common(ctx, exposure, gamma, saturation, stars_rgb_or_null, argb_out, argb_shadow_out);
}
int gamer_b200_tonemap_device(gamer_ctx *ctx, double exposure, double gamma, double saturation) {
    return tonemap_common(ctx, exposure, gamma, saturation, nullptr, nullptr, nullptr);
}

// Rasterizer::RenderStars (rasterizer.cpp:215-225): K4 stamps the list into the device star layer.
int gamer_b200_render_stars(gamer_ctx *ctx, const GamerStar *stars, int32_t n_stars, float *stars_rgb_out_or_null) {
    if (!ctx || n_stars < 0 || (n_stars > 0 && !stars)) return GAMER_E_ARG;
    std::lock_guard<std::mutex> lock(ctx->mu);
    if (!ctx->uploaded) return set_err(ctx, GAMER_E_STATE, "render_stars before upload");
    const int W = ctx->frame.width;
    for (Dev &d : ctx->devs) {
        d.layer_valid = false;
        if (d.rows <= 0) continue;
        CK(cudaSetDevice(d.device));
        const size_t px = (size_t)d.rows * W;
        if (d.cap_stars < px) {
            if (d.stars) cudaFree(d.stars);
            d.stars = nullptr; d.cap_stars = 0;
            CK(cudaMalloc(&d.stars, px * 3 * sizeof(float)));
            d.cap_stars = px;
        }
        if (d.cap_star_list < (size_t)n_stars) {
            if (d.star_list) cudaFree(d.star_list);
            d.star_list = nullptr; d.cap_star_list = 0;
            CK(cudaMalloc(&d.star_list, (size_t)n_stars * sizeof(GamerStar)));
            d.cap_star_list = (size_t)n_stars;
        }
        CK(cudaMemsetAsync(d.stars, 0, px * 3 * sizeof(float), d.st));                     // fill(QVector3D(0,0,0))
        if (n_stars > 0) {
            CK(cudaMemcpyAsync(d.star_list, stars, (size_t)n_stars * sizeof(GamerStar), cudaMemcpyHostToDevice, d.st));
            CK(gk::launch_k4(d.star_list, n_stars, d.stars, d.scene, d.st));
            ctx->launches += 1;
        }
        d.layer_valid = true;
    }
    int rc = sync_all(ctx);
    if (rc) return rc;
    if (stars_rgb_out_or_null) { rc = gather_to_host(ctx, (char *)stars_rgb_out_or_null, 3 * sizeof(float), 3); if (rc) return rc; }
    return GAMER_OK;
}

// Rasterizer::getPercentDone (rasterizer.h:112-118): fraction of pixel blocks finished. Any thread.
int gamer_b200_progress(gamer_ctx *ctx, double *fraction) {
    if (!ctx || !fraction) return GAMER_E_ARG;
    if (!ctx->rendering.load()) { *fraction = ctx->rendered ? 1.0 : 0.0; return GAMER_OK; }
    double done = 0, total = 0;
    for (Dev &d : ctx->devs) {
        if (d.rows <= 0 || d.n_blocks == 0) continue;
        if (cudaSetDevice(d.device) != cudaSuccess) continue;
        if (cudaMemcpyAsync(d.progress_host, d.progress, sizeof(unsigned int), cudaMemcpyDeviceToHost, d.ctrl) != cudaSuccess) continue;
        if (cudaStreamSynchronize(d.ctrl) != cudaSuccess) continue;
        done += *d.progress_host;
        total += d.n_blocks;
    }
    *fraction = total > 0 ? done / total : 0.0;
    return GAMER_OK;
}

// Rasterizer::Abort (rasterizer.h:91-98): blocks that have not started skip their pixels. Any thread.
int gamer_b200_abort(gamer_ctx *ctx) {
    if (!ctx) return GAMER_E_ARG;
    ctx->abort_req.store(1);           // sticky: consumed by the render in flight or, if none is, by the next one
    if (!ctx->rendering.load()) return GAMER_OK;
    static const int one = 1;
    for (Dev &d : ctx->devs) {
        if (cudaSetDevice(d.device) != cudaSuccess) continue;
        cudaMemcpyAsync(d.abort_flag, &one, sizeof(int), cudaMemcpyHostToDevice, d.ctrl);
        cudaStreamSynchronize(d.ctrl);
    }
    return GAMER_OK;
}

int gamer_b200_device_buffers(gamer_ctx *ctx, int32_t which, void **rgb_dev, void **argb_dev, void **argb_shadow_dev,
                              int32_t *rows_out) {
    if (!ctx || which < 0 || which >= (int)ctx->devs.size()) return GAMER_E_ARG;
    Dev &d = ctx->devs[which];
    if (rgb_dev) *rgb_dev = d.rgb;
    if (argb_dev) *argb_dev = d.argb;
    if (argb_shadow_dev) *argb_shadow_dev = d.shadow;
    if (rows_out) *rows_out = d.rows;
    return GAMER_OK;
}

// This context's tile, copied to its place in a full_size x full_size HOST image (which_buf: 0 float RGB, 1 ARGB32,
// 2 shadow ARGB32). One process per GPU, each with the bands of one frame (GamerFrame x0/y0/band_rows/band_stride),
// writes them into ONE host frame the processes share (shared memory, ideally cudaHostRegister'ed): the final gather of
// the tile partition without a device-side collective and without a serial copy by one rank.
int gamer_b200_copy_bands(gamer_ctx *ctx, int32_t which_buf, void *host_full_frame) {
    if (!ctx || !host_full_frame || which_buf < 0 || which_buf > 2) return GAMER_E_ARG;
    std::lock_guard<std::mutex> lock(ctx->mu);
    if (!ctx->rendered) return set_err(ctx, GAMER_E_STATE, "copy_bands before render");
    if (ctx->devs.size() != 1) return set_err(ctx, GAMER_E_STATE, "copy_bands is for single-device contexts (a multi-device context gathers by itself)");
    Dev &d = ctx->devs[0];
    const GamerFrame &f = ctx->frame;
    const size_t elem = which_buf == 0 ? 3 * sizeof(float) : sizeof(uint32_t);
    const char *src = which_buf == 0 ? (const char *)d.rgb : which_buf == 1 ? (const char *)d.argb : (const char *)d.shadow;
    if (f.x0 < 0 || f.y0 < 0 || f.x0 + f.width > f.full_size) return set_err(ctx, GAMER_E_ARG, "tile outside the full image");
    CK(cudaSetDevice(d.device));
    const int run = f.band_rows > 0 ? f.band_rows : f.height;          // local rows that are contiguous in the image
    const size_t src_pitch = (size_t)f.width * elem, dst_pitch = (size_t)f.full_size * elem;
    for (int L = 0; L < f.height; L += run) {
        const int rows = (L + run <= f.height) ? run : f.height - L;
        const int y = f.band_rows > 0 ? f.y0 + (L / f.band_rows) * f.band_stride : f.y0 + L;
        if (y + rows > f.full_size) return set_err(ctx, GAMER_E_ARG, "tile outside the full image");
        char *dst = (char *)host_full_frame + ((size_t)y * f.full_size + f.x0) * elem;
        CK(cudaMemcpy2DAsync(dst, dst_pitch, src + (size_t)L * src_pitch, src_pitch, src_pitch, rows, cudaMemcpyDeviceToHost, d.st));
    }
    CK(cudaStreamSynchronize(d.st));
    return GAMER_OK;
}

int gamer_b200_stream(gamer_ctx *ctx, int32_t which, void **stream_out) {
    if (!ctx || !stream_out || which < 0 || which >= (int)ctx->devs.size()) return GAMER_E_ARG;
    *stream_out = (void *)ctx->devs[which].st;
    return GAMER_OK;
}

int gamer_b200_last_kernel_ms(gamer_ctx *ctx, float *k1_ms, float *k2_ms) {
    if (!ctx) return GAMER_E_ARG;
    if (k1_ms) *k1_ms = ctx->k1_ms;
    if (k2_ms) *k2_ms = ctx->k2_ms;
    return GAMER_OK;
}

int gamer_b200_launch_count(gamer_ctx *ctx, uint64_t *launches) {
    if (!ctx || !launches) return GAMER_E_ARG;
    *launches = ctx->launches;
    return GAMER_OK;
}

int gamer_b200_host_alloc(void **ptr, uint64_t bytes) {
    if (!ptr) return GAMER_E_ARG;
    cudaError_t e = cudaHostAlloc(ptr, bytes, cudaHostAllocDefault);
    if (e != cudaSuccess) { *ptr = nullptr; return cuda_fail(nullptr, e, "cudaHostAlloc"); }
    return GAMER_OK;
}
int gamer_b200_host_free(void *ptr) {
    if (!ptr) return GAMER_OK;
    return cudaFreeHost(ptr) == cudaSuccess ? GAMER_OK : GAMER_E_CUDA;
}

// ---- probes ------------------------------------------------------------------------------------
}  // extern "C"
namespace {
struct DevBuf {                    // freed on every path out of a probe
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    template <class T> T *as() const { return static_cast<T *>(p); }
};
int probe_io(gamer_ctx *ctx, const float *xyz, int32_t n, DevBuf &xyz_dev, DevBuf &val_dev) {
    Dev &d = ctx->devs[0];
    CK(cudaSetDevice(d.device));
    CK(cudaMalloc(&xyz_dev.p, (size_t)n * 3 * sizeof(float)));
    CK(cudaMalloc(&val_dev.p, (size_t)n * sizeof(float)));
    CK(cudaMemcpyAsync(xyz_dev.p, xyz, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice, d.st));
    return GAMER_OK;
}
}  // namespace
extern "C" {

int gamer_b200_probe_simplex(gamer_ctx *ctx, const float *xyz, int32_t n, double frequency, int32_t *lattice_out, float *value_out) {
    if (!ctx || !xyz || n <= 0) return GAMER_E_ARG;
    std::lock_guard<std::mutex> lock(ctx->mu);
    Dev &d = ctx->devs[0];
    DevBuf xd, vd, ld;
    int rc = probe_io(ctx, xyz, n, xd, vd);
    if (rc) return rc;
    CK(cudaMalloc(&ld.p, (size_t)n * 10 * sizeof(int32_t)));
    CK(gk::launch_probe_simplex(xd.as<float>(), n, frequency, ld.as<int32_t>(), vd.as<float>(), d.st));
    ctx->launches += 1;
    if (value_out) CK(cudaMemcpyAsync(value_out, vd.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, d.st));
    if (lattice_out) CK(cudaMemcpyAsync(lattice_out, ld.p, (size_t)n * 10 * sizeof(int32_t), cudaMemcpyDeviceToHost, d.st));
    CK(cudaStreamSynchronize(d.st));
    return GAMER_OK;
}

int gamer_b200_probe_fbm(gamer_ctx *ctx, const float *xyz, int32_t n, int32_t octaves, double persistence, double scale, float *value_out) {
    if (!ctx || !xyz || !value_out || n <= 0 || octaves <= 0 || octaves > 32) return GAMER_E_ARG;
    std::lock_guard<std::mutex> lock(ctx->mu);
    Dev &d = ctx->devs[0];
    DevBuf xd, vd;
    int rc = probe_io(ctx, xyz, n, xd, vd);
    if (rc) return rc;
    double amp = 1, maxamp = 0;
    for (int o = 0; o < octaves; o++) { maxamp += amp; amp *= persistence; }
    CK(gk::launch_probe_fbm(xd.as<float>(), n, octaves, (float)persistence, (float)(1.0 / maxamp), scale, vd.as<float>(), d.st));
    ctx->launches += 1;
    CK(cudaMemcpyAsync(value_out, vd.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, d.st));
    CK(cudaStreamSynchronize(d.st));
    return GAMER_OK;
}

int gamer_b200_probe_ridged(gamer_ctx *ctx, const float *xyz, int32_t n, double frequency, int32_t octaves, double lacunarity,
                            double offset, double gain, float *value_out) {
    if (!ctx || !xyz || !value_out || n <= 0 || octaves <= 0 || octaves > 16) return GAMER_E_ARG;
    std::lock_guard<std::mutex> lock(ctx->mu);
    Dev &d = ctx->devs[0];
    DevBuf xd, vd;
    int rc = probe_io(ctx, xyz, n, xd, vd);
    if (rc) return rc;
    float rw[16];
    const double w = -0.05f;
    for (int o = 0; o < octaves; o++) { rw[o] = (float)pow(frequency, w); frequency *= lacunarity; }
    CK(gk::launch_probe_ridged(xd.as<float>(), n, octaves, (float)lacunarity, (float)offset, (float)gain, rw, vd.as<float>(), d.st));
    ctx->launches += 1;
    CK(cudaMemcpyAsync(value_out, vd.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, d.st));
    CK(cudaStreamSynchronize(d.st));
    return GAMER_OK;
}

// ---- allocation-free variant of the fBm probe (include/gamer_b200.h) ------------------------------------------------
int gamer_b200_fbm_static_max_points(void) { return gk::fbm_static_max_points(); }

int gamer_b200_fbm_static(const float *xyz, int32_t n, int32_t octaves, double persistence, double scale, float *value_out) {
    if (!xyz || !value_out || n <= 0 || octaves <= 0 || octaves > 32)
        return set_err(nullptr, GAMER_E_ARG, "gamer_b200_fbm_static: null pointer, n <= 0 or octaves outside 1..32");
    if (n > gk::fbm_static_max_points())
        return set_err(nullptr, GAMER_E_ARG, "gamer_b200_fbm_static: n = " + std::to_string(n) + " points exceed the static scratch (" +
                       std::to_string(gk::fbm_static_max_points()) + " points per call): split the call");
    double amp = 1, maxamp = 0;
    for (int o = 0; o < octaves; o++) { maxamp += amp; amp *= persistence; }
    cudaError_t e = gk::fbm_static(xyz, n, octaves, (float)persistence, (float)(1.0 / maxamp), scale, value_out);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "gamer_b200_fbm_static");
    return GAMER_OK;
}

}  // extern "C"
