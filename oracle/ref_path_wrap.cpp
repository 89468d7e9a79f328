// TEST INFRASTRUCTURE — a C door onto the reference's OWN render path, compiled in place.
// `make ref` builds oracle/_ref/libgamer_ref_path.so from the reference's unmodified sources where they lie
// under $(REF) (rasterizer.cpp, galaxycomponent(s).cpp, galaxy.cpp, galaxyinstance.cpp, gamercamera.cpp,
// renderingparams.cpp, buffer2d.cpp, util.cpp, noise/*.cpp ...) against the Qt stand-ins of oracle/qtshim
// (value-class arithmetic = oracle/qt5_math.h's restatement of Qt 5.15; QThread::start runs synchronously;
// files, streams, images and messages are small host stand-ins). Nothing of the reference is copied here:
// this file only CALLS its public API in the order consolerenderer.cpp:60-118 does.
// Only tests/golden/make_fixtures.py and tests/ load this library; it exists where /root/reference exists.
#include <cstdint>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "source/galaxy/rasterizer.h"
#include "source/galaxy/galaxycomponent.h"
#include "source/util/util.h"

namespace {
struct RefPath {
    RenderingParams rp;
    Rasterizer *ras = nullptr;
};
QVector3D v3(const float *p) { return QVector3D(p[0], p[1], p[2]); }
}  // namespace

extern "C" {

void *refpath_create(int size, double ray_step, const float *cam, const float *target, const float *up, double fov) {
    RefPath *h = new RefPath;
    h->rp.spectra().PopulateSpectra();                 // consolerenderer.cpp:66
    h->rp.setSize(size);
    h->rp.setRayStep(ray_step);
    h->rp.camera().setCamera(v3(cam));
    h->rp.camera().setTarget(v3(target));
    h->rp.camera().setUp(v3(up));
    h->rp.camera().setPerspective(fov);
    h->ras = new Rasterizer(&h->rp);
    return h;
}

int refpath_add_galaxy(void *hp, const char *gax_path, const float *pos, const float *orient, double iscale, double redshift) {
    RefPath *h = static_cast<RefPath *>(hp);
    Galaxy *g = new Galaxy();
    if (!g->Load(QString(gax_path))) { delete g; return 1; }
    h->ras->AddGalaxy(new GalaxyInstance(g, QString(gax_path), v3(pos), v3(orient), iscale, redshift));
    return 0;
}

// Render() = Prepare + RenderPixels on the (synchronous) thread; preview = RenderDirect (min step 0.01).
int refpath_render(void *hp, int preview, float *rgb_out) {
    RefPath *h = static_cast<RefPath *>(hp);
    if (preview) h->ras->RenderDirect(); else h->ras->Render();
    Buffer2D *b = h->ras->getRenderBuffer();
    if (!b) return 1;
    const int n = h->rp.size();
    for (int j = 0; j < n; j++)
        for (int i = 0; i < n; i++) {
            QVector3D v = b->Get(i, j);
            float *o = rgb_out + (size_t(j) * n + i) * 3;
            o[0] = v.x(); o[1] = v.y(); o[2] = v.z();
        }
    return 0;
}

// Timing door (bench.py --impl reference, cpu_baseline): rows row0, row0+stride, ... (n_rows of them) of the
// frame through the reference's own per-pixel code — Rasterizer::Prepare once, then setupCamera + renderPixel per
// pixel, OpenMP over pixels like RenderPixelsOMP (rasterizer.cpp:283-313) when built with -fopenmp.
// prepare != 0 runs Prepare() (+ one warm-up pixel) first; a caller timing the pixel loop passes it once, untimed.
// out = n_rows * size * 3 floats. Returns the number of threads used.
int refpath_render_rows(void *hp, int row0, int stride, int n_rows, int threads, int prepare, float *out) {
    RefPath *h = static_cast<RefPath *>(hp);
    const int n = h->rp.size();
    if (prepare) {
        h->ras->Prepare();
        // one pixel on this thread first: the reference allocates some tables lazily (noise.cpp:88-89)
        RasterPixel warm = h->ras->renderPixel(h->ras->setupCamera((n / 2) * n + n / 2), h->ras->getGalaxies());
        (void)warm;
    }
    int used = 1;
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
    used = threads;
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads)
#endif
    for (int k = 0; k < n_rows * n; k++) {
        const int j = row0 + (k / n) * stride, i = k % n;
        RasterPixel p = h->ras->renderPixel(h->ras->setupCamera(i + j * n), h->ras->getGalaxies());
        out[size_t(k) * 3 + 0] = p.I().x();
        out[size_t(k) * 3 + 1] = p.I().y();
        out[size_t(k) * 3 + 2] = p.I().z();
    }
    return used;
}

// One ray through Rasterizer::setupCamera + renderPixel (after a render has run Prepare()).
void refpath_pixel(void *hp, int idx, float *rgb) {
    RefPath *h = static_cast<RefPath *>(hp);
    QVector3D dir = h->ras->setupCamera(idx);
    RasterPixel p = h->ras->renderPixel(dir, h->ras->getGalaxies());
    rgb[0] = p.I().x(); rgb[1] = p.I().y(); rgb[2] = p.I().z();
}

void refpath_ray(void *hp, int idx, float *dir) {
    RefPath *h = static_cast<RefPath *>(hp);
    QVector3D d = h->ras->setupCamera(idx);
    dir[0] = d.x(); dir[1] = d.y(); dir[2] = d.z();
}

// Debug aid: the march of one pixel driven from here, calling the reference's component methods sample by
// sample, recording (p.xyz, step, I.xyz) after every sample. Returns the number of samples of galaxy `gal`.
int refpath_trace(void *hp, int idx, int gal, int max_samples, float *out7) {
    RefPath *h = static_cast<RefPath *>(hp);
    QVector3D dir = h->ras->setupCamera(idx);
    GalaxyInstance *gi = h->ras->getGalaxies()[gal];
    Galaxy *g = gi->GetGalaxy();
    QVector3D isp1, isp2;
    double t1, t2;
    QVector3D cam = h->rp.camera().camera() - gi->position();
    bool hit = Util::IntersectSphere(cam, dir, g->galaxyParams().axis(), isp1, isp2, t1, t2);
    if (t2 > 0) isp2 = cam;
    if (t1 > 0 && t2 > 0) hit = false;
    if (!hit) return 0;
    RasterPixel rp;
    double length = (isp1 - isp2).length();
    QVector3D d = (isp1 - isp2).normalized(), origin = isp1, p = origin, ll = (isp2 - origin).normalized();
    double step = h->rp.rayStep();
    rp.scale = step;
    int n = 0;
    while (QVector3D::dotProduct(p - origin, ll) < length + step) {
        step = Util::clamp(pow((p - cam).length(), 1) * h->rp.rayStep(), 0.001, 0.01);
        rp.step = step;
        for (GalaxyComponent *gc : g->components())
            if (gc->getComponentParams().active() == 1) {
                rp.radius = gc->getRadius(p, rp.P, rp.z, gi);
                rp.z = gc->getHeightModulation(rp.z);
                gc->calculateIntensity(&rp, p, gi, step * 200.0);
            }
        if (n < max_samples) {
            float *o = out7 + size_t(n) * 7;
            o[0] = p.x(); o[1] = p.y(); o[2] = p.z(); o[3] = float(step);
            o[4] = rp.I().x(); o[5] = rp.I().y(); o[6] = rp.I().z();
        }
        p = p - d * step;
        n++;
        rp.setI(Util::floor(rp.I()));
    }
    return n;
}

int refpath_assemble(void *hp, double exposure, double gamma, double saturation, uint32_t *image, uint32_t *shadow) {
    RefPath *h = static_cast<RefPath *>(hp);
    h->rp.m_exposure = exposure; h->rp.m_gamma = gamma; h->rp.m_saturation = saturation;
    h->ras->AssembleImage();
    const int n = h->rp.size();
    QImage *a = h->ras->getBuffer(), *s = h->ras->getImageShadowBuffer();
    if (!a || !s) return 1;
    for (int j = 0; j < n; j++)
        for (int i = 0; i < n; i++) {
            image[size_t(j) * n + i] = a->pixel(i, j);
            shadow[size_t(j) * n + i] = s->pixel(i, j);
        }
    return 0;
}

// The background star layer's stamping: the reference's own Buffer2D::RenderGaussian (util/buffer2d.cpp:224-243) for every
// star of a list, on a zeroed size x size buffer — what Buffer2D::RenderStars (:140-173) does once it has drawn a star's
// parameters. The parameters themselves come from libc rand() there and cannot be reproduced; the stamping can.
// stars: n records {int32 x, y, w; float cs[3]} (the layout of GamerStar, include/gamer_b200.h). layer_out = size*size*3.
struct RefStar { int32_t x, y, w; float cs[3]; };
int refpath_stamp_stars(int size, const void *stars, int n, float *layer_out) {
    Buffer2D buf(size);
    buf.fill(QVector3D(0, 0, 0));
    const RefStar *st = static_cast<const RefStar *>(stars);
    for (int k = 0; k < n; k++) buf.RenderGaussian(st[k].x, st[k].y, st[k].w, QVector3D(st[k].cs[0], st[k].cs[1], st[k].cs[2]));
    for (int j = 0; j < size; j++)
        for (int i = 0; i < size; i++) {
            const QVector3D v = buf.Get(i, j);
            float *o = layer_out + (size_t(j) * size + i) * 3;
            o[0] = v.x(); o[1] = v.y(); o[2] = v.z();
        }
    return 0;
}

void refpath_destroy(void *hp) {
    RefPath *h = static_cast<RefPath *>(hp);
    delete h->ras;
    delete h;
}

}  // extern "C"
