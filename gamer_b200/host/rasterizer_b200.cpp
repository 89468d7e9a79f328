// See rasterizer_b200.hpp. Host-only code: no CUDA here, everything device-side is behind the C ABI.
#include "rasterizer_b200.hpp"

#include <algorithm>
#include <cctype>
#include <cmath>
#include <map>
#include <stdexcept>

namespace gamer {

// spectrum.h:50-72 — the six named RGB triples the CLI installs (consolerenderer.cpp:68)
void Spectra::PopulateSpectra() {
    m_spectra = {{"Red", {1, 0.6f, 0.4f}},   {"Yellow", {1, 0.9f, 0.45f}}, {"Blue", {0.4f, 0.6f, 1.0f}},
                 {"White", {1, 1, 1}},       {"Cyan", {0.3f, 0.7f, 1.0f}}, {"Purple", {1.0f, 0.3f, 0.8f}}};
}

static std::string lower(std::string s) {
    for (char& c : s) c = char(std::tolower((unsigned char)c));
    return s;
}

const ComponentSpectrum* Spectra::FindSpectrum(const std::string& name) const {
    const std::string n = lower(name);
    for (const ComponentSpectrum& s : m_spectra)
        if (lower(s.m_name) == n) return &s;
    return nullptr;
}

int GamerCamera::setupViewmatrix(float inv_vp[16]) const {
    const float c[3] = {m_camera.x, m_camera.y, m_camera.z}, t[3] = {m_target.x, m_target.y, m_target.z},
                u[3] = {m_up.x, m_up.y, m_up.z};
    return gamer_b200_camera_inv_vp(c, t, u, double(m_perspective), inv_vp);
}

RasterizerB200::RasterizerB200(const std::vector<int32_t>& device_ids) {
    int rc = gamer_b200_create(&m_ctx, device_ids.data(), int32_t(device_ids.size()));
    if (rc != GAMER_OK) {  // no CPU fallback: without a usable B200 context there is no rasterizer
        const char* msg = gamer_b200_last_error(nullptr);
        throw std::runtime_error(std::string("gamer_b200_create failed: ") + (msg ? msg : "?"));
    }
}

RasterizerB200::~RasterizerB200() {
    if (m_thread.joinable()) {
        gamer_b200_abort(m_ctx);
        m_thread.join();
    }
    gamer_b200_destroy(m_ctx);
}

static int class_id_of(const std::string& name) {  // factory strings of galaxy.cpp:15-42
    static const std::map<std::string, int> ids = {{"bulge", GAMER_CLASS_BULGE},
                                                   {"disk", GAMER_CLASS_DISK},
                                                   {"dust", GAMER_CLASS_DUST},
                                                   {"dust2", GAMER_CLASS_DUST2},
                                                   {"dust positive", GAMER_CLASS_DUST_POSITIVE},
                                                   {"stars", GAMER_CLASS_STARS},
                                                   {"stars small", GAMER_CLASS_STARS_SMALL}};
    auto it = ids.find(name);
    return it == ids.end() ? -1 : it->second;
}

// Rasterizer::Prepare (rasterizer.cpp:190-213) + Galaxy::SetupComponents/SetupSpectra (galaxy.cpp:75-109):
// sort far -> near, instantiate components in file order, resolve spectra (white when unknown).
Scene RasterizerB200::Prepare() const {
    if (!m_renderingParams) throw std::runtime_error("RasterizerB200: setRenderingParams() was not called");
    RenderingParams& rp = *m_renderingParams;
    const Vec3 cam = rp.camera().m_camera;
    auto dist = [&](const GalaxyInstance* gi) {
        const float dx = gi->m_position.x - cam.x, dy = gi->m_position.y - cam.y, dz = gi->m_position.z - cam.z;
        return float(std::sqrt(double(dx) * dx + double(dy) * dy + double(dz) * dz));
    };
    std::vector<GalaxyInstance*> ordered = m_galaxies;
    std::stable_sort(ordered.begin(), ordered.end(),
                     [&](const GalaxyInstance* a, const GalaxyInstance* b) { return dist(a) > dist(b); });
    Scene sc;
    for (const GalaxyInstance* gi : ordered) {
        if (!gi || !gi->m_galaxy) throw std::runtime_error("RasterizerB200: galaxy instance without a galaxy");
        const GalaxyParams& gp = gi->m_galaxy->galaxyParams;
        GamerGalaxy g{};
        const float pos[3] = {gi->m_position.x, gi->m_position.y, gi->m_position.z};
        const float ori[3] = {gi->m_orientation.x, gi->m_orientation.y, gi->m_orientation.z};
        for (int i = 0; i < 3; ++i) {
            g.pos[i] = pos[i];
            g.orient[i] = ori[i];
            g.axis[i] = float(gp.axis[i]);
        }
        gamer_b200_rotation_to_orientation(ori, g.rotmat);  // setupQuaternions, galaxyinstance.cpp:69-71
        g.intensity_scale = gi->m_intensityScale;
        g.winding_b = gp.windingB;
        g.winding_n = gp.windingN;
        g.no_arms = gp.noArms;
        g.arm[0] = gp.arm1; g.arm[1] = gp.arm2; g.arm[2] = gp.arm3; g.arm[3] = gp.arm4;
        g.comp_begin = int32_t(sc.components.size());
        for (const ComponentParams& cp : gi->m_galaxy->componentParams) {
            const int cid = class_id_of(cp.className);
            if (cid < 0) continue;  // unknown class: the reference reports it and skips (galaxy.cpp:38-41)
            GamerComponent c{};
            c.class_id = cid;
            c.active = cp.active;
            c.strength = cp.strength; c.arm = cp.arm; c.z0 = cp.z0; c.r0 = cp.r0; c.inner = cp.inner;
            c.delta = cp.delta; c.winding = cp.winding; c.scale = cp.scale;
            c.noise_offset = cp.noiseOffset; c.noise_tilt = cp.noiseTilt; c.ks = cp.ks;
            const ComponentSpectrum* s = rp.spectra().FindSpectrum(cp.spectrum);
            const Vec3 rgb = s ? s->m_spectrum : Vec3{1, 1, 1};
            c.spec[0] = rgb.x; c.spec[1] = rgb.y; c.spec[2] = rgb.z;
            sc.components.push_back(c);
        }
        g.comp_count = int32_t(sc.components.size()) - g.comp_begin;
        sc.galaxies.push_back(g);
    }
    GamerFrame& f = sc.frame;
    f.width = f.height = f.full_size = rp.size();
    if (rp.camera().setupViewmatrix(f.inv_vp) != GAMER_OK) throw std::runtime_error("camera matrix is singular");
    f.cam[0] = cam.x; f.cam[1] = cam.y; f.cam[2] = cam.z;
    f.ray_step = double(rp.rayStep());
    f.min_ray_step = m_isPreview ? 0.01 : 0.001;  // rasterizer.cpp:437-441
    f.flags = m_noRadialGate ? GAMER_FRAME_NO_RADIAL_GATE : 0;
    f.n_galaxies = int32_t(sc.galaxies.size());
    f.n_components = int32_t(sc.components.size());
    return sc;
}

int RasterizerB200::RenderPixels() {
    Scene sc = Prepare();
    m_width = sc.frame.width;
    m_height = sc.frame.height;
    m_renderBuffer.assign(size_t(m_width) * m_height * 3, 0.0f);
    // Far -> near passes of at most GAMER_MAX_GALAXIES galaxies / GAMER_MAX_COMPONENTS active components;
    // the unscaled accumulation stays in the device framebuffer between passes (gamer_b200.h).
    static const GamerGalaxy no_gal{};
    static const GamerComponent no_comp{};
    const size_t ng = sc.galaxies.size();
    size_t first = 0;
    int rc = GAMER_OK;
    do {
        size_t last = first, active = 0;
        while (last < ng && last - first < GAMER_MAX_GALAXIES) {
            size_t a = 0;
            const GamerGalaxy& g = sc.galaxies[last];
            for (int k = g.comp_begin; k < g.comp_begin + g.comp_count; ++k) a += sc.components[size_t(k)].active == 1.0;
            if (last > first && active + a > GAMER_MAX_COMPONENTS) break;
            active += a;
            ++last;
        }
        std::vector<GamerGalaxy> gals(sc.galaxies.begin() + long(first), sc.galaxies.begin() + long(last));
        std::vector<GamerComponent> comps;
        for (GamerGalaxy& g : gals) {
            const int begin = int(comps.size());
            comps.insert(comps.end(), sc.components.begin() + g.comp_begin, sc.components.begin() + g.comp_begin + g.comp_count);
            g.comp_begin = begin;
        }
        GamerFrame f = sc.frame;
        f.n_galaxies = int32_t(gals.size());
        f.n_components = int32_t(comps.size());
        f.flags = (sc.frame.flags & GAMER_FRAME_NO_RADIAL_GATE) | (first > 0 ? GAMER_FRAME_CONTINUE : 0) |
                  (last < ng ? GAMER_FRAME_PARTIAL : 0);
        rc = gamer_b200_upload(m_ctx, &f, gals.empty() ? &no_gal : gals.data(), comps.empty() ? &no_comp : comps.data());
        if (rc != GAMER_OK) return rc;
        rc = last < ng ? gamer_b200_render_device(m_ctx) : gamer_b200_render(m_ctx, m_renderBuffer.data());
        if (rc != GAMER_OK) break;
        first = last;
    } while (first < ng);
    if (rc == GAMER_E_ABORTED) std::fill(m_renderBuffer.begin(), m_renderBuffer.end(), 0.0f);
    return rc;
}

void RasterizerB200::run() {
    int rc;
    try {
        rc = RenderPixels();
    } catch (const std::exception&) {
        rc = GAMER_E_ARG;
    }
    m_status.store(rc);
    m_state.store(done);
}

void RasterizerB200::Render() {
    if (m_thread.joinable()) m_thread.join();
    m_isPreview = false;
    m_state.store(rendering);
    m_thread = std::thread(&RasterizerB200::run, this);
}

int RasterizerB200::RenderDirect() {
    if (m_thread.joinable()) m_thread.join();
    m_isPreview = true;
    m_state.store(rendering);
    run();
    return m_status.load();
}

void RasterizerB200::Abort() {
    gamer_b200_abort(m_ctx);
    if (m_thread.joinable()) m_thread.join();
    m_state.store(aborting);
}

double RasterizerB200::getPercentDone() const {
    double f = 0;
    gamer_b200_progress(m_ctx, &f);
    return f;
}

const char* RasterizerB200::lastError() const { return gamer_b200_last_error(m_ctx); }

float RasterizerB200::lastKernelMs() const {
    float a = 0, b = 0;
    gamer_b200_last_kernel_ms(m_ctx, &a, &b);
    return a;
}

int RasterizerB200::AssembleImage() {
    if (m_thread.joinable()) m_thread.join();
    RenderingParams& rp = *m_renderingParams;
    const size_t px = size_t(m_width) * m_height;
    m_imageBuffer.assign(px, 0);
    m_imageShadowBuffer.assign(px, 0);
    const float* stars = (!m_isPreview && m_starsBuffer.size() == px * 3) ? m_starsBuffer.data() : nullptr;
    if (!m_isPreview && m_deviceStars && rp.m_noStars > 0) {
        std::vector<GamerStar> list(size_t(rp.m_noStars));
        int rc = gamer_b200_star_list(rp.size(), rp.m_noStars, rp.m_starSize, rp.m_starSizeSpread, rp.m_starStrength, m_starSeed,
                                      list.data());
        if (rc == GAMER_OK) rc = gamer_b200_render_stars(m_ctx, list.data(), int32_t(list.size()), nullptr);
        if (rc != GAMER_OK) return rc;
        stars = GAMER_STARS_LAYER;
    }
    return gamer_b200_tonemap(m_ctx, double(rp.m_exposure), double(rp.m_gamma), double(rp.m_saturation), stars,
                              m_imageBuffer.data(), m_imageShadowBuffer.data());
}

}  // namespace gamer
