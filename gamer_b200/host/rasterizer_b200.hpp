// Qt-free host side above the C ABI: the reference's caller-facing model for the render path and a
// Rasterizer-shaped adapter whose RenderPixels()/AssembleImage() are single calls into libgamer_b200.
//   Spectra / ComponentSpectrum   source/galaxy/spectrum.h:9-72
//   GamerCamera                   source/galaxy/gamercamera.h:7-45, gamercamera.cpp:185-217
//   RenderingParams               source/galaxy/renderingparams.h:19-64
//   GalaxyInstance                source/galaxy/galaxyinstance.h:8-45
//   RasterizerB200                source/galaxy/rasterizer.h:20-127 (same method names and call order)
// In a Qt build of the reference the same adapter derives from Rasterizer.
#pragma once
#include <atomic>
#include <cstdint>
#include <string>
#include <thread>
#include <vector>

#include "../../include/gamer_b200.h"
#include "gax.hpp"

namespace gamer {

struct Vec3 {
    float x = 0, y = 0, z = 0;
};

struct ComponentSpectrum {
    std::string m_name;
    Vec3 m_spectrum;
};

class Spectra {  // spectrum.h:40-72
public:
    std::vector<ComponentSpectrum> m_spectra;
    void PopulateSpectra();
    const ComponentSpectrum* FindSpectrum(const std::string& name) const;  // case-insensitive
};

class GamerCamera {
public:
    Vec3 m_camera{3, 3, 3}, m_target{0, 0, 0}, m_up{0, 1, 0};
    float m_perspective = 70;
    void setCamera(Vec3 v) { m_camera = v; }
    void setTarget(Vec3 v) { m_target = v; }
    void setUp(Vec3 v) { m_up = v; }
    void setPerspective(float f) { m_perspective = f; }
    int setupViewmatrix(float inv_vp[16]) const;  // gamercamera.cpp:185-208
};

class RenderingParams {
public:
    int m_size = 128;
    float m_exposure = 1, m_gamma = 1, m_saturation = 1, m_rayStep = 0.001f;
    int m_noStars = 0;
    double m_starSize = 1, m_starSizeSpread = 1, m_starStrength = 1;   // renderingparams.h:28-30
    // carried for file round trips only (renderingparams.h:20-37)
    int m_previewSize = 64, m_nside = 32;
    double m_detailLevel = 0.01, m_noiseDetail = 1;
    std::string m_renderType, m_galaxyDirectory = "galaxies/", m_sceneDirectory = "scenes/", m_imageDirectory = "images/",
        m_currentGalaxy, m_sceneMode = "galaxy";
    GamerCamera m_camera;
    Spectra m_spectra;
    GamerCamera& camera() { return m_camera; }
    Spectra& spectra() { return m_spectra; }
    int size() const { return m_size; }
    float rayStep() const { return m_rayStep; }
    void setSize(int s) { m_size = s; }
    void setExposure(float v) { m_exposure = v; }
    void setGamma(float v) { m_gamma = v; }
    void setSaturation(float v) { m_saturation = v; }
    void setRayStep(float v) { m_rayStep = v; }
};

class GalaxyInstance {
public:
    const Galaxy* m_galaxy = nullptr;
    std::string m_galaxyName;
    Vec3 m_position, m_orientation{0, 1, 0};
    double m_intensityScale = 1, m_redshift = 0;
    GalaxyInstance() = default;
    GalaxyInstance(const Galaxy* g, std::string name, Vec3 pos, Vec3 orient, double iscale, double redshift)
        : m_galaxy(g), m_galaxyName(std::move(name)), m_position(pos), m_orientation(orient), m_intensityScale(iscale),
          m_redshift(redshift) {}
};

// The POD scene of include/gamer_b200.h, owning its arrays.
struct Scene {
    GamerFrame frame{};
    std::vector<GamerGalaxy> galaxies;
    std::vector<GamerComponent> components;
};

class RasterizerB200 {
public:
    enum State { idle, rendering, done, aborting };

    explicit RasterizerB200(const std::vector<int32_t>& device_ids = {0});
    ~RasterizerB200();
    RasterizerB200(const RasterizerB200&) = delete;
    RasterizerB200& operator=(const RasterizerB200&) = delete;

    // inputs (rasterizer.cpp:48-51, 229-247, 126-136)
    void setRenderingParams(RenderingParams* rp) { m_renderingParams = rp; }
    RenderingParams* getRenderingParams() { return m_renderingParams; }
    void AddGalaxy(GalaxyInstance* gi) { m_galaxies.push_back(gi); }
    void Clear() { m_galaxies.clear(); }
    void CopyFrom(const RasterizerB200* from) { m_galaxies = from->m_galaxies; }
    void setNewSize(int s) { m_renderingParams->setSize(s); }
    // gamer_b200.h GAMER_FRAME_NO_RADIAL_GATE: reproduce a g++/clang -O1+ build of the reference, in which the
    // `intensity > 0.001` test of galaxycomponent.cpp:62 is compiled away (default: the test as written)
    void setNoRadialGate(bool on) { m_noRadialGate = on; }

    // render (rasterizer.cpp:190-213, 283-313, 350-370)
    Scene Prepare() const;
    int RenderPixels();       // upload + K1 + D2H into m_renderBuffer; returns a GAMER_* code
    void Render();            // asynchronous on the rasterizer's own thread, like QThread::start
    int RenderDirect();       // synchronous preview
    void Abort();
    State getState() const { return m_state.load(); }
    double getPercentDone() const;
    int lastStatus() const { return m_status.load(); }
    const char* lastError() const;

    // post-process (rasterizer.cpp:315-327)
    void RenderStars(const std::vector<float>& stars_rgb) { m_starsBuffer = stars_rgb; m_deviceStars = false; }
    // rasterizer.cpp:215-225: layer from noStars/starSize/starSizeSpread/starStrength, stamped on the device (K4)
    void RenderStars(uint64_t seed = 0) { m_deviceStars = true; m_starSeed = seed; }
    int AssembleImage();

    // outputs
    const std::vector<float>& getRenderBuffer() const { return m_renderBuffer; }          // linear float RGB, x + y*N
    const std::vector<uint32_t>& getBuffer() const { return m_imageBuffer; }              // ARGB32, as shown by the GUI
    const std::vector<uint32_t>& getImageShadowBuffer() const { return m_imageShadowBuffer; }  // ARGB32, as saved to PNG
    int width() const { return m_width; }
    int height() const { return m_height; }
    float lastKernelMs() const;

private:
    void run();
    gamer_ctx* m_ctx = nullptr;
    std::vector<GalaxyInstance*> m_galaxies;
    RenderingParams* m_renderingParams = nullptr;
    std::vector<float> m_renderBuffer, m_starsBuffer;
    std::vector<uint32_t> m_imageBuffer, m_imageShadowBuffer;
    std::atomic<State> m_state{idle};
    std::atomic<int> m_status{0};
    bool m_isPreview = false;
    bool m_deviceStars = false;
    bool m_noRadialGate = false;
    uint64_t m_starSeed = 0;
    int m_width = 0, m_height = 0;
    std::thread m_thread;
};

}  // namespace gamer
