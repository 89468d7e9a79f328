// Qt-free reader of GAMER's RenderingParams files (RenderParams.dat) — see gamer_b200/renderparams.py for
// the layout (source/galaxy/renderingparams.h:41-64, gamercamera.h:31-39, spectrum.h:23-31,74-93).
// Older files end after the spectra (no nside / renderType): accepted, like the reference's shipped file.
#pragma once
#include <cstdio>
#include <string>
#include <vector>

#include "gax.hpp"
#include "rasterizer_b200.hpp"

namespace gamer {

inline bool LoadRenderingParams(const std::string& path, RenderingParams& rp) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::vector<uint8_t> data;
    uint8_t buf[4096];
    size_t n;
    while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) data.insert(data.end(), buf, buf + n);
    std::fclose(f);
    detail::ByteReader r(data);
    auto v3 = [&]() { Vec3 v; v.x = float(r.f64()); v.y = float(r.f64()); v.z = float(r.f64()); return v; };
    rp.m_camera.m_camera = v3();
    rp.m_camera.m_target = v3();
    rp.m_camera.m_up = v3();
    rp.m_camera.m_perspective = float(r.f64());
    rp.m_size = int32_t(r.u32());
    rp.m_previewSize = int32_t(r.u32());
    rp.m_exposure = float(r.f64());
    rp.m_gamma = float(r.f64());
    rp.m_saturation = float(r.f64());
    rp.m_detailLevel = r.f64();
    rp.m_noiseDetail = r.f64();
    rp.m_noStars = int32_t(r.u32());
    rp.m_starSize = r.f64();
    rp.m_starSizeSpread = r.f64();
    rp.m_starStrength = r.f64();
    rp.m_rayStep = float(r.f64());
    rp.m_galaxyDirectory = r.qstring();
    rp.m_sceneDirectory = r.qstring();
    rp.m_currentGalaxy = r.qstring();
    rp.m_sceneMode = r.qstring();
    rp.m_imageDirectory = r.qstring();
    const uint32_t ns = r.u32();
    if (ns > 4096) throw std::runtime_error("RenderingParams: implausible spectrum count");
    rp.m_spectra.m_spectra.clear();
    for (uint32_t i = 0; i < ns; ++i) {
        ComponentSpectrum cs;
        cs.m_name = r.qstring();
        cs.m_spectrum = v3();
        rp.m_spectra.m_spectra.push_back(cs);
    }
    if (r.o < data.size()) {
        rp.m_nside = int32_t(r.u32());
        rp.m_renderType = r.qstring();
    }
    if (r.o != data.size()) throw std::runtime_error("RenderingParams: trailing bytes");
    return true;
}

}  // namespace gamer
