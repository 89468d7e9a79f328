// Qt-free reader/writer of GAMER's .gax galaxy files.
// Mirrors the reference's parameter classes by name:
//   ComponentParams  source/galaxy/componentparams.h:7-44
//   GalaxyParams     source/galaxy/galaxyparams.h:10-43
//   Galaxy           source/galaxy/galaxy.h:39-68 (stream operators), galaxy.cpp:173-206 (Save/Load)
// A .gax file is a raw big-endian QDataStream dump: QString = u32 byte length (0xFFFFFFFF = null)
// + UTF-16BE code units; double = IEEE f64; QVector3D = three floats widened to f64 (QDataStream's
// default DoublePrecision); QVector<T> = u32 count + items.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace gamer {

struct ComponentParams {
    std::string className = "bulge";
    double strength = 1.0;
    std::string spectrum = "White";
    double arm = 0.5, z0 = 0.02, r0 = 0.5, active = 1.0, delta = 0.0, winding = 0.1, scale = 1.0,
           noiseOffset = 0.0, noiseTilt = 1.0, ks = 1.0, inner = 0.0;
    std::string name = "New component";
};

struct GalaxyParams {
    std::string name = "NewGalaxy";
    double axis[3] = {1, 1, 1};
    double bulgeDust = 0.025;
    double bulgeAxis[3] = {1, 1, 1};
    double windingB = 0.5, windingN = 4, noArms = 2, arm1 = 0, arm2 = 3.141592653589793,
           arm3 = 2 * 3.141592653589793, arm4 = 3 * 3.141592653589793, innerTwirl = 0, warpAmplitude = 0,
           warpScale = 0;
};

namespace detail {

struct ByteReader {
    const std::vector<uint8_t>& d;
    size_t o = 0;
    explicit ByteReader(const std::vector<uint8_t>& data) : d(data) {}
    void need(size_t n) const {
        if (o + n > d.size()) throw std::runtime_error(".gax: truncated file");
    }
    uint32_t u32() {
        need(4);
        uint32_t v = (uint32_t(d[o]) << 24) | (uint32_t(d[o + 1]) << 16) | (uint32_t(d[o + 2]) << 8) | d[o + 3];
        o += 4;
        return v;
    }
    double f64() {
        need(8);
        uint64_t b = 0;
        for (int i = 0; i < 8; ++i) b = (b << 8) | d[o + i];
        o += 8;
        double v;
        std::memcpy(&v, &b, 8);
        return v;
    }
    // UTF-16BE -> UTF-8 (surrogate pairs handled; the shipped files are ASCII)
    std::string qstring() {
        uint32_t n = u32();
        if (n == 0xFFFFFFFFu) return std::string();
        need(n);
        std::string s;
        for (uint32_t i = 0; i + 1 < n; i += 2) {
            uint32_t c = (uint32_t(d[o + i]) << 8) | d[o + i + 1];
            if (c >= 0xD800 && c < 0xDC00 && i + 3 < n) {
                uint32_t lo = (uint32_t(d[o + i + 2]) << 8) | d[o + i + 3];
                c = 0x10000 + ((c - 0xD800) << 10) + (lo - 0xDC00);
                i += 2;
            }
            if (c < 0x80) s += char(c);
            else if (c < 0x800) { s += char(0xC0 | (c >> 6)); s += char(0x80 | (c & 0x3F)); }
            else if (c < 0x10000) { s += char(0xE0 | (c >> 12)); s += char(0x80 | ((c >> 6) & 0x3F)); s += char(0x80 | (c & 0x3F)); }
            else { s += char(0xF0 | (c >> 18)); s += char(0x80 | ((c >> 12) & 0x3F)); s += char(0x80 | ((c >> 6) & 0x3F)); s += char(0x80 | (c & 0x3F)); }
        }
        o += n;
        return s;
    }
};

struct ByteWriter {
    std::vector<uint8_t> d;
    void u32(uint32_t v) { for (int s = 24; s >= 0; s -= 8) d.push_back(uint8_t(v >> s)); }
    void f64(double v) {
        uint64_t b;
        std::memcpy(&b, &v, 8);
        for (int s = 56; s >= 0; s -= 8) d.push_back(uint8_t(b >> s));
    }
    void qstring(const std::string& s) {
        std::vector<uint16_t> u;
        for (size_t i = 0; i < s.size();) {
            uint32_t c = uint8_t(s[i]);
            int extra = c < 0x80 ? 0 : c < 0xE0 ? 1 : c < 0xF0 ? 2 : 3;
            c = extra == 0 ? c : c & (0x3F >> extra);
            for (int k = 1; k <= extra && i + k < s.size(); ++k) c = (c << 6) | (uint8_t(s[i + k]) & 0x3F);
            i += extra + 1;
            if (c >= 0x10000) { c -= 0x10000; u.push_back(uint16_t(0xD800 + (c >> 10))); u.push_back(uint16_t(0xDC00 + (c & 0x3FF))); }
            else u.push_back(uint16_t(c));
        }
        u32(uint32_t(u.size() * 2));
        for (uint16_t c : u) { d.push_back(uint8_t(c >> 8)); d.push_back(uint8_t(c)); }
    }
};

}  // namespace detail

class Galaxy {
public:
    std::string displayName;
    GalaxyParams galaxyParams;
    std::vector<ComponentParams> componentParams;

    static Galaxy fromBytes(const std::vector<uint8_t>& data) {
        detail::ByteReader r(data);
        Galaxy g;
        g.displayName = r.qstring();
        GalaxyParams& p = g.galaxyParams;
        p.name = r.qstring();
        for (double& v : p.axis) v = r.f64();
        p.bulgeDust = r.f64();
        for (double& v : p.bulgeAxis) v = r.f64();
        double* tail[] = {&p.windingB, &p.windingN, &p.noArms, &p.arm1, &p.arm2, &p.arm3, &p.arm4, &p.innerTwirl,
                          &p.warpAmplitude, &p.warpScale};
        for (double* v : tail) *v = r.f64();
        uint32_t n = r.u32();
        if (n > 4096) throw std::runtime_error(".gax: implausible component count");
        for (uint32_t i = 0; i < n; ++i) {
            ComponentParams c;
            c.className = r.qstring();
            c.strength = r.f64();
            c.spectrum = r.qstring();
            double* f[] = {&c.arm, &c.z0, &c.r0, &c.active, &c.delta, &c.winding, &c.scale, &c.noiseOffset, &c.noiseTilt,
                           &c.ks, &c.inner};
            for (double* v : f) *v = r.f64();
            c.name = r.qstring();
            g.componentParams.push_back(c);
        }
        if (r.o != data.size()) throw std::runtime_error(".gax: trailing bytes");
        return g;
    }

    std::vector<uint8_t> toBytes() const {
        detail::ByteWriter w;
        w.qstring(displayName);
        const GalaxyParams& p = galaxyParams;
        w.qstring(p.name);
        for (double v : p.axis) w.f64(v);
        w.f64(p.bulgeDust);
        for (double v : p.bulgeAxis) w.f64(v);
        for (double v : {p.windingB, p.windingN, p.noArms, p.arm1, p.arm2, p.arm3, p.arm4, p.innerTwirl, p.warpAmplitude,
                         p.warpScale})
            w.f64(v);
        w.u32(uint32_t(componentParams.size()));
        for (const ComponentParams& c : componentParams) {
            w.qstring(c.className);
            w.f64(c.strength);
            w.qstring(c.spectrum);
            for (double v : {c.arm, c.z0, c.r0, c.active, c.delta, c.winding, c.scale, c.noiseOffset, c.noiseTilt, c.ks, c.inner})
                w.f64(v);
            w.qstring(c.name);
        }
        return w.d;
    }

    // Galaxy::Load / Save (galaxy.cpp:173-206). Load returns false when the file cannot be opened,
    // like the reference; a malformed file throws.
    bool Load(const std::string& path) {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) return false;
        std::vector<uint8_t> data;
        uint8_t buf[4096];
        size_t n;
        while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) data.insert(data.end(), buf, buf + n);
        std::fclose(f);
        *this = fromBytes(data);
        return true;
    }

    bool Save(const std::string& path) const {
        FILE* f = std::fopen(path.c_str(), "wb");
        if (!f) return false;
        std::vector<uint8_t> d = toBytes();
        bool ok = std::fwrite(d.data(), 1, d.size(), f) == d.size();
        std::fclose(f);
        return ok;
    }
};

}  // namespace gamer
