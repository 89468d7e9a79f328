// FITS export of the linear render buffer, the C++ twin of gamer_b200/fits.py.
//
// The reference saves one plane per colour channel as BITPIX = -64 (FitsIO::Savedouble, source/util/fitsio.h:18-56,
// called from mainwindow.cpp:890-892 with <base>_red.fits / _green.fits / _blue.fits) and rows bottom-up
// (Buffer2D::toQByteArray, util/buffer2d.cpp:175-185: Get(j, size-1-i)). Its byte "flip" swaps groups of FOUR bytes
// over the first quarter of an array of 8-byte doubles and its header cards are not padded to 80 columns, so the
// files it writes are not valid FITS. This writer keeps the reference's content and orientation and writes the
// standard instead of copying the two bugs: 80-column cards, 2880-byte blocks, big-endian IEEE doubles.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace gamer {
namespace fits_detail {

constexpr size_t kBlock = 2880;

inline void card(std::vector<uint8_t>& out, const char* key, const std::string& value, const char* comment) {
    char line[161];
    if (std::strcmp(key, "END") == 0) std::snprintf(line, sizeof line, "END");
    else if (comment && *comment) std::snprintf(line, sizeof line, "%-8s= %20s / %s", key, value.c_str(), comment);
    else std::snprintf(line, sizeof line, "%-8s= %20s", key, value.c_str());
    size_t n = std::strlen(line);
    if (n > 80) n = 80;
    out.insert(out.end(), line, line + n);
    out.insert(out.end(), 80 - n, uint8_t(' '));
}

inline void pad(std::vector<uint8_t>& out, uint8_t fill) {
    out.insert(out.end(), (kBlock - out.size() % kBlock) % kBlock, fill);
}

}  // namespace fits_detail

// One plane: channel `no` (0..2) of an interleaved float RGB buffer of `width` x `height` pixels, image row 0 at the
// top (the layout of Rasterizer's render buffer), written bottom-up like the reference.
inline std::vector<uint8_t> encode_fits_plane(const float* rgb, int width, int height, int no) {
    using namespace fits_detail;
    std::vector<uint8_t> out;
    out.reserve(kBlock + size_t(width) * height * 8 + kBlock);
    card(out, "SIMPLE", "T", "conforms to FITS");
    card(out, "BITPIX", "-64", "IEEE double");
    card(out, "NAXIS", "2", "");
    card(out, "NAXIS1", std::to_string(width), "");
    card(out, "NAXIS2", std::to_string(height), "");
    card(out, "END", "", "");
    pad(out, uint8_t(' '));
    for (int y = height - 1; y >= 0; --y)
        for (int x = 0; x < width; ++x) {
            const double v = double(rgb[3 * (size_t(y) * width + x) + no]);
            uint64_t bits;
            std::memcpy(&bits, &v, 8);
            for (int s = 56; s >= 0; s -= 8) out.push_back(uint8_t(bits >> s));
        }
    pad(out, 0);
    return out;
}

// FitsIO::Savedouble(filename, no, buffer)
inline bool save_fits_double(const std::string& filename, int no, const float* rgb, int width, int height) {
    if (no < 0 || no > 2 || width <= 0 || height <= 0 || !rgb) return false;
    const std::vector<uint8_t> bytes = encode_fits_plane(rgb, width, height, no);
    FILE* f = std::fopen(filename.c_str(), "wb");
    if (!f) return false;
    const bool ok = std::fwrite(bytes.data(), 1, bytes.size(), f) == bytes.size();
    return std::fclose(f) == 0 && ok;
}

// mainwindow.cpp:890-892: <base>_red.fits, <base>_green.fits, <base>_blue.fits
inline bool save_fits_rgb(const std::string& base, const float* rgb, int width, int height) {
    static const char* suffix[3] = {"_red.fits", "_green.fits", "_blue.fits"};
    for (int no = 0; no < 3; ++no)
        if (!save_fits_double(base + suffix[no], no, rgb, width, height)) return false;
    return true;
}

}  // namespace gamer
