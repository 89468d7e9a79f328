// ref_noise_class_wrap.cpp — TEST INFRASTRUCTURE. extern "C" doors onto the reference's OWN noise CLASSES,
// compiled in place from /root/reference/source/noise/{noise,simplex,simplexnoise}.cpp against the minimal
// Qt stand-ins of oracle/qtshim (QObject: empty; QVector: unused; QVector3D: three floats and `* float`).
// The instance is built the way Rasterizer builds it: new Simplex(1,1,1,1) (galaxy/rasterizer.h:60-72).
// Used by tests/golden/make_fixtures.py to pin the oracle's restated fBm and ridged multifractal against
// Noise::octave_noise_3d (noise.cpp:162-180) and Noise::getRidgedMf (noise.cpp:81-128). Built only into
// oracle/_ref/ (git-ignored); no reference source is copied into this repo.
#include "source/noise/simplex.h"

extern "C" {
void ref_class_octave_noise_3d(const float *xyz, int n, double octaves, double persistence, double scale, double *out) {
    Simplex noise(1, 1, 1, 1);
    for (int i = 0; i < n; i++)
        out[i] = noise.octave_noise_3d(octaves, persistence, scale, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
}
void ref_class_ridged(const float *xyz, int n, double frequency, int octaves, double lacunarity, double offset, double gain,
                      double *out) {
    Simplex noise(1, 1, 1, 1);
    for (int i = 0; i < n; i++)
        out[i] = noise.getRidgedMf(QVector3D(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]), frequency, octaves, lacunarity, offset, gain);
}
void ref_class_raw_3d(const double *xyz, int n, double *out) {
    Simplex noise(1, 1, 1, 1);
    for (int i = 0; i < n; i++) out[i] = noise.raw_3d(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
}
}
