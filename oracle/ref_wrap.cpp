// ref_wrap.cpp — TEST INFRASTRUCTURE. extern "C" doors onto the reference's OWN noise code,
// compiled in place from /root/reference/source/noise/simplexnoise.cpp (the one file of the hot
// path that has no Qt dependency). Built only into oracle/_ref/ (git-ignored); used by tests to
// pin the oracle's restated simplex/fBm bit-for-bit. No reference source is copied into this repo.
extern double raw_noise_3d(const double x, const double y, const double z);          // simplexnoise.cpp:243
extern double octave_noise_3d(const double octaves, const double persistence, const double scale,
                              const double x, const double y, const double z);       // simplexnoise.cpp:78
extern "C" {
double ref_raw_noise_3d(double x, double y, double z) { return raw_noise_3d(x, y, z); }
double ref_octave_noise_3d(double octaves, double persistence, double scale, double x, double y, double z) {
    return octave_noise_3d(octaves, persistence, scale, x, y, z);
}
void ref_raw_noise_3d_batch(const double *xyz, int n, double *out) {
    for (int i = 0; i < n; i++) out[i] = raw_noise_3d(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
}
}
