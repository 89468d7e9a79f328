// kernels.h — host-callable launchers of the sm_100a kernels (kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "scene_dev.h"
#include "../../include/gamer_b200.h"

namespace gk {
cudaError_t prepare_device();   // per device, before the first launch
cudaError_t launch_k1(const DScene &sc, float *out_dev, unsigned long long *counters_dev, const int *abort_dev,
                      unsigned int *progress_dev /* [0] tiles done, [1] ticket; both zeroed by the caller */, bool counted,
                      int sm_count, cudaStream_t st, unsigned int *n_tiles);
cudaError_t launch_k2(const float *rgb, const float *stars, uint32_t *argb, uint32_t *shadow, size_t n, double exposure,
                      double gamma, double saturation, int sm_count, cudaStream_t st, int *n_launches = nullptr);
cudaError_t launch_k4(const GamerStar *stars, int n, float *layer, const DScene &sc, cudaStream_t st);
cudaError_t launch_probe_simplex(const float *xyz, int n, double frequency, int32_t *lat, float *val, cudaStream_t st);
cudaError_t launch_probe_fbm(const float *xyz, int n, int octaves, float pers, float inv_maxamp, double scale, float *out,
                             cudaStream_t st);
cudaError_t launch_probe_ridged(const float *xyz, int n, int octaves, float lac, float offset, float gain, const float *rw,
                                float *out, cudaStream_t st);
// allocation-free fBm evaluation on the current device (scratch in a static __device__ array of kernels.cu)
int fbm_static_max_points();
cudaError_t fbm_static(const float *xyz_host, int n, int octaves, float pers, float inv_maxamp, double scale, float *out_host);
}  // namespace gk
