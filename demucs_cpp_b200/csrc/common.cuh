// Shared declarations for the demucs_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>

namespace dmx
{

// ---- model constants (reference: src/model.hpp:19-24,619-625; src/dsp.hpp:15-18) ----
constexpr int SEG = 343980;      // int(7.8 * 44100)
constexpr int SEG_STRIDE = 257985;
constexpr int MAX_SHIFT = 22050;
constexpr int NFFT = 4096;
constexpr int HOP = 1024;
constexpr int LE = 336;          // kept STFT frames
constexpr int NBINS = 2048;      // kept bins (Nyquist dropped)
constexpr int PAD = 1536;
constexpr int PAD_END = 1620;
constexpr int PADDED = SEG + PAD + PAD_END;      // 347136
constexpr int OLA_LEN = NFFT + HOP * (PADDED / HOP); // 340 frames -> 351232
constexpr int TLEN0 = 85995, TLEN1 = 21499, TLEN2 = 5375, TLEN3 = 1344;

void set_error(const std::string &msg);

#define DMX_CUDA(call)                                                              \
    do                                                                              \
    {                                                                               \
        cudaError_t e__ = (call);                                                   \
        if (e__ != cudaSuccess)                                                     \
        {                                                                           \
            char buf__[512];                                                        \
            snprintf(buf__, sizeof buf__, "%s:%d: %s: %s", __FILE__, __LINE__, #call, \
                     cudaGetErrorString(e__));                                      \
            ::dmx::set_error(buf__);                                                \
            return -1;                                                              \
        }                                                                           \
    } while (0)

#define DMX_LAUNCH_CHECK() DMX_CUDA(cudaGetLastError())

__device__ __forceinline__ float gelu_erf(float x)
{
    // exact-erf GELU (reference: src/conv.hpp:203-204)
    return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

__device__ __forceinline__ float sigmoid_ref(float v)
{
    // reference: src/layers.cpp:115-116
    return 1.0f / (1.0f + expf(-v));
}

// mean / rstd from double (sum, sumsq) with the reference's UNBIASED variance
// (src/layers.hpp:76-95) and eps inside the sqrt (src/layers.cpp:41,77).
__device__ __forceinline__ void stats_to_mean_rstd(const double *st, double count, float eps,
                                                   float &mean, float &rstd)
{
    double s = st[0], ss = st[1];
    double m = s / count;
    double var = (ss - s * m) / (count - 1.0);
    if (var < 0.0)
        var = 0.0;
    mean = (float)m;
    rstd = (float)(1.0 / sqrt(var + (double)eps));
}

} // namespace dmx
