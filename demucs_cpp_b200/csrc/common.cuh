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

__device__ __forceinline__ float fast_ex2(float x)
{
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_rcp(float x)
{
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ float gelu_erf(float x)
{
    // erf GELU, 0.5 x (1 + erf(x / sqrt 2)) (reference: src/conv.hpp:203-204), through
    // erfc(z) = t (a1 + t (a2 + t (a3 + t (a4 + t a5)))) exp(-z^2), t = 1 / (1 + p z), z >= 0  (Abramowitz-Stegun 7.1.26):
    //   x >= 0: x (1 - erfc(z) / 2)      x < 0: x erfc(z) / 2       z = |x| / sqrt 2
    // |error| <= 4.3e-7 absolute over the whole line -- the same as evaluating the formula with a float erff
    // (4.5e-7, both measured against float64) -- at a third of the instructions; the GELU sits in the epilogue
    // / prologue of the widest bandwidth-bound layers, where instruction issue is what limits them.
    const float z = fabsf(x) * 0.70710678118654752440f;
    const float t = fast_rcp(fmaf(0.3275911f, z, 1.0f));
    const float e = fast_ex2(-1.4426950408889634f * z * z);
    float q = fmaf(1.061405429f, t, -1.453152027f);
    q = fmaf(q, t, 1.421413741f);
    q = fmaf(q, t, -0.284496736f);
    q = fmaf(q, t, 0.254829592f);
    const float hq = 0.5f * q * t * e; // erfc(z) / 2
    return x * (x >= 0.f ? 1.0f - hq : hq);
}

__device__ __forceinline__ float sigmoid_ref(float v)
{
    // reference: src/layers.cpp:115-116; exp and reciprocal on the SFU (2 ulp each)
    return fast_rcp(1.0f + fast_ex2(-1.4426950408889634f * v));
}

// n / d for any 32-bit n through one 64-bit high multiply: mul = 2^64 / d rounded up (0 encodes d == 1)
__device__ __forceinline__ int fast_div(int n, unsigned long long mul)
{
    return mul ? (int)__umul64hi((unsigned long long)(unsigned)n, mul) : n;
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
