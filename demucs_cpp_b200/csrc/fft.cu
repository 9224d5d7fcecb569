// STFT / iSTFT for the 7.8 s segment forward, as shared-memory radix-16 Stockham FFTs.
//
// Reference semantics reproduced here (see oracle/demucs_oracle.py for the CPU restatement):
//   * reflect_padding 1536/1620 with the edge sample included  (src/model_inference.cpp:22-46)
//   * demucscpp::stft: hop 1024, periodic Hann, rFFT-4096, 1/sqrt(4096); frames 2..337 and
//     bins 0..2047 kept, complex-as-channels packing            (src/dsp.cpp:51-86,119-149;
//                                                                src/model_inference.cpp:75-99)
//   * demucscpp::istft: sqrt(4096) * unscaled inverse, window / 4096 / (wss + 1e-8) overlap-add
//                                                               (src/dsp.cpp:88-117,151-185)
//   * per-source reassembly: x*std+mean, zero Nyquist bin, two zero frames each side, crop 1536,
//     plus de-normalised time branch                            (src/model_inference.cpp:380-474)
//
// Both stereo channels ride one complex FFT (z = L + iR); the two real spectra are separated by
// Hermitian symmetry.  Layouts: mix [B][SEG][2]; spectrogram activations [B][F=2048][T=336][C].
#include "common.cuh"
#include "kernels.cuh"

namespace dmx
{

constexpr int FFT_THREADS = 256;
constexpr int FFT_YLEN = NFFT + NFFT / 16;                         // second buffer: padded layout of the first pass
constexpr int FFT_SMEM = (NFFT + FFT_YLEN) * (int)sizeof(float2); // ping-pong, 66 KB
constexpr int ISTFT_SMEM = (2 * NFFT + FFT_YLEN) * (int)sizeof(float2); // two sources' spectra + the second buffer

__device__ __forceinline__ float2 cmul(float2 a, float2 b)
{
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

// ------------------------------------------------------------------------------------------------
// 4096-point complex FFT over shared memory: three radix-16 Stockham passes, one 16-point butterfly per thread per
// pass, the butterfly in registers (4 x 4 decomposition), two block barriers.  tw[k] = exp(-2*pi*i*k/4096); INV
// conjugates the twiddles and the +-i rotations (unscaled inverse).
//   pass A: sub-length 4096, stride 1    x[p + 256 k]            -> y[pad(16 p + k)]      * w^(k p)
//   pass B: sub-length  256, stride 16   y[pad(q + 16 p + 256 k)] -> x[q + 256 p + 16 k]   * w^(16 k p)
//   pass C: sub-length   16, stride 256  x[q + 256 k]            -> y[q + 256 k]
// pad(i) = i + i/16 keeps pass A's stride-16 stores conflict free.  The result ends in y (natural order, unpadded).
__device__ __forceinline__ int fft_pad(int i) { return i + (i >> 4); }

template <bool INV> __device__ __forceinline__ float2 mul_i(float2 a) // forward: -i a ; inverse: +i a
{
    return INV ? make_float2(-a.y, a.x) : make_float2(a.y, -a.x);
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }

// 4-point DFT: X1 = a0 -+ i a1 - a2 +- i a3 (upper sign forward)
template <bool INV> __device__ __forceinline__ void dft4(float2 a0, float2 a1, float2 a2, float2 a3, float2 &X0, float2 &X1, float2 &X2, float2 &X3)
{
    const float2 s02 = cadd(a0, a2), d02 = csub(a0, a2), s13 = cadd(a1, a3), d13 = mul_i<INV>(csub(a1, a3));
    X0 = cadd(s02, s13);
    X2 = csub(s02, s13);
    X1 = cadd(d02, d13);
    X3 = csub(d02, d13);
}

template <bool INV> __device__ __forceinline__ void dft16(float2 (&v)[16])
{
    // w16^m = exp(-+ 2 pi i m / 16), m = j2 * k1
    constexpr float C1 = 0.92387953251128674f, S1 = 0.38268343236508977f, C2 = 0.70710678118654752f;
    float2 a[4][4];
#pragma unroll
    for (int j2 = 0; j2 < 4; ++j2)
        dft4<INV>(v[j2], v[4 + j2], v[8 + j2], v[12 + j2], a[j2][0], a[j2][1], a[j2][2], a[j2][3]);
    auto tw = [](float2 z, float c, float sn) { // z * (c - i sn) forward, z * (c + i sn) inverse
        const float si = INV ? sn : -sn;
        return make_float2(z.x * c - z.y * si, z.x * si + z.y * c);
    };
    a[1][1] = tw(a[1][1], C1, S1);
    a[1][2] = tw(a[1][2], C2, C2);
    a[1][3] = tw(a[1][3], S1, C1);
    a[2][1] = tw(a[2][1], C2, C2);
    a[2][2] = mul_i<INV>(a[2][2]);          // m = 4
    a[2][3] = tw(a[2][3], -C2, C2);         // m = 6
    a[3][1] = tw(a[3][1], S1, C1);          // m = 3
    a[3][2] = tw(a[3][2], -C2, C2);         // m = 6
    a[3][3] = tw(a[3][3], -C1, -S1);        // m = 9
#pragma unroll
    for (int k1 = 0; k1 < 4; ++k1)
        dft4<INV>(a[0][k1], a[1][k1], a[2][k1], a[3][k1], v[k1], v[k1 + 4], v[k1 + 8], v[k1 + 12]);
}

// x: NFFT entries (input, natural order); y: FFT_YLEN entries; result in y[0..NFFT).  Caller syncs before and after.
template <bool INV> __device__ __forceinline__ void fft4096_r16(float2 *x, float2 *y, const float2 *__restrict__ tw)
{
    static_assert(FFT_THREADS == NFFT / 16, "one radix-16 butterfly per thread per pass");
    const int t = threadIdx.x;
    float2 v[16];
    // ---- pass A ----
#pragma unroll
    for (int k = 0; k < 16; ++k)
        v[k] = x[t + 256 * k];
    dft16<INV>(v);
#pragma unroll
    for (int k = 0; k < 16; ++k)
    {
        float2 o = v[k];
        if (k > 0)
        {
            float2 w = __ldg(&tw[k * t]);
            if (INV)
                w.y = -w.y;
            o = cmul(w, o);
        }
        y[fft_pad(16 * t + k)] = o;
    }
    __syncthreads();
    // ---- pass B ----
    {
        const int p = t >> 4, q = t & 15;
#pragma unroll
        for (int k = 0; k < 16; ++k)
            v[k] = y[fft_pad(q + 16 * p + 256 * k)];
        dft16<INV>(v);
        // x (pass A's input) may be overwritten now: every thread finished pass A before the barrier above
#pragma unroll
        for (int k = 0; k < 16; ++k)
        {
            float2 o = v[k];
            if (k > 0 && p > 0)
            {
                float2 w = __ldg(&tw[16 * k * p]);
                if (INV)
                    w.y = -w.y;
                o = cmul(w, o);
            }
            x[q + 256 * p + 16 * k] = o;
        }
    }
    __syncthreads();
    // ---- pass C ----
#pragma unroll
    for (int k = 0; k < 16; ++k)
        v[k] = x[t + 256 * k];
    dft16<INV>(v);
#pragma unroll
    for (int k = 0; k < 16; ++k)
        y[t + 256 * k] = v[k];
}

// ------------------------------------------------------------------------------------------------
// STFT: grid (336 frames, B).  Writes raw CaC spectrogram spec[b][f][t][4] = (ReL, ImL, ReR, ImR)
// and accumulates sum / sum-of-squares per batch item into stats[b][2] (double) for the global
// normalisation that follows (src/model_inference.cpp:115-124).
__global__ void __launch_bounds__(FFT_THREADS) stft_kernel(const float2 *__restrict__ mix, // [B][SEG]
                                                           const float *__restrict__ window, // [4096]
                                                           const float2 *__restrict__ tw,    // [4096]
                                                           float4 *__restrict__ spec,        // [B][2048][336]
                                                           double *__restrict__ stats)       // [B][2]
{
    extern __shared__ float2 sm[];
    float2 *x = sm, *y = sm + NFFT;
    const int t = blockIdx.x, b = blockIdx.y;
    const float2 *src = mix + (size_t)b * SEG;
    // kept frame t is STFT frame j = t + 2; its first sample sits at index j*1024 - 2048 of the
    // 1536/1620-padded segment, i.e. at (t*1024 - 1536) of the segment before mirroring.
    const int base = t * HOP - PAD;
    for (int n = threadIdx.x; n < NFFT; n += FFT_THREADS)
    {
        int i = base + n;
        if (i < 0)
            i = -1 - i; // edge-including mirror
        else if (i >= SEG)
            i = 2 * SEG - 1 - i;
        const float2 v = __ldg(&src[i]);
        const float w = __ldg(&window[n]);
        x[n] = make_float2(v.x * w, v.y * w);
    }
    __syncthreads();
    fft4096_r16<false>(x, y, tw);
    __syncthreads();
    const float sc = 1.0f / 64.0f; // 1/sqrt(4096)
    float s1 = 0.f, s2 = 0.f;
    for (int k = threadIdx.x; k < NBINS; k += FFT_THREADS)
    {
        const float2 zk = y[k];
        const float2 zm = y[(NFFT - k) & (NFFT - 1)];
        // XL = (Z[k] + conj(Z[N-k]))/2 ; XR = (Z[k] - conj(Z[N-k]))/(2i)
        float4 o;
        o.x = 0.5f * (zk.x + zm.x) * sc;
        o.y = 0.5f * (zk.y - zm.y) * sc;
        o.z = 0.5f * (zk.y + zm.y) * sc;
        o.w = -0.5f * (zk.x - zm.x) * sc;
        spec[((size_t)b * NBINS + k) * LE + t] = o;
        s1 += o.x + o.y + o.z + o.w;
        s2 += o.x * o.x + o.y * o.y + o.z * o.z + o.w * o.w;
    }
    // block reduce -> one double atomic pair per CTA
    __shared__ double red[2][FFT_THREADS / 32];
    double d1 = s1, d2 = s2;
    for (int o = 16; o > 0; o >>= 1)
    {
        d1 += __shfl_xor_sync(0xffffffffu, d1, o);
        d2 += __shfl_xor_sync(0xffffffffu, d2, o);
    }
    if ((threadIdx.x & 31) == 0)
    {
        red[0][threadIdx.x >> 5] = d1;
        red[1][threadIdx.x >> 5] = d2;
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double a = 0, c = 0;
        for (int i = 0; i < FFT_THREADS / 32; ++i)
        {
            a += red[0][i];
            c += red[1][i];
        }
        atomicAdd(&stats[2 * b + 0], a);
        atomicAdd(&stats[2 * b + 1], c);
    }
}

// ------------------------------------------------------------------------------------------------
// iSTFT stage 1: grid (336 frames, S sources, B).  Reads the freq-branch output
// xout[b][f][t][4*S] (channels s*4 + {ReL,ImL,ReR,ImR}), de-normalises with the per-item
// (mean, std) of the input spectrogram, runs the inverse FFT and writes the windowed frame
// (scaled by 64/4096, not yet divided by the window-sum-square) to frames[b][s][t][4096] (L,R).
// A CTA handles a PAIR of sources: for one bin the pair's 8 values are one whole 32-byte sector of the
// (bins x frames x channels) tensor, whose bins lie 21 KB apart -- the kernel is bound by the rate of those
// scattered sector requests, and a single source would use only half of every sector it pulls.
__global__ void __launch_bounds__(FFT_THREADS) istft_frames_kernel(const float *__restrict__ xout, int S,
                                                                   const float2 *__restrict__ meanstd, // [B] (mean,std)
                                                                   const float *__restrict__ window,
                                                                   const float2 *__restrict__ tw,
                                                                   float2 *__restrict__ frames, const size_t sk, const size_t st_)
{
    extern __shared__ float2 sm[];
    float2 *xa = sm, *xb = sm + NFFT, *y = sm + 2 * NFFT;
    const int t = blockIdx.x, s0 = 2 * blockIdx.y, b = blockIdx.z;
    const float2 ms = meanstd[b];
    const int C = 4 * S;
    const bool pair = s0 + 1 < S; // an odd source count leaves the last CTA row with one source
    // the eight bins of a thread are 21 KB apart in memory: issue all loads before touching any of them
    float4 va[NBINS / FFT_THREADS], vb[NBINS / FFT_THREADS];
#pragma unroll
    for (int i = 0; i < NBINS / FFT_THREADS; ++i)
    {
        // frame-major input (production): the bins of frame t are 64 contiguous bytes apart, a warp reads 2 KB in one piece;
        // bin-major (the stand-alone dmx_istft_segment): 21 KB apart
        const float *src = xout + (size_t)b * NBINS * LE * C + (size_t)(threadIdx.x + i * FFT_THREADS) * sk + (size_t)t * st_ + 4 * s0;
        va[i] = __ldg(reinterpret_cast<const float4 *>(src));
        vb[i] = pair ? __ldg(reinterpret_cast<const float4 *>(src + 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < NBINS / FFT_THREADS; ++i)
    {
        const int k = threadIdx.x + i * FFT_THREADS;
#pragma unroll
        for (int h = 0; h < 2; ++h)
        {
            const float4 v = h ? vb[i] : va[i];
            float2 *x = h ? xb : xa;
            const float lr = v.x * ms.y + ms.x, li = v.y * ms.y + ms.x;
            const float rr = v.z * ms.y + ms.x, ri = v.w * ms.y + ms.x;
            if (k == 0)
            {
                // real inverse ignores the imaginary part of DC; Nyquist bin is zero-padded
                x[0] = make_float2(lr, rr);
                x[NFFT / 2] = make_float2(0.f, 0.f);
            }
            else
            {
                // Z[k] = XL + i XR ; Z[N-k] = conj(XL) + i conj(XR)
                x[k] = make_float2(lr - ri, li + rr);
                x[NFFT - k] = make_float2(lr + ri, rr - li);
            }
        }
    }
    __syncthreads();
    const float sc = 64.0f / 4096.0f; // sqrt(4096) / 4096
#pragma unroll 1
    for (int h = 0; h < (pair ? 2 : 1); ++h)
    {
        fft4096_r16<true>(h ? xb : xa, y, tw);
        __syncthreads();
        float2 *dst = frames + (((size_t)b * S + s0 + h) * LE + t) * NFFT;
        for (int n = threadIdx.x; n < NFFT; n += FFT_THREADS)
        {
            const float w = __ldg(&window[n]) * sc;
            const float2 v = y[n];
            dst[n] = make_float2(v.x * w, v.y * w);
        }
        __syncthreads(); // y is rewritten by the second transform
    }
}

// iSTFT stage 2: overlap-add the four frames covering each sample, each divided by (wss + 1e-8)
// in increasing frame order like the reference's accumulation loop, crop, and add the
// de-normalised time branch.  out[b][s][n] (L,R).  xt_out: [B][SEG][2*S] channel s*2 + ch.
__global__ void __launch_bounds__(256) istft_ola_kernel(const float2 *__restrict__ frames, int S,
                                                        const float *__restrict__ wss,       // [OLA_LEN]
                                                        const float *__restrict__ xt_out,
                                                        const float2 *__restrict__ meanstd_t, // [B]
                                                        float2 *__restrict__ out)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const int s = blockIdx.y, b = blockIdx.z;
    if (n >= SEG)
        return;
    const int q = n + PAD + NFFT / 2; // index in the 351232-long OLA buffer
    const float inv = 1.0f / (__ldg(&wss[q]) + 1e-8f);
    const int jhi = q / HOP;
    float2 acc = make_float2(0.f, 0.f);
    const float2 *fb = frames + ((size_t)b * S + s) * LE * (size_t)NFFT;
#pragma unroll
    for (int d = 3; d >= 0; --d)
    {
        const int j = jhi - d;  // STFT frame number, nonzero frames are 2..337
        const int t = j - 2;
        if (t >= 0 && t < LE)
        {
            const float2 v = __ldg(&fb[(size_t)t * NFFT + (q - j * HOP)]);
            acc.x += v.x * inv;
            acc.y += v.y * inv;
        }
    }
    const float2 ms = meanstd_t[b];
    const float2 xt = *reinterpret_cast<const float2 *>(xt_out + ((size_t)b * SEG + n) * (2 * S) + 2 * s);
    acc.x += xt.x * ms.y + ms.x;
    acc.y += xt.y * ms.y + ms.x;
    out[((size_t)b * S + s) * SEG + n] = acc;
}

// ------------------------------------------------------------------------------------------------
// Opts the FFT kernels into 64 KB of dynamic shared memory; call once per device with it current.
int fft_init_device()
{
    DMX_CUDA(cudaFuncSetAttribute(stft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FFT_SMEM));
    DMX_CUDA(cudaFuncSetAttribute(istft_frames_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ISTFT_SMEM));
    return 0;
}

int launch_stft(const float *mix, int B, const float *window, const float *tw, float *spec, double *stats,
                cudaStream_t st)
{
    stft_kernel<<<dim3(LE, B), FFT_THREADS, FFT_SMEM, st>>>((const float2 *)mix, window, (const float2 *)tw,
                                                            (float4 *)spec, stats);
    DMX_LAUNCH_CHECK();
    return 0;
}

int launch_istft(const float *xout, int B, int S, const float *meanstd, const float *window, const float *tw,
                 const float *wss, const float *xt_out, const float *meanstd_t, float *frames_scratch,
                 float *out, cudaStream_t st, int frame_major)
{
    const size_t C = 4 * (size_t)S;
    istft_frames_kernel<<<dim3(LE, (S + 1) / 2, B), FFT_THREADS, ISTFT_SMEM, st>>>(
        xout, S, (const float2 *)meanstd, window, (const float2 *)tw, (float2 *)frames_scratch, frame_major ? C : (size_t)LE * C,
        frame_major ? (size_t)NBINS * C : C);
    DMX_LAUNCH_CHECK();
    istft_ola_kernel<<<dim3((SEG + 255) / 256, S, B), 256, 0, st>>>(
        (const float2 *)frames_scratch, S, wss, xt_out, (const float2 *)meanstd_t, (float2 *)out);
    DMX_LAUNCH_CHECK();
    return 0;
}

} // namespace dmx
