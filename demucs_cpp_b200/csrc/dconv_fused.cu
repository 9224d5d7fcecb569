// Fused DConv residual branch for the wide frequency-branch levels (C = 48, 96): one CTA keeps one
// frequency row x[T=336][C] of one batch item resident in shared memory and runs BOTH sub-blocks on it
//     h = conv_k3_dil(x) ; g = GELU(GN1(h)) ; y = W2 g + b2 ; x += scale * GLU(GN2(y))
// (reference: demucscpp::apply_dconv, src/layers.cpp:152-375; v3 twin :723-894), so a DConv costs one HBM read
// and one write of x instead of ~12 passes over x / h / y through separate GEMM and pointwise launches.
// GroupNorm(1) statistics are exact block reductions over the row (a GroupNorm item IS a frequency row,
// src/layers.cpp:23-46), in double, unbiased variance.  y (2C wide) is never stored: it is computed once for
// the statistics and once more for the normalise-GLU-residual pass (6..24 MACs per value).
// All arithmetic fp32 on CUDA cores: these layers have K or N of 6..24 and are bandwidth / latency bound, not
// tensor-core shaped; used by both precision modes.
#include "kernels.cuh"

namespace dmx
{

constexpr int DF_T = LE;          // 336 frames
constexpr int DF_HALF = DF_T / 2; // each thread owns frames t and t + 168: every weight fetched from shared memory
constexpr int DF_THREADS = 192;   // feeds two positions (halves the shared-memory traffic per FMA); 168 live threads

// Packed parameter block per DConv (floats), for sub-block d = 0, 1 at offset d * stride:
//   W1[3][C][HID] | b1[HID] | g1w[HID] | g1b[HID] | W2[2C][HID] | b2[2C] | g2w[2C] | g2b[2C] | scale[C]
__host__ __device__ constexpr int df_stride(int C, int HID) { return 3 * C * HID + 3 * HID + 2 * C * HID + 6 * C + C; }

__device__ __forceinline__ void block_sum2(double &a, double &b, double *red)
{
    for (int o = 16; o > 0; o >>= 1)
    {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads(); // red may still be read from the previous reduction
    if (lane == 0)
    {
        red[2 * warp] = a;
        red[2 * warp + 1] = b;
    }
    __syncthreads();
    double sa = 0.0, sb = 0.0;
    for (int w = 0; w < DF_THREADS / 32; ++w)
    {
        sa += red[2 * w];
        sb += red[2 * w + 1];
    }
    a = sa;
    b = sb;
}

template <int C, int HID>
__global__ void __launch_bounds__(DF_THREADS) dconv_fused_kernel(float *__restrict__ x, const float *__restrict__ params)
{
    constexpr int LD = C + 1; // odd row stride: a warp reading one channel of 32 consecutive frames is conflict free
    constexpr int STRIDE = df_stride(C, HID);
    extern __shared__ __align__(16) float sm[];
    float *xs = sm;                      // [T][LD]
    float *ps = xs + DF_T * LD;          // parameters of the current sub-block
    __shared__ double red[2 * (DF_THREADS / 32)];

    const int tid = threadIdx.x;
    float *xg = x + (size_t)blockIdx.x * DF_T * C;
    for (int i = tid; i < DF_T * C / 4; i += DF_THREADS)
    {
        const float4 v = *reinterpret_cast<const float4 *>(xg + 4 * i);
        const int t = (4 * i) / C, c = (4 * i) % C;
        float *d = xs + t * LD + c;
        d[0] = v.x;
        d[1] = v.y;
        d[2] = v.z;
        d[3] = v.w;
    }
    const bool live = tid < DF_HALF;
    const int tp[2] = {tid, tid + DF_HALF};

#pragma unroll 1
    for (int sub = 0; sub < 2; ++sub)
    {
        const int dil = 1 << sub;
        __syncthreads(); // xs complete (initial load / previous sub-block's update); ps free
        for (int i = tid; i < STRIDE; i += DF_THREADS)
            ps[i] = __ldg(&params[sub * STRIDE + i]);
        __syncthreads();
        const float *W1 = ps, *b1 = W1 + 3 * C * HID, *g1w = b1 + HID, *g1b = g1w + HID;
        const float *W2 = g1b + HID, *b2 = W2 + 2 * C * HID, *g2w = b2 + 2 * C, *g2b = g2w + 2 * C, *scale = g2b + 2 * C;

        // ---- conv k3, dilation dil, zero padded: h[p][j] for this thread's two frames ----
        float h[2][HID];
#pragma unroll
        for (int p = 0; p < 2; ++p)
#pragma unroll
            for (int j = 0; j < HID; ++j)
                h[p][j] = b1[j];
        if (live)
        {
#pragma unroll
            for (int tap = 0; tap < 3; ++tap)
            {
                const int t0 = tp[0] + (tap - 1) * dil, t1 = tp[1] + (tap - 1) * dil;
                const bool v0 = t0 >= 0 && t0 < DF_T, v1 = t1 >= 0 && t1 < DF_T;
                const float *x0 = xs + (v0 ? t0 : 0) * LD, *x1 = xs + (v1 ? t1 : 0) * LD;
                const float *w = W1 + tap * C * HID;
#pragma unroll 4
                for (int c = 0; c < C; ++c)
                {
                    const float a0 = v0 ? x0[c] : 0.f, a1 = v1 ? x1[c] : 0.f;
#pragma unroll
                    for (int j = 0; j < HID; ++j)
                    {
                        const float wv = w[c * HID + j];
                        h[0][j] = fmaf(wv, a0, h[0][j]);
                        h[1][j] = fmaf(wv, a1, h[1][j]);
                    }
                }
            }
        }
        double s1 = 0.0, s2 = 0.0;
        if (live)
        {
            float a = 0.f, b = 0.f;
#pragma unroll
            for (int p = 0; p < 2; ++p)
#pragma unroll
                for (int j = 0; j < HID; ++j)
                {
                    a += h[p][j];
                    b += h[p][j] * h[p][j];
                }
            s1 = a;
            s2 = b;
        }
        block_sum2(s1, s2, red);
        {
            const double n = (double)DF_T * HID;
            const double m = s1 / n;
            double var = (s2 - s1 * m) / (n - 1.0);
            var = var < 0.0 ? 0.0 : var;
            const float mean = (float)m, rstd = (float)(1.0 / sqrt(var + 1e-5));
#pragma unroll
            for (int p = 0; p < 2; ++p)
#pragma unroll
                for (int j = 0; j < HID; ++j)
                    h[p][j] = gelu_erf((h[p][j] - mean) * rstd * g1w[j] + g1b[j]);
        }
        // ---- 1x1 conv to 2C channels, pass A: statistics only ----
        s1 = 0.0;
        s2 = 0.0;
        if (live)
        {
            float a = 0.f, b = 0.f;
#pragma unroll 4
            for (int n = 0; n < 2 * C; ++n)
            {
                float y0 = b2[n], y1 = y0;
#pragma unroll
                for (int j = 0; j < HID; ++j)
                {
                    const float wv = W2[n * HID + j];
                    y0 = fmaf(wv, h[0][j], y0);
                    y1 = fmaf(wv, h[1][j], y1);
                }
                a += y0 + y1;
                b += y0 * y0 + y1 * y1;
            }
            s1 = a;
            s2 = b;
        }
        block_sum2(s1, s2, red);
        // ---- pass B: recompute, GroupNorm, GLU, LayerScale, residual into the resident rows ----
        if (live)
        {
            const double n = (double)DF_T * 2 * C;
            const double m = s1 / n;
            double var = (s2 - s1 * m) / (n - 1.0);
            var = var < 0.0 ? 0.0 : var;
            const float mean = (float)m, rstd = (float)(1.0 / sqrt(var + 1e-5));
            float *xr0 = xs + tp[0] * LD, *xr1 = xs + tp[1] * LD;
#pragma unroll 2
            for (int c = 0; c < C; ++c)
            {
                float ya0 = b2[c], yg0 = b2[C + c], ya1 = ya0, yg1 = yg0;
#pragma unroll
                for (int j = 0; j < HID; ++j)
                {
                    const float wa = W2[c * HID + j], wg = W2[(C + c) * HID + j];
                    ya0 = fmaf(wa, h[0][j], ya0);
                    yg0 = fmaf(wg, h[0][j], yg0);
                    ya1 = fmaf(wa, h[1][j], ya1);
                    yg1 = fmaf(wg, h[1][j], yg1);
                }
                const float ga = rstd * g2w[c], gg = rstd * g2w[C + c], sc = scale[c];
                const float va0 = (ya0 - mean) * ga + g2b[c], vg0 = (yg0 - mean) * gg + g2b[C + c];
                const float va1 = (ya1 - mean) * ga + g2b[c], vg1 = (yg1 - mean) * gg + g2b[C + c];
                xr0[c] += sc * (va0 * sigmoid_ref(vg0));
                xr1[c] += sc * (va1 * sigmoid_ref(vg1));
            }
        }
    }
    __syncthreads();
    for (int i = tid; i < DF_T * C / 4; i += DF_THREADS)
    {
        const int tt = (4 * i) / C, c = (4 * i) % C;
        const float *s = xs + tt * LD + c;
        *reinterpret_cast<float4 *>(xg + 4 * i) = make_float4(s[0], s[1], s[2], s[3]);
    }
}

template <int C, int HID> static int df_smem() { return (DF_T * (C + 1) + df_stride(C, HID)) * (int)sizeof(float); }

template <int C, int HID> static int df_launch(float *x, const float *params, int items, cudaStream_t st)
{
    dconv_fused_kernel<C, HID><<<items, DF_THREADS, df_smem<C, HID>(), st>>>(x, params);
    DMX_LAUNCH_CHECK();
    return 0;
}

bool dconv_fused_supported(int C, int hid, int V)
{
    return V == DF_T && ((C == 48 && (hid == 6 || hid == 12)) || (C == 96 && (hid == 12 || hid == 24)));
}

int dconv_fused_param_floats(int C, int hid) { return 2 * df_stride(C, hid); }

int dconv_fused_init_device()
{
    DMX_CUDA(cudaFuncSetAttribute(dconv_fused_kernel<48, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, df_smem<48, 6>()));
    DMX_CUDA(cudaFuncSetAttribute(dconv_fused_kernel<48, 12>, cudaFuncAttributeMaxDynamicSharedMemorySize, df_smem<48, 12>()));
    DMX_CUDA(cudaFuncSetAttribute(dconv_fused_kernel<96, 12>, cudaFuncAttributeMaxDynamicSharedMemorySize, df_smem<96, 12>()));
    DMX_CUDA(cudaFuncSetAttribute(dconv_fused_kernel<96, 24>, cudaFuncAttributeMaxDynamicSharedMemorySize, df_smem<96, 24>()));
    return 0;
}

// x: [items][336][C] in place; items = B * F frequency rows
int launch_dconv_fused(float *x, const float *params, int items, int C, int hid, cudaStream_t st)
{
    if (C == 48 && hid == 6)
        return df_launch<48, 6>(x, params, items, st);
    if (C == 48 && hid == 12)
        return df_launch<48, 12>(x, params, items, st);
    if (C == 96 && hid == 12)
        return df_launch<96, 12>(x, params, items, st);
    if (C == 96 && hid == 24)
        return df_launch<96, 24>(x, params, items, st);
    set_error("dconv_fused: unsupported shape");
    return -1;
}

} // namespace dmx
