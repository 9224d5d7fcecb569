// HBM-bound normalisation / element-wise kernels of the segment forward.  Every variance is the
// reference's UNBIASED one, statistics are carried as double (sum, sumsq).
#include "kernels.cuh"

#include <cuda_bf16.h>

namespace dmx
{

// ------------------------------------------------------------------------------------------------
// sum / sumsq over n_per_item floats of each batch item.  grid (chunks, B).
__global__ void __launch_bounds__(256) sum_stats_kernel(const float *__restrict__ x, size_t n_per_item,
                                                        double *__restrict__ stats)
{
    const int b = blockIdx.y;
    const float *xb = x + (size_t)b * n_per_item;
    float s = 0.f, ss = 0.f;
    double ds = 0.0, dss = 0.0;
    int cnt = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_per_item; i += (size_t)gridDim.x * blockDim.x)
    {
        const float v = __ldg(&xb[i]);
        s += v;
        ss += v * v;
        if (++cnt == 64)
        { // spill fp32 partials into double regularly to bound rounding
            ds += s;
            dss += ss;
            s = ss = 0.f;
            cnt = 0;
        }
    }
    ds += s;
    dss += ss;
    for (int o = 16; o > 0; o >>= 1)
    {
        ds += __shfl_xor_sync(0xffffffffu, ds, o);
        dss += __shfl_xor_sync(0xffffffffu, dss, o);
    }
    __shared__ double red[2][8];
    if ((threadIdx.x & 31) == 0)
    {
        red[0][threadIdx.x >> 5] = ds;
        red[1][threadIdx.x >> 5] = dss;
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double a = 0, c = 0;
        for (int i = 0; i < 8; ++i)
        {
            a += red[0][i];
            c += red[1][i];
        }
        atomicAdd(&stats[2 * b + 0], a);
        atomicAdd(&stats[2 * b + 1], c);
    }
}

int launch_sum_stats(const float *x, int B, size_t n_per_item, double *stats, cudaStream_t st)
{
    int chunks = (int)((n_per_item + 256 * 16 - 1) / (256 * 16));
    if (chunks > 592)
        chunks = 592; // 4 CTAs per SM
    if (chunks < 1)
        chunks = 1;
    sum_stats_kernel<<<dim3(chunks, B), 256, 0, st>>>(x, n_per_item, stats);
    DMX_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// x = (x - mean) / (std + 1e-5), std = sqrt(sum((x-mean)^2)/(n-1))   (src/model_inference.cpp:115-144)
__global__ void __launch_bounds__(256) normalize_input_kernel(float *__restrict__ x, size_t n_per_item,
                                                              const double *__restrict__ stats,
                                                              float2 *__restrict__ meanstd)
{
    const int b = blockIdx.y;
    const double s = stats[2 * b], ss = stats[2 * b + 1];
    const double n = (double)n_per_item;
    const double m = s / n;
    double var = (ss - s * m) / (n - 1.0);
    if (var < 0)
        var = 0;
    const float mean = (float)m, stdv = (float)sqrt(var);
    if (blockIdx.x == 0 && threadIdx.x == 0)
        meanstd[b] = make_float2(mean, stdv);
    const float inv = 1.0f / (stdv + 1e-5f);
    float *xb = x + (size_t)b * n_per_item;
    const size_t n4 = n_per_item / 4;
    float4 *x4 = reinterpret_cast<float4 *>(xb);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
    {
        float4 v = x4[i];
        v.x = (v.x - mean) * inv;
        v.y = (v.y - mean) * inv;
        v.z = (v.z - mean) * inv;
        v.w = (v.w - mean) * inv;
        x4[i] = v;
    }
    for (size_t i = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_per_item;
         i += (size_t)gridDim.x * blockDim.x)
        xb[i] = (xb[i] - mean) * inv;
}

int launch_normalize_input(float *x, int B, size_t n_per_item, const double *stats, float *meanstd, cudaStream_t st)
{
    if ((n_per_item % 4) != 0 || (reinterpret_cast<uintptr_t>(x) & 15))
    {
        set_error("normalize_input: item size must be a multiple of 4 floats and 16-byte aligned");
        return -1;
    }
    int chunks = (int)((n_per_item / 4 + 256 * 4 - 1) / (256 * 4));
    if (chunks > 1184)
        chunks = 1184;
    if (chunks < 1)
        chunks = 1;
    normalize_input_kernel<<<dim3(chunks, B), 256, 0, st>>>(x, n_per_item, stats, (float2 *)meanstd);
    DMX_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// DConv tail (src/layers.cpp:240-253, 348-374): y is the raw 1x1-conv output [rows][2C]; GroupNorm(1)
// with per-item statistics, GLU over channels, LayerScale, residual add into x [rows][C].
__global__ void __launch_bounds__(256) gn_glu_residual_kernel(const float *__restrict__ y, float *__restrict__ x,
                                                              int rows_per_b, int V, int C, int U,
                                                              const double *__restrict__ stats, double count,
                                                              int per_u, const float *__restrict__ gw,
                                                              const float *__restrict__ gb,
                                                              const float *__restrict__ scale, size_t total4,
                                                              __nv_bfloat16 *__restrict__ img, int img_ld)
{
    const int c4n = C / 4;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (size_t)gridDim.x * blockDim.x)
    {
        const size_t row = i / c4n;
        const int c = (int)(i - row * c4n) * 4;
        const int b = (int)(row / rows_per_b);
        const int r = (int)(row - (size_t)b * rows_per_b);
        const int item = per_u ? b * U + r / V : b;
        float mean, rstd;
        stats_to_mean_rstd(stats + 2 * (size_t)item, count, 1e-5f, mean, rstd);
        const float4 a = __ldg(reinterpret_cast<const float4 *>(y + row * (size_t)(2 * C) + c));
        const float4 g = __ldg(reinterpret_cast<const float4 *>(y + row * (size_t)(2 * C) + C + c));
        const float4 wa = __ldg(reinterpret_cast<const float4 *>(gw + c)), ba = __ldg(reinterpret_cast<const float4 *>(gb + c));
        const float4 wg = __ldg(reinterpret_cast<const float4 *>(gw + C + c)), bg = __ldg(reinterpret_cast<const float4 *>(gb + C + c));
        const float4 sc = __ldg(reinterpret_cast<const float4 *>(scale + c));
        float4 xv = *reinterpret_cast<float4 *>(x + row * (size_t)C + c);
        xv.x += sc.x * (((a.x - mean) * rstd * wa.x + ba.x) * sigmoid_ref((g.x - mean) * rstd * wg.x + bg.x));
        xv.y += sc.y * (((a.y - mean) * rstd * wa.y + ba.y) * sigmoid_ref((g.y - mean) * rstd * wg.y + bg.y));
        xv.z += sc.z * (((a.z - mean) * rstd * wa.z + ba.z) * sigmoid_ref((g.z - mean) * rstd * wg.z + bg.z));
        xv.w += sc.w * (((a.w - mean) * rstd * wa.w + ba.w) * sigmoid_ref((g.w - mean) * rstd * wg.w + bg.w));
        *reinterpret_cast<float4 *>(x + row * (size_t)C + c) = xv;
        if (img)
        {
            // the finished values also leave as the bf16 channels-last image the next convolution's tensor-map loads read
            __nv_bfloat162 lo = __floats2bfloat162_rn(xv.x, xv.y), hi = __floats2bfloat162_rn(xv.z, xv.w);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t *>(&lo);
            pk.y = *reinterpret_cast<uint32_t *>(&hi);
            *reinterpret_cast<uint2 *>(img + row * (size_t)img_ld + c) = pk;
        }
    }
}

int launch_gn_glu_residual(const float *y, float *x, int B, int U, int V, int C, const double *stats, int per_u,
                           const float *gw, const float *gb, const float *scale, cudaStream_t st, void *img, int img_ld)
{
    if (C % 4)
    {
        set_error("gn_glu_residual: C % 4 != 0");
        return -1;
    }
    const size_t rows = (size_t)B * U * V;
    const size_t total4 = rows * (C / 4);
    const double count = per_u ? (double)V * 2 * C : (double)U * V * 2 * C;
    size_t blocks = (total4 + 255) / 256;
    if (blocks > 148 * 16)
        blocks = 148 * 16;
    gn_glu_residual_kernel<<<(unsigned)blocks, 256, 0, st>>>(y, x, U * V, V, C, U, stats, count, per_u, gw, gb, scale,
                                                             total4, static_cast<__nv_bfloat16 *>(img), img_ld);
    DMX_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// LayerNorm over C with unbiased variance (src/layers.cpp:121-150); one warp per row, C <= 512,
// optional additive table (positional embedding, src/crosstransformer.cpp:244-263).
__global__ void __launch_bounds__(256) layer_norm_kernel(const float *__restrict__ x, float *__restrict__ y, int rows,
                                                         int C, const float *__restrict__ w, const float *__restrict__ b,
                                                         const float *__restrict__ add, int add_rows)
{
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows)
        return;
    const float *xr = x + (size_t)row * C;
    float v[16];
    const int per = C / 32; // <= 16
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i)
        if (i < per)
        {
            v[i] = __ldg(&xr[lane + 32 * i]);
            s += v[i];
        }
    for (int o = 16; o > 0; o >>= 1)
        s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s / (float)C;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i)
        if (i < per)
        {
            const float d = v[i] - mean;
            q += d * d;
        }
    for (int o = 16; o > 0; o >>= 1)
        q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = 1.0f / sqrtf(q / (float)(C - 1) + 1e-5f);
    float *yr = y + (size_t)row * C;
    const float *ar = add ? add + (size_t)(row % add_rows) * C : nullptr;
#pragma unroll
    for (int i = 0; i < 16; ++i)
        if (i < per)
        {
            const int c = lane + 32 * i;
            float o = (v[i] - mean) * rstd * __ldg(&w[c]) + __ldg(&b[c]);
            if (ar)
                o += __ldg(&ar[c]);
            yr[c] = o;
        }
}

int launch_layer_norm(const float *x, float *y, int rows, int C, const float *w, const float *b, const float *add,
                      int add_rows, cudaStream_t st)
{
    if (C % 32 || C > 512)
    {
        set_error("layer_norm: C must be a multiple of 32 and <= 512");
        return -1;
    }
    layer_norm_kernel<<<(rows + 7) / 8, 256, 0, st>>>(x, y, rows, C, w, b, add, add_rows > 0 ? add_rows : 1);
    DMX_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// LayerNorm whose output is the tcgen05 GEMM's A-operand image: bf16, tile-packed [row/128][C/8][row%128][8].
// One CTA = 32 consecutive rows (never straddling a 128-row tile).  Each warp normalises 4 rows (lane l owns the
// 8-channel chunks l, l+32; C <= 512) into a shared tile [chunk][32 rows][16 B]; the tile is then written out one
// chunk per warp pass, lane = row: 512 contiguous bytes per store instead of 32 scattered 16-byte pieces.
constexpr int LNP_ROWS = 32;
constexpr int LNP_CHS = LNP_ROWS * 16 + 16; // chunk stride in bytes, padded: conflict-free 16-byte stores by lane = chunk
__global__ void __launch_bounds__(256) layer_norm_pack_kernel(const float *__restrict__ x, uint4 *__restrict__ ypk, int rows,
                                                              int C, const float *__restrict__ w, const float *__restrict__ b)
{
    __shared__ __align__(16) uint8_t tile[64 * LNP_CHS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row0 = blockIdx.x * LNP_ROWS;
    const int nch = C / 8;
#pragma unroll 1
    for (int rr = 0; rr < 4; ++rr)
    {
        const int rl = warp * 4 + rr, row = row0 + rl;
        if (row >= rows)
            break;
        const float *xr = x + (size_t)row * C;
        float v[2][8];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 2; ++i)
        {
            const int ch = lane + 32 * i;
            if (ch < nch)
            {
                const float4 a = __ldg(reinterpret_cast<const float4 *>(xr + ch * 8));
                const float4 c = __ldg(reinterpret_cast<const float4 *>(xr + ch * 8 + 4));
                v[i][0] = a.x; v[i][1] = a.y; v[i][2] = a.z; v[i][3] = a.w;
                v[i][4] = c.x; v[i][5] = c.y; v[i][6] = c.z; v[i][7] = c.w;
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    s += v[i][e];
            }
        }
        for (int o = 16; o > 0; o >>= 1)
            s += __shfl_xor_sync(0xffffffffu, s, o);
        const float mean = s / (float)C;
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < 2; ++i)
            if (lane + 32 * i < nch)
            {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                {
                    const float d = v[i][e] - mean;
                    q += d * d;
                }
            }
        for (int o = 16; o > 0; o >>= 1)
            q += __shfl_xor_sync(0xffffffffu, q, o);
        const float rstd = 1.0f / sqrtf(q / (float)(C - 1) + 1e-5f);
#pragma unroll
        for (int i = 0; i < 2; ++i)
        {
            const int ch = lane + 32 * i;
            if (ch < nch)
            {
                const float4 w0 = __ldg(reinterpret_cast<const float4 *>(w + ch * 8)), w1 = __ldg(reinterpret_cast<const float4 *>(w + ch * 8 + 4));
                const float4 b0 = __ldg(reinterpret_cast<const float4 *>(b + ch * 8)), b1 = __ldg(reinterpret_cast<const float4 *>(b + ch * 8 + 4));
                const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
                const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                float o8[8];
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    o8[e] = (v[i][e] - mean) * rstd * ww[e] + bb[e];
                uint4 pk;
                __nv_bfloat162 h0 = __floats2bfloat162_rn(o8[0], o8[1]), h1 = __floats2bfloat162_rn(o8[2], o8[3]),
                               h2 = __floats2bfloat162_rn(o8[4], o8[5]), h3 = __floats2bfloat162_rn(o8[6], o8[7]);
                pk.x = *reinterpret_cast<uint32_t *>(&h0);
                pk.y = *reinterpret_cast<uint32_t *>(&h1);
                pk.z = *reinterpret_cast<uint32_t *>(&h2);
                pk.w = *reinterpret_cast<uint32_t *>(&h3);
                *reinterpret_cast<uint4 *>(tile + ch * LNP_CHS + rl * 16) = pk;
            }
        }
    }
    __syncthreads();
    // row0 is a multiple of 32: the 32 rows share one 128-row tile, at rows (row0 & 127) .. +31 of it
    const int row = row0 + lane;
    for (int ch = warp; ch < nch; ch += 8)
        if (row < rows)
            ypk[((size_t)(row0 >> 7) * nch + ch) * 128 + (row & 127)] = *reinterpret_cast<const uint4 *>(tile + ch * LNP_CHS + lane * 16);
}

int launch_layer_norm_pack(const float *x, void *ypk, int rows, int C, const float *w, const float *b, cudaStream_t st)
{
    if (C % 8 || C > 512)
    {
        set_error("layer_norm_pack: C must be a multiple of 8 and <= 512");
        return -1;
    }
    layer_norm_pack_kernel<<<(rows + LNP_ROWS - 1) / LNP_ROWS, 256, 0, st>>>(x, (uint4 *)ypk, rows, C, w, b);
    DMX_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// MyGroupNorm (one group over (T, C) per item, src/layers.cpp:517-530) applied in place.
__global__ void __launch_bounds__(256) group_norm_apply_kernel(float *__restrict__ x, int rows_per_item, int C,
                                                               const double *__restrict__ stats,
                                                               const float *__restrict__ w, const float *__restrict__ b,
                                                               size_t total4)
{
    const int c4n = C / 4;
    const double count = (double)rows_per_item * C;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (size_t)gridDim.x * blockDim.x)
    {
        const size_t row = i / c4n;
        const int c = (int)(i - row * c4n) * 4;
        const int item = (int)(row / rows_per_item);
        float mean, rstd;
        stats_to_mean_rstd(stats + 2 * (size_t)item, count, 1e-5f, mean, rstd);
        float4 v = *reinterpret_cast<float4 *>(x + row * (size_t)C + c);
        const float4 ww = __ldg(reinterpret_cast<const float4 *>(w + c)), bb = __ldg(reinterpret_cast<const float4 *>(b + c));
        v.x = (v.x - mean) * rstd * ww.x + bb.x;
        v.y = (v.y - mean) * rstd * ww.y + bb.y;
        v.z = (v.z - mean) * rstd * ww.z + bb.z;
        v.w = (v.w - mean) * rstd * ww.w + bb.w;
        *reinterpret_cast<float4 *>(x + row * (size_t)C + c) = v;
    }
}

int launch_group_norm_apply(float *x, int B, int rows_per_item, int C, const double *stats, const float *w,
                            const float *b, cudaStream_t st)
{
    const size_t total4 = (size_t)B * rows_per_item * (C / 4);
    size_t blocks = (total4 + 255) / 256;
    if (blocks > 148 * 16)
        blocks = 148 * 16;
    group_norm_apply_kernel<<<(unsigned)blocks, 256, 0, st>>>(x, rows_per_item, C, stats, w, b, total4);
    DMX_LAUNCH_CHECK();
    return 0;
}

} // namespace dmx
