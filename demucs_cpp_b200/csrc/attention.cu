// Multi-head attention softmax(Q K^T / sqrt(dh)) V, streaming-softmax (flash) formulation so the
// T x S score matrix (231 MB per layer per segment at T=S=2688, 8 heads) never reaches HBM.
// fp32 CUDA-core variant for the full-precision mode.  Reference: src/layers.cpp:442-482
// (heads are contiguous dh-wide column blocks; max-subtracted softmax over keys).
#include "kernels.cuh"

namespace dmx
{

constexpr int AT_BQ = 64; // queries per CTA
constexpr int AT_BK = 64; // keys per tile
constexpr int AT_THREADS = 256;

template <int DH>
__global__ void __launch_bounds__(AT_THREADS) attention_kernel(const float *__restrict__ Q, int ldq,
                                                               const float *__restrict__ K, int ldk,
                                                               const float *__restrict__ V, int ldv,
                                                               float *__restrict__ O, int ldo, int T, int S,
                                                               float scale)
{
    constexpr int DC = DH / 16; // output columns per thread (4 for dh=64, 3 for dh=48)
    extern __shared__ __align__(16) float smem[];
    float *Qt = smem;                  // [DH][AT_BQ]
    float *Kt = Qt + DH * AT_BQ;       // [DH][AT_BK]
    float *Vs = Kt + DH * AT_BK;       // [AT_BK][DH]
    float *Ps = Vs + AT_BK * DH;       // [AT_BQ][AT_BK]

    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int q0 = blockIdx.x * AT_BQ;
    const int h = blockIdx.y, b = blockIdx.z;
    const float *Qb = Q + (size_t)b * T * ldq + h * DH;
    const float *Kb = K + (size_t)b * S * ldk + h * DH;
    const float *Vb = V + (size_t)b * S * ldv + h * DH;

    // load the query tile transposed: lane -> query row keeps shared stores conflict free
    for (int idx = tid; idx < AT_BQ * (DH / 4); idx += AT_THREADS)
    {
        const int r = idx % AT_BQ, dq = idx / AT_BQ;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (q0 + r < T)
            v = __ldg(reinterpret_cast<const float4 *>(Qb + (size_t)(q0 + r) * ldq + dq * 4));
        Qt[(dq * 4 + 0) * AT_BQ + r] = v.x;
        Qt[(dq * 4 + 1) * AT_BQ + r] = v.y;
        Qt[(dq * 4 + 2) * AT_BQ + r] = v.z;
        Qt[(dq * 4 + 3) * AT_BQ + r] = v.w;
    }

    float m_run[4], l_run[4], o[4][DC];
#pragma unroll
    for (int i = 0; i < 4; ++i)
    {
        m_run[i] = -INFINITY;
        l_run[i] = 0.f;
#pragma unroll
        for (int c = 0; c < DC; ++c)
            o[i][c] = 0.f;
    }

    for (int k0 = 0; k0 < S; k0 += AT_BK)
    {
        __syncthreads(); // previous tile fully consumed (also orders the Qt stores on the first pass)
        for (int idx = tid; idx < AT_BK * (DH / 4); idx += AT_THREADS)
        {
            const int r = idx % AT_BK, dq = idx / AT_BK;
            float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
            if (k0 + r < S)
            {
                kv = __ldg(reinterpret_cast<const float4 *>(Kb + (size_t)(k0 + r) * ldk + dq * 4));
                vv = __ldg(reinterpret_cast<const float4 *>(Vb + (size_t)(k0 + r) * ldv + dq * 4));
            }
            Kt[(dq * 4 + 0) * AT_BK + r] = kv.x;
            Kt[(dq * 4 + 1) * AT_BK + r] = kv.y;
            Kt[(dq * 4 + 2) * AT_BK + r] = kv.z;
            Kt[(dq * 4 + 3) * AT_BK + r] = kv.w;
            *reinterpret_cast<float4 *>(&Vs[r * DH + dq * 4]) = vv;
        }
        __syncthreads();

        // scores: rows 4ty..4ty+3, keys 4tx..4tx+3
        float s[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j)
                s[i][j] = 0.f;
#pragma unroll 8
        for (int d = 0; d < DH; ++d)
        {
            const float4 a = *reinterpret_cast<const float4 *>(&Qt[d * AT_BQ + 4 * ty]);
            const float4 k = *reinterpret_cast<const float4 *>(&Kt[d * AT_BK + 4 * tx]);
            const float av[4] = {a.x, a.y, a.z, a.w};
            const float kv[4] = {k.x, k.y, k.z, k.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    s[i][j] = fmaf(av[i], kv[j], s[i][j]);
        }
        float corr[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
        {
            float mx = -INFINITY;
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                s[i][j] = (k0 + 4 * tx + j < S) ? s[i][j] * scale : -INFINITY;
                mx = fmaxf(mx, s[i][j]);
            }
#pragma unroll
            for (int o2 = 8; o2 > 0; o2 >>= 1)
                mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o2));
            const float m_new = fmaxf(m_run[i], mx);
            corr[i] = (m_run[i] == -INFINITY) ? 0.f : expf(m_run[i] - m_new);
            float rs = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                const float pv = (s[i][j] == -INFINITY) ? 0.f : expf(s[i][j] - m_new);
                s[i][j] = pv;
                rs += pv;
            }
#pragma unroll
            for (int o2 = 8; o2 > 0; o2 >>= 1)
                rs += __shfl_xor_sync(0xffffffffu, rs, o2);
            l_run[i] = l_run[i] * corr[i] + rs;
            m_run[i] = m_new;
            *reinterpret_cast<float4 *>(&Ps[(4 * ty + i) * AT_BK + 4 * tx]) = make_float4(s[i][0], s[i][1], s[i][2], s[i][3]);
#pragma unroll
            for (int c = 0; c < DC; ++c)
                o[i][c] *= corr[i];
        }
        __syncthreads();
        // O[4ty+i][tx*DC + c] += sum_j P[4ty+i][j] * V[j][tx*DC + c]
#pragma unroll 4
        for (int j = 0; j < AT_BK; ++j)
        {
            float vv[DC];
#pragma unroll
            for (int c = 0; c < DC; ++c)
                vv[c] = Vs[j * DH + tx * DC + c];
#pragma unroll
            for (int i = 0; i < 4; ++i)
            {
                const float pv = Ps[(4 * ty + i) * AT_BK + j];
#pragma unroll
                for (int c = 0; c < DC; ++c)
                    o[i][c] = fmaf(pv, vv[c], o[i][c]);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
    {
        const int r = q0 + 4 * ty + i;
        if (r < T)
        {
            const float inv = 1.0f / l_run[i];
            float *dst = O + ((size_t)b * T + r) * ldo + h * DH + tx * DC;
#pragma unroll
            for (int c = 0; c < DC; ++c)
                dst[c] = o[i][c] * inv;
        }
    }
}

int launch_attention(const float *Q, int ldq, const float *K, int ldk, const float *V, int ldv, float *O, int ldo,
                     int B, int T, int S, int heads, int dh, cudaStream_t st)
{
    const float scale = 1.0f / sqrtf((float)dh);
    dim3 grid((T + AT_BQ - 1) / AT_BQ, heads, B);
    if ((ldq | ldk | ldv) % 4)
    {
        set_error("attention: leading dimensions must be multiples of 4");
        return -1;
    }
    if (dh == 64)
    {
        const int smem = (64 * AT_BQ + 64 * AT_BK + AT_BK * 64 + AT_BQ * AT_BK) * (int)sizeof(float);
        DMX_CUDA(cudaFuncSetAttribute(attention_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attention_kernel<64><<<grid, AT_THREADS, smem, st>>>(Q, ldq, K, ldk, V, ldv, O, ldo, T, S, scale);
    }
    else if (dh == 48)
    {
        const int smem = (48 * AT_BQ + 48 * AT_BK + AT_BK * 48 + AT_BQ * AT_BK) * (int)sizeof(float);
        DMX_CUDA(cudaFuncSetAttribute(attention_kernel<48>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attention_kernel<48><<<grid, AT_THREADS, smem, st>>>(Q, ldq, K, ldk, V, ldv, O, ldo, T, S, scale);
    }
    else
    {
        set_error("attention: unsupported head dim");
        return -1;
    }
    DMX_LAUNCH_CHECK();
    return 0;
}

} // namespace dmx
