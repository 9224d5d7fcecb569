// fp32 tap-GEMM (see kernels.cuh for the operator definition).  This is the full-precision path
// (north_star "fp32 mode", per-layer tolerance 1e-4): CUDA-core FMA, 128 x BN output tile per CTA,
// BK = 16, double-buffered shared memory with register prefetch, fused prologues (skip add,
// GroupNorm+GELU) and epilogues (bias, GELU, GLU, LayerScale residual, GroupNorm statistics,
// transposed-conv scatter + crop).
#include "kernels.cuh"

namespace dmx
{

constexpr int G_BM = 128;
constexpr int G_BK = 16;
constexpr int G_THREADS = 256;

struct RowCoord
{
    int b, uo, vo;
};

__device__ __forceinline__ RowCoord decode_row(int m, int Uo, int Vo)
{
    RowCoord rc;
    const int per = Uo * Vo;
    rc.b = m / per;
    const int r = m - rc.b * per;
    rc.uo = r / Vo;
    rc.vo = r - rc.uo * Vo;
    return rc;
}

template <int BN, int TM, int TN, bool VECA>
__global__ void __launch_bounds__(G_THREADS) gemm_kernel(const GemmParams p)
{
    constexpr int NTX = BN / TN;
    constexpr int NTY = G_THREADS / NTX;
    static_assert(NTY * TM == G_BM, "tile shape");
    constexpr int B_F4 = G_BK * BN / 4;                          // float4 per B tile
    constexpr int B_PER_THREAD = (B_F4 + G_THREADS - 1) / G_THREADS;

    __shared__ __align__(16) float As[2][G_BK][G_BM];
    __shared__ __align__(16) float Bs[2][G_BK][BN];

    const int tid = threadIdx.x;
    const int tx = tid % NTX, ty = tid / NTX;
    const int m0 = blockIdx.x * G_BM, n0 = blockIdx.y * BN;

    // ---- A loader assignment: one row, two float4 k-slots per thread ----
    const int la = tid & (G_BM - 1);
    const int kq0 = tid / G_BM; // 0..1
    const int lm = m0 + la;
    const bool lvalid = lm < p.M;
    RowCoord lrc = decode_row(lvalid ? lm : 0, p.Uo, p.Vo);
    float pmean = 0.f, prstd = 1.f;
    if (p.pro == 1 && lvalid)
    {
        const int ui = lrc.uo * p.su + p.du[0];
        const int item = p.pro_per_u ? lrc.b * p.Ui + ui : lrc.b;
        stats_to_mean_rstd(p.pro_stats + 2 * (size_t)item, p.pro_count, 1e-5f, pmean, prstd);
    }

    auto load_a = [&](int kt, float4 (&ra)[2]) {
#pragma unroll
        for (int i = 0; i < 2; ++i)
        {
            const int k = kt * G_BK + (kq0 + 2 * i) * 4;
            float v[4] = {0.f, 0.f, 0.f, 0.f};
            if (lvalid && k < p.K)
            {
                if (VECA)
                {
                    const int tap = k / p.Cin;
                    const int c = k - tap * p.Cin;
                    const int ui = lrc.uo * p.su + p.du[tap];
                    const int vi = lrc.vo * p.sv + p.dv[tap];
                    if (ui >= 0 && ui < p.Ui && vi >= 0 && vi < p.Vi)
                    {
                        const size_t off = (((size_t)lrc.b * p.Ui + ui) * p.Vi + vi) * p.lda + c;
                        float4 t = __ldg(reinterpret_cast<const float4 *>(p.A + off));
                        if (p.A2)
                        {
                            const float4 t2 = __ldg(reinterpret_cast<const float4 *>(p.A2 + off));
                            t.x += t2.x;
                            t.y += t2.y;
                            t.z += t2.z;
                            t.w += t2.w;
                        }
                        v[0] = t.x;
                        v[1] = t.y;
                        v[2] = t.z;
                        v[3] = t.w;
                        if (p.pro == 1)
                        {
#pragma unroll
                            for (int e = 0; e < 4; ++e)
                                v[e] = gelu_erf((v[e] - pmean) * prstd * __ldg(&p.pro_w[c + e]) + __ldg(&p.pro_b[c + e]));
                        }
                    }
                }
                else
                {
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                    {
                        const int ke = k + e;
                        if (ke < p.K)
                        {
                            const int tap = ke / p.Cin;
                            const int c = ke - tap * p.Cin;
                            const int ui = lrc.uo * p.su + p.du[tap];
                            const int vi = lrc.vo * p.sv + p.dv[tap];
                            if (ui >= 0 && ui < p.Ui && vi >= 0 && vi < p.Vi)
                            {
                                const size_t off = (((size_t)lrc.b * p.Ui + ui) * p.Vi + vi) * p.lda + c;
                                float t = __ldg(p.A + off);
                                if (p.A2)
                                    t += __ldg(p.A2 + off);
                                if (p.pro == 1)
                                    t = gelu_erf((t - pmean) * prstd * __ldg(&p.pro_w[c]) + __ldg(&p.pro_b[c]));
                                v[e] = t;
                            }
                        }
                    }
                }
            }
            ra[i] = make_float4(v[0], v[1], v[2], v[3]);
        }
    };
    auto store_a = [&](int buf, const float4 (&ra)[2]) {
#pragma unroll
        for (int i = 0; i < 2; ++i)
        {
            const int kk = (kq0 + 2 * i) * 4;
            As[buf][kk + 0][la] = ra[i].x;
            As[buf][kk + 1][la] = ra[i].y;
            As[buf][kk + 2][la] = ra[i].z;
            As[buf][kk + 3][la] = ra[i].w;
        }
    };
    auto load_b = [&](int kt, float4 (&rb)[B_PER_THREAD]) {
#pragma unroll
        for (int i = 0; i < B_PER_THREAD; ++i)
        {
            const int idx = tid + i * G_THREADS;
            if (idx < B_F4)
            {
                const int kk = idx / (BN / 4), nq = idx % (BN / 4);
                rb[i] = __ldg(reinterpret_cast<const float4 *>(p.W + (size_t)(kt * G_BK + kk) * p.ldw + n0 + nq * 4));
            }
        }
    };
    auto store_b = [&](int buf, const float4 (&rb)[B_PER_THREAD]) {
#pragma unroll
        for (int i = 0; i < B_PER_THREAD; ++i)
        {
            const int idx = tid + i * G_THREADS;
            if (idx < B_F4)
            {
                const int kk = idx / (BN / 4), nq = idx % (BN / 4);
                *reinterpret_cast<float4 *>(&Bs[buf][kk][nq * 4]) = rb[i];
            }
        }
    };

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j)
            acc[i][j] = 0.f;

    const int nk = (p.K + G_BK - 1) / G_BK;
    float4 ra[2];
    float4 rb[B_PER_THREAD];
    load_a(0, ra);
    load_b(0, rb);
    store_a(0, ra);
    store_b(0, rb);
    __syncthreads();
    for (int kt = 0; kt < nk; ++kt)
    {
        const int cur = kt & 1;
        if (kt + 1 < nk)
        {
            load_a(kt + 1, ra);
            load_b(kt + 1, rb);
        }
#pragma unroll
        for (int kk = 0; kk < G_BK; ++kk)
        {
            float a[TM], bb[TN];
#pragma unroll
            for (int i = 0; i < TM; i += (TM >= 4 ? 4 : TM))
            {
                if constexpr (TM >= 4)
                {
                    const float4 t = *reinterpret_cast<const float4 *>(&As[cur][kk][ty * TM + i]);
                    a[i] = t.x;
                    a[i + 1] = t.y;
                    a[i + 2] = t.z;
                    a[i + 3] = t.w;
                }
                else
                {
#pragma unroll
                    for (int e = 0; e < TM; ++e)
                        a[e] = As[cur][kk][ty * TM + e];
                }
            }
#pragma unroll
            for (int j = 0; j < TN; j += 4)
            {
                const float4 t = *reinterpret_cast<const float4 *>(&Bs[cur][kk][tx * TN + j]);
                bb[j] = t.x;
                bb[j + 1] = t.y;
                bb[j + 2] = t.z;
                bb[j + 3] = t.w;
            }
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j)
                    acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
        }
        if (kt + 1 < nk)
        {
            store_a(cur ^ 1, ra);
            store_b(cur ^ 1, rb);
        }
        __syncthreads();
    }

    // ---------------------------------------------------------------- epilogue
    const int nbase = n0 + tx * TN;
    float bias[TN], gam[TN];
#pragma unroll
    for (int j = 0; j < TN; ++j)
    {
        bias[j] = p.bias ? __ldg(&p.bias[nbase + j]) : 0.f; // bias is zero padded like W
        gam[j] = 0.f;
    }
    int cur_item = -1;
    double st_s = 0.0, st_ss = 0.0;
    const int nlimit = p.glu ? p.N / 2 : p.N;   // valid OUTPUT columns
    const int ncol0 = p.glu ? nbase / 2 : nbase; // first output column held by this thread
    const int nout = p.glu ? TN / 2 : TN;        // output columns held by this thread
    if (p.gamma)
    {
#pragma unroll
        for (int j = 0; j < TN; ++j)
            if (ncol0 + j < nlimit)
                gam[j] = __ldg(&p.gamma[ncol0 + j]);
    }

#pragma unroll
    for (int i = 0; i < TM; ++i)
    {
        const int m = m0 + ty * TM + i;
        const bool rvalid = m < p.M;
        const RowCoord rc = decode_row(rvalid ? m : 0, p.Uo, p.Vo);
        float val[TN];
#pragma unroll
        for (int j = 0; j < TN; ++j)
        {
            float v = acc[i][j] + bias[j];
            if (p.act == 1)
                v = gelu_erf(v);
            val[j] = v;
        }
        if (p.glu)
        {
#pragma unroll
            for (int j = 0; j < TN / 2; ++j)
                val[j] = val[2 * j] * sigmoid_ref(val[2 * j + 1]);
        }
        float rs = 0.f, rss = 0.f;
        if (rvalid)
        {
#pragma unroll
            for (int j0 = 0; j0 < TN; j0 += 4)
            {
                if (j0 >= nout)
                    break;
                const int gs = (nout - j0) < 4 ? (nout - j0) : 4; // 4, or 2 for GLU with TN == 4
                const int nc0 = ncol0 + j0;
                if (nc0 >= nlimit)
                    break;
                const int r = nc0 / p.Nco;
                const int co0 = nc0 - r * p.Nco;
                bool done = false;
                if (co0 + gs <= p.Nco && nc0 + gs <= nlimit)
                {
                    const int uu = rc.uo * p.ous + r * p.our + p.ouo;
                    const int vv = rc.vo * p.ovs + r * p.ovr + p.ovo;
                    done = true; // whole group shares one destination pixel
                    if (uu >= 0 && uu < p.Uout && vv >= 0 && vv < p.Vout)
                    {
                        float *dst = p.C + (p.out_vu ? ((size_t)rc.b * p.Vout + vv) * p.Uout + uu : ((size_t)rc.b * p.Uout + uu) * p.Vout + vv) * p.ldc + co0;
                        float o[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            o[e] = (e < gs) ? val[j0 + e] : 0.f;
                        if (p.rowadd)
                        {
                            const float *radd = p.rowadd + (size_t)rc.uo * p.rowadd_ld + co0;
#pragma unroll
                            for (int e = 0; e < 4; ++e)
                                if (e < gs)
                                    o[e] += __ldg(radd + e);
                        }
                        if (p.addC)
                        {
                            const float *ad = p.addC + (dst - p.C);
#pragma unroll
                            for (int e = 0; e < 4; ++e)
                                if (e < gs)
                                    o[e] += __ldg(ad + e);
                        }
                        const bool al16 = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
                        const bool al8 = (reinterpret_cast<uintptr_t>(dst) & 7) == 0;
                        if (gs == 4 && al16)
                        {
                            if (p.gamma)
                            {
                                const float4 old = *reinterpret_cast<const float4 *>(dst);
                                o[0] = old.x + gam[j0 + 0] * o[0];
                                o[1] = old.y + gam[j0 + 1] * o[1];
                                o[2] = old.z + gam[j0 + 2] * o[2];
                                o[3] = old.w + gam[j0 + 3] * o[3];
                            }
                            *reinterpret_cast<float4 *>(dst) = make_float4(o[0], o[1], o[2], o[3]);
                        }
                        else if (gs == 2 && al8)
                        {
                            if (p.gamma)
                            {
                                const float2 old = *reinterpret_cast<const float2 *>(dst);
                                o[0] = old.x + gam[j0 + 0] * o[0];
                                o[1] = old.y + gam[j0 + 1] * o[1];
                            }
                            *reinterpret_cast<float2 *>(dst) = make_float2(o[0], o[1]);
                        }
                        else
                        {
#pragma unroll
                            for (int e = 0; e < 4; ++e)
                                if (e < gs)
                                {
                                    if (p.gamma)
                                        o[e] = dst[e] + gam[j0 + e] * o[e];
                                    dst[e] = o[e];
                                }
                        }
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            if (e < gs)
                            {
                                rs += o[e];
                                rss += o[e] * o[e];
                            }
                    }
                }
                if (!done)
                {
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                    {
                        const int nc = nc0 + e;
                        if (e < gs && nc < nlimit)
                        {
                            const int r2 = nc / p.Nco;
                            const int co = nc - r2 * p.Nco;
                            const int u2 = rc.uo * p.ous + r2 * p.our + p.ouo;
                            const int v2 = rc.vo * p.ovs + r2 * p.ovr + p.ovo;
                            if (u2 >= 0 && u2 < p.Uout && v2 >= 0 && v2 < p.Vout)
                            {
                                float *dst = p.C + (p.out_vu ? ((size_t)rc.b * p.Vout + v2) * p.Uout + u2 : ((size_t)rc.b * p.Uout + u2) * p.Vout + v2) * p.ldc + co;
                                float o = val[j0 + e];
                                if (p.rowadd)
                                    o += __ldg(p.rowadd + (size_t)rc.uo * p.rowadd_ld + co);
                                if (p.addC)
                                    o += __ldg(p.addC + (dst - p.C));
                                if (p.gamma)
                                    o = *dst + gam[j0 + e] * o;
                                *dst = o;
                                rs += o;
                                rss += o * o;
                            }
                        }
                    }
                }
            }
        }
        if (p.stats)
        {
            // reduce this row's partial sums across the NTX threads that share the row
#pragma unroll
            for (int o = NTX / 2; o > 0; o >>= 1)
            {
                rs += __shfl_xor_sync(0xffffffffu, rs, o);
                rss += __shfl_xor_sync(0xffffffffu, rss, o);
            }
            if (tx == 0 && rvalid)
            {
                const int item = p.stat_per_u ? rc.b * p.Uo + rc.uo : rc.b;
                if (item != cur_item)
                {
                    if (cur_item >= 0)
                    {
                        atomicAdd(&p.stats[2 * (size_t)cur_item + 0], st_s);
                        atomicAdd(&p.stats[2 * (size_t)cur_item + 1], st_ss);
                    }
                    cur_item = item;
                    st_s = 0.0;
                    st_ss = 0.0;
                }
                st_s += (double)rs;
                st_ss += (double)rss;
            }
        }
    }
    if (p.stats && tx == 0 && cur_item >= 0)
    {
        atomicAdd(&p.stats[2 * (size_t)cur_item + 0], st_s);
        atomicAdd(&p.stats[2 * (size_t)cur_item + 1], st_ss);
    }
}

template <int BN, int TM, int TN>
static int launch_cfg(const GemmParams &p, cudaStream_t st)
{
    dim3 grid((p.M + G_BM - 1) / G_BM, (p.N + BN - 1) / BN);
    const bool vec = (p.Cin % 4 == 0) && (p.lda % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.A) & 15) == 0) &&
                     (!p.A2 || (reinterpret_cast<uintptr_t>(p.A2) & 15) == 0);
    if (vec)
        gemm_kernel<BN, TM, TN, true><<<grid, G_THREADS, 0, st>>>(p);
    else
        gemm_kernel<BN, TM, TN, false><<<grid, G_THREADS, 0, st>>>(p);
    DMX_LAUNCH_CHECK();
    return 0;
}

int launch_gemm(const GemmParams &p, cudaStream_t st)
{
    if (p.M <= 0 || p.N <= 0 || p.K <= 0)
    {
        set_error("launch_gemm: empty problem");
        return -1;
    }
    if (p.glu && (p.N % 2))
    {
        set_error("launch_gemm: GLU needs an even N");
        return -1;
    }
    // the stats shuffle-reduction assumes the NTX threads of a row sit in one warp (true for all cfgs)
    if (p.N > 64)
        return launch_cfg<128, 8, 8>(p, st);
    if (p.N > 32)
        return launch_cfg<64, 8, 4>(p, st);
    if (p.N > 16)
        return launch_cfg<32, 4, 4>(p, st);
    return launch_cfg<16, 2, 4>(p, st);
}

} // namespace dmx
