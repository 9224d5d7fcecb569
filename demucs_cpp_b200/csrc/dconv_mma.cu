// bf16-mode twin of dconv_fused.cu: the same fused DConv residual branch (one CTA keeps one frequency row
// x[T=336][C] resident and runs both sub-blocks on it; reference demucscpp::apply_dconv, src/layers.cpp:152-375,
// v3 twin :723-894), with the two convolutions of each sub-block on the tensor cores through warp-level
// mma.sync (bf16 operands, fp32 accumulate) instead of CUDA-core FMAs.
//
// Why mma.sync and not tcgen05 here: the GEMMs are 16..32 wide (hidden = C/8 or C/4 channels) and 336 rows long,
// their operands already sit in this CTA's shared memory and their results are consumed in registers by the
// GroupNorm / GELU / GLU math of the same thread -- exactly the m16n8k{8,16} fragment shapes.  The accumulator
// fragment of the k3 conv (16 positions x 8 hidden) IS the A fragment of the following 1x1 conv, so the hidden
// activations never leave registers.  The SIMT kernel was bound by FMA issue (ncu: 428 M instructions, 72 % issue
// utilisation for C = 48); here a warp instruction does 2048 / 4096 MACs.
//
//     h = conv_k3_dil(x) ; g = GELU(GN1(h)) ; y = W2 g + b2 ; x += scale * GLU(GN2(y))
// GroupNorm statistics: exact block reductions over the row in double, unbiased variance.  y is evaluated twice
// (statistics, then normalise-GLU-residual) and never stored.  The residual stream x stays fp32 (updated in place in
// global memory, in the accumulator-fragment layout, while the row is still L2 resident); only the MMA operands (x as
// the conv input, g, the weights) are rounded to bf16, like every other GEMM of the bf16 mode.
#include "kernels.cuh"

#include <cuda_bf16.h>

#include <cstring>

namespace dmx
{
namespace dm
{

constexpr int T = LE;             // 336 frames = 21 m-tiles of 16
constexpr int WARPS = 7;          // 3 m-tiles per warp
constexpr int THREADS = WARPS * 32;
constexpr int MT_PER_WARP = 3;
constexpr int HALO = 2;           // zero rows on both sides of the bf16 copy (dilations 1 and 2)

__host__ __device__ constexpr int hid_pad(int hid) { return hid <= 8 ? 8 : (hid <= 16 ? 16 : 32); }
// element strides of the bf16 operand arrays, chosen so that the 8 rows x 4 words of a fragment load fall into
// 32 different banks: word stride mod 32 must be an odd multiple of 4
__host__ __device__ constexpr int xld(int C) { return C + 8; }                    // 28 / 52 words
__host__ __device__ constexpr int k1ld(int C) { return 3 * C + 8; }               // 76 / 148 words
__host__ __device__ constexpr int w2ld(int hp) { return hp == 8 ? 8 : hp + 8; }   // 4 / 12 / 20 words
__host__ __device__ constexpr int nvec(int C, int hp) { return 3 * hp + 7 * C; }  // b1 g1w g1b | b2 g2w g2b (2C each) | scale
__host__ __device__ constexpr size_t smem_bytes(int C, int hid)
{
    return (size_t)(T + 2 * HALO) * xld(C) * 2 + (size_t)hid_pad(hid) * k1ld(C) * 2 +
           (size_t)2 * C * w2ld(hid_pad(hid)) * 2 + (size_t)nvec(C, hid_pad(hid)) * 4 + 64;
}
// fp32 parameter block shared with dconv_fused.cu, per sub-block:
//   W1[3][C][HID] | b1[HID] | g1w[HID] | g1b[HID] | W2[2C][HID] | b2[2C] | g2w[2C] | g2b[2C] | scale[C]
__host__ __device__ constexpr int p_stride(int C, int HID) { return 3 * C * HID + 3 * HID + 2 * C * HID + 6 * C + C; }

// Pre-packed parameters (built once at model finalize by dconv_mma_pack): per sub-block, exactly the shared-memory
// images the kernels use -- w1b [HP][K1LD] bf16 (k = tap*C + c, zero padded) | w2b [2C][W2LD] bf16 | vec fp32
// (b1 g1w g1b [HP each] | b2 g2w g2b [2C each] | scale [C], rounded up to 4 floats) -- so a CTA fetches them with
// 16-byte async copies instead of converting fp32 weights element by element (that conversion, a dependent L2 load
// per element, was most of the run time of the short tiled kernels).
__host__ __device__ constexpr int pk_w1_bytes(int C, int hp) { return hp * k1ld(C) * 2; }
__host__ __device__ constexpr int pk_w2_bytes(int C, int hp) { return 2 * C * w2ld(hp) * 2; }
__host__ __device__ constexpr int pk_vec_bytes(int C, int hp) { return ((nvec(C, hp) + 3) & ~3) * 4; }
__host__ __device__ constexpr int pk_sub_bytes(int C, int hp) { return pk_w1_bytes(C, hp) + pk_w2_bytes(C, hp) + pk_vec_bytes(C, hp); }

__host__ __device__ constexpr int rw_bytes(int C) { return 2 * C * xld(C) * 2 + 2 * C * 4; }
__host__ __device__ constexpr int cin_bytes(int C) { return C * 40 * 2 + C * 4; }

__device__ __forceinline__ void copy16_async(void *sdst, const void *gsrc, int bytes)
{
    for (int i = threadIdx.x * 16; i < bytes; i += THREADS * 16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(
                         (uint32_t)__cvta_generic_to_shared(reinterpret_cast<uint8_t *>(sdst) + i)),
                     "l"(reinterpret_cast<const uint8_t *>(gsrc) + i)
                     : "memory");
}
__device__ __forceinline__ void async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ void mma_k16(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                        uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
                 "{%0, %1, %2, %3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_k8(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0)
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%0, %1, %2, %3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi)
{
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t *>(&h);
}
__device__ __forceinline__ uint32_t ld32(const __nv_bfloat16 *p) { return *reinterpret_cast<const uint32_t *>(p); }

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
    for (int w = 0; w < WARPS; ++w)
    {
        sa += red[2 * w];
        sb += red[2 * w + 1];
    }
    a = sa;
    b = sb;
}

// RW: the encoder's 1x1 "rewrite" conv + GLU (+ frequency embedding) that follows the DConv (src/encdec.cpp:118-164) runs
// on the row while it is still here: out[row][c] = (Wa x + ba) * sigmoid(Wg x + bg) + rowadd[u][c]; the DConv result
// itself is then not written back (nothing else reads it).  rw: Wr bf16 [2C][C + 8] | bias fp32 [2C].
// CIN: the encoder's strided input conv (k8 s4 p2 along frequency, 4 CaC channels, + GELU; src/encdec.cpp:24-50) that
// PRECEDES the DConv is computed here too: the row's A fragments are gathered straight from the spectrogram
// [B][4F][336][4] (8 taps x 4 channels = K 32), so x is produced on the SM (written once for the residual updates)
// instead of being written by a GEMM launch and read back.  cin: Wc bf16 [C][40] (k = tap*4 + ch) | bias fp32 [C].
template <int C, int HID, bool RW, bool CIN>
__global__ void __launch_bounds__(THREADS, (C == 48 && HID == 6) ? 4 : 1)
    dconv_mma_kernel(float *__restrict__ x, const uint8_t *__restrict__ params, const uint8_t *__restrict__ rw,
                     float *__restrict__ out, const float *__restrict__ rowadd, const int F, const float *__restrict__ spec,
                     const uint8_t *__restrict__ cin)
{
    constexpr int HP = hid_pad(HID);
    constexpr int NT1 = HP / 8;       // n-tiles of the k3 conv
    constexpr int NT2 = 2 * C / 8;    // n-tiles of the 1x1 conv
    constexpr int XLD = xld(C), K1LD = k1ld(C), W2LD = w2ld(HP);
    extern __shared__ __align__(16) uint8_t smraw[];
    float *vec = reinterpret_cast<float *>(smraw);                                // small fp32 vectors
    __nv_bfloat16 *xb = reinterpret_cast<__nv_bfloat16 *>(vec + ((nvec(C, HP) + 3) & ~3)); // [T + 4][XLD]
    __nv_bfloat16 *w1b = xb + (T + 2 * HALO) * XLD;                               // [HP][K1LD], k = tap*C + c
    __nv_bfloat16 *w2b = w1b + HP * K1LD;                                         // [2C][W2LD]
    __nv_bfloat16 *wr = w2b + 2 * C * W2LD;                                       // RW: [2C][XLD] | bias fp32 [2C]
    float *b1 = vec, *g1w = b1 + HP, *g1b = g1w + HP, *b2 = g1b + HP, *g2w = b2 + 2 * C, *g2b = g2w + 2 * C,
          *scale = g2b + 2 * C;
    __shared__ double red[2 * WARPS];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    float *xg = x + (size_t)blockIdx.x * T * C;
    if (RW)
        copy16_async(wr, rw, rw_bytes(C)); // waited for together with the first sub-block's parameters

    // ---- load the row: fp32 copy + bf16 operand copy (halo rows zero) ----
    for (int i = tid; i < 2 * HALO * XLD / 2; i += THREADS)
    {
        const int r = i / (XLD / 2), c2 = i % (XLD / 2);
        const int row = r < HALO ? r : T + r; // rows 0,1 and T+2,T+3
        reinterpret_cast<uint32_t *>(xb + row * XLD)[c2] = 0u;
    }
    if (CIN)
    {
        constexpr int CLD = 40; // 32 + 8: 20 words per row, conflict-free fragment loads
        __nv_bfloat16 *wc = wr + (RW ? rw_bytes(C) / 2 : 0);
        copy16_async(wc, cin, cin_bytes(C));
        async_wait_all();
        __syncthreads();
        const float *cb = reinterpret_cast<const float *>(wc + C * CLD);
        const int bi = blockIdx.x / F, u = blockIdx.x - bi * F;
        const float *sb = spec + (size_t)bi * (4 * F) * T * 4;
        // one m-tile at a time (24 accumulator registers): this stage must not set the kernel's register count
#pragma unroll 1
        for (int i = 0; i < MT_PER_WARP; ++i)
        {
            const int row = 16 * (warp + WARPS * i) + g;
            float acc[C / 8][4];
#pragma unroll
            for (int nt = 0; nt < C / 8; ++nt)
            {
                acc[nt][0] = acc[nt][2] = cb[8 * nt + 2 * t];
                acc[nt][1] = acc[nt][3] = cb[8 * nt + 2 * t + 1];
            }
#pragma unroll
            for (int ks = 0; ks < 2; ++ks)
            {
                // k = tap*4 + ch: this lane's k pair (2t, 2t+1) is channels (t&1)*2, +1 of tap 4ks + (t>>1); k + 8 = two taps on
                const int tap0 = 4 * ks + (t >> 1), ch0 = (t & 1) * 2;
                const int f0 = 4 * u - 2 + tap0, f1 = f0 + 2;
                float2 p0 = make_float2(0.f, 0.f), p1 = p0, p2 = p0, p3 = p0;
                if (f0 >= 0 && f0 < 4 * F)
                {
                    p0 = __ldg(reinterpret_cast<const float2 *>(sb + ((size_t)f0 * T + row) * 4 + ch0));
                    p1 = __ldg(reinterpret_cast<const float2 *>(sb + ((size_t)f0 * T + row + 8) * 4 + ch0));
                }
                if (f1 >= 0 && f1 < 4 * F)
                {
                    p2 = __ldg(reinterpret_cast<const float2 *>(sb + ((size_t)f1 * T + row) * 4 + ch0));
                    p3 = __ldg(reinterpret_cast<const float2 *>(sb + ((size_t)f1 * T + row + 8) * 4 + ch0));
                }
                const uint32_t a0 = pack2(p0.x, p0.y), a1 = pack2(p1.x, p1.y), a2 = pack2(p2.x, p2.y), a3 = pack2(p3.x, p3.y);
#pragma unroll
                for (int nt = 0; nt < C / 8; ++nt)
                {
                    const __nv_bfloat16 *wp = wc + (8 * nt + g) * CLD + 16 * ks + 2 * t;
                    mma_k16(acc[nt], a0, a1, a2, a3, ld32(wp), ld32(wp + 8));
                }
            }
#pragma unroll
            for (int nt = 0; nt < C / 8; ++nt)
            {
                const int c = 8 * nt + 2 * t;
                const float2 x0 = make_float2(gelu_erf(acc[nt][0]), gelu_erf(acc[nt][1]));
                const float2 x1 = make_float2(gelu_erf(acc[nt][2]), gelu_erf(acc[nt][3]));
                *reinterpret_cast<float2 *>(xg + (size_t)row * C + c) = x0;
                *reinterpret_cast<float2 *>(xg + (size_t)(row + 8) * C + c) = x1;
                *reinterpret_cast<uint32_t *>(xb + (row + HALO) * XLD + c) = pack2(x0.x, x0.y);
                *reinterpret_cast<uint32_t *>(xb + (row + 8 + HALO) * XLD + c) = pack2(x1.x, x1.y);
            }
        }
    }
    // the conv operand copy: fp32 row -> bf16 (the fp32 residual itself stays in global memory / L2 and is updated
    // there in the fragment layout, see pass B)
#pragma unroll 6
    for (int i = tid; i < (CIN ? 0 : T * C / 4); i += THREADS)
    {
        const float4 v = *reinterpret_cast<const float4 *>(xg + 4 * i);
        const int tt = (4 * i) / C, c = (4 * i) % C;
        uint2 pk;
        pk.x = pack2(v.x, v.y);
        pk.y = pack2(v.z, v.w);
        *reinterpret_cast<uint2 *>(xb + (tt + HALO) * XLD + c) = pk;
    }

#pragma unroll 1
    for (int sub = 0; sub < 2; ++sub)
    {
        const int dil = 1 << sub;
        const uint8_t *P = params + (size_t)sub * pk_sub_bytes(C, HP);
        __syncthreads(); // previous sub-block done with the weights; xs / xb complete
        // ---- parameters of this sub-block: pre-packed shared-memory images ----
        copy16_async(w1b, P, pk_w1_bytes(C, HP));
        copy16_async(w2b, P + pk_w1_bytes(C, HP), pk_w2_bytes(C, HP));
        copy16_async(vec, P + pk_w1_bytes(C, HP) + pk_w2_bytes(C, HP), pk_vec_bytes(C, HP));
        async_wait_all();
        __syncthreads();

        // ---- conv k3, dilation dil: h[16 positions][HP] per m-tile, accumulators in the C-fragment layout ----
        float acc[MT_PER_WARP][NT1][4];
#pragma unroll
        for (int i = 0; i < MT_PER_WARP; ++i)
#pragma unroll
            for (int nt = 0; nt < NT1; ++nt)
            {
                acc[i][nt][0] = acc[i][nt][2] = b1[8 * nt + 2 * t];
                acc[i][nt][1] = acc[i][nt][3] = b1[8 * nt + 2 * t + 1];
            }
#pragma unroll
        for (int tap = 0; tap < 3; ++tap)
        {
            const int shift = (tap - 1) * dil + HALO;
#pragma unroll 2
            for (int kc = 0; kc < C / 16; ++kc)
            {
                uint32_t bf[NT1][2];
#pragma unroll
                for (int nt = 0; nt < NT1; ++nt)
                {
                    const __nv_bfloat16 *wp = w1b + (8 * nt + g) * K1LD + tap * C + 16 * kc + 2 * t;
                    bf[nt][0] = ld32(wp);
                    bf[nt][1] = ld32(wp + 8);
                }
#pragma unroll
                for (int i = 0; i < MT_PER_WARP; ++i)
                {
                    const int row0 = 16 * (warp + WARPS * i) + shift + g;
                    const __nv_bfloat16 *ap = xb + row0 * XLD + 16 * kc + 2 * t;
                    const uint32_t a0 = ld32(ap), a1 = ld32(ap + 8 * XLD), a2 = ld32(ap + 8), a3 = ld32(ap + 8 * XLD + 8);
#pragma unroll
                    for (int nt = 0; nt < NT1; ++nt)
                        mma_k16(acc[i][nt], a0, a1, a2, a3, bf[nt][0], bf[nt][1]);
                }
            }
        }
        // ---- GroupNorm(1) statistics of h (padded hidden channels are exactly zero) ----
        double s1 = 0.0, s2 = 0.0;
        {
            float a = 0.f, b = 0.f;
#pragma unroll
            for (int i = 0; i < MT_PER_WARP; ++i)
#pragma unroll
                for (int nt = 0; nt < NT1; ++nt)
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                    {
                        a += acc[i][nt][e];
                        b = fmaf(acc[i][nt][e], acc[i][nt][e], b);
                    }
            s1 = a;
            s2 = b;
        }
        block_sum2(s1, s2, red);
        // ---- GN1 + GELU; the C fragments become the A fragments of the 1x1 conv ----
        uint32_t af[MT_PER_WARP][NT1][2];
        {
            const double n = (double)T * HID;
            const double m = s1 / n;
            double var = (s2 - s1 * m) / (n - 1.0);
            var = var < 0.0 ? 0.0 : var;
            const float mean = (float)m, rstd = (float)(1.0 / sqrt(var + 1e-5));
#pragma unroll
            for (int nt = 0; nt < NT1; ++nt)
            {
                const int j = 8 * nt + 2 * t;
                const float w0 = rstd * g1w[j], w1 = rstd * g1w[j + 1], o0 = g1b[j], o1 = g1b[j + 1];
#pragma unroll
                for (int i = 0; i < MT_PER_WARP; ++i)
                {
                    af[i][nt][0] = pack2(gelu_erf((acc[i][nt][0] - mean) * w0 + o0), gelu_erf((acc[i][nt][1] - mean) * w1 + o1));
                    af[i][nt][1] = pack2(gelu_erf((acc[i][nt][2] - mean) * w0 + o0), gelu_erf((acc[i][nt][3] - mean) * w1 + o1));
                }
            }
        }
        // y tile nt (8 output channels) of the three m-tiles: d[i] = b2 + W2[8nt.., :] g
        auto conv2 = [&](int nt, float (&d)[MT_PER_WARP][4]) {
            const float bb0 = b2[8 * nt + 2 * t], bb1 = b2[8 * nt + 2 * t + 1];
#pragma unroll
            for (int i = 0; i < MT_PER_WARP; ++i)
            {
                d[i][0] = d[i][2] = bb0;
                d[i][1] = d[i][3] = bb1;
            }
            const __nv_bfloat16 *wp = w2b + (8 * nt + g) * W2LD + 2 * t;
            if (HP == 8)
            {
                const uint32_t b0 = ld32(wp);
#pragma unroll
                for (int i = 0; i < MT_PER_WARP; ++i)
                    mma_k8(d[i], af[i][0][0], af[i][0][1], b0);
            }
            else
            {
#pragma unroll
                for (int s = 0; s < HP / 16; ++s)
                {
                    const uint32_t b0 = ld32(wp + 16 * s), b1v = ld32(wp + 16 * s + 8);
#pragma unroll
                    for (int i = 0; i < MT_PER_WARP; ++i)
                        mma_k16(d[i], af[i][(2 * s) % NT1][0], af[i][(2 * s) % NT1][1], af[i][(2 * s + 1) % NT1][0],
                                af[i][(2 * s + 1) % NT1][1], b0, b1v);
                }
            }
        };
        // ---- 1x1 conv to 2C channels, pass A: statistics only ----
        s1 = 0.0;
        s2 = 0.0;
        {
            float a = 0.f, b = 0.f;
#pragma unroll 4
            for (int nt = 0; nt < NT2; ++nt)
            {
                float d[MT_PER_WARP][4];
                conv2(nt, d);
#pragma unroll
                for (int i = 0; i < MT_PER_WARP; ++i)
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                    {
                        a += d[i][e];
                        b = fmaf(d[i][e], d[i][e], b);
                    }
            }
            s1 = a;
            s2 = b;
        }
        block_sum2(s1, s2, red);
        // ---- pass B: recompute, GroupNorm, GLU, LayerScale, residual into the resident fp32 rows ----
        {
            const double n = (double)T * 2 * C;
            const double m = s1 / n;
            double var = (s2 - s1 * m) / (n - 1.0);
            var = var < 0.0 ? 0.0 : var;
            const float mean = (float)m, rstd = (float)(1.0 / sqrt(var + 1e-5));
#pragma unroll 2
            for (int nt = 0; nt < C / 8; ++nt)
            {
                float da[MT_PER_WARP][4], dg[MT_PER_WARP][4];
                conv2(nt, da);
                conv2(nt + C / 8, dg);
                const int c = 8 * nt + 2 * t;
                const float ga0 = rstd * g2w[c], ga1 = rstd * g2w[c + 1], oa0 = g2b[c], oa1 = g2b[c + 1];
                const float gg0 = rstd * g2w[C + c], gg1 = rstd * g2w[C + c + 1], og0 = g2b[C + c], og1 = g2b[C + c + 1];
                const float sc0 = scale[c], sc1 = scale[c + 1];
#pragma unroll
                for (int i = 0; i < MT_PER_WARP; ++i)
                {
                    // residual update straight on global memory (8 rows x 32 contiguous bytes per request: whole
                    // sectors; the row was read moments ago, so it comes from L2) and, after the first sub-block, the
                    // refreshed bf16 operand for the second one (nobody reads xb between the k3 conv and here)
                    const int row = 16 * (warp + WARPS * i) + g;
                    float2 *r0 = reinterpret_cast<float2 *>(xg + (size_t)row * C + c);
                    float2 *r1 = reinterpret_cast<float2 *>(xg + (size_t)(row + 8) * C + c);
                    float2 x0 = *r0, x1 = *r1;
                    x0.x += sc0 * (((da[i][0] - mean) * ga0 + oa0) * sigmoid_ref((dg[i][0] - mean) * gg0 + og0));
                    x0.y += sc1 * (((da[i][1] - mean) * ga1 + oa1) * sigmoid_ref((dg[i][1] - mean) * gg1 + og1));
                    x1.x += sc0 * (((da[i][2] - mean) * ga0 + oa0) * sigmoid_ref((dg[i][2] - mean) * gg0 + og0));
                    x1.y += sc1 * (((da[i][3] - mean) * ga1 + oa1) * sigmoid_ref((dg[i][3] - mean) * gg1 + og1));
                    if (!RW || sub == 0)
                    {
                        *r0 = x0;
                        *r1 = x1;
                    }
                    if (RW || sub == 0)
                    {
                        *reinterpret_cast<uint32_t *>(xb + (row + HALO) * XLD + c) = pack2(x0.x, x0.y);
                        *reinterpret_cast<uint32_t *>(xb + (row + 8 + HALO) * XLD + c) = pack2(x1.x, x1.y);
                    }
                }
            }
        }
    }
    if (RW)
    {
        // ---- rewrite 1x1 + GLU on the finished row (bf16 copy in xb), straight to the encoder output ----
        __syncthreads();
        const float *rb = reinterpret_cast<const float *>(wr + 2 * C * XLD);
        const int u = blockIdx.x % F;
        float *og = out + (size_t)blockIdx.x * T * C;
#pragma unroll 1
        for (int nt = 0; nt < C / 8; ++nt)
        {
            const int c = 8 * nt + 2 * t;
            float da[MT_PER_WARP][4], dg[MT_PER_WARP][4];
#pragma unroll
            for (int i = 0; i < MT_PER_WARP; ++i)
            {
                da[i][0] = da[i][2] = rb[c];
                da[i][1] = da[i][3] = rb[c + 1];
                dg[i][0] = dg[i][2] = rb[C + c];
                dg[i][1] = dg[i][3] = rb[C + c + 1];
            }
#pragma unroll
            for (int kc = 0; kc < C / 16; ++kc)
            {
                const __nv_bfloat16 *wa = wr + (8 * nt + g) * XLD + 16 * kc + 2 * t;
                const __nv_bfloat16 *wg = wa + C * XLD;
                const uint32_t ba0 = ld32(wa), ba1 = ld32(wa + 8), bg0 = ld32(wg), bg1 = ld32(wg + 8);
#pragma unroll
                for (int i = 0; i < MT_PER_WARP; ++i)
                {
                    const __nv_bfloat16 *ap = xb + (16 * (warp + WARPS * i) + g + HALO) * XLD + 16 * kc + 2 * t;
                    const uint32_t a0 = ld32(ap), a1 = ld32(ap + 8 * XLD), a2 = ld32(ap + 8), a3 = ld32(ap + 8 * XLD + 8);
                    mma_k16(da[i], a0, a1, a2, a3, ba0, ba1);
                    mma_k16(dg[i], a0, a1, a2, a3, bg0, bg1);
                }
            }
            float e0 = 0.f, e1 = 0.f;
            if (rowadd)
            {
                e0 = __ldg(rowadd + (size_t)u * C + c);
                e1 = __ldg(rowadd + (size_t)u * C + c + 1);
            }
#pragma unroll
            for (int i = 0; i < MT_PER_WARP; ++i)
            {
                const int row = 16 * (warp + WARPS * i) + g;
                *reinterpret_cast<float2 *>(og + (size_t)row * C + c) =
                    make_float2(da[i][0] * sigmoid_ref(dg[i][0]) + e0, da[i][1] * sigmoid_ref(dg[i][1]) + e1);
                *reinterpret_cast<float2 *>(og + (size_t)(row + 8) * C + c) =
                    make_float2(da[i][2] * sigmoid_ref(dg[i][2]) + e0, da[i][3] * sigmoid_ref(dg[i][3]) + e1);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Time branch (one GroupNorm item = a whole batch element of L samples): the same arithmetic on 336-position tiles
// of the signal, in three launches per sub-block because the statistics span the item:
//   phase 0: h = conv_k3_dil(x) + b1 -> hbuf (fp32 [B][L][hs]), (sum, sumsq) of h -> stats1[b]
//   phase 1: g = GELU(GN1(h)); y = W2 g + b2 in registers only,  (sum, sumsq) of y -> stats2[b]
//   phase 2: recompute g and y; x += scale * GLU(GN2(y))
// x is read twice and written once, h (C/8 wide) written once and read twice, y (2C wide) never stored -- against
// x / h / y passing through three GEMM + pointwise launches with y materialised (src/layers.cpp:152-375).
template <int C, int HID, int PHASE, int MTW>
__global__ void __launch_bounds__(THREADS)
    dconv_time_kernel(float *__restrict__ x, float *__restrict__ hbuf, const uint8_t *__restrict__ P, const int L, const int hs,
                      const int dil, double *__restrict__ stats1, double *__restrict__ stats2)
{
    constexpr int HP = hid_pad(HID);
    constexpr int NT1 = HP / 8;
    constexpr int NT2 = 2 * C / 8;
    constexpr int TT = 16 * WARPS * MTW; // positions per tile: MTW m-tiles of 16 per warp
    constexpr int XLD = xld(C), K1LD = k1ld(C), W2LD = w2ld(HP);
    extern __shared__ __align__(16) uint8_t smraw[];
    // phase 0: xb | w1b | b1      phase 1: w2b | vec      phase 2: xs | w2b | vec
    __shared__ double red[2 * WARPS];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const int b = blockIdx.y;
    const int p0 = blockIdx.x * TT; // first position of the tile
    float *xg = x + (size_t)b * L * C;
    float *hg = hbuf + (size_t)b * L * hs;
    if (PHASE == 0)
    {
        __nv_bfloat16 *xb = reinterpret_cast<__nv_bfloat16 *>(smraw); // [T + 4][XLD], row r <-> position p0 + r - 2
        __nv_bfloat16 *w1b = xb + (TT + 2 * HALO) * XLD;
        float *b1 = reinterpret_cast<float *>(w1b + HP * K1LD);
#pragma unroll 6
        for (int i = tid; i < (TT + 2 * HALO) * C / 4; i += THREADS)
        {
            const int r = (4 * i) / C, c = (4 * i) % C;
            const int pos = p0 + r - HALO;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (pos >= 0 && pos < L)
                v = *reinterpret_cast<const float4 *>(xg + (size_t)pos * C + c);
            uint2 pk;
            pk.x = pack2(v.x, v.y);
            pk.y = pack2(v.z, v.w);
            *reinterpret_cast<uint2 *>(xb + r * XLD + c) = pk;
        }
        copy16_async(w1b, P, pk_w1_bytes(C, HP));
        copy16_async(b1, P + pk_w1_bytes(C, HP) + pk_w2_bytes(C, HP), HP * 4);
        async_wait_all();
        __syncthreads();
        float acc[MTW][NT1][4];
#pragma unroll
        for (int i = 0; i < MTW; ++i)
#pragma unroll
            for (int nt = 0; nt < NT1; ++nt)
            {
                acc[i][nt][0] = acc[i][nt][2] = b1[8 * nt + 2 * t];
                acc[i][nt][1] = acc[i][nt][3] = b1[8 * nt + 2 * t + 1];
            }
#pragma unroll
        for (int tap = 0; tap < 3; ++tap)
        {
            const int shift = (tap - 1) * dil + HALO;
#pragma unroll 2
            for (int kc = 0; kc < C / 16; ++kc)
            {
                uint32_t bf[NT1][2];
#pragma unroll
                for (int nt = 0; nt < NT1; ++nt)
                {
                    const __nv_bfloat16 *wp = w1b + (8 * nt + g) * K1LD + tap * C + 16 * kc + 2 * t;
                    bf[nt][0] = ld32(wp);
                    bf[nt][1] = ld32(wp + 8);
                }
#pragma unroll
                for (int i = 0; i < MTW; ++i)
                {
                    const int row0 = 16 * (warp + WARPS * i) + shift + g;
                    const __nv_bfloat16 *ap = xb + row0 * XLD + 16 * kc + 2 * t;
                    const uint32_t a0 = ld32(ap), a1 = ld32(ap + 8 * XLD), a2 = ld32(ap + 8), a3 = ld32(ap + 8 * XLD + 8);
#pragma unroll
                    for (int nt = 0; nt < NT1; ++nt)
                        mma_k16(acc[i][nt], a0, a1, a2, a3, bf[nt][0], bf[nt][1]);
                }
            }
        }
        float sa = 0.f, sb = 0.f;
#pragma unroll
        for (int i = 0; i < MTW; ++i)
        {
            const int q0 = p0 + 16 * (warp + WARPS * i) + g, q1 = q0 + 8;
#pragma unroll
            for (int nt = 0; nt < NT1; ++nt)
            {
                const int j = 8 * nt + 2 * t;
                if (j < hs)
                {
                    if (q0 < L)
                    {
                        *reinterpret_cast<float2 *>(hg + (size_t)q0 * hs + j) = make_float2(acc[i][nt][0], acc[i][nt][1]);
                        sa += acc[i][nt][0] + acc[i][nt][1];
                        sb += acc[i][nt][0] * acc[i][nt][0] + acc[i][nt][1] * acc[i][nt][1];
                    }
                    if (q1 < L)
                    {
                        *reinterpret_cast<float2 *>(hg + (size_t)q1 * hs + j) = make_float2(acc[i][nt][2], acc[i][nt][3]);
                        sa += acc[i][nt][2] + acc[i][nt][3];
                        sb += acc[i][nt][2] * acc[i][nt][2] + acc[i][nt][3] * acc[i][nt][3];
                    }
                }
            }
        }
        double s1 = sa, s2 = sb;
        block_sum2(s1, s2, red);
        if (tid == 0)
        {
            atomicAdd(&stats1[2 * b], s1);
            atomicAdd(&stats1[2 * b + 1], s2);
        }
        return;
    }

    // ---- phases 1 and 2: hidden activations back from hbuf, normalised, as A fragments ----
    float *vec = reinterpret_cast<float *>(smraw);
    __nv_bfloat16 *w2b = reinterpret_cast<__nv_bfloat16 *>(vec + ((nvec(C, HP) + 3) & ~3));
    float *b1 = vec, *g1w = b1 + HP, *g1b = g1w + HP, *b2 = g1b + HP, *g2w = b2 + 2 * C, *g2b = g2w + 2 * C,
          *scale = g2b + 2 * C;
    (void)b1;
    copy16_async(w2b, P + pk_w1_bytes(C, HP), pk_w2_bytes(C, HP));
    copy16_async(vec, P + pk_w1_bytes(C, HP) + pk_w2_bytes(C, HP), pk_vec_bytes(C, HP));
    async_wait_all();
    __syncthreads();

    uint32_t af[MTW][NT1][2];
    {
        float mean, rstd;
        stats_to_mean_rstd(stats1 + 2 * b, (double)L * HID, 1e-5f, mean, rstd);
#pragma unroll
        for (int nt = 0; nt < NT1; ++nt)
        {
            const int j = 8 * nt + 2 * t;
            const float w0 = rstd * g1w[j], w1 = rstd * g1w[j + 1], o0 = g1b[j], o1 = g1b[j + 1];
#pragma unroll
            for (int i = 0; i < MTW; ++i)
            {
                const int q0 = p0 + 16 * (warp + WARPS * i) + g, q1 = q0 + 8;
                float2 h0 = make_float2(0.f, 0.f), h1 = h0;
                if (j < hs)
                {
                    if (q0 < L)
                        h0 = *reinterpret_cast<const float2 *>(hg + (size_t)q0 * hs + j);
                    if (q1 < L)
                        h1 = *reinterpret_cast<const float2 *>(hg + (size_t)q1 * hs + j);
                }
                af[i][nt][0] = pack2(gelu_erf((h0.x - mean) * w0 + o0), gelu_erf((h0.y - mean) * w1 + o1));
                af[i][nt][1] = pack2(gelu_erf((h1.x - mean) * w0 + o0), gelu_erf((h1.y - mean) * w1 + o1));
            }
        }
    }
    auto conv2 = [&](int nt, float (&d)[MTW][4]) {
        const float bb0 = b2[8 * nt + 2 * t], bb1 = b2[8 * nt + 2 * t + 1];
#pragma unroll
        for (int i = 0; i < MTW; ++i)
        {
            d[i][0] = d[i][2] = bb0;
            d[i][1] = d[i][3] = bb1;
        }
        const __nv_bfloat16 *wp = w2b + (8 * nt + g) * W2LD + 2 * t;
        if (HP == 8)
        {
            const uint32_t b0 = ld32(wp);
#pragma unroll
            for (int i = 0; i < MTW; ++i)
                mma_k8(d[i], af[i][0][0], af[i][0][1], b0);
        }
        else
        {
#pragma unroll
            for (int s = 0; s < HP / 16; ++s)
            {
                const uint32_t b0 = ld32(wp + 16 * s), b1v = ld32(wp + 16 * s + 8);
#pragma unroll
                for (int i = 0; i < MTW; ++i)
                    mma_k16(d[i], af[i][(2 * s) % NT1][0], af[i][(2 * s) % NT1][1], af[i][(2 * s + 1) % NT1][0],
                            af[i][(2 * s + 1) % NT1][1], b0, b1v);
            }
        }
    };
    bool v0[MTW], v1[MTW];
#pragma unroll
    for (int i = 0; i < MTW; ++i)
    {
        const int q0 = p0 + 16 * (warp + WARPS * i) + g;
        v0[i] = q0 < L;
        v1[i] = q0 + 8 < L;
    }
    if (PHASE == 1)
    {
        float a = 0.f, bq = 0.f;
#pragma unroll 4
        for (int nt = 0; nt < NT2; ++nt)
        {
            float d[MTW][4];
            conv2(nt, d);
#pragma unroll
            for (int i = 0; i < MTW; ++i)
            {
                if (v0[i])
                {
                    a += d[i][0] + d[i][1];
                    bq += d[i][0] * d[i][0] + d[i][1] * d[i][1];
                }
                if (v1[i])
                {
                    a += d[i][2] + d[i][3];
                    bq += d[i][2] * d[i][2] + d[i][3] * d[i][3];
                }
            }
        }
        double s1 = a, s2 = bq;
        block_sum2(s1, s2, red);
        if (tid == 0)
        {
            atomicAdd(&stats2[2 * b], s1);
            atomicAdd(&stats2[2 * b + 1], s2);
        }
        return;
    }
    {
        float mean, rstd;
        stats_to_mean_rstd(stats2 + 2 * b, (double)L * 2 * C, 1e-5f, mean, rstd);
#pragma unroll 2
        for (int nt = 0; nt < C / 8; ++nt)
        {
            float da[MTW][4], dg[MTW][4];
            conv2(nt, da);
            conv2(nt + C / 8, dg);
            const int c = 8 * nt + 2 * t;
            const float ga0 = rstd * g2w[c], ga1 = rstd * g2w[c + 1], oa0 = g2b[c], oa1 = g2b[c + 1];
            const float gg0 = rstd * g2w[C + c], gg1 = rstd * g2w[C + c + 1], og0 = g2b[C + c], og1 = g2b[C + c + 1];
            const float sc0 = scale[c], sc1 = scale[c + 1];
#pragma unroll
            for (int i = 0; i < MTW; ++i)
            {
                // the residual goes straight to global memory in the fragment layout: per request 8 rows x 32
                // contiguous bytes (whole sectors), no shared-memory copy of x and no block barrier in this phase
                float2 *r0 = reinterpret_cast<float2 *>(xg + (size_t)(p0 + 16 * (warp + WARPS * i) + g) * C + c);
                float2 *r1 = reinterpret_cast<float2 *>(reinterpret_cast<float *>(r0) + 8 * C);
                if (v0[i])
                {
                    float2 xv = *r0;
                    xv.x += sc0 * (((da[i][0] - mean) * ga0 + oa0) * sigmoid_ref((dg[i][0] - mean) * gg0 + og0));
                    xv.y += sc1 * (((da[i][1] - mean) * ga1 + oa1) * sigmoid_ref((dg[i][1] - mean) * gg1 + og1));
                    *r0 = xv;
                }
                if (v1[i])
                {
                    float2 xv = *r1;
                    xv.x += sc0 * (((da[i][2] - mean) * ga0 + oa0) * sigmoid_ref((dg[i][2] - mean) * gg0 + og0));
                    xv.y += sc1 * (((da[i][3] - mean) * ga1 + oa1) * sigmoid_ref((dg[i][3] - mean) * gg1 + og1));
                    *r1 = xv;
                }
            }
        }
    }
}

__host__ __device__ constexpr size_t time_smem(int C, int hid, int phase, int mtw)
{
    return phase == 0 ? (size_t)(16 * WARPS * mtw + 2 * HALO) * xld(C) * 2 + (size_t)hid_pad(hid) * k1ld(C) * 2 + hid_pad(hid) * 4 + 16
                      : (size_t)((nvec(C, hid_pad(hid)) + 3) & ~3) * 4 + (size_t)2 * C * w2ld(hid_pad(hid)) * 2 + 16;
}

// m-tiles per warp: 3 (336-position tiles) for the long 48- / 96-channel signals, 1 (112 positions) for the
// 192-channel level, whose items are short (336 frames, 5375 samples): three times the CTAs to fill the SMs
__host__ __device__ constexpr int time_mtw(int C) { return C >= 192 ? 1 : 3; }

template <int C, int HID> static int time_set_attr()
{
    constexpr int M = time_mtw(C);
    DMX_CUDA(cudaFuncSetAttribute(dconv_time_kernel<C, HID, 0, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)time_smem(C, HID, 0, M)));
    DMX_CUDA(cudaFuncSetAttribute(dconv_time_kernel<C, HID, 1, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)time_smem(C, HID, 1, M)));
    DMX_CUDA(cudaFuncSetAttribute(dconv_time_kernel<C, HID, 2, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)time_smem(C, HID, 2, M)));
    return 0;
}

template <int C, int HID>
static int time_launch(float *x, float *hbuf, const uint8_t *params, int B, int L, int hs, double *const *stats, cudaStream_t st)
{
    constexpr int M = time_mtw(C);
    constexpr int TT = 16 * WARPS * M;
    const dim3 grid((L + TT - 1) / TT, B);
    for (int sub = 0; sub < 2; ++sub)
    {
        const uint8_t *P = params + (size_t)sub * pk_sub_bytes(C, hid_pad(HID));
        double *s1 = stats[2 * sub], *s2 = stats[2 * sub + 1];
        dconv_time_kernel<C, HID, 0, M><<<grid, THREADS, time_smem(C, HID, 0, M), st>>>(x, hbuf, P, L, hs, 1 << sub, s1, s2);
        dconv_time_kernel<C, HID, 1, M><<<grid, THREADS, time_smem(C, HID, 1, M), st>>>(x, hbuf, P, L, hs, 1 << sub, s1, s2);
        dconv_time_kernel<C, HID, 2, M><<<grid, THREADS, time_smem(C, HID, 2, M), st>>>(x, hbuf, P, L, hs, 1 << sub, s1, s2);
    }
    DMX_LAUNCH_CHECK();
    return 0;
}

template <int C, int HID> static int set_attr()
{
    DMX_CUDA(cudaFuncSetAttribute(dconv_mma_kernel<C, HID, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)smem_bytes(C, HID)));
    return 0;
}
template <int C, int HID> static int launch(float *x, const uint8_t *params, int items, cudaStream_t st)
{
    dconv_mma_kernel<C, HID, false, false><<<items, THREADS, smem_bytes(C, HID), st>>>(x, params, nullptr, nullptr, nullptr, 1,
                                                                                    nullptr, nullptr);
    DMX_LAUNCH_CHECK();
    return 0;
}
template <int C, int HID> static int set_attr_rw()
{
    DMX_CUDA(cudaFuncSetAttribute(dconv_mma_kernel<C, HID, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)smem_bytes(C, HID) + rw_bytes(C)));
    DMX_CUDA(cudaFuncSetAttribute(dconv_mma_kernel<C, HID, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)smem_bytes(C, HID) + rw_bytes(C) + cin_bytes(C)));
    return 0;
}
template <int C, int HID>
static int launch_rw(float *x, const uint8_t *params, const uint8_t *rw, float *out, const float *rowadd, int F, int items,
                     const float *spec, const uint8_t *cin, cudaStream_t st)
{
    if (spec)
        dconv_mma_kernel<C, HID, true, true><<<items, THREADS, smem_bytes(C, HID) + rw_bytes(C) + cin_bytes(C), st>>>(
            x, params, rw, out, rowadd, F, spec, cin);
    else
        dconv_mma_kernel<C, HID, true, false><<<items, THREADS, smem_bytes(C, HID) + rw_bytes(C), st>>>(x, params, rw, out, rowadd,
                                                                                                  F, nullptr, nullptr);
    DMX_LAUNCH_CHECK();
    return 0;
}

} // namespace dm

bool dconv_mma_supported(int C, int hid, int V)
{
    return V == dm::T && ((C == 48 && (hid == 6 || hid == 12)) || (C == 96 && (hid == 12 || hid == 24)));
}

int dconv_mma_init_device()
{
    return dm::set_attr_rw<48, 6>() || dm::set_attr<48, 6>() || dm::set_attr<48, 12>() || dm::set_attr<96, 12>() || dm::set_attr<96, 24>() ||
           dm::time_set_attr<48, 6>() ||
           dm::time_set_attr<48, 12>() || dm::time_set_attr<96, 12>() || dm::time_set_attr<96, 24>() ||
           dm::time_set_attr<192, 24>();
}

bool dconv_time_mma_supported(int C, int hid, int hs)
{
    return (hs % 2) == 0 && hs >= hid &&
           ((C == 48 && (hid == 6 || hid == 12)) || (C == 96 && (hid == 12 || hid == 24)) || (C == 192 && hid == 24));
}

// Encoder level with C = 48, hidden 6 (htdemucs encoder 0): DConv + the following 1x1 rewrite / GLU / frequency
// embedding in one launch.  x [items][336][48] (DConv input, left holding the first sub-block's update only),
// out [items][336][48]; rw = block built by dconv_mma_rw_pack; rowadd [F][48] or null, u = item % F.
bool dconv_mma_rw_supported(int C, int hid, int V) { return V == dm::T && C == 48 && hid == 6; }
int dconv_mma_rw_pack_floats(int C) { return dm::rw_bytes(C) / 4; }
void dconv_mma_rw_pack(int C, const float *W, const float *bias, float *outp)
{
    uint8_t *o = reinterpret_cast<uint8_t *>(outp);
    memset(o, 0, (size_t)dm::rw_bytes(C));
    __nv_bfloat16 *w = reinterpret_cast<__nv_bfloat16 *>(o);
    const int ld = dm::xld(C);
    for (int n = 0; n < 2 * C; ++n)
        for (int k = 0; k < C; ++k)
            w[n * ld + k] = __float2bfloat16(W[(size_t)n * C + k]);
    memcpy(o + (size_t)2 * C * ld * 2, bias, (size_t)2 * C * 4);
}
// Input conv block (Cin = 4, k8 along frequency): W [C][4][8] (PyTorch conv weight order) -> bf16 [C][40], k = tap*4 + ch
int dconv_mma_cin_pack_floats(int C) { return dm::cin_bytes(C) / 4; }
void dconv_mma_cin_pack(int C, const float *W, const float *bias, float *outp)
{
    uint8_t *o = reinterpret_cast<uint8_t *>(outp);
    memset(o, 0, (size_t)dm::cin_bytes(C));
    __nv_bfloat16 *w = reinterpret_cast<__nv_bfloat16 *>(o);
    for (int n = 0; n < C; ++n)
        for (int ch = 0; ch < 4; ++ch)
            for (int tap = 0; tap < 8; ++tap)
                w[n * 40 + tap * 4 + ch] = __float2bfloat16(W[((size_t)n * 4 + ch) * 8 + tap]);
    memcpy(o + (size_t)C * 40 * 2, bias, (size_t)C * 4);
}
// spec / cin non-null: x is first computed from the spectrogram [B][4F][336][4] by the input conv + GELU
int launch_dconv_mma_rw(float *x, const void *packed, const void *rw, float *out, const float *rowadd, int F, int items,
                        int C, int hid, const float *spec, const void *cin, cudaStream_t st)
{
    if (C == 48 && hid == 6)
        return dm::launch_rw<48, 6>(x, static_cast<const uint8_t *>(packed), static_cast<const uint8_t *>(rw), out, rowadd, F,
                                    items, spec, static_cast<const uint8_t *>(cin), st);
    set_error("dconv_mma_rw: unsupported shape");
    return -1;
}

// Floats (4-byte units) of the packed parameter block of one DConv (two sub-blocks); 0 = shape not handled here.
int dconv_mma_pack_floats(int C, int hid)
{
    if (!dconv_mma_supported(C, hid, dm::T) && !dconv_time_mma_supported(C, hid, hid + (hid & 1)))
        return 0;
    return 2 * dm::pk_sub_bytes(C, dm::hid_pad(hid)) / 4;
}

// blk: the fp32 block of dconv_fused.cu (per sub-block W1[3][C][hid] | b1 g1w g1b | W2[2C][hid] | b2 g2w g2b | scale)
// -> out: the shared-memory images of dconv_mma.cu, weights rounded to bf16 (round to nearest even)
void dconv_mma_pack(int C, int hid, const float *blk, float *out)
{
    const int hp = dm::hid_pad(hid), K1 = dm::k1ld(C), W2 = dm::w2ld(hp);
    const int sub = dm::pk_sub_bytes(C, hp);
    uint8_t *o = reinterpret_cast<uint8_t *>(out);
    memset(o, 0, (size_t)2 * sub);
    for (int s = 0; s < 2; ++s)
    {
        const float *P = blk + (size_t)s * dm::p_stride(C, hid);
        const float *Pb1 = P + 3 * C * hid, *Pg1w = Pb1 + hid, *Pg1b = Pg1w + hid, *PW2 = Pg1b + hid;
        const float *Pb2 = PW2 + 2 * C * hid, *Pg2w = Pb2 + 2 * C, *Pg2b = Pg2w + 2 * C, *Psc = Pg2b + 2 * C;
        __nv_bfloat16 *w1 = reinterpret_cast<__nv_bfloat16 *>(o + (size_t)s * sub);
        __nv_bfloat16 *w2 = reinterpret_cast<__nv_bfloat16 *>(o + (size_t)s * sub + dm::pk_w1_bytes(C, hp));
        float *vec = reinterpret_cast<float *>(o + (size_t)s * sub + dm::pk_w1_bytes(C, hp) + dm::pk_w2_bytes(C, hp));
        for (int n = 0; n < hid; ++n)
            for (int k = 0; k < 3 * C; ++k)
                w1[n * K1 + k] = __float2bfloat16(P[(size_t)k * hid + n]);
        for (int n = 0; n < 2 * C; ++n)
            for (int k = 0; k < hid; ++k)
                w2[n * W2 + k] = __float2bfloat16(PW2[(size_t)n * hid + k]);
        float *b1 = vec, *g1w = b1 + hp, *g1b = g1w + hp, *b2 = g1b + hp, *g2w = b2 + 2 * C, *g2b = g2w + 2 * C,
              *scale = g2b + 2 * C;
        for (int i = 0; i < hid; ++i)
        {
            b1[i] = Pb1[i];
            g1w[i] = Pg1w[i];
            g1b[i] = Pg1b[i];
        }
        for (int i = 0; i < 2 * C; ++i)
        {
            b2[i] = Pb2[i];
            g2w[i] = Pg2w[i];
            g2b[i] = Pg2b[i];
        }
        for (int i = 0; i < C; ++i)
            scale[i] = Psc[i];
    }
}

// Tiled form: x [B][L][C] in place (B GroupNorm items of L positions: batch elements of the time branch, or
// frequency rows with L = 336), hbuf [B][L][hs] scratch, stats: four zeroed (sum, sumsq)[B] buffers (sub-block 0 GN1,
// GN2, sub-block 1 GN1, GN2); six launches.
int launch_dconv_time_mma(float *x, float *hbuf, const void *packed, int B, int L, int C, int hid, int hs,
                          double *const *stats, cudaStream_t st)
{
    const uint8_t *params = static_cast<const uint8_t *>(packed);
    if (C == 48 && hid == 6)
        return dm::time_launch<48, 6>(x, hbuf, params, B, L, hs, stats, st);
    if (C == 48 && hid == 12)
        return dm::time_launch<48, 12>(x, hbuf, params, B, L, hs, stats, st);
    if (C == 96 && hid == 12)
        return dm::time_launch<96, 12>(x, hbuf, params, B, L, hs, stats, st);
    if (C == 96 && hid == 24)
        return dm::time_launch<96, 24>(x, hbuf, params, B, L, hs, stats, st);
    if (C == 192 && hid == 24)
        return dm::time_launch<192, 24>(x, hbuf, params, B, L, hs, stats, st);
    set_error("dconv_time_mma: unsupported shape");
    return -1;
}

// x: [items][336][C] in place; items = B * F frequency rows; packed: the block built by dconv_mma_pack
int launch_dconv_mma(float *x, const void *packed, int items, int C, int hid, cudaStream_t st)
{
    const uint8_t *params = static_cast<const uint8_t *>(packed);
    if (C == 48 && hid == 6)
        return dm::launch<48, 6>(x, params, items, st);
    if (C == 48 && hid == 12)
        return dm::launch<48, 12>(x, params, items, st);
    if (C == 96 && hid == 12)
        return dm::launch<96, 12>(x, params, items, st);
    if (C == 96 && hid == 24)
        return dm::launch<96, 24>(x, params, items, st);
    set_error("dconv_mma: unsupported shape");
    return -1;
}

} // namespace dmx
