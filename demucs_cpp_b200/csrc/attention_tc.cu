// bf16 tensor-core attention for sm_100a (bf16 mode): softmax(Q K^T / sqrt(dh)) V per head with the
// streaming-softmax formulation, both GEMMs on tcgen05 with accumulators in TMEM.
//
// One CTA = 128 queries of one (batch item, head).  Per 128-key tile:
//     S  = Q K^T   tcgen05.mma  M=128, N=128, K=dh   (Q, K: K-major canonical smem tiles)     -> TMEM cols [0,128)
//     each thread owns one query row: tcgen05.ld its S row, scale, running max / sum in fp32, P = exp(S - m)
//     written as bf16 into the canonical K-major A-operand layout (conflict-free 16-byte stores)
//     PV = P V     tcgen05.mma  M=128, N=dh,  K=128  (V: MN-major canonical tile, i.e. stored as it lies in
//                                                      memory, dh contiguous)                   -> TMEM cols [128,128+dh)
//     O  = O * exp(m_old - m_new) + PV   in registers (fp32)
// Q/K/V arrive as bf16 (the QKV projection's epilogue writes them that way); the output is fp32 because it
// feeds the out-projection's fp32 A-gather.  The T x S score matrix never leaves the SM.
// Reference: src/layers.cpp:442-482 (heads = contiguous dh-wide column blocks, max-subtracted softmax).
#include "kernels.cuh"

#include <cuda_bf16.h>

namespace dmx
{
namespace atc
{

constexpr int BQ = 128;   // queries per CTA (UMMA M)
constexpr int BKEY = 128; // keys per tile
constexpr int THREADS = 128;
constexpr int PAD = 16;   // bytes added to the K-direction core stride: conflict-free 16-byte stores

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
                 "selp.u32 %0, 1, 0, p;\n"
                 "}\n"
                 : "=r"(ok)
                 : "r"(smem_u32(bar)), "r"(parity)
                 : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t n = 0;
    while (!mbar_try_wait(bar, parity))
        if (++n > (1u << 22))
            __trap();
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo)
{
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
// kind::f16, D = F32, A = B = BF16, A K-major; b_mn = 1 selects an MN-major B operand
__device__ __forceinline__ uint32_t make_idesc(int n, int b_mn)
{
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)b_mn << 16) | ((uint32_t)(n >> 3) << 17) |
           ((uint32_t)(BQ >> 4) << 24);
}
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum)
{
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "setp.ne.b32 p, %4, 0;\n"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
                 "}\n" ::"r"(tmem_d),
                 "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
                 : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t *bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float *v)
{
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i)
        v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi)
{
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t *>(&h);
}

// DH = head dim (64 or 48).  smem tiles:
//   Qs: K-major, rows = queries, k = dh   : chunk (row, kc) at kc*LBO_R + row*16,  LBO_R = 128*16 + PAD
//   Ks: K-major, rows = keys,    k = dh   : same geometry
//   Ps: K-major, rows = queries, k = keys : chunk (row, kc) at kc*LBO_R + row*16,  16 chunks
//   Vs: MN-major, mn = dh, k = keys       : chunk (key, dg) at dg*SBO_V + key*16,  SBO_V = 128*16 + PAD
//       (k-groups of 8 keys are 128 B apart = LBO_V)
template <int DH>
__global__ void __launch_bounds__(THREADS) attention_tc_kernel(const __nv_bfloat16 *__restrict__ Q, int ldq,
                                                               const __nv_bfloat16 *__restrict__ K, int ldk,
                                                               const __nv_bfloat16 *__restrict__ V, int ldv,
                                                               float *__restrict__ O, int ldo, int T, int S, float scale,
                                                               uint4 *__restrict__ Opk, int heads)
{
    constexpr int DCH = DH / 8;                    // 16-byte chunks per row of Q/K/V
    constexpr uint32_t LBO_R = BQ * 16 + PAD;      // 2064
    constexpr uint32_t SBO_V = BKEY * 16 + PAD;    // 2064
    constexpr uint32_t Q_BYTES = DCH * LBO_R, K_BYTES = DCH * LBO_R, V_BYTES = DCH * SBO_V, P_BYTES = 16 * LBO_R;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *Qs = smem, *Ks = Qs + Q_BYTES, *Vs = Ks + K_BYTES, *Ps = Vs + V_BYTES;
    __shared__ __align__(8) uint64_t mma_bar;
    __shared__ uint32_t tmem_slot;
    (void)P_BYTES;

    const int tid = threadIdx.x, warp = tid >> 5;
    const int q0 = blockIdx.x * BQ;
    const int h = blockIdx.y, b = blockIdx.z;
    const __nv_bfloat16 *Qb = Q + (size_t)b * T * ldq + h * DH;
    const __nv_bfloat16 *Kb = K + (size_t)b * S * ldk + h * DH;
    const __nv_bfloat16 *Vb = V + (size_t)b * S * ldv + h * DH;

    if (tid == 0)
    {
        mbar_init(&mma_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                     "r"(256u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // Q tile: chunk (row, kc): 8 lanes of a row read DH*2 contiguous bytes
    for (int idx = tid; idx < BQ * DCH; idx += THREADS)
    {
        const int row = idx / DCH, kc = idx % DCH;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (q0 + row < T)
            v = __ldg(reinterpret_cast<const uint4 *>(Qb + (size_t)(q0 + row) * ldq + kc * 8));
        *reinterpret_cast<uint4 *>(Qs + kc * LBO_R + row * 16) = v;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_S = tmem_slot, tmem_PV = tmem_slot + 128;
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
    const uint32_t idesc_s = make_idesc(BKEY, 0), idesc_pv = make_idesc(DH, 1);

    float o[DH];
#pragma unroll
    for (int i = 0; i < DH; ++i)
        o[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;
    uint32_t phase = 0;

    for (int k0 = 0; k0 < S; k0 += BKEY)
    {
        // ---- K and V tiles (previous tile's MMAs have completed: every thread waited on mma_bar) ----
        for (int idx = tid; idx < BKEY * DCH; idx += THREADS)
        {
            const int key = idx / DCH, c = idx % DCH;
            uint4 kv = make_uint4(0u, 0u, 0u, 0u), vv = kv;
            if (k0 + key < S)
            {
                kv = __ldg(reinterpret_cast<const uint4 *>(Kb + (size_t)(k0 + key) * ldk + c * 8));
                vv = __ldg(reinterpret_cast<const uint4 *>(Vb + (size_t)(k0 + key) * ldv + c * 8));
            }
            *reinterpret_cast<uint4 *>(Ks + c * LBO_R + key * 16) = kv;
            *reinterpret_cast<uint4 *>(Vs + c * SBO_V + key * 16) = vv;
        }
        fence_proxy_async();
        __syncthreads();
        // ---- S = Q K^T ----
        if (tid == 0)
        {
            tc_fence_after();
            const uint32_t qa = smem_u32(Qs), ka = smem_u32(Ks);
#pragma unroll
            for (int j = 0; j < DH / 16; ++j)
                umma(tmem_S, make_desc(qa + j * 2 * LBO_R, LBO_R, 128), make_desc(ka + j * 2 * LBO_R, LBO_R, 128), idesc_s,
                     j > 0 ? 1u : 0u);
            umma_commit(&mma_bar);
        }
        mbar_wait(&mma_bar, phase);
        phase ^= 1u;
        tc_fence_after();
        // ---- streaming softmax on this thread's row ----
        // two passes over the TMEM row (TMEM reads are cheap; keeping all 128 scores live would spill)
        float corr;
        {
            float mx = -INFINITY;
#pragma unroll
            for (int c = 0; c < BKEY; c += 16)
            {
                float s[16];
                tmem_ld16(tmem_S + lane_off + (uint32_t)c, s);
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (k0 + c + j < S)
                        mx = fmaxf(mx, s[j]);
            }
            mx *= scale; // scale > 0: max commutes with it
            const float m_new = fmaxf(m_run, mx);
            corr = (m_run == -INFINITY) ? 0.f : __expf(m_run - m_new);
            float rs = 0.f;
#pragma unroll
            for (int c = 0; c < BKEY; c += 16)
            {
                float s[16];
                tmem_ld16(tmem_S + lane_off + (uint32_t)c, s);
                float pv[16];
#pragma unroll
                for (int j = 0; j < 16; ++j)
                {
                    pv[j] = (k0 + c + j < S) ? __expf(s[j] * scale - m_new) : 0.f;
                    rs += pv[j];
                }
                uint4 pk;
                pk.x = pack_bf16x2(pv[0], pv[1]);
                pk.y = pack_bf16x2(pv[2], pv[3]);
                pk.z = pack_bf16x2(pv[4], pv[5]);
                pk.w = pack_bf16x2(pv[6], pv[7]);
                *reinterpret_cast<uint4 *>(Ps + (c / 8) * LBO_R + tid * 16) = pk;
                pk.x = pack_bf16x2(pv[8], pv[9]);
                pk.y = pack_bf16x2(pv[10], pv[11]);
                pk.z = pack_bf16x2(pv[12], pv[13]);
                pk.w = pack_bf16x2(pv[14], pv[15]);
                *reinterpret_cast<uint4 *>(Ps + (c / 8 + 1) * LBO_R + tid * 16) = pk;
            }
            l_run = l_run * corr + rs;
            m_run = m_new;
        }
        tc_fence_before();
        fence_proxy_async();
        __syncthreads();
        // ---- PV = P V ----
        if (tid == 0)
        {
            tc_fence_after();
            const uint32_t pa = smem_u32(Ps), va = smem_u32(Vs);
#pragma unroll
            for (int j = 0; j < BKEY / 16; ++j)
                umma(tmem_PV, make_desc(pa + j * 2 * LBO_R, LBO_R, 128), make_desc(va + j * 256, 128, SBO_V), idesc_pv,
                     j > 0 ? 1u : 0u);
            umma_commit(&mma_bar);
        }
        mbar_wait(&mma_bar, phase);
        phase ^= 1u;
        tc_fence_after();
        {
            float pv[DH];
#pragma unroll
            for (int c = 0; c < DH; c += 16)
                tmem_ld16(tmem_PV + lane_off + (uint32_t)c, &pv[c]);
#pragma unroll
            for (int i = 0; i < DH; ++i)
                o[i] = o[i] * corr + pv[i];
        }
        tc_fence_before();
    }
    const int row = q0 + tid;
    if (row < T)
    {
        const float inv = 1.0f / l_run;
        if (Opk)
        {
            // tile-packed bf16: the out-projection GEMM's A image, [row/128][d/8][row%128][8]
            const size_t grow = (size_t)b * T + row;
            const int nch = heads * DH / 8;
#pragma unroll
            for (int i = 0; i < DH; i += 8)
            {
                uint4 pk;
                pk.x = pack_bf16x2(o[i] * inv, o[i + 1] * inv);
                pk.y = pack_bf16x2(o[i + 2] * inv, o[i + 3] * inv);
                pk.z = pack_bf16x2(o[i + 4] * inv, o[i + 5] * inv);
                pk.w = pack_bf16x2(o[i + 6] * inv, o[i + 7] * inv);
                Opk[((grow >> 7) * nch + (h * DH + i) / 8) * 128 + (grow & 127)] = pk;
            }
        }
        else
        {
            float *dst = O + ((size_t)b * T + row) * ldo + h * DH;
#pragma unroll
            for (int i = 0; i < DH; i += 4)
                *reinterpret_cast<float4 *>(dst + i) = make_float4(o[i] * inv, o[i + 1] * inv, o[i + 2] * inv, o[i + 3] * inv);
        }
    }
    __syncthreads();
    if (warp == 0)
    {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_slot), "r"(256u) : "memory");
    }
}

template <int DH> static int smem_bytes()
{
    constexpr int DCH = DH / 8;
    return DCH * (BQ * 16 + PAD) * 2 + DCH * (BKEY * 16 + PAD) + 16 * (BQ * 16 + PAD);
}

} // namespace atc

int attention_tc_init_device()
{
    DMX_CUDA(cudaFuncSetAttribute(atc::attention_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  atc::smem_bytes<64>()));
    DMX_CUDA(cudaFuncSetAttribute(atc::attention_tc_kernel<48>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  atc::smem_bytes<48>()));
    return 0;
}

// Q/K/V: bf16, O: fp32.  Leading dimensions in elements; all must be multiples of 8 (16-byte rows).
int launch_attention_tc(const void *Q, int ldq, const void *K, int ldk, const void *V, int ldv, float *O, int ldo, int B,
                        int T, int S, int heads, int dh, cudaStream_t st, void *Opk)
{
    if ((ldq | ldk | ldv) % 8 || ldo % 4)
    {
        set_error("attention_tc: leading dimensions must be multiples of 8 (bf16) / 4 (fp32)");
        return -1;
    }
    const float scale = 1.0f / sqrtf((float)dh);
    dim3 grid((T + atc::BQ - 1) / atc::BQ, heads, B);
    const __nv_bfloat16 *q = static_cast<const __nv_bfloat16 *>(Q), *k = static_cast<const __nv_bfloat16 *>(K),
                        *v = static_cast<const __nv_bfloat16 *>(V);
    if (dh == 64)
        atc::attention_tc_kernel<64><<<grid, atc::THREADS, atc::smem_bytes<64>(), st>>>(q, ldq, k, ldk, v, ldv, O, ldo, T, S,
                                                                                          scale, (uint4 *)Opk, heads);
    else if (dh == 48)
        atc::attention_tc_kernel<48><<<grid, atc::THREADS, atc::smem_bytes<48>(), st>>>(q, ldq, k, ldk, v, ldv, O, ldo, T, S,
                                                                                          scale, (uint4 *)Opk, heads);
    else
    {
        set_error("attention_tc: unsupported head dim");
        return -1;
    }
    DMX_LAUNCH_CHECK();
    return 0;
}

} // namespace dmx
