// bf16 tensor-core tap-GEMM for sm_100a: tcgen05.mma (one elected thread issues), fp32 accumulators in
// TMEM, operands in shared memory in the canonical no-swizzle K-major core-matrix layout.
//
//   * A operand (activations, fp32 channels-last in HBM) is gathered per output row by 128 producer
//     threads -- one tile row each -- with the tap / zero-padding predicates of the implicit GEMM and the
//     fused prologues (skip add, GroupNorm+GELU) applied in registers, converted to bf16 and stored as
//     16-byte core-matrix rows (a warp writes 512 contiguous bytes: conflict free).
//   * B operand (weights) is pre-packed on the host as the exact shared-memory image of every
//     (n-tile, k-block) and fetched with ONE 1-D bulk async copy (cp.async.bulk ... mbarrier
//     complete_tx) per stage -- no tensor map needed.
//   * A 3/4-stage mbarrier ring couples producers and the MMA thread (full[] : 128 row arrivals + the
//     bulk copy's transaction bytes; empty[] : tcgen05.commit).  The same 128 threads then run the
//     epilogue: tcgen05.ld of their own TMEM lane (= their output row), bias / GELU / GLU / LayerScale
//     residual / GroupNorm statistics / transposed-conv scatter, fp32 stores.
//
// Tile: 128 rows x BN columns (BN = N rounded to 16, <= 256, chosen per layer), BK = 64.
//
// HALO mode (stride-1 convolutions with >= 3 taps: the 3x3 / k3 decoder rewrites and the dilated k3 convs of the
// DConv branches): a tile is 128 consecutive positions v of ONE (item, u) line, and the producers stage the
// input patch -- (u taps) planes x (128 + v-tap span) rows x Cin channels -- ONCE, as bf16 in the canonical
// layout [plane][Cin/8][row][8].  Every tap's A operand is then just a descriptor into that patch (plane
// selects du, a 16-byte start-address shift per row selects dv), so each input row crosses L2 -> SM once per
// tile instead of once per tap (the gather of the 3x3 rewrite ran at the L2 bandwidth limit: 6 GB for a 0.5 GB
// tensor).  Only the weights still stream through the k-block ring.
#include "kernels.cuh"

#include <cuda.h>
#include <cuda_bf16.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace dmx
{
namespace tc
{

constexpr int BM = 128;
constexpr int BK = 64;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
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
// Bounded wait: a protocol bug traps (launch error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t n = 0;
    while (!mbar_try_wait(bar, parity))
        if (++n > (1u << 22))
            __trap();
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// Shared-memory matrix descriptor, SWIZZLE_NONE, K-major: core matrix = 8 rows x 16 bytes;
// LBO = byte distance between the two K-adjacent core matrices of one UMMA_K=16 slice,
// SBO = byte distance between 8-row groups.  (bit layout: cute/arch/mma_sm100_desc.hpp SmemDescriptor)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo)
{
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46; // descriptor version for sm_100
    return d;               // base_offset = 0, lbo_mode = 0, layout_type = SWIZZLE_NONE (0)
}

// Instruction descriptor for kind::f16: D = F32, A = B = BF16, both K-major, M = 128, N = bn.
// The same for kind::tf32: A = B = TF32 (fp32 words in shared memory), K = 8 per instruction.
__device__ __forceinline__ uint32_t make_idesc_tf32(int bn)
{
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}

__device__ __forceinline__ uint32_t make_idesc(int bn)
{
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
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
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16])
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

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32])
{
    uint32_t r[32];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]),
                   "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]),
                   "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i)
        v[i] = __uint_as_float(r[i]);
}

struct RowCoord
{
    int b, uo, vo;
};
__device__ __forceinline__ RowCoord decode_row(int m, const GemmParams &p)
{
    RowCoord rc;
    rc.b = fast_div(m, p.div_per);
    const int r = m - rc.b * (p.Uo * p.Vo);
    rc.uo = fast_div(r, p.div_vo);
    rc.vo = r - rc.uo * p.Vo;
    return rc;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi)
{
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t *>(&h);
}

// Per-row bookkeeping for the epilogue.
struct RowInfo
{
    long long outoff; // element offset in C of (b, uo*ous + ouo, vo*ovs + ovo), channel 0 (may be "virtual")
    int uu0, vv0;     // that output position (for the bounds checks of the scatter mapping)
    int uo, item;     // input-row frequency index (row-add table) and statistics item (-1: row invalid)
    int prow, orow;   // row index in the tile-packed output space (out_pack); output position index outoff / ldc
};

constexpr int PROD_WARPS = 8;
constexpr int PROD_THREADS = PROD_WARPS * 32;   // 256: A producers, then epilogue
constexpr int THREADS = PROD_THREADS + 64;      // + warp 8 (B loader) + warp 9 (MMA issuer)
constexpr int A_LBO = BM * 16 + 16;             // 2064 B: +16 B pad makes the producers' 16-byte stores conflict free
constexpr int A_STAGE = 8 * A_LBO;              // 16512 B per stage
constexpr int EPI_LD = 36;                      // floats per staged row (32 + 4 pad)
constexpr int EPI_WARP = 5120;                   // per-warp epilogue buffer: 32 x EPI_LD floats, or two 2 KB store tiles (1 KB aligned)
constexpr int EPI_BYTES = PROD_WARPS * EPI_WARP;

// Store 4 consecutive OUTPUT columns nc0..nc0+3 of one row (general mapping: transposed-conv scatter).
__device__ __forceinline__ void store_group4_scatter(const GemmParams &p, const RowInfo &ri, int nc0, int nlimit,
                                                     float (&o)[4], float &rs, float &rss)
{
#pragma unroll
    for (int e = 0; e < 4; ++e)
    {
        const int nc = nc0 + e;
        if (nc < nlimit)
        {
            const int r = nc / p.Nco;
            const int co = nc - r * p.Nco;
            const int uu = ri.uu0 + r * p.our, vv = ri.vv0 + r * p.ovr;
            if (uu >= 0 && uu < p.Uout && vv >= 0 && vv < p.Vout)
            {
                float *dst = p.C + ri.outoff +
                             (p.out_vu ? (long long)r * p.our + (long long)r * p.ovr * p.Uout
                                       : (long long)r * p.our * p.Vout + (long long)r * p.ovr) * p.ldc + co;
                float v = o[e];
                if (p.rowadd)
                    v += __ldg(p.rowadd + (size_t)ri.uo * p.rowadd_ld + co);
                if (p.addC)
                    v += __ldg(p.addC + (dst - p.C));
                if (p.gamma)
                    v = *dst + __ldg(p.gamma + nc) * v;
                *dst = v;
                rs += v;
                rss += v * v;
            }
        }
    }
}

// The same for mappings whose channel groups never straddle a phase (Nco % 4 == 0, aligned rows): one phase
// decode, one bounds check and one 16-byte store per group.
__device__ __forceinline__ void store_group4_scatter_vec(const GemmParams &p, const RowInfo &ri, int nc0, float (&o)[4],
                                                         float &rs, float &rss)
{
    const int r = fast_div(nc0, p.div_nco);
    const int co = nc0 - r * p.Nco;
    const int uu = ri.uu0 + r * p.our, vv = ri.vv0 + r * p.ovr;
    if ((unsigned)uu < (unsigned)p.Uout && (unsigned)vv < (unsigned)p.Vout)
    {
        float *dst = p.C + ri.outoff +
                     (p.out_vu ? (long long)r * p.our + (long long)r * p.ovr * p.Uout
                               : (long long)r * p.our * p.Vout + (long long)r * p.ovr) * p.ldc + co;
        if (p.rowadd)
        {
            const float4 ra = __ldg(reinterpret_cast<const float4 *>(p.rowadd + (size_t)ri.uo * p.rowadd_ld + co));
            o[0] += ra.x;
            o[1] += ra.y;
            o[2] += ra.z;
            o[3] += ra.w;
        }
        if (p.addC)
        {
            const float4 ad = __ldg(reinterpret_cast<const float4 *>(p.addC + (dst - p.C)));
            o[0] += ad.x;
            o[1] += ad.y;
            o[2] += ad.z;
            o[3] += ad.w;
        }
        if (p.gamma)
        {
            const float4 g = __ldg(reinterpret_cast<const float4 *>(p.gamma + nc0));
            const float4 old = *reinterpret_cast<const float4 *>(dst);
            o[0] = old.x + g.x * o[0];
            o[1] = old.y + g.y * o[1];
            o[2] = old.z + g.z * o[2];
            o[3] = old.w + g.w * o[3];
        }
        if (!p.skip_c)
            *reinterpret_cast<float4 *>(dst) = make_float4(o[0], o[1], o[2], o[3]);
        if (p.Cbf)
        {
            uint2 pk2;
            pk2.x = pack_bf16x2(o[0], o[1]);
            pk2.y = pack_bf16x2(o[2], o[3]);
            *reinterpret_cast<uint2 *>(reinterpret_cast<__nv_bfloat16 *>(p.Cbf) +
                                       ((long long)ri.orow + (long long)r * p.our * p.Vout + (long long)r * p.ovr) * p.Cbf_ld + co) = pk2;
        }
        rs = o[0] + o[1] + o[2] + o[3];
        rss = o[0] * o[0] + o[1] * o[1] + o[2] * o[2] + o[3] * o[3];
    }
}

// HALO mode: stage the input patch of tile (item hb, line hu, first position hv0) as bf16 into a_smem.
// unit = (patch row, 8-channel chunk), chunk fastest: consecutive lanes read consecutive 32-byte pieces of the
// channels-last input (a whole plane of the patch is one contiguous span of memory).
__device__ __forceinline__ void halo_stage_patch(const GemmParams &p, const int tid, const int c8n, const uint32_t h_lbo,
                                                 const int hb, const int hu, const int hv0, uint8_t *a_smem)
{
    const int rows = p.halo_rows;
    const int total = p.halo_nu * rows * c8n;
    const long long a2d = p.A2 ? (p.A2 - p.A) : 0;
    for (int u0 = tid; u0 < total; u0 += 4 * PROD_THREADS)
    {
        float4 lo[4], hi[4];
        int dsto[4];
        const float *src[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
        {
            const int u = u0 + i * PROD_THREADS;
            dsto[i] = -1;
            src[i] = nullptr;
            if (u < total)
            {
                const int prow = fast_div(u, p.div_c8);
                const int kcc = u - prow * c8n;
                const int plane = fast_div(prow, p.div_rows);
                const int r = prow - plane * rows;
                const int ui = hu + p.halo_dumin + plane, vi = hv0 + p.halo_dvmin + r;
                dsto[i] = (plane * c8n + kcc) * (int)h_lbo + r * 16;
                if ((unsigned)ui < (unsigned)p.Ui && (unsigned)vi < (unsigned)p.Vi)
                    src[i] = p.A + (((long long)hb * p.Ui + ui) * p.Vi + vi) * p.lda + kcc * 8;
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
        {
            lo[i] = hi[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (src[i])
            {
                lo[i] = __ldg(reinterpret_cast<const float4 *>(src[i]));
                hi[i] = __ldg(reinterpret_cast<const float4 *>(src[i] + 4));
            }
        }
        if (p.A2)
        {
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (src[i])
                {
                    const float4 t0 = __ldg(reinterpret_cast<const float4 *>(src[i] + a2d));
                    const float4 t1 = __ldg(reinterpret_cast<const float4 *>(src[i] + a2d + 4));
                    lo[i].x += t0.x;
                    lo[i].y += t0.y;
                    lo[i].z += t0.z;
                    lo[i].w += t0.w;
                    hi[i].x += t1.x;
                    hi[i].y += t1.y;
                    hi[i].z += t1.z;
                    hi[i].w += t1.w;
                }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (dsto[i] >= 0)
            {
                uint4 pk;
                pk.x = pack_bf16x2(lo[i].x, lo[i].y);
                pk.y = pack_bf16x2(lo[i].z, lo[i].w);
                pk.z = pack_bf16x2(hi[i].x, hi[i].y);
                pk.w = pack_bf16x2(hi[i].z, hi[i].w);
                *reinterpret_cast<uint4 *>(a_smem + dsto[i]) = pk;
            }
    }
}

// Gather bookkeeping of a producer thread: base pointer and valid-tap mask of its four tile rows r0 + 32 i.
__device__ __forceinline__ void producer_rows(const GemmParams &p, const int m0, const int r0, const float *(&rp)[4],
                                              uint32_t (&rmask)[4])
{
#pragma unroll
    for (int i = 0; i < 4; ++i)
    {
        const int m = m0 + r0 + 32 * i;
        rmask[i] = 0;
        rp[i] = p.A;
        if (m < p.M)
        {
            const RowCoord rc = decode_row(m, p);
            const int ui0 = rc.uo * p.su, vi0 = rc.vo * p.sv;
            rp[i] = p.A + (((long long)rc.b * p.Ui + ui0) * p.Vi + vi0) * p.lda;
            for (int t = 0; t < p.ntaps; ++t)
                if ((unsigned)(ui0 + p.du[t]) < (unsigned)p.Ui && (unsigned)(vi0 + p.dv[t]) < (unsigned)p.Vi)
                    rmask[i] |= 1u << t;
        }
    }
}

// One k-block of the A operand for a producer thread: chunk column k..k+7 of its four rows, with the tap /
// zero-padding predicates, the skip add and the GroupNorm+GELU prologue applied in registers, as packed bf16.
template <bool VECA>
__device__ __forceinline__ void gather_kblock(const GemmParams &p, const int k, const int r0, const float *const (&rp)[4],
                                              const uint32_t (&rmask)[4], const long long a2d, const float2 *rowstat,
                                              uint4 (&pk)[4])
{
    if (VECA)
    {
        const bool h0 = k < p.K, h1 = k + 4 < p.K;
        int tap0 = 0, tap1 = 0, c0 = 0, c1 = 0;
        long long toff0 = 0, toff1 = 0;
        if (h0)
        {
            tap0 = k / p.Cin;
            c0 = k - tap0 * p.Cin;
            toff0 = ((long long)p.du[tap0] * p.Vi + p.dv[tap0]) * p.lda + c0;
        }
        if (h1)
        {
            tap1 = (k + 4) / p.Cin;
            c1 = (k + 4) - tap1 * p.Cin;
            toff1 = ((long long)p.du[tap1] * p.Vi + p.dv[tap1]) * p.lda + c1;
        }
        float4 lo[4], hi[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
        {
            const bool v0 = h0 && ((rmask[i] >> tap0) & 1u), v1 = h1 && ((rmask[i] >> tap1) & 1u);
            lo[i] = v0 ? __ldg(reinterpret_cast<const float4 *>(rp[i] + toff0)) : make_float4(0.f, 0.f, 0.f, 0.f);
            hi[i] = v1 ? __ldg(reinterpret_cast<const float4 *>(rp[i] + toff1)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (p.A2)
        {
#pragma unroll
            for (int i = 0; i < 4; ++i)
            {
                const bool v0 = h0 && ((rmask[i] >> tap0) & 1u), v1 = h1 && ((rmask[i] >> tap1) & 1u);
                if (v0)
                {
                    const float4 t = __ldg(reinterpret_cast<const float4 *>(rp[i] + a2d + toff0));
                    lo[i].x += t.x;
                    lo[i].y += t.y;
                    lo[i].z += t.z;
                    lo[i].w += t.w;
                }
                if (v1)
                {
                    const float4 t = __ldg(reinterpret_cast<const float4 *>(rp[i] + a2d + toff1));
                    hi[i].x += t.x;
                    hi[i].y += t.y;
                    hi[i].z += t.z;
                    hi[i].w += t.w;
                }
            }
        }
        if (p.pro == 1)
        {
            float4 w0 = make_float4(0.f, 0.f, 0.f, 0.f), b0 = w0, w1 = w0, b1 = w0;
            if (h0)
            {
                w0 = __ldg(reinterpret_cast<const float4 *>(p.pro_w + c0));
                b0 = __ldg(reinterpret_cast<const float4 *>(p.pro_b + c0));
            }
            if (h1)
            {
                w1 = __ldg(reinterpret_cast<const float4 *>(p.pro_w + c1));
                b1 = __ldg(reinterpret_cast<const float4 *>(p.pro_b + c1));
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
            {
                const float2 ms = rowstat[r0 + 32 * i];
                if (h0 && ((rmask[i] >> tap0) & 1u))
                {
                    lo[i].x = gelu_erf((lo[i].x - ms.x) * ms.y * w0.x + b0.x);
                    lo[i].y = gelu_erf((lo[i].y - ms.x) * ms.y * w0.y + b0.y);
                    lo[i].z = gelu_erf((lo[i].z - ms.x) * ms.y * w0.z + b0.z);
                    lo[i].w = gelu_erf((lo[i].w - ms.x) * ms.y * w0.w + b0.w);
                }
                if (h1 && ((rmask[i] >> tap1) & 1u))
                {
                    hi[i].x = gelu_erf((hi[i].x - ms.x) * ms.y * w1.x + b1.x);
                    hi[i].y = gelu_erf((hi[i].y - ms.x) * ms.y * w1.y + b1.y);
                    hi[i].z = gelu_erf((hi[i].z - ms.x) * ms.y * w1.z + b1.z);
                    hi[i].w = gelu_erf((hi[i].w - ms.x) * ms.y * w1.w + b1.w);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
        {
            pk[i].x = pack_bf16x2(lo[i].x, lo[i].y);
            pk[i].y = pack_bf16x2(lo[i].z, lo[i].w);
            pk[i].z = pack_bf16x2(hi[i].x, hi[i].y);
            pk[i].w = pack_bf16x2(hi[i].z, hi[i].w);
        }
    }
    else
    {
#pragma unroll
        for (int i = 0; i < 4; ++i)
        {
            float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            const float2 ms = rowstat[r0 + 32 * i];
#pragma unroll
            for (int e = 0; e < 8; ++e)
            {
                const int ke = k + e;
                if (ke < p.K)
                {
                    const int tap = ke / p.Cin;
                    const int c = ke - tap * p.Cin;
                    if ((rmask[i] >> tap) & 1u)
                    {
                        const long long off = ((long long)p.du[tap] * p.Vi + p.dv[tap]) * p.lda + c;
                        float t = __ldg(rp[i] + off);
                        if (p.A2)
                            t += __ldg(rp[i] + a2d + off);
                        if (p.pro == 1)
                            t = gelu_erf((t - ms.x) * ms.y * __ldg(&p.pro_w[c]) + __ldg(&p.pro_b[c]));
                        v[e] = t;
                    }
                }
            }
            pk[i].x = pack_bf16x2(v[0], v[1]);
            pk[i].y = pack_bf16x2(v[2], v[3]);
            pk[i].z = pack_bf16x2(v[4], v[5]);
            pk[i].w = pack_bf16x2(v[6], v[7]);
        }
    }
}

__device__ __forceinline__ void bulk_s2g(void *gdst, const void *ssrc, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

__device__ __forceinline__ bool bulk_ok(const GemmParams &p, const int mode)
{
    return mode == 1 && !p.addC && !p.rowadd && !p.out_pack && !p.out_bf16 && !p.Cbf && (p.N % (p.glu ? 8 : 4)) == 0;
}

// Epilogue for row-major fp32 outputs with the identity mapping (the convolutions and rewrites of the encoders /
// decoders, the DConv GEMMs, the LayerScale-residual linears of the transformer): a lane keeps the row of its TMEM lane, applies bias / GELU /
// GLU in registers, writes the finished piece of its row (32 raw columns) into its own 144-byte slot of the warp's
// staging buffer and hands it to the bulk-copy engine (cp.async.bulk shared -> global): the global stores cost one
// instruction per row piece, are fully contiguous per request, and nothing is exchanged between lanes.  The
// transposing form below spent ~35 instructions per 4 outputs, and these layers are bound by instruction issue.
__device__ __forceinline__ void epilogue_tile_bulk(const GemmParams &p, const int bn, const int n0, const uint32_t lane_addr,
                                                   const RowInfo &ri, float *stage, const int c_begin, const int c_step,
                                                   const int lane)
{
    const int ntile_end = (n0 + bn) < p.N ? (n0 + bn) : p.N;
    const bool valid = ri.item >= 0;
    float *slot = stage + lane * EPI_LD;
    float *crow = p.C + ri.outoff;
    double st_s = 0.0, st_ss = 0.0;
    for (int c0 = c_begin; c0 < bn; c0 += c_step)
    {
        const int rawc0 = n0 + c0;
        if (rawc0 >= ntile_end)
            break;
        float v[32];
        tmem_ld32(lane_addr + (uint32_t)c0, v);
        const int nraw = (ntile_end - rawc0) < 32 ? (ntile_end - rawc0) : 32; // multiple of 4 (8 with GLU)
        if (p.bias)
        {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
                if (j < nraw)
                {
                    const float4 b = __ldg(reinterpret_cast<const float4 *>(p.bias + rawc0 + j));
                    v[j] += b.x;
                    v[j + 1] += b.y;
                    v[j + 2] += b.z;
                    v[j + 3] += b.w;
                }
        }
        if (p.act == 1)
        {
#pragma unroll
            for (int j = 0; j < 32; ++j)
                v[j] = gelu_erf(v[j]);
        }
        int nout = nraw;
        if (p.glu)
        {
#pragma unroll
            for (int j = 0; j < 16; ++j)
                v[j] = v[2 * j] * sigmoid_ref(v[2 * j + 1]);
            nout = nraw >> 1;
        }
        if (p.gamma && valid)
        {
            // LayerScale residual: the lane reads the old piece of its own row (the residual stream is L2 resident)
            const float *orow = crow + (p.glu ? (rawc0 >> 1) : rawc0);
            const float *grow = p.gamma + (p.glu ? (rawc0 >> 1) : rawc0);
#pragma unroll
            for (int j = 0; j < 32; j += 4)
                if (j < nout)
                {
                    const float4 o = *reinterpret_cast<const float4 *>(orow + j);
                    const float4 gm = __ldg(reinterpret_cast<const float4 *>(grow + j));
                    v[j] = o.x + gm.x * v[j];
                    v[j + 1] = o.y + gm.y * v[j + 1];
                    v[j + 2] = o.z + gm.z * v[j + 2];
                    v[j + 3] = o.w + gm.w * v[j + 3];
                }
        }
        float rs = 0.f, rss = 0.f;
        bulk_wait_read(); // the previous piece has left the slot
#pragma unroll
        for (int j = 0; j < 32; j += 4)
            if (j < nout)
            {
                *reinterpret_cast<float4 *>(slot + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                rs += v[j] + v[j + 1] + v[j + 2] + v[j + 3];
                rss += v[j] * v[j] + v[j + 1] * v[j + 1] + v[j + 2] * v[j + 2] + v[j + 3] * v[j + 3];
            }
        fence_proxy_async();
        if (valid)
        {
            bulk_s2g(crow + (p.glu ? (rawc0 >> 1) : rawc0), slot, (uint32_t)nout * 4u);
            bulk_commit();
            st_s += (double)rs;
            st_ss += (double)rss;
        }
    }
    bulk_wait_read(); // the staging buffer is reused by the next tile / freed at exit
    if (p.stats)
    {
        // one item per warp in the common case: reduce across lanes before touching global memory
        const int item0 = __shfl_sync(0xffffffffu, ri.item, 0);
        if (__all_sync(0xffffffffu, ri.item == item0))
        {
            for (int o = 16; o > 0; o >>= 1)
            {
                st_s += __shfl_xor_sync(0xffffffffu, st_s, o);
                st_ss += __shfl_xor_sync(0xffffffffu, st_ss, o);
            }
            if (lane == 0 && item0 >= 0)
            {
                atomicAdd(&p.stats[2 * (size_t)item0 + 0], st_s);
                atomicAdd(&p.stats[2 * (size_t)item0 + 1], st_ss);
            }
        }
        else if (valid)
        {
            atomicAdd(&p.stats[2 * (size_t)ri.item + 0], st_s);
            atomicAdd(&p.stats[2 * (size_t)ri.item + 1], st_ss);
        }
    }
}

__device__ __forceinline__ bool scatter_lean(const GemmParams &p, const int plain_out)
{
    return plain_out == 2 && !p.glu && !p.gamma && !p.rowadd && !p.stats && (p.N & 3) == 0 && !p.out_bf16 && !p.out_pack &&
           p.dbg_epi != 4;
}

// The transposed convolutions' epilogue (scatter mapping with 16-byte aligned 4-channel groups, no GLU / LayerScale /
// row-add / statistics): the same staging as epilogue_tile below, but a lane's column group -- hence its phase r, its
// channel and every phase-dependent offset -- is fixed for the whole chunk and decoded ONCE, and the skip operand of all
// rows a lane serves is loaded four rows at a time before the first one is used.  (Measured: the generic path spent 1.9 of
// the 2.9 ms these layers take per 16 segments here, 1.2 of them in the load -> add -> store chain of one row at a time;
// now 0.9.  Requesting the skip rows into L2 while the tile's MMAs still run changed nothing.)
__device__ __forceinline__ void epilogue_tile_scatter(const GemmParams &p, const int bn, const int n0, const uint32_t lane_addr,
                                                      const RowInfo *rowinfo, float *stage, const int q, const int c_begin,
                                                      const int c_step, const int lane)
{
    const int ntile_end = (n0 + bn) < p.N ? (n0 + bn) : p.N;
    const int g = lane & 7, rsub = lane >> 3;
    for (int c0 = c_begin; c0 < bn; c0 += c_step)
    {
        float val[32];
        tmem_ld16(lane_addr + (uint32_t)c0, *reinterpret_cast<float(*)[16]>(&val[0]));
        if (c0 + 16 < bn)
            tmem_ld16(lane_addr + (uint32_t)(c0 + 16), *reinterpret_cast<float(*)[16]>(&val[16]));
        else
        {
#pragma unroll
            for (int j = 16; j < 32; ++j)
                val[j] = 0.f;
        }
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<float4 *>(&stage[lane * EPI_LD + j]) = make_float4(val[j], val[j + 1], val[j + 2], val[j + 3]);
        __syncwarp();
        const int outc = n0 + c0 + 4 * g;
        const bool have = outc + 4 <= ntile_end; // N % 4 == 0: the group is whole or absent
        const int r = fast_div(outc, p.div_nco);
        const int co = outc - r * p.Nco;
        const int du = r * p.our, dv = r * p.ovr;
        const long long poff = (p.out_vu ? (long long)du + (long long)dv * p.Uout : (long long)du * p.Vout + (long long)dv) * p.ldc + co;
        const float *addc = p.addC ? p.addC + poff : nullptr;
        float *cdst = p.C + poff;
        __nv_bfloat16 *bdst =
            p.Cbf ? reinterpret_cast<__nv_bfloat16 *>(p.Cbf) + ((long long)du * p.Vout + (long long)dv) * p.Cbf_ld + co : nullptr;
        float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p.bias)
            b4 = __ldg(reinterpret_cast<const float4 *>(p.bias + outc));
        const RowInfo *rq = rowinfo + q * 32 + rsub;
        // two batches of four rows: a batch's skip loads are all in flight before the first one is consumed
#pragma unroll
        for (int hb = 0; hb < 2; ++hb)
        {
            float4 ad[4];
            long long off[4];
            unsigned okmask = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i)
            {
                const RowInfo &ri = rq[(hb * 4 + i) * 4];
                const bool ok = have && ri.item >= 0 && (unsigned)(ri.uu0 + du) < (unsigned)p.Uout &&
                                (unsigned)(ri.vv0 + dv) < (unsigned)p.Vout;
                okmask |= ok ? (1u << i) : 0u;
                off[i] = ri.outoff;
                ad[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (ok && addc)
                    ad[i] = __ldg(reinterpret_cast<const float4 *>(addc + off[i]));
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
            {
                if (!((okmask >> i) & 1u))
                    continue;
                const int it = hb * 4 + i;
                const int rl = it * 4 + rsub;
                const float4 t0 = *reinterpret_cast<const float4 *>(&stage[rl * EPI_LD + 4 * g]);
                float o0 = t0.x + b4.x, o1 = t0.y + b4.y, o2 = t0.z + b4.z, o3 = t0.w + b4.w;
                if (p.act == 1)
                {
                    o0 = gelu_erf(o0);
                    o1 = gelu_erf(o1);
                    o2 = gelu_erf(o2);
                    o3 = gelu_erf(o3);
                }
                o0 += ad[i].x;
                o1 += ad[i].y;
                o2 += ad[i].z;
                o3 += ad[i].w;
                if (!p.skip_c)
                    *reinterpret_cast<float4 *>(cdst + off[i]) = make_float4(o0, o1, o2, o3);
                if (bdst)
                {
                    uint2 pk2;
                    pk2.x = pack_bf16x2(o0, o1);
                    pk2.y = pack_bf16x2(o2, o3);
                    *reinterpret_cast<uint2 *>(bdst + (long long)rq[it * 4].orow * p.Cbf_ld) = pk2;
                }
            }
        }
        __syncwarp(); // converged again before the next chunk's aligned TMEM loads
    }
}

// Epilogue of one 128 x bn tile for the TMEM lane quadrant q (tile rows 32q..32q+31), column chunks c_begin,
// c_begin + c_step, ...  Per 32-column chunk each lane pulls its own row's raw accumulators from TMEM and stages
// them in shared memory (stage: this warp's private 32 x EPI_LD floats); the warp then re-reads them transposed so
// that 8 lanes write 128 contiguous bytes of one output row: coalesced stores / residual reads.  Bias, GELU, GLU,
// row-add, skip add, LayerScale residual and the statistics are applied on the transposed side.
__device__ __forceinline__ void epilogue_tile(const GemmParams &p, const int bn, const int plain_out, const int n0,
                                              const uint32_t lane_addr, const RowInfo *rowinfo, float *stage, const int q,
                                              const int c_begin, const int c_step, const int lane)
{
    if (scatter_lean(p, plain_out))
    {
        epilogue_tile_scatter(p, bn, n0, lane_addr, rowinfo, stage, q, c_begin, c_step, lane);
        return;
    }
    const int ntile_end = (n0 + bn) < p.N ? (n0 + bn) : p.N; // never beyond N nor into the next column tile
    const int nlimit = p.glu ? (ntile_end >> 1) : ntile_end;   // valid OUTPUT columns
    const int gpr = p.glu ? 4 : 8;                             // 4-column output groups per row per chunk
    const int rows_per_it = 32 / gpr;
    const int g = lane % gpr, rsub = lane / gpr;
    int cur_item = -1;
    double st_s = 0.0, st_ss = 0.0;
    for (int c0 = c_begin; c0 < bn; c0 += c_step)
    {
        float val[32];
        tmem_ld16(lane_addr + (uint32_t)c0, *reinterpret_cast<float(*)[16]>(&val[0]));
        if (c0 + 16 < bn)
            tmem_ld16(lane_addr + (uint32_t)(c0 + 16), *reinterpret_cast<float(*)[16]>(&val[16]));
        else
        {
#pragma unroll
            for (int j = 16; j < 32; ++j)
                val[j] = 0.f;
        }
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<float4 *>(&stage[lane * EPI_LD + j]) = make_float4(val[j], val[j + 1], val[j + 2], val[j + 3]);
        __syncwarp();
        const int nbase = n0 + c0;                          // first RAW column of the chunk
        // this lane's raw columns: plain 4g..4g+3, GLU 8g..8g+7 (-> outputs 4g..4g+3)
        const int rawc = nbase + (p.glu ? 8 * g : 4 * g);
        const int outc = (p.glu ? (nbase >> 1) : nbase) + 4 * g;
        float bias[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (p.bias)
        {
            const float4 b0 = __ldg(reinterpret_cast<const float4 *>(p.bias + rawc)); // zero padded to ldw
            bias[0] = b0.x;
            bias[1] = b0.y;
            bias[2] = b0.z;
            bias[3] = b0.w;
            if (p.glu)
            {
                const float4 b1 = __ldg(reinterpret_cast<const float4 *>(p.bias + rawc + 4));
                bias[4] = b1.x;
                bias[5] = b1.y;
                bias[6] = b1.z;
                bias[7] = b1.w;
            }
        }
        float4 gam = make_float4(0.f, 0.f, 0.f, 0.f);
        const bool full4 = outc + 4 <= nlimit;
        if (p.gamma && full4)
            gam = __ldg(reinterpret_cast<const float4 *>(p.gamma + outc));
        // Residual / skip operands of the identity mapping: pull all of this lane's rows towards L1 first, so the
        // loop below does not expose one full memory latency per row (load -> fma -> store chains).
        const float *pre_src = p.gamma ? p.C : p.addC;
        if (plain_out == 1 && full4 && pre_src != nullptr)
        {
            for (int it = 0; it < gpr; ++it)
            {
                const RowInfo &ri = rowinfo[q * 32 + it * rows_per_it + rsub];
                if (ri.item >= 0)
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(pre_src + ri.outoff + outc));
            }
        }
        for (int it = 0; it < gpr; ++it)
        {
            const int rl = it * rows_per_it + rsub;
            const RowInfo ri = rowinfo[q * 32 + rl];
            float rs = 0.f, rss = 0.f;
            if (ri.item >= 0) // row inside the problem
            {
                float o[4];
                if (p.glu)
                {
                    const float4 t0 = *reinterpret_cast<const float4 *>(&stage[rl * EPI_LD + 8 * g]);
                    const float4 t1 = *reinterpret_cast<const float4 *>(&stage[rl * EPI_LD + 8 * g + 4]);
                    float r8[8] = {t0.x + bias[0], t0.y + bias[1], t0.z + bias[2], t0.w + bias[3],
                                   t1.x + bias[4], t1.y + bias[5], t1.z + bias[6], t1.w + bias[7]};
                    if (p.act == 1)
                    {
#pragma unroll
                        for (int e = 0; e < 8; ++e)
                            r8[e] = gelu_erf(r8[e]);
                    }
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        o[e] = r8[2 * e] * sigmoid_ref(r8[2 * e + 1]);
                }
                else
                {
                    const float4 t0 = *reinterpret_cast<const float4 *>(&stage[rl * EPI_LD + 4 * g]);
                    o[0] = t0.x + bias[0];
                    o[1] = t0.y + bias[1];
                    o[2] = t0.z + bias[2];
                    o[3] = t0.w + bias[3];
                    if (p.act == 1)
                    {
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            o[e] = gelu_erf(o[e]);
                    }
                }
                if (plain_out == 2 && full4)
                    store_group4_scatter_vec(p, ri, outc, o, rs, rss);
                else if (plain_out == 1 && full4)
                {
                    // identity mapping: output row == GEMM row, 16-byte aligned groups
                    float *dst = p.C + ri.outoff + outc;
                    if (p.rowadd)
                    {
                        const float4 ra = __ldg(reinterpret_cast<const float4 *>(p.rowadd + (size_t)ri.uo * p.rowadd_ld + outc));
                        o[0] += ra.x;
                        o[1] += ra.y;
                        o[2] += ra.z;
                        o[3] += ra.w;
                    }
                    if (p.addC)
                    {
                        const float4 ad = *reinterpret_cast<const float4 *>(p.addC + ri.outoff + outc);
                        o[0] += ad.x;
                        o[1] += ad.y;
                        o[2] += ad.z;
                        o[3] += ad.w;
                    }
                    if (p.gamma)
                    {
                        const float4 old = *reinterpret_cast<const float4 *>(dst);
                        o[0] = old.x + gam.x * o[0];
                        o[1] = old.y + gam.y * o[1];
                        o[2] = old.z + gam.z * o[2];
                        o[3] = old.w + gam.w * o[3];
                    }
                    if (p.out_pack)
                    {
                        // tile-packed bf16 for a consumer GEMM with K = ldc: [row/128][col/8][row%128][col%8]
                        const int mrow = ri.prow;
                        const size_t off = ((size_t)(mrow >> 7) * (p.ldc >> 3) + (outc >> 3)) * 1024 + (size_t)(mrow & 127) * 8 +
                                           (outc & 7);
                        uint2 pk2;
                        pk2.x = pack_bf16x2(o[0], o[1]);
                        pk2.y = pack_bf16x2(o[2], o[3]);
                        *reinterpret_cast<uint2 *>(reinterpret_cast<__nv_bfloat16 *>(p.C) + off) = pk2;
                    }
                    else if (p.out_bf16)
                    {
                        uint2 pk2;
                        pk2.x = pack_bf16x2(o[0], o[1]);
                        pk2.y = pack_bf16x2(o[2], o[3]);
                        *reinterpret_cast<uint2 *>(reinterpret_cast<__nv_bfloat16 *>(p.C) + ri.outoff + outc) = pk2;
                    }
                    else
                    {
                        *reinterpret_cast<float4 *>(dst) = make_float4(o[0], o[1], o[2], o[3]);
                        if (p.Cbf)
                        {
                            uint2 pk2;
                            pk2.x = pack_bf16x2(o[0], o[1]);
                            pk2.y = pack_bf16x2(o[2], o[3]);
                            *reinterpret_cast<uint2 *>(reinterpret_cast<__nv_bfloat16 *>(p.Cbf) + (long long)ri.orow * p.Cbf_ld + outc) = pk2;
                        }
                    }
                    rs = o[0] + o[1] + o[2] + o[3];
                    rss = o[0] * o[0] + o[1] * o[1] + o[2] * o[2] + o[3] * o[3];
                }
                else if (outc < nlimit)
                    store_group4_scatter(p, ri, outc, nlimit, o, rs, rss);
            }
            if (p.stats)
            {
                if (ri.item != cur_item)
                {
                    if (cur_item >= 0)
                    {
                        atomicAdd(&p.stats[2 * (size_t)cur_item + 0], st_s);
                        atomicAdd(&p.stats[2 * (size_t)cur_item + 1], st_ss);
                    }
                    cur_item = ri.item;
                    st_s = 0.0;
                    st_ss = 0.0;
                }
                st_s += (double)rs;
                st_ss += (double)rss;
            }
        }
    }
    if (p.stats)
    {
        // one item per warp in the common case: reduce across lanes before touching global memory
        const int item0 = __shfl_sync(0xffffffffu, cur_item, 0);
        if (__all_sync(0xffffffffu, cur_item == item0))
        {
            for (int o = 16; o > 0; o >>= 1)
            {
                st_s += __shfl_xor_sync(0xffffffffu, st_s, o);
                st_ss += __shfl_xor_sync(0xffffffffu, st_ss, o);
            }
            if (lane == 0 && item0 >= 0)
            {
                atomicAdd(&p.stats[2 * (size_t)item0 + 0], st_s);
                atomicAdd(&p.stats[2 * (size_t)item0 + 1], st_ss);
            }
        }
        else if (cur_item >= 0)
        {
            atomicAdd(&p.stats[2 * (size_t)cur_item + 0], st_s);
            atomicAdd(&p.stats[2 * (size_t)cur_item + 1], st_ss);
        }
    }
}

__device__ __forceinline__ void tma_load_4d(void *smem_dst, const CUtensorMap *tm, uint64_t *bar, int c0, int c1, int c2, int c3)
{
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                 ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap *tm, const void *smem_src, int c0, int c1, int c2)
{
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                     reinterpret_cast<uint64_t>(tm)),
                 "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }

// Shared-memory matrix descriptor of a swizzled K-major operand whose rows are one swizzle span wide (pitch bytes):
// 8-row groups lie SBO = 8 * pitch apart; the leading-dimension offset is not used.  layout: 2 = 128B, 4 = 64B, 6 = 32B.
__device__ __forceinline__ uint64_t make_desc_sw(uint32_t saddr, uint32_t sbo, uint32_t layout)
{
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(layout & 7u) << 61;
    return d;
}

// Epilogue for row-major fp32 outputs with the identity mapping: per 32 raw columns a lane applies bias / GELU / GLU (and the
// LayerScale residual, reading the old piece of its own row) to its own row, the warp assembles 32 rows x 16 columns in
// shared memory -- 16-byte pieces XOR-placed as the tensor map's SWIZZLE_64B expects, so the stores do not collide on
// banks -- and lane 0 hands the tile to ONE tensor-map store; rows past the end of the line (and columns past N) are
// clipped by the map.  Two 2 KB tiles per warp alternate.  (One cp.async.bulk per lane and piece, as epilogue_tile_bulk
// does, runs at the copy engine's ~12 ns per operation: 4.5 us for a 128 x 96 tile.)
__device__ __forceinline__ void epilogue_tile_tmast(const GemmParams &p, const CUtensorMap *tmC, const int bn, const int n0,
                                                    const uint32_t lane_addr, const RowInfo &ri, uint8_t *stage, const int c_begin,
                                                    const int c_step, const int lane, const int v_row0, const int line, int &parity)
{
    const int ntile_end = (n0 + bn) < p.N ? (n0 + bn) : p.N;
    const bool valid = ri.item >= 0;
    double st_s = 0.0, st_ss = 0.0;
    const int sw = (lane >> 1) & 3; // SWIZZLE_64B phase of this lane's 64-byte row
    const float *crow = p.C + ri.outoff;
    for (int c0 = c_begin; c0 < bn; c0 += c_step)
    {
        const int rawc0 = n0 + c0;
        if (rawc0 >= ntile_end)
            break;
        float v[32];
        tmem_ld32(lane_addr + (uint32_t)c0, v);
        const int nraw = (ntile_end - rawc0) < 32 ? (ntile_end - rawc0) : 32;
        if (p.bias)
        {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
                if (j < nraw)
                {
                    const float4 b = __ldg(reinterpret_cast<const float4 *>(p.bias + rawc0 + j));
                    v[j] += b.x;
                    v[j + 1] += b.y;
                    v[j + 2] += b.z;
                    v[j + 3] += b.w;
                }
        }
        if (p.act == 1)
        {
#pragma unroll
            for (int j = 0; j < 32; ++j)
                v[j] = gelu_erf(v[j]);
        }
        int nout = nraw;
        if (p.glu)
        {
#pragma unroll
            for (int j = 0; j < 16; ++j)
                v[j] = v[2 * j] * sigmoid_ref(v[2 * j + 1]);
            nout = nraw >> 1;
        }
        const int outc0 = p.glu ? (rawc0 >> 1) : rawc0;
        if (p.gamma && valid)
        {
            const float *orow = crow + outc0;
            const float *grow = p.gamma + outc0;
#pragma unroll
            for (int j = 0; j < 32; j += 4)
                if (j < nout)
                {
                    const float4 o = *reinterpret_cast<const float4 *>(orow + j);
                    const float4 gm = __ldg(reinterpret_cast<const float4 *>(grow + j));
                    v[j] = o.x + gm.x * v[j];
                    v[j + 1] = o.y + gm.y * v[j + 1];
                    v[j + 2] = o.z + gm.z * v[j + 2];
                    v[j + 3] = o.w + gm.w * v[j + 3];
                }
        }
        float rs = 0.f, rss = 0.f;
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (16 * h < nout)
            {
                uint8_t *buf = stage + parity * 2048;
                parity ^= 1;
                if (lane == 0)
                    bulk_wait_read1(); // the store issued two tiles ago has drained this buffer
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 4; ++j)
                {
                    const int e = 16 * h + 4 * j;
                    *reinterpret_cast<float4 *>(buf + lane * 64 + ((j ^ sw) << 4)) = make_float4(v[e], v[e + 1], v[e + 2], v[e + 3]);
                    if (e < nout)
                    {
                        rs += v[e] + v[e + 1] + v[e + 2] + v[e + 3];
                        rss += v[e] * v[e] + v[e + 1] * v[e + 1] + v[e + 2] * v[e + 2] + v[e + 3] * v[e + 3];
                    }
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0)
                {
                    tma_store_3d(tmC, buf, outc0 + 16 * h, v_row0, line);
                    bulk_commit();
                }
            }
        if (valid)
        {
            st_s += (double)rs;
            st_ss += (double)rss;
        }
    }
    if (p.stats)
    {
        const int item0 = __shfl_sync(0xffffffffu, ri.item, 0);
        if (__all_sync(0xffffffffu, ri.item == item0))
        {
            for (int o = 16; o > 0; o >>= 1)
            {
                st_s += __shfl_xor_sync(0xffffffffu, st_s, o);
                st_ss += __shfl_xor_sync(0xffffffffu, st_ss, o);
            }
            if (lane == 0 && item0 >= 0)
            {
                atomicAdd(&p.stats[2 * (size_t)item0 + 0], st_s);
                atomicAdd(&p.stats[2 * (size_t)item0 + 1], st_ss);
            }
        }
        else if (valid)
        {
            atomicAdd(&p.stats[2 * (size_t)ri.item + 0], st_s);
            atomicAdd(&p.stats[2 * (size_t)ri.item + 1], st_ss);
        }
    }
}

// Epilogue for row-major fp32 outputs with the identity mapping whose tile rows are CONSECUTIVE output rows (flat GEMM rows, or
// positions of one line): per 32 raw columns a lane finishes its own row in registers (bias / GELU / GLU, statistics), the
// warp turns the 32 x 32 (or 32 x 16) piece around through its private buffer, and every store instruction then writes whole
// 128-byte (64-byte) row segments -- with the LayerScale residual read the same coalesced way.  No row table, no per-row
// address arithmetic.  (The tensor-map store above moves such a piece with one instruction, but the copy engine handles a
// tile row by row: for outputs with a 2 KB row pitch -- the transformer's 512-wide residual stream -- it sustained only
// ~9 ns per 64-byte row, 17 us per 128 x 256 tile against 2 us of MMAs.)
__device__ __forceinline__ void epilogue_tile_direct(const GemmParams &p, const int bn, const int n0, const uint32_t lane_addr,
                                                     const int item, float *crow0, const int nvalid, float *stage, const int c_begin,
                                                     const int c_step, const int lane)
{
    // crow0: output row of tile row q*32 (this warp's first row); nvalid: how many of the warp's 32 rows exist
    const int ntile_end = (n0 + bn) < p.N ? (n0 + bn) : p.N;
    const bool valid = lane < nvalid;
    double st_s = 0.0, st_ss = 0.0;
    const int gpr = p.glu ? 4 : 8;       // 4-column groups per row of a piece
    const int rpi = 32 / gpr;            // rows per store instruction
    const int g = lane % gpr, rsub = lane / gpr;
    for (int c0 = c_begin; c0 < bn; c0 += c_step)
    {
        const int rawc0 = n0 + c0;
        if (rawc0 >= ntile_end)
            break;
        float v[32];
        tmem_ld32(lane_addr + (uint32_t)c0, v);
        const int nraw = (ntile_end - rawc0) < 32 ? (ntile_end - rawc0) : 32;
        if (p.bias)
        {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
                if (j < nraw)
                {
                    const float4 b = __ldg(reinterpret_cast<const float4 *>(p.bias + rawc0 + j));
                    v[j] += b.x;
                    v[j + 1] += b.y;
                    v[j + 2] += b.z;
                    v[j + 3] += b.w;
                }
        }
        if (p.act == 1)
        {
#pragma unroll
            for (int j = 0; j < 32; ++j)
                v[j] = gelu_erf(v[j]);
        }
        int nout = nraw;
        if (p.glu)
        {
#pragma unroll
            for (int j = 0; j < 16; ++j)
                v[j] = v[2 * j] * sigmoid_ref(v[2 * j + 1]);
            nout = nraw >> 1;
        }
        const int outc0 = p.glu ? (rawc0 >> 1) : rawc0;
        __syncwarp(); // the previous piece has been read back
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (j < gpr)
                *reinterpret_cast<float4 *>(&stage[lane * EPI_LD + 4 * j]) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        __syncwarp();
        const int col = outc0 + 4 * g;
        const bool colok = 4 * g < nout;
        float4 gam = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p.gamma && colok)
            gam = __ldg(reinterpret_cast<const float4 *>(p.gamma + col));
        float rs = 0.f, rss = 0.f;
#pragma unroll
        for (int it = 0; it < 8; ++it)
            if (it < gpr)
            {
                const int r = it * rpi + rsub;
                if (r < nvalid && colok)
                {
                    float4 o = *reinterpret_cast<const float4 *>(&stage[r * EPI_LD + 4 * g]);
                    float *dst = crow0 + (size_t)r * p.ldc + col;
                    if (p.gamma)
                    {
                        const float4 old = *reinterpret_cast<const float4 *>(dst);
                        o.x = old.x + gam.x * o.x;
                        o.y = old.y + gam.y * o.y;
                        o.z = old.z + gam.z * o.z;
                        o.w = old.w + gam.w * o.w;
                    }
                    *reinterpret_cast<float4 *>(dst) = o;
                    rs += o.x + o.y + o.z + o.w;
                    rss += o.x * o.x + o.y * o.y + o.z * o.z + o.w * o.w;
                }
            }
        st_s += (double)rs;
        st_ss += (double)rss;
    }
    (void)valid;
    if (p.stats)
    {
        // the warp's rows belong to one statistics item in the common case (checked by the caller: item >= 0), else -1
        for (int o = 16; o > 0; o >>= 1)
        {
            st_s += __shfl_xor_sync(0xffffffffu, st_s, o);
            st_ss += __shfl_xor_sync(0xffffffffu, st_ss, o);
        }
        if (lane == 0 && item >= 0)
        {
            atomicAdd(&p.stats[2 * (size_t)item + 0], st_s);
            atomicAdd(&p.stats[2 * (size_t)item + 1], st_ss);
        }
    }
}

template <int STAGES, bool VECA, bool HALO>
__global__ void __launch_bounds__(THREADS, 3) gemm_tc_kernel(const __grid_constant__ CUtensorMap tmC, const GemmParams p,
                                                          const __nv_bfloat16 *__restrict__ Wb, const int bn,
                                                          const int plain_out, const int tma_store)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[STAGES];
    __shared__ __align__(8) uint64_t empty_bar[STAGES];
    __shared__ __align__(8) uint64_t tmem_full_bar;
    __shared__ __align__(8) uint64_t a_full_bar; // HALO: the whole patch is staged
    __shared__ uint32_t tmem_slot;
    __shared__ RowInfo rowinfo[BM];
    __shared__ float2 rowstat[BM]; // (mean, rstd) of the GroupNorm prologue per row

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.x * BM;
    const int ntile = blockIdx.y;
    const int n0 = ntile * bn;
    const int nk = (p.K + BK - 1) / BK;
    const uint32_t b_stage_bytes = (uint32_t)bn * BK * 2;
    // HALO: tile -> (item hb, line hu, first position hv0); the patch replaces the A ring
    int hb = 0, hu = 0, hv0 = 0;
    const int c8n = p.Cin >> 3;
    const uint32_t h_lbo = (uint32_t)(p.halo_rows | 1) * 16; // odd row count: conflict-free 16-byte producer stores
    if (HALO)
    {
        const int bu = fast_div((int)blockIdx.x, p.div_tpr);
        hv0 = ((int)blockIdx.x - bu * p.halo_tpr) * BM;
        hb = fast_div(bu, p.div_uo);
        hu = bu - hb * p.Uo;
    }
    const uint32_t a_bytes = HALO ? (((uint32_t)p.halo_nu * c8n * h_lbo + 127u) & ~127u) : (uint32_t)(STAGES * A_STAGE);
    uint8_t *a_smem = smem;                       // STAGES x A_STAGE, or the HALO patch
    uint8_t *b_smem = smem + a_bytes;              // STAGES x bn*128 B
    uint32_t tmem_cols = 32;
    while ((int)tmem_cols < bn)
        tmem_cols <<= 1;

    if (tid == 0)
    {
        for (int s = 0; s < STAGES; ++s)
        {
            mbar_init(&full_bar[s], (p.Apk || HALO) ? 1u : (uint32_t)(PROD_THREADS + 1));
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&tmem_full_bar, 1);
        mbar_init(&a_full_bar, PROD_THREADS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < BM)
    {
        int m = m0 + tid;
        bool valid = m < p.M;
        RowCoord rc;
        if (HALO)
        {
            valid = hv0 + tid < p.Vo;
            rc.b = hb;
            rc.uo = hu;
            rc.vo = valid ? hv0 + tid : 0;
            m = (hb * p.Uo + hu) * p.Vo + rc.vo;
        }
        else
            rc = decode_row(valid ? m : 0, p);
        RowInfo ri;
        ri.uu0 = rc.uo * p.ous + p.ouo;
        ri.vv0 = rc.vo * p.ovs + p.ovo;
        ri.outoff = (p.out_vu ? ((long long)rc.b * p.Vout + ri.vv0) * p.Uout + ri.uu0 : ((long long)rc.b * p.Uout + ri.uu0) * p.Vout + ri.vv0) * p.ldc;
        ri.uo = rc.uo;
        ri.item = valid ? (p.stat_per_u ? rc.b * p.Uo + rc.uo : rc.b) : -1;
        ri.prow = p.pack_T ? rc.b * p.pack_T + rc.uo * p.Vo + rc.vo : m; // per-item padding to whole 128-row tiles
        ri.orow = (rc.b * p.Uout + ri.uu0) * p.Vout + ri.vv0;
        rowinfo[tid] = ri;
        float pmean = 0.f, prstd = 1.f;
        if (p.pro == 1 && valid)
        {
            const int ui = rc.uo * p.su + p.du[0];
            const int item = p.pro_per_u ? rc.b * p.Ui + ui : rc.b;
            stats_to_mean_rstd(p.pro_stats + 2 * (size_t)item, p.pro_count, 1e-5f, pmean, prstd);
        }
        rowstat[tid] = make_float2(pmean, prstd);
    }
    if (warp == PROD_WARPS)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                     "r"(tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp < PROD_WARPS)
    {
        // ================= A producers =================
        // A k-block is 128 rows x 8 chunks of 8 values.  Thread t owns chunk column kc = t % 8 of rows
        // t/8 + 32*i (i = 0..3): the 8 lanes of a row read 256 contiguous bytes (coalesced) and a warp's
        // 16-byte bf16 stores hit 8 distinct bank groups thanks to the padded LBO.  Per row the thread
        // keeps a base pointer and a bit mask of the taps that fall inside the input grid, so the inner
        // loop is: tap offset (once per k-block), predicated 16-byte loads, 4 cvt, one 16-byte store.
        if (HALO)
        {
            halo_stage_patch(p, tid, c8n, h_lbo, hb, hu, hv0, a_smem);
            fence_proxy_async();
            mbar_arrive(&a_full_bar);
        }
        else
        {
        const int kc = tid & 7;
        const int r0 = tid >> 3;
        const float *rp[4];
        uint32_t rmask[4];
        producer_rows(p, m0, r0, rp, rmask);
        const long long a2d = p.A2 ? (p.A2 - p.A) : 0;
        for (int kb = 0; kb < (p.Apk ? 0 : nk); ++kb) // packed-A mode: the operand is bulk-copied, nothing to gather
        {
            const int s = kb % STAGES;
            const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
            const int k = kb * BK + kc * 8;
            uint4 pk[4];
            gather_kblock<VECA>(p, k, r0, rp, rmask, a2d, rowstat, pk);
            mbar_wait(&empty_bar[s], ph ^ 1u);
            uint8_t *dst = a_smem + s * A_STAGE + kc * A_LBO + r0 * 16; // core (row/8, kc): kc*LBO + row*16
#pragma unroll
            for (int i = 0; i < 4; ++i)
                *reinterpret_cast<uint4 *>(dst + i * (32 * 16)) = pk[i];
            fence_proxy_async(); // make the generic-proxy stores visible to the tensor core's async proxy
            mbar_arrive(&full_bar[s]);
        }

        } // !HALO

        // ================= epilogue =================
        // Warps w and w+4 share TMEM lanes / tile rows 32*(w%4)..+31 and alternate over the 32-column chunks;
        // the staging buffers are the drained pipeline buffers.
        const int q = warp & 3, hsel = warp >> 2;
        mbar_wait(&tmem_full_bar, 0);
        tc_fence_after();
        static_assert((EPI_LD * 4) % 16 == 0, "bulk copies need 16-byte aligned row slots");
        if (tma_store && bulk_ok(p, plain_out) && !p.no_bulk)
        {
            // tensor-map store: flat rows (line 0), or one (item, u) line per tile in HALO mode
            int parity = 0;
            epilogue_tile_tmast(p, &tmC, bn, n0, tmem_base + ((uint32_t)(q * 32) << 16), rowinfo[q * 32 + lane],
                                smem + warp * EPI_WARP, 32 * hsel, 64, lane, (HALO ? hv0 : m0) + q * 32, HALO ? hb * p.Uo + hu : 0,
                                parity);
            if (lane == 0)
                bulk_wait_read();
        }
        else if (bulk_ok(p, plain_out) && !p.no_bulk)
            epilogue_tile_bulk(p, bn, n0, tmem_base + ((uint32_t)(q * 32) << 16), rowinfo[q * 32 + lane],
                               reinterpret_cast<float *>(smem + warp * EPI_WARP), 32 * hsel, 64, lane);
        else
            epilogue_tile(p, bn, plain_out, n0, tmem_base + ((uint32_t)(q * 32) << 16), rowinfo,
                          reinterpret_cast<float *>(smem + warp * EPI_WARP), q, 32 * hsel, 64, lane);
        tc_fence_before();
    }
    else if (warp == PROD_WARPS)
    {
        // ================= B loader: one bulk copy per stage =================
        if (lane == 0)
        {
            const uint8_t *src = reinterpret_cast<const uint8_t *>(Wb) + (size_t)ntile * nk * b_stage_bytes;
            for (int kb = 0; kb < nk; ++kb)
            {
                const int s = kb % STAGES;
                const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
                mbar_wait(&empty_bar[s], ph ^ 1u);
                if (p.Apk)
                {
                    // tile-packed bf16 activations: the k-block's A operand is one contiguous 16 KB image
                    const uint8_t *asrc = reinterpret_cast<const uint8_t *>(p.Apk) +
                                          ((size_t)blockIdx.x * nk + kb) * (size_t)(BM * BK * 2);
                    mbar_arrive_expect_tx(&full_bar[s], b_stage_bytes + BM * BK * 2);
                    bulk_g2s(a_smem + s * A_STAGE, asrc, BM * BK * 2, &full_bar[s]);
                }
                else
                    mbar_arrive_expect_tx(&full_bar[s], b_stage_bytes);
                bulk_g2s(b_smem + (size_t)s * b_stage_bytes, src + (size_t)kb * b_stage_bytes, b_stage_bytes, &full_bar[s]);
            }
        }
    }
    else
    {
        // ================= MMA issuer: a single thread =================
        if (lane == 0)
        {
            const uint32_t idesc = make_idesc(bn);
            const uint32_t b_lbo = (uint32_t)bn * 16, sbo = 128;
            const uint32_t a_lbo = p.Apk ? (uint32_t)(BM * 16) : (uint32_t)A_LBO;
            for (int kb = 0; kb < nk; ++kb)
            {
                const int s = kb % STAGES;
                const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
                mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                const uint32_t a_base = smem_u32(a_smem + (HALO ? 0 : s * A_STAGE));
                const uint32_t b_base = smem_u32(b_smem + (size_t)s * b_stage_bytes);
                if (HALO)
                {
                    if (kb == 0)
                    {
                        mbar_wait(&a_full_bar, 0);
                        tc_fence_after();
                    }
                    // k = tap*Cin + c, Cin % 16 == 0: a 16-wide k step lies inside one tap; its A operand is the patch
                    // plane of that tap's du, chunks c/8 and c/8+1, rows shifted by the tap's dv
                    for (int j = 0; j < BK / 16; ++j)
                    {
                        const int k = kb * BK + j * 16;
                        if (k >= p.K)
                            break;
                        const int tap = k / p.Cin, c = k - tap * p.Cin;
                        const uint32_t aoff = (uint32_t)((p.du[tap] - p.halo_dumin) * c8n + (c >> 3)) * h_lbo +
                                              (uint32_t)(p.dv[tap] - p.halo_dvmin) * 16u;
                        const uint64_t ad = make_desc(a_base + aoff, h_lbo, sbo);
                        const uint64_t bd = make_desc(b_base + j * 2 * b_lbo, b_lbo, sbo);
                        umma(tmem_base, ad, bd, idesc, (kb > 0 || j > 0) ? 1u : 0u);
                    }
                }
                else
                {
#pragma unroll
                    for (int j = 0; j < BK / 16; ++j)
                    {
                        const uint64_t ad = make_desc(a_base + j * 2 * a_lbo, a_lbo, sbo);
                        const uint64_t bd = make_desc(b_base + j * 2 * b_lbo, b_lbo, sbo);
                        umma(tmem_base, ad, bd, idesc, (kb > 0 || j > 0) ? 1u : 0u);
                    }
                }
                umma_commit(&empty_bar[s]); // frees the stage once these MMAs have read it
            }
            umma_commit(&tmem_full_bar);    // accumulator complete
        }
    }
    __syncthreads();
    if (warp == PROD_WARPS)
    {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
    }
}


// Epilogue for tile-packed bf16 outputs [row/128][col/8][row%128][8] (bias, optional GELU; no residual / GLU / stats):
// a lane keeps its own row (= its TMEM lane), so 8 consecutive columns are one 16-byte store and the 32 lanes of a
// warp write 512 contiguous bytes -- no transposition through shared memory, no row table.  prow < 0: row outside M.
__device__ __forceinline__ void epilogue_tile_packed(const GemmParams &p, const int bn, const int n0, const uint32_t lane_addr,
                                                     const int prow, const int c_begin, const int c_step)
{
    const int ntile_end = (n0 + bn) < p.N ? (n0 + bn) : p.N;
    uint4 *Cp = reinterpret_cast<uint4 *>(p.C);
    const size_t rowbase = (size_t)(prow < 0 ? 0 : (prow >> 7)) * (size_t)(p.ldc >> 3) * 128 + (size_t)(prow & 127);
    for (int c0 = c_begin; c0 < bn; c0 += c_step)
    {
        if (n0 + c0 >= ntile_end)
            break;
        float v[32];
        tmem_ld32(lane_addr + (uint32_t)c0, v);
#pragma unroll
        for (int g = 0; g < 4; ++g)
        {
            const int col = n0 + c0 + 8 * g;
            if (col < ntile_end) // N % 8 == 0: the whole group is inside
            {
                float o[8];
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    o[e] = v[8 * g + e];
                if (p.bias)
                {
                    const float4 b0 = __ldg(reinterpret_cast<const float4 *>(p.bias + col));
                    const float4 b1 = __ldg(reinterpret_cast<const float4 *>(p.bias + col + 4));
                    o[0] += b0.x;
                    o[1] += b0.y;
                    o[2] += b0.z;
                    o[3] += b0.w;
                    o[4] += b1.x;
                    o[5] += b1.y;
                    o[6] += b1.z;
                    o[7] += b1.w;
                }
                if (p.act == 1)
                {
#pragma unroll
                    for (int e = 0; e < 8; ++e)
                        o[e] = gelu_erf(o[e]);
                }
                uint4 pk;
                pk.x = pack_bf16x2(o[0], o[1]);
                pk.y = pack_bf16x2(o[2], o[3]);
                pk.z = pack_bf16x2(o[4], o[5]);
                pk.w = pack_bf16x2(o[6], o[7]);
                if (prow >= 0)
                    Cp[rowbase + (size_t)(col >> 3) * 128] = pk;
            }
        }
    }
}

// =====================================================================================================================
// Persistent, warp-specialised form of the same operator: one CTA per SM loops over output tiles, with TWO
// accumulators in TMEM and the operand ring running across tile boundaries, so the gather / weight copies / MMAs of
// tile i+1 overlap the epilogue of tile i inside the CTA (the one-tile-per-CTA kernel above leaves the SM idle through
// each CTA's serial setup -> load -> MMA -> epilogue chain; ncu: tensor pipe 12-20 % active on the transformer GEMMs).
//   warps 0-7    producers: A gather (or HALO patch staging, double buffered); absent when A arrives tile-packed
//   next 16      epilogue: four per TMEM lane quadrant, each taking every fourth 32-column chunk
//   then         bulk-copy loader warp (weights, tile-packed A) and the MMA issuer warp
// Barriers: full/empty per ring stage (as above, phases counted over the CTA's whole k-block sequence),
// acc_full/acc_empty per accumulator, patch_full/patch_empty per HALO patch buffer.
constexpr int P_EPI_WARPS = 16;
constexpr int P_MAX_STAGES = 6;
constexpr int P_EPI_WARP = 5120;                                    // per warp: 32 x EPI_LD transposition buffer / two 2 KB store tiles
constexpr int P_EPI_BYTES = P_EPI_WARPS * P_EPI_WARP + 1024;        // + alignment slack
constexpr int P_ROW_BYTES = P_EPI_WARPS * 32 * (int)sizeof(RowInfo); // private row tables
__host__ __device__ constexpr int p_threads(bool prod) { return ((prod ? PROD_WARPS : 0) + P_EPI_WARPS + 2) * 32; }

template <bool VECA, bool HALO, bool PROD>
__global__ void __launch_bounds__(p_threads(PROD), 1)
    gemm_tc_persist_kernel(const __grid_constant__ CUtensorMap tmC, const GemmParams p, const __nv_bfloat16 *__restrict__ Wb,
                           const int bn, const int plain_out, const int stages, const int mtiles, const int ntiles,
                           const int tma_store)
{
    constexpr int PW = PROD ? PROD_WARPS : 0; // first epilogue warp; PW % 4 == 0 keeps warp % 4 == lane quadrant
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[P_MAX_STAGES], empty_bar[P_MAX_STAGES];
    __shared__ __align__(8) uint64_t acc_full[2], acc_empty[2], patch_full[2], patch_empty[2];
    __shared__ uint32_t tmem_slot;
    __shared__ float2 rowstat[2][BM]; // producers only: GroupNorm prologue (mean, rstd) per tile row, by tile parity

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int nk = (p.K + BK - 1) / BK;
    const int total_tiles = mtiles * ntiles;
    const uint32_t b_stage_bytes = (uint32_t)bn * BK * 2;
    const int c8n = p.Cin >> 3;
    const uint32_t h_lbo = (uint32_t)(p.halo_rows | 1) * 16;
    const uint32_t patch_bytes = HALO ? (((uint32_t)p.halo_nu * c8n * h_lbo + 127u) & ~127u) : 0u;
    const uint32_t a_bytes = HALO ? 2u * patch_bytes : (uint32_t)(stages * A_STAGE);
    uint8_t *a_smem = smem;
    uint8_t *b_smem = smem + a_bytes;
    // epilogue buffers: 1024-byte aligned (the tensor-map store tiles are swizzled), P_EPI_WARP bytes per warp
    uint8_t *epi_base = b_smem + (size_t)stages * b_stage_bytes;
    epi_base += (1024u - (smem_u32(epi_base) & 1023u)) & 1023u;
    float *epi_smem = reinterpret_cast<float *>(epi_base);
    RowInfo *row_smem = reinterpret_cast<RowInfo *>(epi_base + P_EPI_WARPS * P_EPI_WARP);
    uint32_t acc_cols = 32;
    while ((int)acc_cols < bn)
        acc_cols <<= 1;

    if (tid == 0)
    {
        for (int s = 0; s < stages; ++s)
        {
            mbar_init(&full_bar[s], (p.Apk || HALO) ? 1u : (uint32_t)(PROD_THREADS + 1));
            mbar_init(&empty_bar[s], 1);
        }
        for (int a = 0; a < 2; ++a)
        {
            mbar_init(&acc_full[a], 1);
            mbar_init(&acc_empty[a], P_EPI_WARPS);
            mbar_init(&patch_full[a], PROD_THREADS);
            mbar_init(&patch_empty[a], 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == PW + P_EPI_WARPS)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                     "r"(2u * acc_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    // tile -> (row tile mt, column tile nt): column tiles of one row tile run on neighbouring CTAs at the same time,
    // so the A operand they share is fetched from HBM once
    auto halo_coords = [&](int mt, int &hb, int &hu, int &hv0) {
        const int bu = fast_div(mt, p.div_tpr);
        hv0 = (mt - bu * p.halo_tpr) * BM;
        hb = fast_div(bu, p.div_uo);
        hu = bu - hb * p.Uo;
    };

    if (PROD && warp < PW)
    {
        // ================= A producers =================
        const int kc = tid & 7, r0 = tid >> 3;
        const long long a2d = p.A2 ? (p.A2 - p.A) : 0;
        int it = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it)
        {
            const int mt = tile / ntiles;
            if (HALO)
            {
                int hb, hu, hv0;
                halo_coords(mt, hb, hu, hv0);
                const int pb = it & 1;
                mbar_wait(&patch_empty[pb], (uint32_t)((it >> 1) & 1) ^ 1u);
                halo_stage_patch(p, tid, c8n, h_lbo, hb, hu, hv0, a_smem + pb * patch_bytes);
                fence_proxy_async();
                mbar_arrive(&patch_full[pb]);
            }
            else
            {
                const int m0 = mt * BM;
                const float2 *rs = rowstat[it & 1];
                if (p.pro == 1)
                {
                    if (tid < BM)
                    {
                        float pmean = 0.f, prstd = 1.f;
                        const int m = m0 + tid;
                        if (m < p.M)
                        {
                            const RowCoord rc = decode_row(m, p);
                            const int ui = rc.uo * p.su + p.du[0];
                            const int item = p.pro_per_u ? rc.b * p.Ui + ui : rc.b;
                            stats_to_mean_rstd(p.pro_stats + 2 * (size_t)item, p.pro_count, 1e-5f, pmean, prstd);
                        }
                        rowstat[it & 1][tid] = make_float2(pmean, prstd);
                    }
                    asm volatile("bar.sync 1, %0;" ::"n"(PROD_THREADS) : "memory");
                }
                const float *rp[4];
                uint32_t rmask[4];
                producer_rows(p, m0, r0, rp, rmask);
                for (int kb = 0; kb < nk; ++kb)
                {
                    const int gk = it * nk + kb;
                    const int s = gk % stages;
                    const uint32_t ph = (uint32_t)(gk / stages) & 1u;
                    uint4 pk[4];
                    gather_kblock<VECA>(p, kb * BK + kc * 8, r0, rp, rmask, a2d, rs, pk);
                    mbar_wait(&empty_bar[s], ph ^ 1u);
                    uint8_t *dst = a_smem + s * A_STAGE + kc * A_LBO + r0 * 16;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        *reinterpret_cast<uint4 *>(dst + i * (32 * 16)) = pk[i];
                    fence_proxy_async();
                    mbar_arrive(&full_bar[s]);
                }
            }
        }
    }
    else if (warp < PW + P_EPI_WARPS)
    {
        // ================= epilogue =================
        const int e = warp - PW;
        const int q = e & 3;      // == warp % 4: the TMEM lane quadrant this warp may read
        const int csel = e >> 2;  // which of every four 32-column chunks
        float *stage = epi_smem + e * (P_EPI_WARP / 4);
        RowInfo *rowinfo = row_smem + e * 32 - q * 32; // private table, indexed by tile row q*32 + r like the shared one
        const bool use_tmast = tma_store && !HALO && bulk_ok(p, plain_out) && !p.no_bulk;
        int st_parity = 0;
        const bool packed_direct = p.out_pack && plain_out == 1 && !p.glu && !p.gamma && !p.addC && !p.rowadd && !p.stats &&
                                   (p.N % 8) == 0;
        int it = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it)
        {
            const int mt = tile / ntiles, nt = tile - mt * ntiles;
            {
                const int row = q * 32 + lane;
                int m = mt * BM + row;
                bool valid = m < p.M;
                RowCoord rc;
                if (HALO)
                {
                    int hb, hu, hv0;
                    halo_coords(mt, hb, hu, hv0);
                    valid = hv0 + row < p.Vo;
                    rc.b = hb;
                    rc.uo = hu;
                    rc.vo = valid ? hv0 + row : 0;
                    m = (hb * p.Uo + hu) * p.Vo + rc.vo;
                }
                else
                    rc = decode_row(valid ? m : 0, p);
                RowInfo ri;
                ri.uu0 = rc.uo * p.ous + p.ouo;
                ri.vv0 = rc.vo * p.ovs + p.ovo;
                ri.outoff = (p.out_vu ? ((long long)rc.b * p.Vout + ri.vv0) * p.Uout + ri.uu0 : ((long long)rc.b * p.Uout + ri.uu0) * p.Vout + ri.vv0) * p.ldc;
                ri.uo = rc.uo;
                ri.item = valid ? (p.stat_per_u ? rc.b * p.Uo + rc.uo : rc.b) : -1;
                ri.prow = p.pack_T ? rc.b * p.pack_T + rc.uo * p.Vo + rc.vo : m;
                ri.orow = (rc.b * p.Uout + ri.uu0) * p.Vout + ri.vv0;
                rowinfo[row] = ri;
            }
            const int my_prow = rowinfo[q * 32 + lane].item >= 0 ? rowinfo[q * 32 + lane].prow : -1;
            __syncwarp();
            const int acc = it & 1;
            mbar_wait(&acc_full[acc], (uint32_t)((it >> 1) & 1));
            tc_fence_after();
            const uint32_t lane_addr = tmem_base + (uint32_t)acc * acc_cols + ((uint32_t)(q * 32) << 16);
            if (packed_direct)
                epilogue_tile_packed(p, bn, nt * bn, lane_addr, my_prow, 32 * csel, 32 * (P_EPI_WARPS / 4));
            else if (use_tmast)
                epilogue_tile_tmast(p, &tmC, bn, nt * bn, lane_addr, rowinfo[q * 32 + lane], reinterpret_cast<uint8_t *>(stage),
                                    32 * csel, 32 * (P_EPI_WARPS / 4), lane, mt * BM + q * 32, 0, st_parity);
            else if (bulk_ok(p, plain_out) && !p.no_bulk)
                epilogue_tile_bulk(p, bn, nt * bn, lane_addr, rowinfo[q * 32 + lane], stage, 32 * csel,
                                   32 * (P_EPI_WARPS / 4), lane);
            else
                epilogue_tile(p, bn, plain_out, nt * bn, lane_addr, rowinfo, stage, q, 32 * csel, 32 * (P_EPI_WARPS / 4), lane);
            tc_fence_before();
            __syncwarp();
            if (lane == 0)
                mbar_arrive(&acc_empty[acc]);
        }
        if (use_tmast && lane == 0)
            bulk_wait_read(); // the store tiles live in this CTA's shared memory: drain before exit
    }
    else if (warp == PW + P_EPI_WARPS)
    {
        // ================= loader: one bulk copy (two with tile-packed A) per stage =================
        if (lane == 0)
        {
            int it = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it)
            {
                const int mt = tile / ntiles, nt = tile - mt * ntiles;
                const uint8_t *src = reinterpret_cast<const uint8_t *>(Wb) + (size_t)nt * nk * b_stage_bytes;
                for (int kb = 0; kb < nk; ++kb)
                {
                    const int gk = it * nk + kb;
                    const int s = gk % stages;
                    const uint32_t ph = (uint32_t)(gk / stages) & 1u;
                    mbar_wait(&empty_bar[s], ph ^ 1u);
                    if (p.Apk)
                    {
                        const uint8_t *asrc = reinterpret_cast<const uint8_t *>(p.Apk) +
                                              ((size_t)mt * nk + kb) * (size_t)(BM * BK * 2);
                        mbar_arrive_expect_tx(&full_bar[s], b_stage_bytes + BM * BK * 2);
                        bulk_g2s(a_smem + s * A_STAGE, asrc, BM * BK * 2, &full_bar[s]);
                    }
                    else
                        mbar_arrive_expect_tx(&full_bar[s], b_stage_bytes);
                    bulk_g2s(b_smem + (size_t)s * b_stage_bytes, src + (size_t)kb * b_stage_bytes, b_stage_bytes, &full_bar[s]);
                }
            }
        }
    }
    else
    {
        // ================= MMA issuer: a single thread =================
        if (lane == 0)
        {
            const uint32_t idesc = make_idesc(bn);
            const uint32_t b_lbo = (uint32_t)bn * 16, sbo = 128;
            const uint32_t a_lbo = p.Apk ? (uint32_t)(BM * 16) : (uint32_t)A_LBO;
            int it = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it)
            {
                const int acc = it & 1;
                const uint32_t d_tmem = tmem_base + (uint32_t)acc * acc_cols;
                mbar_wait(&acc_empty[acc], (uint32_t)((it >> 1) & 1) ^ 1u); // epilogue of tile it-2 has drained it
                tc_fence_after();
                if (HALO)
                {
                    mbar_wait(&patch_full[it & 1], (uint32_t)((it >> 1) & 1));
                    tc_fence_after();
                }
                for (int kb = 0; kb < nk; ++kb)
                {
                    const int gk = it * nk + kb;
                    const int s = gk % stages;
                    const uint32_t ph = (uint32_t)(gk / stages) & 1u;
                    mbar_wait(&full_bar[s], ph);
                    tc_fence_after();
                    const uint32_t b_base = smem_u32(b_smem + (size_t)s * b_stage_bytes);
                    if (HALO)
                    {
                        const uint32_t a_base = smem_u32(a_smem + (it & 1) * patch_bytes);
                        for (int j = 0; j < BK / 16; ++j)
                        {
                            const int k = kb * BK + j * 16;
                            if (k >= p.K)
                                break;
                            const int tap = k / p.Cin, c = k - tap * p.Cin;
                            const uint32_t aoff = (uint32_t)((p.du[tap] - p.halo_dumin) * c8n + (c >> 3)) * h_lbo +
                                                  (uint32_t)(p.dv[tap] - p.halo_dvmin) * 16u;
                            umma(d_tmem, make_desc(a_base + aoff, h_lbo, sbo), make_desc(b_base + j * 2 * b_lbo, b_lbo, sbo),
                                 idesc, (kb > 0 || j > 0) ? 1u : 0u);
                        }
                    }
                    else
                    {
                        const uint32_t a_base = smem_u32(a_smem + s * A_STAGE);
#pragma unroll
                        for (int j = 0; j < BK / 16; ++j)
                            umma(d_tmem, make_desc(a_base + j * 2 * a_lbo, a_lbo, sbo),
                                 make_desc(b_base + j * 2 * b_lbo, b_lbo, sbo), idesc, (kb > 0 || j > 0) ? 1u : 0u);
                    }
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(&acc_full[acc]);
                if (HALO)
                    umma_commit(&patch_empty[it & 1]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == PW + P_EPI_WARPS)
    {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2u * acc_cols) : "memory");
    }
}

// =====================================================================================================================
// TMA-fed persistent implicit-GEMM convolution.  The A operand is a bf16 channels-last IMAGE [B][U][V][Cin] in global
// memory (written by the producing layer's epilogue) and reaches shared memory through a tensor map: one
// cp.async.bulk.tensor per BOX = (plane u, rows v0 + dv0 .. + rows - 1, a chunk of cw channels), landed in the tensor
// core's swizzled K-major layout (cw * 2 = 128 / 64 / 32-byte rows, SWIZZLE_128B / 64B / 32B), zero fill outside the
// grid = the convolution's zero padding.  No thread touches the operand.  Every tap that reads the plane uses the SAME
// box: its A descriptor starts dv - dv0 rows further down (start address + shift * row pitch), so a 3 x 3 convolution
// moves each input row L2 -> SM three times less than nine separate tap loads would.
//
// What bounds these layers (measured with dmx_debug_conv on B200): (1) the weights -- with one 128-row tile per pass every
// tile re-streams the whole K x N weight image L2 -> SM at ~30 GB/s per SM; so G = 2 row tiles share every weight block
// (two accumulators fed from one block), and a weight image that fits stays resident in shared memory for the whole
// kernel; (2) the store side -- one cp.async.bulk per lane and 32-column piece (384 per tile) runs at the copy engine's
// ~12 ns per operation, 4.5 us per tile; here a warp stages its 32 rows x 32 columns in a swizzled tile and ONE tensor-map
// store (clipped at the line end by the map's bounds) moves it.
//   warps [0, ew)   epilogue (ew % 4 == 0, so warp % 4 == TMEM lane quadrant)
//   warp ew         A loader (TMA boxes)      warp ew + 1   weight loader (bulk copies)      warp ew + 2   MMA issuer
// Rings: boxes (a_full / a_empty) and 64-row weight blocks (b_full / b_empty); accumulators acc_full / acc_empty.
constexpr int T_MAX_ASLOTS = 16;
constexpr int T_MAX_BSTAGES = 16;
constexpr int T_MAX_EW = 16;
constexpr int T_MAX_ACC = 4;
constexpr int T_EPI_WARP_BYTES = 5120; // one 32 x 128-byte or two 32 x 64-byte store tiles (or the 32 x EPI_LD transposition buffer)

// umma with the descriptors given as (lo, hi) words: the issuing thread only adds to the lo words between instructions
__device__ __forceinline__ void umma_lh(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc,
                                        uint32_t accum)
{
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 ".reg .b64 da, db;\n"
                 "mov.b64 da, {%1, %2};\n"
                 "mov.b64 db, {%3, %4};\n"
                 "setp.ne.b32 p, %6, 0;\n"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n"
                 "}\n" ::"r"(tmem_d),
                 "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accum)
                 : "memory");
}

__device__ __forceinline__ void umma_lh_tf32(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                             uint32_t idesc, uint32_t accum)
{
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 ".reg .b64 da, db;\n"
                 "mov.b64 da, {%1, %2};\n"
                 "mov.b64 db, {%3, %4};\n"
                 "setp.ne.b32 p, %6, 0;\n"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %5, p;\n"
                 "}\n" ::"r"(tmem_d),
                 "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accum)
                 : "memory");
}

__global__ void __launch_bounds__((T_MAX_EW + 4) * 32, 1)
    gemm_tc_tma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmC, const GemmParams p,
                       const __nv_bfloat16 *__restrict__ Wt, const int bn, const int plain_out, const int a_slots, const int b_stages,
                       const int mtiles, const int ntiles, const int ew, const int G, const int nacc, const int b_resident,
                       const int tma_store, const int epi_warp_bytes)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t a_full[T_MAX_ASLOTS], a_empty[T_MAX_ASLOTS], b_full[T_MAX_BSTAGES], b_empty[T_MAX_BSTAGES];
    __shared__ __align__(8) uint64_t acc_full[T_MAX_ACC], acc_empty[T_MAX_ACC];
    __shared__ uint32_t tmem_slot;

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int kblk = p.t_tf32 ? 32 : BK; // K per weight block: 128 bytes per weight row either way
    const int nkb = (p.K + kblk - 1) / kblk;
    const int npairs = (mtiles + G - 1) / G;
    const int total_items = npairs * ntiles;
    const int ewg = ew / G; // epilogue warps per tile of a pair
    const uint32_t b_stage_bytes = (uint32_t)bn * BK * 2;
    uint8_t *a_smem = smem;
    const uint32_t slot_bytes = (uint32_t)p.t_cps * (uint32_t)p.t_box_bytes; // a ring slot holds the cps chunk boxes of a plane
    // per-warp epilogue buffers (1024-aligned): nothing for the direct tile-packed epilogue, 4 KB of store tiles for the
    // tensor-map store, T_EPI_WARP_BYTES + a 32-row table for the transposing epilogue
    uint8_t *epi_smem = smem + (size_t)a_slots * slot_bytes;
    uint8_t *b_smem = epi_smem + (size_t)ew * epi_warp_bytes;
    const bool flat = p.Apk != nullptr; // tile-packed A: tiles are 128 consecutive GEMM rows, no line structure
    uint32_t acc_cols = 32;
    while ((int)acc_cols < bn)
        acc_cols <<= 1;

    if (tid == 0)
    {
        for (int s = 0; s < a_slots; ++s)
        {
            mbar_init(&a_full[s], 1);
            mbar_init(&a_empty[s], 1);
        }
        for (int s = 0; s < b_stages; ++s)
        {
            mbar_init(&b_full[s], p.Apk ? 2u : 1u); // tile-packed A travels in the weight ring's stages (both loaders arrive)
            mbar_init(&b_empty[s], (uint32_t)G);   // every MMA issuer of a pair releases the block
        }
        for (int a = 0; a < nacc; ++a)
        {
            mbar_init(&acc_full[a], 1);
            mbar_init(&acc_empty[a], (uint32_t)ewg);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == ew + 2)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                     "r"((uint32_t)nacc * acc_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    auto tile_coords = [&](int mt, int &hb, int &hu, int &hv0) {
        const int bu = fast_div(mt, p.div_tpr);
        hv0 = (mt - bu * p.halo_tpr) * BM;
        hb = fast_div(bu, p.div_uo);
        hu = bu - hb * p.Uo;
    };

    if (warp < ew)
    {
        // ================= epilogue: warp group g = warp / ewg drains tile g of every pair =================
        const int g = warp / ewg, wg = warp - g * ewg;
        const int q = warp & 3; // ewg % 4 == 0, so this is also wg & 3: the TMEM lane quadrant this warp may read
        const int csel = wg >> 2, nsel = ewg >> 2;
        uint8_t *stage = epi_smem + (size_t)warp * epi_warp_bytes;
        RowInfo *rowinfo = reinterpret_cast<RowInfo *>(stage + T_EPI_WARP_BYTES) - q * 32; // transposing epilogue only
        const bool packed_direct = p.out_pack && plain_out == 1 && !p.glu && !p.gamma && !p.addC && !p.rowadd && !p.stats &&
                                   (p.N % 8) == 0;
        const bool use_tmast = tma_store && bulk_ok(p, plain_out);
        // coalesced direct stores where the output rows are far apart (row pitch >= 1 KB: the tensor-map store is slow there)
        const bool use_direct = (tma_store & 2) && bulk_ok(p, plain_out);
        const bool need_table = !packed_direct && !use_tmast && !(bulk_ok(p, plain_out) && !p.no_bulk);
        int parity = 0, itm = 0;
        for (int item = blockIdx.x; item < total_items; item += gridDim.x, ++itm)
        {
            const int pr = item / ntiles, nt = item - pr * ntiles;
            const int mt = pr * G + g;
            if (mt >= mtiles)
                break;
            const int tl = itm * G + g;
            int hb = 0, hu = 0, hv0 = 0;
            RowInfo ri;
            {
                const int row = q * 32 + lane;
                RowCoord rc;
                bool valid;
                int m;
                if (flat)
                {
                    m = mt * BM + row;
                    valid = m < p.M;
                    rc = decode_row(valid ? m : 0, p);
                }
                else
                {
                    tile_coords(mt, hb, hu, hv0);
                    valid = hv0 + row < p.Vo;
                    rc.b = hb;
                    rc.uo = hu;
                    rc.vo = valid ? hv0 + row : 0;
                    m = (hb * p.Uo + hu) * p.Vo + rc.vo;
                }
                ri.uu0 = rc.uo * p.ous + p.ouo;
                ri.vv0 = rc.vo * p.ovs + p.ovo;
                ri.outoff = (p.out_vu ? ((long long)rc.b * p.Vout + ri.vv0) * p.Uout + ri.uu0 : ((long long)rc.b * p.Uout + ri.uu0) * p.Vout + ri.vv0) * p.ldc;
                ri.uo = rc.uo;
                ri.item = valid ? (p.stat_per_u ? rc.b * p.Uo + rc.uo : rc.b) : -1;
                ri.prow = p.pack_T ? rc.b * p.pack_T + rc.uo * p.Vo + rc.vo : m;
                ri.orow = (rc.b * p.Uout + ri.uu0) * p.Vout + ri.vv0;
                if (need_table)
                    rowinfo[row] = ri;
            }
            const int my_prow = ri.item >= 0 ? ri.prow : -1;
            __syncwarp();
            const int acc = tl % nacc;
            mbar_wait(&acc_full[acc], (uint32_t)(tl / nacc) & 1u);
            tc_fence_after();
            const uint32_t lane_addr = tmem_base + (uint32_t)acc * acc_cols + ((uint32_t)(q * 32) << 16);
            // rows of this warp: consecutive output rows?  (always within a line; flat rows only while they stay in one item)
            const int item0 = __shfl_sync(0xffffffffu, ri.item, 0);
            const int nvalid = __popc(__ballot_sync(0xffffffffu, ri.item >= 0));
            const bool one_item = __all_sync(0xffffffffu, ri.item < 0 || ri.item == item0);
            const long long off0 = __shfl_sync(0xffffffffu, ri.outoff, 0);
            if (use_direct && (one_item || !p.stats) && nvalid > 0)
                epilogue_tile_direct(p, bn, nt * bn, lane_addr, one_item ? item0 : -1, p.C + off0, nvalid,
                                     reinterpret_cast<float *>(stage), 32 * csel, 32 * nsel, lane);
            else if (use_tmast)
                epilogue_tile_tmast(p, &tmC, bn, nt * bn, lane_addr, ri, stage, 32 * csel, 32 * nsel, lane,
                                    flat ? mt * BM + q * 32 : hv0 + q * 32, flat ? 0 : hb * p.Uo + hu, parity);
            else if (packed_direct)
                epilogue_tile_packed(p, bn, nt * bn, lane_addr, my_prow, 32 * csel, 32 * nsel);
            else if (bulk_ok(p, plain_out) && !p.no_bulk)
                epilogue_tile_bulk(p, bn, nt * bn, lane_addr, ri, reinterpret_cast<float *>(stage), 32 * csel, 32 * nsel, lane);
            else
                epilogue_tile(p, bn, plain_out, nt * bn, lane_addr, rowinfo, reinterpret_cast<float *>(stage), q, 32 * csel, 32 * nsel,
                              lane);
            tc_fence_before();
            __syncwarp();
            if (lane == 0)
                mbar_arrive(&acc_empty[acc]);
        }
        if (use_tmast && lane == 0)
            bulk_wait_read(); // the store tiles live in this CTA's shared memory: drain before exit
    }
    else if (warp == ew)
    {
        // ================= A loader: one tensor-map copy per box =================
        if (lane == 0)
        {
            asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
            const uint32_t box_tx = (uint32_t)p.t_rows * (uint32_t)p.t_pitch * (uint32_t)p.t_cps;
            const int ngrp = p.t_nchunk / p.t_cps;
            int slot = 0;
            uint32_t ph = 0;
            for (int item = blockIdx.x; item < total_items; item += gridDim.x)
            {
                const int pr = item / ntiles;
                int hb[2] = {0, 0}, hu[2] = {0, 0}, hv0[2] = {0, 0}, ng = 0;
                for (int g = 0; g < G && pr * G + g < mtiles; ++g, ++ng)
                    if (!flat)
                        tile_coords(pr * G + g, hb[g], hu[g], hv0[g]);
                if (flat)
                {
                    // tile-packed activations: a k-block of A is one contiguous 16 KB image per row tile; it rides in the weight
                    // ring's stage for the same k-block (one full / empty barrier pair per k-block for the MMA issuers)
                    for (int kb = 0; kb < ngrp; ++kb)
                    {
                        mbar_wait(&b_empty[slot], ph ^ 1u);
                        mbar_arrive_expect_tx(&b_full[slot], (uint32_t)ng * BM * BK * 2);
                        for (int g = 0; g < ng; ++g)
                            bulk_g2s(a_smem + ((size_t)slot * G + g) * (BM * BK * 2),
                                     reinterpret_cast<const uint8_t *>(p.Apk) + ((size_t)(pr * G + g) * ngrp + kb) * (size_t)(BM * BK * 2),
                                     BM * BK * 2, &b_full[slot]);
                        if (++slot == b_stages)
                        {
                            slot = 0;
                            ph ^= 1u;
                        }
                    }
                    continue;
                }
                for (int qg = 0; qg < ngrp; ++qg)
                    for (int b = 0; b < p.t_nbox; ++b)
                        for (int g = 0; g < ng; ++g)
                        {
                            mbar_wait(&a_empty[slot], ph ^ 1u);
                            mbar_arrive_expect_tx(&a_full[slot], box_tx);
                            for (int qi = 0; qi < p.t_cps; ++qi)
                                tma_load_4d(a_smem + (size_t)slot * slot_bytes + (size_t)qi * p.t_box_bytes, &tmA, &a_full[slot],
                                            (qg * p.t_cps + qi) * p.t_cw, hv0[g] * p.sv + p.t_box_dv[b], hu[g] * p.su + p.t_box_du[b],
                                            hb[g]);
                            if (++slot == a_slots)
                            {
                                slot = 0;
                                ph ^= 1u;
                            }
                        }
            }
        }
    }
    else if (warp == ew + 1)
    {
        // ================= weight loader: one bulk copy per 64-row block (once only when the image is resident) =================
        if (lane == 0)
        {
            int s = 0;
            uint32_t ph = 0;
            for (int item = blockIdx.x; item < total_items; item += gridDim.x)
            {
                const int nt = item % ntiles;
                const uint8_t *src = reinterpret_cast<const uint8_t *>(Wt) + (size_t)nt * nkb * b_stage_bytes;
                for (int kb = 0; kb < nkb; ++kb)
                {
                    if (!b_resident)
                        mbar_wait(&b_empty[s], ph ^ 1u);
                    mbar_arrive_expect_tx(&b_full[s], b_stage_bytes);
                    bulk_g2s(b_smem + (size_t)s * b_stage_bytes, src + (size_t)kb * b_stage_bytes, b_stage_bytes, &b_full[s]);
                    if (++s == b_stages)
                    {
                        s = 0;
                        ph ^= 1u;
                    }
                }
                if (b_resident)
                    break;
            }
        }
    }
    else
    {
        // ================= MMA issuers: one thread per tile of the pair.  The loop between two tcgen05.mma is a handful of
        // integer instructions (descriptor lo-word adds, ring counters): a lone thread issues a dependent instruction every
        // few cycles, and at N = 96 an MMA is only 48 cycles long. =================
        const int g = warp - (ew + 2);
        if (lane == 0 && g < G)
        {
            const bool tf32 = p.t_tf32 != 0;
            const uint32_t idesc = tf32 ? make_idesc_tf32(bn) : make_idesc(bn);
            const uint32_t b_lbo16 = (uint32_t)bn;                                  // (bn * 16 bytes) >> 4
            // A descriptors: swizzled K-major boxes (32 bytes per K = 16 step), or the canonical no-swizzle image of tile-packed
            // activations (two 2 KB core-matrix columns per step)
            const uint64_t a_proto = flat ? make_desc(0, (uint32_t)(BM * 16), 128) : make_desc_sw(0, 8u * (uint32_t)p.t_pitch, (uint32_t)p.t_layout);
            const uint32_t a_step16 = flat ? (uint32_t)(2 * BM) : 2u;
            const uint64_t b_proto = make_desc(0, (uint32_t)bn * 16, 128);
            const uint32_t a_hi = (uint32_t)(a_proto >> 32), a_lo0 = (uint32_t)a_proto;
            const uint32_t b_hi = (uint32_t)(b_proto >> 32), b_lo0 = (uint32_t)b_proto;
            const uint32_t a_smem16 = smem_u32(a_smem) >> 4, box16 = (uint32_t)p.t_box_bytes >> 4, slot16 = slot_bytes >> 4;
            const uint32_t b_smem16 = smem_u32(b_smem) >> 4, bstage16 = b_stage_bytes >> 4;
            const int spc = tf32 ? (p.t_kw >> 3) : (p.t_kw >> 4), cps = p.t_cps, tpb = p.t_tpb, nbox = p.t_nbox,
                      ngrp = p.t_nchunk / p.t_cps;
            const uint32_t shift16 = (uint32_t)p.t_shift16;
            const int slots_per_item = ngrp * nbox * G;
            int itm = 0;
            for (int item = blockIdx.x; item < total_items; item += gridDim.x, ++itm)
            {
                const int pr = item / ntiles;
                const int ng = (mtiles - pr * G) < G ? (mtiles - pr * G) : G;
                if (g >= ng)
                {
                    // the last, incomplete pair (always this CTA's last item): no tile for this issuer, but the weight
                    // blocks still need its release
                    if (!b_resident)
                    {
                        const int gb = itm * nkb;
                        int bs = gb % b_stages;
                        uint32_t bph = (uint32_t)(gb / b_stages) & 1u;
                        for (int kb = 0; kb < nkb; ++kb)
                        {
                            mbar_wait(&b_full[bs], bph);
                            mbar_arrive(&b_empty[bs]);
                            if (++bs == b_stages)
                            {
                                bs = 0;
                                bph ^= 1u;
                            }
                        }
                    }
                    break;
                }
                const int tl = itm * G + g, acc = tl % nacc;
                const uint32_t d_tmem = tmem_base + (uint32_t)acc * acc_cols;
                // ring positions at the start of this item (all earlier items of this CTA were full pairs)
                const int ga = itm * slots_per_item + g;
                int slot = ga % a_slots;
                uint32_t aph = (uint32_t)(ga / a_slots) & 1u;
                int bs = 0;
                uint32_t bph = 0;
                if (!b_resident)
                {
                    const int gb = itm * nkb;
                    bs = gb % b_stages;
                    bph = (uint32_t)(gb / b_stages) & 1u;
                }
                mbar_wait(&acc_empty[acc], ((uint32_t)(tl / nacc) & 1u) ^ 1u);
                tc_fence_after();
                uint32_t accum = 0, bstep = 0, b_lo = 0;
                if (flat)
                {
                    // one stage = this k-block's weights + the pair's A images: one wait, four MMAs, one release
                    for (int kb = 0; kb < nkb; ++kb)
                    {
                        mbar_wait(&b_full[bs], bph);
                        tc_fence_after();
                        const uint32_t a_lo = a_lo0 + a_smem16 + (uint32_t)(bs * G + g) * (uint32_t)(BM * BK * 2 / 16);
                        b_lo = b_lo0 + b_smem16 + (uint32_t)bs * bstage16;
#pragma unroll
                        for (int j = 0; j < BK / 16; ++j)
                        {
                            umma_lh(d_tmem, a_lo + (uint32_t)j * a_step16, a_hi, b_lo + (uint32_t)j * 2u * b_lbo16, b_hi, idesc, accum);
                            accum = 1;
                        }
                        umma_commit(&b_empty[bs]);
                        if (++bs == b_stages)
                        {
                            bs = 0;
                            bph ^= 1u;
                        }
                    }
                    umma_commit(&acc_full[acc]);
                    continue;
                }
                for (int qg = 0; qg < ngrp; ++qg)
                    for (int b = 0; b < nbox; ++b)
                    {
                        mbar_wait(&a_full[slot], aph);
                        tc_fence_after();
                        uint32_t a_chunk = a_lo0 + a_smem16 + (uint32_t)slot * slot16;
                        for (int qi = 0; qi < cps; ++qi, a_chunk += box16)
                        {
                            uint32_t a_tap = a_chunk;
                            for (int t = 0; t < tpb; ++t, a_tap += shift16)
                            {
                                uint32_t a_lo = a_tap;
                                for (int s = 0; s < spc; ++s, a_lo += a_step16)
                                {
                                    if (bstep == 0)
                                    {
                                        if (!b_resident || itm == 0)
                                        {
                                            mbar_wait(&b_full[bs], bph);
                                            tc_fence_after();
                                        }
                                        b_lo = b_lo0 + b_smem16 + (uint32_t)bs * bstage16;
                                    }
                                    if (tf32)
                                        umma_lh_tf32(d_tmem, a_lo, a_hi, b_lo, b_hi, idesc, accum);
                                    else
                                        umma_lh(d_tmem, a_lo, a_hi, b_lo, b_hi, idesc, accum);
                                    accum = 1;
                                    b_lo += 2 * b_lbo16;
                                    if (++bstep == 4)
                                    {
                                        bstep = 0;
                                        if (!b_resident)
                                            umma_commit(&b_empty[bs]);
                                        if (++bs == b_stages)
                                        {
                                            bs = 0;
                                            bph ^= 1u;
                                        }
                                    }
                                }
                            }
                        }
                        umma_commit(&a_empty[slot]);
                        slot += ng;
                        if (slot >= a_slots)
                        {
                            slot -= a_slots;
                            aph ^= 1u;
                        }
                    }
                if (bstep != 0 && !b_resident)
                    umma_commit(&b_empty[bs]); // the last, partial weight block of the tile
                umma_commit(&acc_full[acc]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == ew + 2)
    {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)nacc * acc_cols)
                     : "memory");
    }
}

} // namespace tc

// Column-tile width for a layer with N output columns: N itself (rounded up to 16) when it fits one
// tile, otherwise the largest multiple-of-16 divisor of N that is <= 256.
int tc_tile_n(int N)
{
    const int n16 = (N + 15) / 16 * 16;
    if (n16 <= 256)
        return n16;
    for (int bn = 256; bn >= 64; bn -= 16)
        if (N % bn == 0)
            return bn;
    return 256;
}

size_t tc_packed_elems(int K, int N)
{
    const int bn = tc_tile_n(N);
    const int nkb = (K + tc::BK - 1) / tc::BK, nt = (N + bn - 1) / bn;
    return (size_t)nt * nkb * bn * tc::BK;
}

// Host-side index of element (k, n) in the packed bf16 image: [n-tile][k-block][kc = (k%64)/8][n%bn][k%8]
size_t tc_packed_index(int K, int N, int k, int n)
{
    const int bn = tc_tile_n(N);
    const int nkb = (K + tc::BK - 1) / tc::BK;
    const int nt = n / bn, nn = n % bn, kb = k / tc::BK, kr = k % tc::BK;
    return (((size_t)nt * nkb + kb) * 8 + kr / 8) * ((size_t)bn * 8) + (size_t)nn * 8 + (kr % 8);
}

// ---- tf32 variant of the plan: fp32 activations in 32-channel chunks (128-byte rows, SWIZZLE_128B) ----
int tc_tma_plan32(int ntaps, const int *du, const int *dv, int Cin, bool share_dv, TmaPlan *pl)
{
    // the bf16 planner with a multiple-of-64 channel count gives the box / tap structure; the chunking is replaced
    if (Cin % 4 || tc_tma_plan(ntaps, du, dv, 64, share_dv, pl))
        return -1;
    pl->cw = pl->kw = 32;
    pl->nchunk = (Cin + 31) / 32;
    pl->pitch = 128;
    pl->layout = 2;
    pl->cpitch = Cin;
    pl->box_bytes = (pl->rows * pl->pitch + 1023) / 1024 * 1024;
    pl->cps = 1;
    return 0;
}

int tc_tma_k32(const TmaPlan &pl) { return pl.nchunk * 32 * pl.ntap; }

int tc_tma_korig32(const TmaPlan &pl, int Cin, int kp)
{
    const int cc = kp % 32;
    int r = kp / 32;
    const int k = r % pl.tpb;
    r /= pl.tpb;
    const int b = r % pl.nbox, q = r / pl.nbox;
    const int c = q * 32 + cc;
    return c < Cin ? pl.tap_id[b * pl.tpb + k] * Cin + c : -1;
}

// fp32 weight image for kind::tf32: [n-tile][k-block of 32][kc = (k % 32) / 4][n % bn][k % 4]
size_t tc_packed_elems32(int K, int N)
{
    const int bn = tc_tile_n(N);
    return (size_t)((N + bn - 1) / bn) * ((K + 31) / 32) * bn * 32;
}

size_t tc_packed_index32(int K, int N, int k, int n)
{
    const int bn = tc_tile_n(N);
    const int nkb = (K + 31) / 32;
    const int nt = n / bn, nn = n % bn, kb = k / 32, kr = k % 32;
    return (((size_t)nt * nkb + kb) * 8 + kr / 4) * ((size_t)bn * 4) + (size_t)nn * 4 + (kr % 4);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn tma_encoder();
// Tensor map of a row-major fp32 output for the store epilogue: (columns, positions of a line, lines), tiles of 32 rows x 16
// columns, SWIZZLE_64B.  Returns 0 and sets *ok when the mapping qualifies.
static int make_store_map(const GemmParams &p, int nout, int vout, long long lines, CUtensorMap *tm, int *ok)
{
    *ok = 0;
    memset(tm, 0, sizeof *tm);
    static const bool off = getenv("DMX_TMA_NO_STORE") != nullptr;
    EncodeTiledFn enc = tma_encoder();
    if (off || !enc || p.addC || p.rowadd || p.out_pack || p.out_bf16 || p.Cbf || (p.N % (p.glu ? 8 : 4)) != 0 || (p.ldc % 4) != 0)
        return 0;
    const cuuint64_t dims[3] = {(cuuint64_t)nout, (cuuint64_t)vout, (cuuint64_t)lines};
    const cuuint64_t strides[2] = {(cuuint64_t)p.ldc * 4, (cuuint64_t)vout * p.ldc * 4};
    const cuuint32_t box[3] = {16, 32, 1};
    const cuuint32_t es[3] = {1, 1, 1};
    const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, p.C, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
    {
        set_error("cuTensorMapEncodeTiled (output) failed (" + std::to_string((int)r) + ")");
        return -1;
    }
    *ok = 1;
    return 0;
}

template <int STAGES> static int tc_set_attr()
{
    const int max_smem = 227 * 1024 - 8 * 1024;
    DMX_CUDA(cudaFuncSetAttribute(tc::gemm_tc_kernel<STAGES, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
    DMX_CUDA(cudaFuncSetAttribute(tc::gemm_tc_kernel<STAGES, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
    if (STAGES == 2)
        DMX_CUDA(cudaFuncSetAttribute(tc::gemm_tc_kernel<2, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
    return 0;
}

template <bool VECA, bool HALO, bool PROD> static int tc_set_attr_persist()
{
    DMX_CUDA(cudaFuncSetAttribute(tc::gemm_tc_persist_kernel<VECA, HALO, PROD>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  227 * 1024 - 8 * 1024));
    return 0;
}

int tc_init_device()
{
    return tc_set_attr<1>() || tc_set_attr<2>() || tc_set_attr<3>() || tc_set_attr_persist<true, false, false>();
}

// Persistent launch: one CTA per SM, each looping over tiles.  Returns 1 when the shape does not fit.
template <bool VECA, bool HALO, bool PROD>
static int tc_launch_persist(const GemmParams &q, const __nv_bfloat16 *w, int bn, int plain, int mtiles, int ntiles,
                             size_t patch, cudaStream_t st)
{
    static int sms = 0;
    if (!sms)
    {
        int dev = 0;
        DMX_CUDA(cudaGetDevice(&dev));
        DMX_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    }
    const size_t stage_bytes = (HALO ? 0 : (size_t)tc::A_STAGE) + (size_t)bn * tc::BK * 2;
    const size_t fixed = tc::P_EPI_BYTES + tc::P_ROW_BYTES + 2 * patch;
    const size_t budget = 216 * 1024;
    if (fixed + 2 * stage_bytes > budget)
        return 1;
    int stages = (int)((budget - fixed) / stage_bytes);
    stages = stages > tc::P_MAX_STAGES ? tc::P_MAX_STAGES : stages;
    const int smem = (int)(fixed + (size_t)stages * stage_bytes);
    const int total = mtiles * ntiles;
    const int grid = total < sms ? total : sms;
    CUtensorMap tmc;
    int tma_store = 0;
    if (plain == 1 && !HALO)
    {
        if (make_store_map(q, q.glu ? q.N / 2 : q.N, q.M, 1, &tmc, &tma_store))
            return -1;
    }
    else
        memset(&tmc, 0, sizeof tmc);
    tc::gemm_tc_persist_kernel<VECA, HALO, PROD>
        <<<grid, tc::p_threads(PROD), smem, st>>>(tmc, q, w, bn, plain, stages, mtiles, ntiles, tma_store);
    DMX_LAUNCH_CHECK();
    return 0;
}

template <int STAGES>
static int tc_launch(const GemmParams &p, const __nv_bfloat16 *w, int bn, bool vec, int plain, cudaStream_t st)
{
    dim3 grid((p.M + tc::BM - 1) / tc::BM, (p.N + bn - 1) / bn);
    int smem = STAGES * (tc::A_STAGE + bn * tc::BK * 2);
    if (smem < tc::EPI_BYTES)
        smem = tc::EPI_BYTES; // the epilogue stages its transposition in the pipeline buffers
    CUtensorMap tmc;
    int tma_store = 0;
    if (plain == 1)
    {
        if (make_store_map(p, p.glu ? p.N / 2 : p.N, p.M, 1, &tmc, &tma_store))
            return -1;
    }
    else
        memset(&tmc, 0, sizeof tmc);
    if (vec)
        tc::gemm_tc_kernel<STAGES, true, false><<<grid, tc::THREADS, smem, st>>>(tmc, p, w, bn, plain, tma_store);
    else
        tc::gemm_tc_kernel<STAGES, false, false><<<grid, tc::THREADS, smem, st>>>(tmc, p, w, bn, plain, tma_store);
    DMX_LAUNCH_CHECK();
    return 0;
}

int launch_gemm_tma_packed(const GemmParams &p, const void *Wb, cudaStream_t st);
bool tc_tma_available();

int launch_gemm_tc(const GemmParams &p, const void *Wb, cudaStream_t st)
{
    if (p.M <= 0 || p.N <= 0 || p.K <= 0 || !Wb)
    {
        set_error("launch_gemm_tc: bad problem");
        return -1;
    }
    if (p.glu && (p.N % 2))
    {
        set_error("launch_gemm_tc: GLU needs an even N");
        return -1;
    }
    if (p.ntaps > 16)
    {
        set_error("launch_gemm_tc: too many taps");
        return -1;
    }
    const int bn = tc_tile_n(p.N);
    const int nout = p.glu ? p.N / 2 : p.N;
    GemmParams q = p;
    auto recip = [](int d) { return d <= 1 ? 0ull : ~0ull / (unsigned long long)d + 1ull; };
    q.div_per = recip(p.Uo * p.Vo);
    q.div_vo = recip(p.Vo);
    q.div_nco = recip(p.Nco);
    static const int nob = getenv("DMX_NO_BULK") ? 1 : 0;
    q.no_bulk = nob;
    static const int dbg_epi = getenv("DMX_DBG_EPI") ? atoi(getenv("DMX_DBG_EPI")) : 0;
    q.dbg_epi = dbg_epi;
    const bool vec = (p.Cin % 4 == 0) && (p.lda % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.A) & 15) == 0) &&
                     (!p.A2 || (reinterpret_cast<uintptr_t>(p.A2) & 15) == 0);
    // identity output mapping with 16-byte aligned rows -> the epilogue's fast path
    const int plain = (p.ous == 1 && p.ovs == 1 && p.our == 0 && p.ovr == 0 && p.ouo == 0 && p.ovo == 0 &&
                       p.Uout == p.Uo && p.Vout == p.Vo && p.Nco >= nout && (p.ldc % 4) == 0 &&
                       (reinterpret_cast<uintptr_t>(p.C) & 15) == 0 && (!p.rowadd || (p.rowadd_ld % 4) == 0) &&
                       (!p.addC || (reinterpret_cast<uintptr_t>(p.addC) & 15) == 0))
                          ? 1
                          : 0;
    // scatter mapping (transposed convolutions) whose 4-column groups stay inside one phase and are 16-byte aligned
    const int vscat = (!plain && (p.Nco % 4) == 0 && (nout % 4) == 0 && (p.ldc % 4) == 0 &&
                       (reinterpret_cast<uintptr_t>(p.C) & 15) == 0 && (!p.rowadd || (p.rowadd_ld % 4) == 0) &&
                       (!p.gamma || (reinterpret_cast<uintptr_t>(p.gamma) & 15) == 0) &&
                       (!p.addC || (reinterpret_cast<uintptr_t>(p.addC) & 15) == 0))
                          ? 2
                          : 0;
    if ((p.out_bf16 || p.out_pack) && (!plain || p.gamma || p.addC || (nout % 4) || (p.out_pack && (p.ldc % 8))))
    {
        set_error("launch_gemm_tc: bf16 / packed output needs the identity mapping, no residual and N % 4 == 0");
        return -1;
    }
    if (p.Apk && (p.K % tc::BK))
    {
        set_error("launch_gemm_tc: packed A needs K % 64 == 0");
        return -1;
    }
    const __nv_bfloat16 *w = static_cast<const __nv_bfloat16 *>(Wb);
    const int nk = (p.K + tc::BK - 1) / tc::BK;
    // ---- halo mode: same-size stride-1 convolution with >= 3 taps whose input patch fits beside the weight ring ----
    if (vec && !p.Apk && p.pro == 0 && p.ntaps >= 3 && p.su == 1 && p.sv == 1 && p.Ui == p.Uo && p.Vi == p.Vo &&
        (p.Cin % 16) == 0 && p.K == p.ntaps * p.Cin && p.lda == p.Cin)
    {
        int dumin = p.du[0], dumax = p.du[0], dvmin = p.dv[0], dvmax = p.dv[0];
        for (int t = 1; t < p.ntaps; ++t)
        {
            dumin = std::min(dumin, p.du[t]);
            dumax = std::max(dumax, p.du[t]);
            dvmin = std::min(dvmin, p.dv[t]);
            dvmax = std::max(dvmax, p.dv[t]);
        }
        const int nu = dumax - dumin + 1, rows = tc::BM + dvmax - dvmin;
        const size_t patch = (((size_t)nu * (p.Cin / 8) * ((rows | 1) * 16)) + 127) & ~(size_t)127;
        const size_t total = patch + 2 * (size_t)bn * tc::BK * 2; // two weight stages: three CTAs share an SM
        if (nu <= 3 && dvmax - dvmin <= 16 && total <= 112 * 1024)
        {
            q.halo_nu = nu;
            q.halo_dumin = dumin;
            q.halo_dvmin = dvmin;
            q.halo_rows = rows;
            q.halo_tpr = (p.Vo + tc::BM - 1) / tc::BM;
            q.div_tpr = recip(q.halo_tpr);
            q.div_uo = recip(p.Uo);
            q.div_c8 = recip(p.Cin / 8);
            q.div_rows = recip(rows);
            const int items = p.M / (p.Uo * p.Vo);
            dim3 grid(items * p.Uo * q.halo_tpr, (p.N + bn - 1) / bn);
            const int smem = (int)std::max(total, (size_t)tc::EPI_BYTES);
            CUtensorMap tmc;
            int tma_store = 0;
            if (plain == 1)
            {
                if (make_store_map(q, nout, p.Vo, (long long)items * p.Uo, &tmc, &tma_store))
                    return -1;
            }
            else
                memset(&tmc, 0, sizeof tmc);
            tc::gemm_tc_kernel<2, true, true><<<grid, tc::THREADS, smem, st>>>(tmc, q, w, bn, plain | vscat, tma_store);
            DMX_LAUNCH_CHECK();
            return 0;
        }
    }
    {
        // tile-packed A (the transformer's linear layers): persistent kernel.  The gather / halo layers stay on the
        // one-tile-per-CTA kernel: they are bound by producer + epilogue instruction issue, which three co-resident
        // CTAs (24 producer warps) serve better than one persistent CTA (measured: 20.3 -> 25.2 ms/step persistent).
        // tile-packed A (the transformer's linear layers): the ring kernel with A riding in the weight ring's stages, one
        // tile per pass, 16 epilogue warps -- 4 stages of 48 KB in flight against the 2 of the persistent kernel below
        // (7.30 -> 6.87 ms/step); DMX_NO_TMA_PACKED=1 goes back to it
        static const bool packed_tma = getenv("DMX_NO_TMA_PACKED") == nullptr;
        if (p.Apk && plain && packed_tma && tc_tma_available())
            return launch_gemm_tma_packed(p, Wb, st);
        if (p.Apk)
        {
            const int mt = (p.M + tc::BM - 1) / tc::BM, nt = (p.N + bn - 1) / bn;
            const int r = tc_launch_persist<true, false, false>(q, w, bn, plain | vscat, mt, nt, 0, st);
            if (r <= 0)
                return r;
        }
    }
    if (nk == 1)
        return tc_launch<1>(q, w, bn, vec, plain | vscat, st);
    // wide tiles: two stages keep the CTA under half of the shared memory so two CTAs share an SM (one's
    // epilogue overlaps the other's main loop)
    if (nk == 2 || bn > 128)
        return tc_launch<2>(q, w, bn, vec, plain | vscat, st);
    return tc_launch<3>(q, w, bn, vec, plain | vscat, st);
}

// ---------------------------------------------------------------------------------------------------------------------
// TMA-fed mode: plan, tensor map and launch
int tc_tma_plan(int ntaps, const int *du, const int *dv, int Cin, bool share_dv, TmaPlan *pl)
{
    if (ntaps < 1 || ntaps > 9 || Cin % 16)
        return -1;
    TmaPlan t;
    t.cw = Cin % 64 == 0 ? 64 : (Cin % 32 == 0 ? 32 : 16);
    t.kw = t.cw;
    t.cpitch = tc_tma_cpitch(Cin);
    if (t.cpitch != Cin)
    {
        t.cw = 64; // one 128-byte row per position, 48 of its 64 channels used
        t.kw = Cin;
    }
    t.nchunk = Cin / t.kw;
    t.pitch = t.cw * 2;
    t.layout = t.cw == 64 ? 2 : (t.cw == 32 ? 4 : 6);
    int dvmin = dv[0], dvmax = dv[0];
    for (int i = 1; i < ntaps; ++i)
    {
        dvmin = std::min(dvmin, dv[i]);
        dvmax = std::max(dvmax, dv[i]);
    }
    t.nbox = 0;
    t.ntap = 0;
    if (share_dv)
    {
        t.rows = tc::BM + dvmax - dvmin;
        // boxes = distinct du in order of first appearance; taps grouped by box, increasing dv inside a box
        for (int i = 0; i < ntaps; ++i)
        {
            bool seen = false;
            for (int k = 0; k < t.nbox; ++k)
                seen = seen || t.box_du[k] == du[i];
            if (!seen)
            {
                t.box_du[t.nbox] = du[i];
                t.box_dv[t.nbox] = dvmin;
                ++t.nbox;
            }
        }
        if (ntaps % t.nbox)
            return -1;
        t.tpb = ntaps / t.nbox;
        t.shift = t.tpb > 1 ? (dvmax - dvmin) / (t.tpb - 1) : 0;
        for (int b = 0; b < t.nbox; ++b)
            for (int k = 0; k < t.tpb; ++k)
            {
                // the tap of this plane at row shift k * shift
                int found = -1;
                for (int i = 0; i < ntaps; ++i)
                    if (du[i] == t.box_du[b] && dv[i] - dvmin == k * t.shift)
                        found = i;
                if (found < 0)
                    return -1; // not a regular tap grid
                t.tap_id[t.ntap++] = found;
            }
    }
    else
    {
        t.rows = tc::BM;
        for (int i = 0; i < ntaps; ++i)
        {
            t.box_du[i] = du[i];
            t.box_dv[i] = dv[i];
            t.tap_id[i] = i;
        }
        t.nbox = t.ntap = ntaps;
        t.tpb = 1;
        t.shift = 0;
    }
    if (t.rows > 256)
        return -1;
    t.box_bytes = (t.rows * t.pitch + 1023) / 1024 * 1024;
    // narrow chunks: all chunks of a plane travel together (one ring slot, one barrier) when that stays small
    t.cps = (t.nchunk > 1 && t.nchunk * t.box_bytes <= 18 * 1024) ? t.nchunk : 1;
    *pl = t;
    return 0;
}

int tc_tma_cpitch(int C) { return C == 48 ? 64 : C; }

int tc_tma_korig(const TmaPlan &pl, int Cin, int kp)
{
    const int cc = kp % pl.kw;
    int r = kp / pl.kw;
    const int k = r % pl.tpb;
    r /= pl.tpb;
    const int qi = r % pl.cps;
    r /= pl.cps;
    const int b = r % pl.nbox, qg = r / pl.nbox;
    return pl.tap_id[b * pl.tpb + k] * Cin + (qg * pl.cps + qi) * pl.kw + cc;
}

static EncodeTiledFn tma_encoder()
{
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried)
    {
        tried = true;
        void *f = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) == cudaSuccess &&
            qr == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(f);
        else
            cudaGetLastError();
    }
    return fn;
}

bool tc_tma_available()
{
    static const bool off = getenv("DMX_NO_TMA") != nullptr;
    return !off && tma_encoder() != nullptr;
}

static int tma_set_attr()
{
    DMX_CUDA(cudaFuncSetAttribute(tc::gemm_tc_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 4 * 1024));
    return 0;
}

static int mtiles_of_lines(int items, int Uo, int tpr) { return items * Uo * tpr; }

// Shared-memory budget, ring depths, pairing and launch of gemm_tc_tma_kernel.  slots_per_tile = ring slots one row tile
// consumes; (store_v, store_lines) = the output viewed as lines of store_v rows for the tensor-map store.
static int tma_launch_core(const GemmParams &q, const CUtensorMap &tm, const __nv_bfloat16 *W, int bn, int plain_out, int nout,
                           int mtiles, size_t slot_bytes, int slots_per_tile, long long store_v, long long store_lines,
                           cudaStream_t st, int g_want = 0)
{
    static int sms = 0;
    if (!sms)
    {
        int dev = 0;
        DMX_CUDA(cudaGetDevice(&dev));
        DMX_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        if (tma_set_attr())
            return -1;
    }
    static const int ew_env = getenv("DMX_TMA_EW") ? atoi(getenv("DMX_TMA_EW")) : 0;
    static const int g_env = getenv("DMX_TMA_G") ? atoi(getenv("DMX_TMA_G")) : 0;
    static const bool no_res = getenv("DMX_TMA_NO_RESIDENT") != nullptr;
    static const bool no_tmast = getenv("DMX_TMA_NO_STORE") != nullptr;
    const size_t b_stage = (size_t)bn * tc::BK * 2; // = bn x 32 fp32 for the tf32 variant
    const size_t budget = 220 * 1024;
    const int kblk = q.t_tf32 ? 32 : tc::BK;
    const int nkb = (q.K + kblk - 1) / kblk;
    const int ntiles = (q.N + bn - 1) / bn;
    int acc_cols = 32;
    while (acc_cols < bn)
        acc_cols <<= 1;
    // tensor map of the output for the store epilogue
    CUtensorMap tmc;
    int tma_store = 0;
    if (plain_out == 1 && !no_tmast)
    {
        if (make_store_map(q, nout, (int)store_v, store_lines, &tmc, &tma_store))
            return -1;
    }
    else
        memset(&tmc, 0, sizeof tmc);
    // per-warp epilogue shared memory by epilogue kind (see the kernel)
    const bool packed_direct = q.out_pack && plain_out == 1 && !q.glu && !q.gamma && !q.addC && !q.rowadd && !q.stats && (q.N % 8) == 0;
    const bool tmast = tma_store && plain_out == 1 && !q.addC && !q.rowadd && !q.out_pack && !q.out_bf16 && !q.Cbf &&
                       (q.N % (q.glu ? 8 : 4)) == 0;
    // coalesced direct stores instead of the tensor-map store: measured no better (transformer GEMMs 7.04 -> 7.39 ms/step
    // when used for the 512-wide residual outputs, conv GEMMs 13.3 -> 13.7 when used everywhere), so only on request
    static const bool direct_env = getenv("DMX_TMA_DIRECT") != nullptr && atoi(getenv("DMX_TMA_DIRECT")) == 1;
    const bool direct = tmast && direct_env;
    if (direct)
        tma_store |= 2;
    const int epi_warp = packed_direct ? 0 : (direct ? tc::T_EPI_WARP_BYTES : (tmast ? 4096 : tc::T_EPI_WARP_BYTES + 32 * (int)sizeof(tc::RowInfo)));
    // G = 2 row tiles share every weight block (one MMA issuer and one group of epilogue warps per tile); four accumulators
    // (one pair draining while the next fills) when they fit in TMEM
    int G = (g_env == 1 || g_env == 2) ? g_env : (g_want ? g_want : 2);
    if (mtiles < 2 * sms && !g_env)
        G = 1; // too few tiles to pair without idling SMs
    const int nacc = (2 * G * acc_cols <= 512) ? 2 * G : G;
    if (nacc * acc_cols > 512)
    {
        set_error("launch_gemm_tma: accumulators do not fit in TMEM");
        return -1;
    }
    auto fixed_for = [&](int e) { return ((size_t)e * epi_warp + 1023) / 1024 * 1024; };
    // epilogue warps: 8 per tile when shared memory allows (the epilogue is a chain of TMEM load -> math -> store tile ->
    // tensor-map store per 32 columns: more warps = more chains in flight), else 4 per tile
    int ew = 8 * G;
    int b_resident = 0, b_stages = 0;
    {
        // preference order: resident weights with 8 then 4 epilogue warps per tile; streamed weights with (8, 3 stages),
        // (4, 3 stages), (4, 2 stages); every variant needs at least 2 boxes per tile of the pair in the ring
        const int min_boxes = 2 * G;
        bool found = false;
        const bool can_res = !no_res && ntiles == 1 && nkb <= tc::T_MAX_BSTAGES;
        // total epilogue warps: 16 (8 per tile of a pair, 16 for a single tile), 8, or 4 per tile
        const int ews[3] = {16, 8, 4 * G};
        for (int i = 0; i < 3 && !found && can_res; ++i)
            if (fixed_for(ews[i]) + (size_t)nkb * b_stage + (size_t)(min_boxes + 1) * slot_bytes <= budget)
            {
                ew = ews[i];
                b_resident = 1;
                b_stages = nkb;
                found = true;
            }
        const int tries[4][2] = {{16, 3}, {8, 3}, {4 * G, 3}, {4 * G, 2}};
        for (int t = 0; t < 4 && !found; ++t)
        {
            const int stg = std::min(tries[t][1], std::max(nkb, 2));
            if (fixed_for(tries[t][0]) + (size_t)stg * b_stage + (size_t)min_boxes * slot_bytes <= budget)
            {
                ew = tries[t][0];
                b_stages = stg;
                found = true;
            }
        }
        if (!found)
        {
            set_error("launch_gemm_tma: tile does not fit in shared memory");
            return -1;
        }
    }
    if (ew_env == 4 * G || ew_env == 8 || ew_env == 16)
        ew = ew_env;
    const size_t fixed = fixed_for(ew);
    if (q.Apk)
    {
        // tile-packed A: a stage holds the k-block's weights and the pair's A images
        const size_t stage_total = b_stage + (size_t)G * slot_bytes;
        int stages = (int)((budget - fixed) / stage_total);
        stages = std::min(stages, std::min(8, std::max(nkb, 2)));
        if (stages < 2)
        {
            set_error("launch_gemm_tma: tile does not fit in shared memory");
            return -1;
        }
        const size_t smem_f = fixed + (size_t)stages * stage_total;
        const int npairs_f = (mtiles + G - 1) / G;
        const int total_f = npairs_f * ntiles;
        tc::gemm_tc_tma_kernel<<<total_f < sms ? total_f : sms, (ew + 2 + G) * 32, smem_f, st>>>(
            tm, tmc, q, W, bn, plain_out, stages * G, stages, mtiles, ntiles, ew, G, nacc, 0, tma_store, epi_warp);
        DMX_LAUNCH_CHECK();
        return 0;
    }
    // Ring depths.  Streamed weights are what these layers wait for: a stage is refilled only after the MMAs that read it
    // have completed and the copy has crossed L2 -> SM (~1.5 us), so the bytes in flight bound the rate.  Boxes first get a
    // modest ring (4 per tile of the pair), then the weight ring takes what is left, then the boxes the remainder.
    const int slots_per_item = G * slots_per_tile;
    int a_max = (int)((budget - fixed - b_stages * b_stage) / slot_bytes);
    a_max = std::min(a_max, tc::T_MAX_ASLOTS);
    if (a_max < 2 * G)
    {
        set_error("launch_gemm_tma: box ring does not fit in shared memory");
        return -1;
    }
    int a_slots = std::min(a_max, std::max(2 * G, std::min(4 * G, slots_per_item)));
    while (!b_resident && b_stages < tc::T_MAX_BSTAGES && b_stages < nkb &&
           fixed + (size_t)(b_stages + 1) * b_stage + (size_t)a_slots * slot_bytes <= budget)
        ++b_stages;
    while (a_slots < tc::T_MAX_ASLOTS && a_slots < 2 * slots_per_item &&
           fixed + (size_t)b_stages * b_stage + (size_t)(a_slots + 1) * slot_bytes <= budget)
        ++a_slots;
    const size_t smem = fixed + (size_t)b_stages * b_stage + (size_t)a_slots * slot_bytes;
    const int npairs = (mtiles + G - 1) / G;
    const int total = npairs * ntiles;
    const int grid = total < sms ? total : sms;
    tc::gemm_tc_tma_kernel<<<grid, (ew + 2 + G) * 32, smem, st>>>(tm, tmc, q, W, bn, plain_out, a_slots, b_stages, mtiles, ntiles, ew,
                                                             G, nacc, b_resident, tma_store, epi_warp);
    DMX_LAUNCH_CHECK();
    return 0;
}

// p: a stride-1 "same" convolution (Ui == Uo * 1, Vi == Vo, su == sv == 1) or any tap set without v shifts, Cin % 16 == 0,
// K == ntaps * Cin, A given as the bf16 image p.Aimg, weights as p.Wt (K' order of tc_tma_plan with share_dv = true unless
// the environment says otherwise).
int launch_gemm_tma(const GemmParams &p, cudaStream_t st)
{
    EncodeTiledFn enc = tma_encoder();
    if (!enc)
    {
        set_error("launch_gemm_tma: cuTensorMapEncodeTiled is not available from this driver");
        return -1;
    }
    const bool tf32 = p.Wt32 != nullptr;
    if ((!tf32 && (!p.Aimg || !p.Wt || p.Cin % 16)) || (tf32 && (!p.A || p.Cin % 4 || p.lda % 4)) || p.M <= 0 || p.N <= 0 ||
        p.K != p.ntaps * p.Cin || p.pro || p.A2 || p.Apk || p.sv != 1)
    {
        set_error("launch_gemm_tma: unsupported problem");
        return -1;
    }
    static const bool share_env = getenv("DMX_TMA_NO_SHARE") == nullptr;
    const bool share = share_env && !p.t_noshare;
    TmaPlan pl;
    if (tf32 ? tc_tma_plan32(p.ntaps, p.du, p.dv, p.Cin, share, &pl) : tc_tma_plan(p.ntaps, p.du, p.dv, p.Cin, share, &pl))
    {
        set_error("launch_gemm_tma: no plan for this tap set");
        return -1;
    }
    const int bn = tc_tile_n(p.N);
    const int nout = p.glu ? p.N / 2 : p.N;
    GemmParams q = p;
    auto recip = [](int d) { return d <= 1 ? 0ull : ~0ull / (unsigned long long)d + 1ull; };
    q.div_per = recip(p.Uo * p.Vo);
    q.div_vo = recip(p.Vo);
    q.div_nco = recip(p.Nco);
    q.halo_tpr = (p.Vo + tc::BM - 1) / tc::BM;
    q.div_tpr = recip(q.halo_tpr);
    q.div_uo = recip(p.Uo);
    q.t_nbox = pl.nbox;
    for (int i = 0; i < 9; ++i)
    {
        q.t_box_du[i] = i < pl.nbox ? pl.box_du[i] : 0;
        q.t_box_dv[i] = i < pl.nbox ? pl.box_dv[i] : 0;
    }
    q.t_tpb = pl.tpb;
    q.t_shift16 = pl.shift * pl.pitch / 16;
    q.t_cps = pl.cps;
    q.t_rows = pl.rows;
    q.t_cw = pl.cw;
    q.t_kw = pl.kw;
    q.t_nchunk = pl.nchunk;
    q.t_pitch = pl.pitch;
    q.t_box_bytes = pl.box_bytes;
    q.t_layout = pl.layout;
    q.t_tf32 = tf32 ? 1 : 0;
    if (tf32)
        q.K = tc_tma_k32(pl); // K' includes the zero-filled channels of the last chunk
    const size_t slot_bytes = (size_t)pl.cps * pl.box_bytes; // one ring slot = the cps chunk boxes of a plane
    const int plain = (p.ous == 1 && p.ovs == 1 && p.our == 0 && p.ovr == 0 && p.ouo == 0 && p.ovo == 0 && p.Uout == p.Uo &&
                       p.Vout == p.Vo && p.Nco >= nout && (p.ldc % 4) == 0 && (reinterpret_cast<uintptr_t>(p.C) & 15) == 0 &&
                       (!p.rowadd || (p.rowadd_ld % 4) == 0) && (!p.addC || (reinterpret_cast<uintptr_t>(p.addC) & 15) == 0))
                          ? 1
                          : 0;
    const int vscat = (!plain && (p.Nco % 4) == 0 && (nout % 4) == 0 && (p.ldc % 4) == 0 &&
                       (reinterpret_cast<uintptr_t>(p.C) & 15) == 0 && (!p.rowadd || (p.rowadd_ld % 4) == 0) &&
                       (!p.gamma || (reinterpret_cast<uintptr_t>(p.gamma) & 15) == 0) &&
                       (!p.addC || (reinterpret_cast<uintptr_t>(p.addC) & 15) == 0))
                          ? 2
                          : 0;
    if (!plain && !vscat)
    {
        set_error("launch_gemm_tma: output mapping not supported");
        return -1;
    }
    // tensor map over the image: dims (C, V, U, B), innermost first
    const int items = p.M / (p.Uo * p.Vo);
    CUtensorMap tm;
    {
        const cuuint64_t cp = (cuuint64_t)(tf32 ? p.lda : (p.Aimg_ld ? p.Aimg_ld : p.Cin));
        if (!tf32 && (int)cp != pl.cpitch)
        {
            set_error("launch_gemm_tma: the image's row pitch does not match the plan");
            return -1;
        }
        const cuuint64_t esz = tf32 ? 4 : 2;
        // tf32: the channel extent is Cin, so the last 32-channel chunk of e.g. 48 channels is zero-filled past channel 47
        const cuuint64_t dims[4] = {tf32 ? (cuuint64_t)p.Cin : cp, (cuuint64_t)p.Vi, (cuuint64_t)p.Ui, (cuuint64_t)items};
        const cuuint64_t strides[3] = {cp * esz, (cuuint64_t)p.Vi * cp * esz, (cuuint64_t)p.Ui * p.Vi * cp * esz};
        const cuuint32_t box[4] = {(cuuint32_t)pl.cw, (cuuint32_t)pl.rows, 1, 1};
        const cuuint32_t es[4] = {1, 1, 1, 1};
        const CUtensorMapSwizzle sw = pl.pitch == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                                      : (pl.pitch == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
        const CUresult r = enc(&tm, tf32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                               const_cast<void *>(tf32 ? static_cast<const void *>(p.A) : p.Aimg), dims, strides, box, es,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS)
        {
            set_error("launch_gemm_tma: cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
            return -1;
        }
    }
    return tma_launch_core(q, tm, static_cast<const __nv_bfloat16 *>(tf32 ? p.Wt32 : p.Wt), bn, plain | vscat, nout,
                           mtiles_of_lines(items, p.Uo, q.halo_tpr),
                           slot_bytes, pl.nbox * (pl.nchunk / pl.cps), p.Vout, (long long)items * p.Uout, st);
}

// The transformer's linear layers on the same kernel: A arrives tile-packed ([row/128][K/8][row%128][8] bf16, written by
// LayerNorm / attention / the previous GEMM), i.e. one contiguous 16 KB image per (row tile, 64-wide k-block), fetched by
// bulk copies into the box ring; W = the ordinary packed weight image.  Same rings, pairing and epilogues as above.
int launch_gemm_tma_packed(const GemmParams &p, const void *Wb, cudaStream_t st)
{
    if (!p.Apk || !Wb || p.M <= 0 || p.N <= 0 || p.K % tc::BK || p.ntaps != 1 || p.pro || p.A2)
    {
        set_error("launch_gemm_tma_packed: unsupported problem");
        return -1;
    }
    const int bn = tc_tile_n(p.N);
    const int nout = p.glu ? p.N / 2 : p.N;
    GemmParams q = p;
    auto recip = [](int d) { return d <= 1 ? 0ull : ~0ull / (unsigned long long)d + 1ull; };
    q.div_per = recip(p.Uo * p.Vo);
    q.div_vo = recip(p.Vo);
    q.div_nco = recip(p.Nco);
    q.t_nbox = 1;
    q.t_tpb = 1;
    q.t_shift16 = 0;
    q.t_cps = 1;
    q.t_cw = q.t_kw = tc::BK;
    q.t_nchunk = p.K / tc::BK;
    q.t_pitch = 128;
    q.t_rows = tc::BM;
    q.t_box_bytes = tc::BM * tc::BK * 2;
    q.t_layout = 0;
    const int plain = (p.ous == 1 && p.ovs == 1 && p.our == 0 && p.ovr == 0 && p.ouo == 0 && p.ovo == 0 && p.Uout == p.Uo &&
                       p.Vout == p.Vo && p.Nco >= nout && (p.ldc % 4) == 0 && (reinterpret_cast<uintptr_t>(p.C) & 15) == 0 &&
                       (!p.rowadd || (p.rowadd_ld % 4) == 0) && (!p.addC || (reinterpret_cast<uintptr_t>(p.addC) & 15) == 0))
                          ? 1
                          : 0;
    if (!plain)
    {
        set_error("launch_gemm_tma_packed: output mapping not supported");
        return -1;
    }
    if ((p.out_bf16 || p.out_pack) && (p.gamma || p.addC || (nout % 4) || (p.out_pack && (p.ldc % 8))))
    {
        set_error("launch_gemm_tma_packed: bf16 / packed output needs no residual and N % 4 == 0");
        return -1;
    }
    CUtensorMap tm;
    memset(&tm, 0, sizeof tm);
    // pairing needs both accumulators of a 256-wide tile pair at once (512 TMEM columns), so a pair's MMAs and epilogues
    // alternate instead of overlapping: worth it only when the MMA phase is long (K >= 1024: FFN down-projection)
    static const int pg_env = getenv("DMX_TMA_PACKED_G") ? atoi(getenv("DMX_TMA_PACKED_G")) : 0;
    int acc_cols = 32;
    while (acc_cols < bn)
        acc_cols <<= 1;
    // measured (transformer GEMMs, ms/step): one tile per pass 6.87, pairs for K >= 1024 only 7.46, pairs everywhere 8.18
    const int g_want = pg_env ? pg_env : ((4 * acc_cols <= 512) ? 2 : 1);
    return tma_launch_core(q, tm, static_cast<const __nv_bfloat16 *>(Wb), bn, plain, nout, (p.M + tc::BM - 1) / tc::BM,
                           (size_t)tc::BM * tc::BK * 2, p.K / tc::BK, p.M, 1, st, g_want);
}

} // namespace dmx
