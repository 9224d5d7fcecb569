// bf16 tensor-core attention for sm_100a (bf16 mode): softmax(Q K^T / sqrt(dh)) V per head with the
// streaming-softmax formulation, both GEMMs on tcgen05 with accumulators in TMEM.
//
// One CTA = 128 queries of one (batch item, head).  Per 128-key tile:
//     S  = Q K^T   tcgen05.mma  M=128, N=128, K=dh   (Q, K: K-major canonical smem tiles)     -> TMEM cols [0,128)
//     256 threads: thread (row, half) owns 64 of the 128 keys of its query row: tcgen05.ld that half row, running
//     max / sum in fp32 (log2 domain), P = exp2(S*c - m) written as bf16 into the canonical K-major A-operand
//     layout (conflict-free 16-byte stores).  The two halves keep SEPARATE (m, l, O) streams, so no per-tile
//     exchange between them is needed: they are merged once at the end through shared memory.
//     PV_h = P_h V_h  tcgen05.mma  M=128, N=dh, K=64 per half (V: MN-major canonical tile, i.e. stored as it lies
//                                                      in memory, dh contiguous)      -> TMEM cols [128+h*dh, ...)
//     O_h accumulates IN TMEM over all key tiles (the PV MMAs run with the accumulate flag).  The maximum used in the
//     exponent may lag the true row maximum by up to 2^8; O_h and the row sum are rescaled (TMEM load / scale / store)
//     only when it is raised -- the first tiles -- so a tile costs one pass over its scores and ONE commit: S of
//     tile j+1 is issued right behind PV of tile j.
// Q/K/V arrive tile-packed bf16 (the QKV projection's epilogue writes them that way, fetched by bulk copies); the
// output is written tile-packed bf16 for the out-projection.  The T x S score matrix never leaves the SM.
// Reference: src/layers.cpp:442-482 (heads = contiguous dh-wide column blocks, max-subtracted softmax).
#include "kernels.cuh"

#include <cuda_bf16.h>

namespace dmx
{
namespace atc
{

constexpr int BQ = 128;   // queries per CTA (UMMA M)
constexpr int BKEY = 128; // keys per tile
constexpr int THREADS = 256; // two threads per query row (one per 64-key half of a tile)

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
// 32 consecutive TMEM columns of this thread's lane.  The wait names the registers as in/out operands so that no
// use of them can be scheduled above it.
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t *r)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]),
                   "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]),
                   "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                   "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]),
                   "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]),
                   "+r"(r[30]), "+r"(r[31])::"memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t *r)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])::"memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t *r)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
                 "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
                 "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
                 "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
                 : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float ex2f(float x)
{
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi)
{
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t *>(&h);
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// DH = head dim (64 or 48).  Q / K / V arrive TILE-PACKED (the projection GEMM's epilogue writes them that way), so
// a (128 rows x DH) tile of one head is ONE contiguous image that is already the tcgen05 operand layout:
//   [DH/8 chunks][128 rows][8 bf16]  = K-major for Q and K (rows = queries / keys), MN-major for V (mn = dh, k = keys)
// and is fetched with a single bulk async copy per tile; K/V tiles are double buffered so the copies of tile j+1 run
// under the MMAs and the softmax of tile j.  No thread touches the operands.
template <int DH>
__global__ void __launch_bounds__(THREADS, 2)
    attention_tc_kernel(const __nv_bfloat16 *__restrict__ Qp, int ldq8, int qc8, const __nv_bfloat16 *__restrict__ Kp,
                        int ldk8, int kc8, const __nv_bfloat16 *__restrict__ Vp, int ldv8, int vc8,
                        uint4 *__restrict__ Opk, int T, int Tpad, int S, int Spad, int heads, float scale_log2)
{
    constexpr int DCH = DH / 8;                   // 16-byte chunks per row
    constexpr uint32_t LBO = BQ * 16;             // 2048: chunk stride inside a packed tile
    constexpr uint32_t TILE = DCH * LBO;          // bytes of a 128 x DH tile
    constexpr int HK = BKEY / 2;                  // keys per thread per tile
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *Qs = smem, *Ks = Qs + TILE, *Vs = Ks + 2 * TILE, *Ps = Vs + 2 * TILE;
    __shared__ __align__(8) uint64_t mma_bar, q_full, kv_full[2];
    __shared__ uint32_t tmem_slot;

    const int tid = threadIdx.x, warp = tid >> 5;
    const int half = warp >> 2;      // which 64 keys of every tile
    const int row = tid & (BQ - 1);  // query row == TMEM lane (warp % 4 selects the lane quadrant)
    const int qt = blockIdx.x;
    const int h = blockIdx.y, b = blockIdx.z;
    const int q0 = qt * BQ;
    const int ntiles = (S + BKEY - 1) / BKEY;
    // packed tile (item b, row tile rt, head h) of a matrix with ld8 chunks per row and first chunk c8:
    auto tile_ptr = [&](const __nv_bfloat16 *base, int ld8, int c8, int rows_pad, int rt) {
        return reinterpret_cast<const uint8_t *>(base) +
               (((size_t)(b * (rows_pad >> 7) + rt) * ld8 + c8 + h * DCH) * (size_t)LBO);
    };

    if (tid == 0)
    {
        mbar_init(&mma_bar, 1);
        mbar_init(&q_full, 1);
        mbar_init(&kv_full[0], 1);
        mbar_init(&kv_full[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                     "r"(256u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (tid == 0)
    {
        mbar_arrive_expect_tx(&q_full, TILE);
        bulk_g2s(Qs, tile_ptr(Qp, ldq8, qc8, Tpad, qt), TILE, &q_full);
        mbar_arrive_expect_tx(&kv_full[0], 2 * TILE);
        bulk_g2s(Ks, tile_ptr(Kp, ldk8, kc8, Spad, 0), TILE, &kv_full[0]);
        bulk_g2s(Vs, tile_ptr(Vp, ldv8, vc8, Spad, 0), TILE, &kv_full[0]);
    }
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tmem_S = tmem_slot + lane_off + (uint32_t)(half * HK);
    const uint32_t tmem_PV = tmem_slot + lane_off + 128u + (uint32_t)(half * DH);
    const uint32_t idesc_s = make_idesc(BKEY, 0), idesc_pv = make_idesc(DH, 1);
    uint8_t *Prow = Ps + (half * (HK / 8)) * LBO + row * 16;

    // The running output O_h stays in TMEM for the whole key loop (PV accumulates into it); the row maximum used in the
    // exponent, m_use, is allowed to lag the true maximum by up to 2^8 (P <= 256 is harmless in bf16 / fp32) and O_h
    // and the row sum are rescaled -- a TMEM load / scale / store by the whole warp -- only when it is raised, which
    // happens in the first tiles only.  The result is the same softmax: P, the sum and O always share one m_use.
    // One commit per tile: S(j+1) is issued right behind PV(j), and its completion implies PV(j)'s.
    float m_use = -INFINITY, l_run = 0.f; // log2 domain
    constexpr float LAG = 8.0f;

    if (tid == 0)
    {
        mbar_wait(&q_full, 0);
        mbar_wait(&kv_full[0], 0);
        tc_fence_after();
        const uint32_t qa = smem_u32(Qs), ka = smem_u32(Ks);
#pragma unroll
        for (int jj = 0; jj < DH / 16; ++jj)
            umma(tmem_slot, make_desc(qa + jj * 2 * LBO, LBO, 128), make_desc(ka + jj * 2 * LBO, LBO, 128), idesc_s,
                 jj > 0 ? 1u : 0u);
        umma_commit(&mma_bar);
    }
    for (int j = 0; j < ntiles; ++j)
    {
        const int buf = j & 1;
        mbar_wait(&mma_bar, (uint32_t)(j & 1)); // S(j) complete, and with it PV(j-1): the other K/V buffer is free
        tc_fence_after();
        if (tid == 0 && j + 1 < ntiles)
        {
            const int nb = (j + 1) & 1;
            mbar_arrive_expect_tx(&kv_full[nb], 2 * TILE);
            bulk_g2s(Ks + nb * TILE, tile_ptr(Kp, ldk8, kc8, Spad, j + 1), TILE, &kv_full[nb]);
            bulk_g2s(Vs + nb * TILE, tile_ptr(Vp, ldv8, vc8, Spad, j + 1), TILE, &kv_full[nb]);
        }
        // ---- this thread's 64 scores of the tile, read once ----
        const int nvalid = min(HK, max(0, S - j * BKEY - half * HK));
        uint32_t s[HK];
        tmem_ld32(tmem_S, s);
        tmem_ld32(tmem_S + 32u, s + 32);
        float mx = -INFINITY;
        if (nvalid == HK)
        {
#pragma unroll
            for (int jj = 0; jj < HK; ++jj)
                mx = fmaxf(mx, __uint_as_float(s[jj]));
        }
        else
        {
#pragma unroll
            for (int jj = 0; jj < HK; ++jj)
                if (jj < nvalid)
                    mx = fmaxf(mx, __uint_as_float(s[jj]));
        }
        mx *= scale_log2; // scale > 0: max commutes with it
        const bool raise = mx > m_use + LAG || (m_use == -INFINITY && mx > -INFINITY);
        if (__any_sync(0xffffffffu, raise))
        {
            const float f = raise ? ((m_use == -INFINITY) ? 0.f : ex2f(m_use - mx)) : 1.0f;
            if (j > 0)
            {
#pragma unroll
                for (int c = 0; c < DH; c += 16)
                {
                    uint32_t ov[16];
                    tmem_ld16(tmem_PV + (uint32_t)c, ov);
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        ov[i] = __float_as_uint(__uint_as_float(ov[i]) * f);
                    tmem_st16(tmem_PV + (uint32_t)c, ov);
                }
                tmem_wait_st();
            }
            l_run *= f;
            if (raise)
                m_use = mx;
        }
        const float m_sub = (m_use == -INFINITY) ? 0.f : m_use; // no real key yet: the p below are all forced to 0
        float rs = 0.f;
#pragma unroll
        for (int c = 0; c < HK; c += 8)
        {
            float pv[8];
#pragma unroll
            for (int e = 0; e < 8; ++e)
            {
                const float p = ex2f(fmaf(__uint_as_float(s[c + e]), scale_log2, -m_sub));
                pv[e] = (nvalid == HK || c + e < nvalid) ? p : 0.f;
                rs += pv[e];
            }
            uint4 pk;
            pk.x = pack_bf16x2(pv[0], pv[1]);
            pk.y = pack_bf16x2(pv[2], pv[3]);
            pk.z = pack_bf16x2(pv[4], pv[5]);
            pk.w = pack_bf16x2(pv[6], pv[7]);
            *reinterpret_cast<uint4 *>(Prow + (c / 8) * LBO) = pk;
        }
        l_run += rs;
        tc_fence_before();
        fence_proxy_async(); // P was written through the generic proxy
        __syncthreads();
        // ---- PV_h += P_h V_h for both key halves, then S of the next tile right behind it ----
        if (tid == 0)
        {
            tc_fence_after();
            const uint32_t pa = smem_u32(Ps), va = smem_u32(Vs + buf * TILE);
#pragma unroll
            for (int jj = 0; jj < BKEY / 16; ++jj)
                umma(tmem_slot + 128u + (uint32_t)((jj / 4) * DH), make_desc(pa + jj * 2 * LBO, LBO, 128),
                     make_desc(va + jj * 256, 128, LBO), idesc_pv, ((jj & 3) || j > 0) ? 1u : 0u);
            if (j + 1 < ntiles)
            {
                const int nb = (j + 1) & 1;
                mbar_wait(&kv_full[nb], (uint32_t)(((j + 1) / 2) & 1));
                tc_fence_after();
                const uint32_t qa = smem_u32(Qs), ka = smem_u32(Ks + nb * TILE);
#pragma unroll
                for (int jj = 0; jj < DH / 16; ++jj)
                    umma(tmem_slot, make_desc(qa + jj * 2 * LBO, LBO, 128), make_desc(ka + jj * 2 * LBO, LBO, 128), idesc_s,
                         jj > 0 ? 1u : 0u);
            }
            umma_commit(&mma_bar);
        }
    }
    mbar_wait(&mma_bar, (uint32_t)(ntiles & 1)); // the last PV
    tc_fence_after();
    float o[DH];
#pragma unroll
    for (int c = 0; c < DH; c += 16)
    {
        uint32_t ov[16];
        tmem_ld16(tmem_PV + (uint32_t)c, ov);
#pragma unroll
        for (int i = 0; i < 16; ++i)
            o[c + i] = __uint_as_float(ov[i]);
    }
    tc_fence_before();
    const float m_run = m_use;
    // ---- merge the two key halves of every row (all MMAs and copies have completed: operand smem is free) ----
    float *mg = reinterpret_cast<float *>(smem); // [DH + 2][128]
    __syncthreads();
    if (half == 1)
    {
        mg[row] = m_run;
        mg[BQ + row] = l_run;
#pragma unroll
        for (int i = 0; i < DH; ++i)
            mg[(2 + i) * BQ + row] = o[i];
    }
    __syncthreads();
    const int grow_i = q0 + row;
    if (half == 0 && grow_i < T)
    {
        const float m1 = mg[row], l1 = mg[BQ + row];
        const float m = fmaxf(m_run, m1); // half 0 always holds real keys: m_run is finite
        const float a0 = ex2f(m_run - m), a1 = (m1 == -INFINITY) ? 0.f : ex2f(m1 - m);
        const float inv = 1.0f / (l_run * a0 + l1 * a1);
        const float w0 = a0 * inv, w1 = a1 * inv;
        // tile-packed bf16 (flat rows): the out-projection GEMM's A image, [row/128][d/8][row%128][8]
        const size_t grow = (size_t)b * T + grow_i;
        const int nch = heads * DCH;
#pragma unroll
        for (int i = 0; i < DH; i += 8)
        {
            float r[8];
#pragma unroll
            for (int e = 0; e < 8; ++e)
                r[e] = o[i + e] * w0 + mg[(2 + i + e) * BQ + row] * w1;
            uint4 pk;
            pk.x = pack_bf16x2(r[0], r[1]);
            pk.y = pack_bf16x2(r[2], r[3]);
            pk.z = pack_bf16x2(r[4], r[5]);
            pk.w = pack_bf16x2(r[6], r[7]);
            Opk[((grow >> 7) * nch + h * DCH + i / 8) * 128 + (grow & 127)] = pk;
        }
    }
    __syncthreads();
    if (warp == 0)
    {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_slot), "r"(256u) : "memory");
    }
}

template <int DH> static int smem_bytes() { return (DH / 8) * (BQ * 16) * 5 + 16 * (BQ * 16); }

} // namespace atc

int attention_tc_init_device()
{
    DMX_CUDA(cudaFuncSetAttribute(atc::attention_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  atc::smem_bytes<64>()));
    DMX_CUDA(cudaFuncSetAttribute(atc::attention_tc_kernel<48>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  atc::smem_bytes<48>()));
    return 0;
}

int launch_attention_tc(const void *Qp, int ldq, int q_col, const void *Kp, int ldk, int k_col, const void *Vp, int ldv,
                        int v_col, void *Opk, int B, int T, int Tpad, int S, int Spad, int heads, int dh, cudaStream_t st)
{
    if ((ldq | ldk | ldv | q_col | k_col | v_col) % 8 || (Tpad | Spad) % 128 || Tpad < T || Spad < S)
    {
        set_error("attention_tc: packed operands need 8-column alignment and 128-row padded items");
        return -1;
    }
    const float scale = 1.4426950408889634f / sqrtf((float)dh); // log2(e) folded in: the kernel works with exp2
    dim3 grid((T + atc::BQ - 1) / atc::BQ, heads, B);
    const __nv_bfloat16 *q = static_cast<const __nv_bfloat16 *>(Qp), *k = static_cast<const __nv_bfloat16 *>(Kp),
                        *v = static_cast<const __nv_bfloat16 *>(Vp);
    if (dh == 64)
        atc::attention_tc_kernel<64><<<grid, atc::THREADS, atc::smem_bytes<64>(), st>>>(
            q, ldq / 8, q_col / 8, k, ldk / 8, k_col / 8, v, ldv / 8, v_col / 8, (uint4 *)Opk, T, Tpad, S, Spad, heads, scale);
    else if (dh == 48)
        atc::attention_tc_kernel<48><<<grid, atc::THREADS, atc::smem_bytes<48>(), st>>>(
            q, ldq / 8, q_col / 8, k, ldk / 8, k_col / 8, v, ldv / 8, v_col / 8, (uint4 *)Opk, T, Tpad, S, Spad, heads, scale);
    else
    {
        set_error("attention_tc: unsupported head dim");
        return -1;
    }
    DMX_LAUNCH_CHECK();
    return 0;
}

} // namespace dmx
