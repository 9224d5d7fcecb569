// Kernels specific to hdemucs_mmi (v3): GroupNorm with G groups, the 2-layer bidirectional LSTM recurrence
// and LocalState attention.  All tensors channels-last fp32: [B][rows][C].
#include "kernels.cuh"

#include <cstdlib>

#include <cooperative_groups.h>

namespace dmx
{

// ------------------------------------------------------------------------------------------------
// GroupNorm(G) statistics: (sum, sumsq) per (batch item, group) over (rows, C/G) jointly
// (reference: groupnorm::generalized_group_norm, src/layers.hpp:130-170 -- statistics span dim 0 too,
// which in our layout is simply all `rows` of the item).  grid (chunks, G, B); stats [B][G][2] zeroed.
__global__ void __launch_bounds__(256) group_stats_kernel(const float *__restrict__ x, int rows, int C, int G,
                                                          double *__restrict__ stats)
{
    const int g = blockIdx.y, b = blockIdx.z;
    const int cg = C / G;
    const float *xb = x + (size_t)b * rows * C + (size_t)g * cg;
    const size_t total = (size_t)rows * cg;
    float s = 0.f, ss = 0.f;
    double ds = 0.0, dss = 0.0;
    int cnt = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x)
    {
        const size_t r = i / cg;
        const int c = (int)(i - r * cg);
        const float v = __ldg(&xb[r * C + c]);
        s += v;
        ss += v * v;
        if (++cnt == 64)
        {
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
        double a = 0, c2 = 0;
        for (int i = 0; i < 8; ++i)
        {
            a += red[0][i];
            c2 += red[1][i];
        }
        atomicAdd(&stats[2 * ((size_t)b * G + g) + 0], a);
        atomicAdd(&stats[2 * ((size_t)b * G + g) + 1], c2);
    }
}

// Apply: y = act(GN(x)) with optional GLU over channels (first half * sigmoid(second half), src/layers.cpp:96-119)
// and an optional row crop: output row r reads input row r + row0.  x: [B][rows_in][C]; y: [B][rows_out][Cy],
// Cy = C (or C/2 with GLU).
__global__ void __launch_bounds__(256) group_norm_g_apply_kernel(const float *__restrict__ x, float *__restrict__ y,
                                                                 int rows_in, int rows_out, int row0, int C, int G,
                                                                 const double *__restrict__ stats,
                                                                 const float *__restrict__ w, const float *__restrict__ bv,
                                                                 int act, int glu, size_t total)
{
    const int Cy = glu ? C / 2 : C;
    const int cg = C / G;
    const double count = (double)rows_in * cg;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x)
    {
        const size_t row = i / Cy;
        const int c = (int)(i - row * Cy);
        const int b = (int)(row / rows_out);
        const int r = (int)(row - (size_t)b * rows_out);
        const float *xr = x + ((size_t)b * rows_in + r + row0) * C;
        float mean, rstd;
        stats_to_mean_rstd(stats + 2 * ((size_t)b * G + c / cg), count, 1e-5f, mean, rstd);
        float v = (__ldg(&xr[c]) - mean) * rstd * __ldg(&w[c]) + __ldg(&bv[c]);
        if (act == 1)
            v = gelu_erf(v);
        if (glu)
        {
            const int c2 = c + Cy;
            float m2, r2;
            stats_to_mean_rstd(stats + 2 * ((size_t)b * G + c2 / cg), count, 1e-5f, m2, r2);
            float gte = (__ldg(&xr[c2]) - m2) * r2 * __ldg(&w[c2]) + __ldg(&bv[c2]);
            if (act == 1)
                gte = gelu_erf(gte);
            v = v * sigmoid_ref(gte);
        }
        y[i] = v;
    }
}

int launch_group_norm_g(const float *x, float *y, int B, int rows_in, int rows_out, int row0, int C, int G,
                        double *stats_zeroed, const float *w, const float *b, int act, int glu, cudaStream_t st)
{
    if (C % G || (glu && (C % 2)))
    {
        set_error("group_norm_g: bad channel/group counts");
        return -1;
    }
    const size_t per = (size_t)rows_in * (C / G);
    int chunks = (int)((per + 256 * 8 - 1) / (256 * 8));
    if (chunks > 64)
        chunks = 64;
    if (chunks < 1)
        chunks = 1;
    group_stats_kernel<<<dim3(chunks, G, B), 256, 0, st>>>(x, rows_in, C, G, stats_zeroed);
    DMX_LAUNCH_CHECK();
    const size_t total = (size_t)B * rows_out * (glu ? C / 2 : C);
    size_t blocks = (total + 255) / 256;
    if (blocks > 148 * 8)
        blocks = 148 * 8;
    group_norm_g_apply_kernel<<<(unsigned)blocks, 256, 0, st>>>(x, y, rows_in, rows_out, row0, C, G, stats_zeroed, w, b, act,
                                                                glu, total);
    DMX_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// LSTM recurrence (reference: demucscpp_v3::lstm_forward, src/lstm.cpp:68-147).  The input projections of
// all steps (W_ih x_t + b_ih + b_hh) are hoisted into one GEMM; this kernel runs the sequential part:
//     gates = pre[t] + W_hh h ;  i,f,o = sigmoid, g = tanh ;  c = f*c + i*g ;  h = o*tanh(c)
// One CTA per (batch item, direction), one thread per hidden unit (its four gate rows).  W_hh is stored
// transposed [k][4h] so that the h threads read consecutive weights; h lives in shared memory.
//   pre : [B][T][2][4h]  (direction-major inside a step row: forward gates, then backward gates)
//   out : [B][T][2h]     (forward hidden | backward hidden), the layout the next layer / Linear consumes
__global__ void lstm_recurrent_kernel(const float *__restrict__ pre, const float *__restrict__ whh_t_fwd,
                                      const float *__restrict__ whh_t_bwd, float *__restrict__ out, int T, int H)
{
    extern __shared__ float hs[]; // [2][H] double-buffered hidden state
    const int dir = blockIdx.x, b = blockIdx.y;
    const int u = threadIdx.x;
    const float *W = dir == 0 ? whh_t_fwd : whh_t_bwd;
    const int G4 = 4 * H;
    float c = 0.f;
    if (u < H)
    {
        hs[u] = 0.f;
        hs[H + u] = 0.f;
    }
    __syncthreads();
    int cur = 0;
    for (int step = 0; step < T; ++step)
    {
        const int t = dir == 0 ? step : T - 1 - step;
        if (u < H)
        {
            const float *pr = pre + (((size_t)b * T + t) * 2 + dir) * G4;
            float gi = __ldg(&pr[u]), gf = __ldg(&pr[H + u]), gg = __ldg(&pr[2 * H + u]), go = __ldg(&pr[3 * H + u]);
            const float *hp = hs + cur * H;
#pragma unroll 4
            for (int k = 0; k < H; ++k)
            {
                const float hk = hp[k];
                const float *wr = W + (size_t)k * G4;
                gi = fmaf(__ldg(&wr[u]), hk, gi);
                gf = fmaf(__ldg(&wr[H + u]), hk, gf);
                gg = fmaf(__ldg(&wr[2 * H + u]), hk, gg);
                go = fmaf(__ldg(&wr[3 * H + u]), hk, go);
            }
            const float i_t = 1.0f / (1.0f + expf(-gi));
            const float f_t = 1.0f / (1.0f + expf(-gf));
            const float g_t = tanhf(gg);
            const float o_t = 1.0f / (1.0f + expf(-go));
            c = f_t * c + i_t * g_t;
            const float h = o_t * tanhf(c);
            hs[(cur ^ 1) * H + u] = h;
            out[((size_t)b * T + t) * (2 * H) + dir * H + u] = h;
        }
        __syncthreads();
        cur ^= 1;
    }
}

// ------------------------------------------------------------------------------------------------
// Cluster-persistent LSTM recurrence.  One thread-block CLUSTER per direction runs the chains of ALL batch
// items: CTA c of the cluster owns 24 hidden units (their 4 x 24 = 96 gate rows of W_hh) and keeps those
// weights in REGISTERS for the whole sequence (thread = (row, quarter of the K range): H/4 weights each).
// Per step every CTA computes its rows for all batch items from the full hidden state in shared memory,
// the 24 x B new hidden values are written straight into every CTA's shared memory over DSMEM, and one
// cluster barrier ends the step.  Cluster size = H / 24: 8 CTAs for H = 192, 16 CTAs for H = 384.
//   whh : W_hh as stored by PyTorch, [4H][H] row-major, for this direction
constexpr int LSTM_UNITS = 24;   // hidden units per CTA
constexpr int LSTM_ROWS = 96;    // gate rows per CTA
constexpr int LSTM_THREADS = 384;
constexpr int LSTM_BMAX = 8;

template <int H>
__global__ void __launch_bounds__(LSTM_THREADS, 1)
    lstm_cluster_kernel(const float *__restrict__ pre, const float *__restrict__ whh_fwd,
                        const float *__restrict__ whh_bwd, float *__restrict__ out, int Btot, int T, int chunk)
{
    constexpr int NC = H / LSTM_UNITS; // cluster size
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    const int crank = (int)cluster.block_rank();
    // cluster ci = (direction, chunk of `chunk` batch items): all chunks of the batch run at the same time on their
    // own clusters instead of one after the other
    const int ci = blockIdx.x / NC;
    const int dir = ci & 1;
    const int b0 = (ci >> 1) * chunk;
    const int B = (Btot - b0) < chunk ? (Btot - b0) : chunk;
    pre += (size_t)b0 * T * 2 * 4 * H;
    out += (size_t)b0 * T * 2 * H;
    const float *W = dir == 0 ? whh_fwd : whh_bwd;

    __shared__ __align__(16) float hbuf[2][LSTM_BMAX][H];       // full hidden state of every batch item, double buffered
    __shared__ float gates[LSTM_BMAX][LSTM_ROWS];               // this CTA's gate pre-activations

    const int tid = threadIdx.x;
    // Mat-vec mapping: warp wi owns 8 of the CTA's 96 gate rows, lane l the K indices l, l+32, ... (KJ of them).
    // Every hidden value is read from shared memory once per warp (conflict free) and reused for 8 rows; the
    // 32 partial sums (8 rows x 4 batch items) of a half-batch are combined with a transposing butterfly
    // (31 shuffles) that leaves lane v = row_local*4 + b_local with the finished dot product.
    constexpr int KJ = H / 32;
    const int wi = tid >> 5, lane = tid & 31;
    float w[8][KJ];
#pragma unroll
    for (int r = 0; r < 8; ++r)
    {
        const int row = wi * 8 + r;
        const int grow = (row / LSTM_UNITS) * H + crank * LSTM_UNITS + (row % LSTM_UNITS);
#pragma unroll
        for (int j = 0; j < KJ; ++j)
            w[r][j] = __ldg(&W[(size_t)grow * H + lane + 32 * j]);
    }
    // this lane finishes (row = wi*8 + lane/4, batch = half*4 + lane%4)
    const int frow = wi * 8 + (lane >> 2);
    const int fgrow = (frow / LSTM_UNITS) * H + crank * LSTM_UNITS + (frow % LSTM_UNITS);
    for (int i = tid; i < 2 * LSTM_BMAX * H; i += LSTM_THREADS)
        (&hbuf[0][0][0])[i] = 0.f;
    // cell state: thread (unit, b) for tid < 24 * B
    const int ub = tid / LSTM_UNITS, uu = tid % LSTM_UNITS;
    float c = 0.f;
    cluster.sync();

    const int G4 = 4 * H;
    int cur = 0;
    for (int step = 0; step < T; ++step)
    {
        const int t = dir == 0 ? step : T - 1 - step;
        // input-projection terms of this step: issued first so their L2 latency hides behind the mat-vec
        float prev[2];
#pragma unroll
        for (int hf = 0; hf < 2; ++hf)
        {
            const int b = hf * 4 + (lane & 3);
            prev[hf] = b < B ? __ldg(&pre[(((size_t)b * T + t) * 2 + dir) * G4 + fgrow]) : 0.f;
        }
#pragma unroll
        for (int hf = 0; hf < 2; ++hf)
        {
            if (hf * 4 >= B)
                break;
            float v[32];
#pragma unroll
            for (int i = 0; i < 32; ++i)
                v[i] = 0.f;
#pragma unroll
            for (int j = 0; j < KJ; ++j)
            {
                float hv[4];
#pragma unroll
                for (int bl = 0; bl < 4; ++bl)
                    hv[bl] = hbuf[cur][hf * 4 + bl][lane + 32 * j];
#pragma unroll
                for (int r = 0; r < 8; ++r)
#pragma unroll
                    for (int bl = 0; bl < 4; ++bl)
                        v[r * 4 + bl] = fmaf(w[r][j], hv[bl], v[r * 4 + bl]);
            }
            // transposing butterfly: after the five rounds lane l holds the full sum of value index l
#pragma unroll
            for (int off = 16, n = 32; off >= 1; off >>= 1, n >>= 1)
            {
                const bool up = (lane & off) != 0;
#pragma unroll
                for (int i = 0; i < 16; ++i)
                {
                    if (i < n / 2)
                    {
                        const float keep = up ? v[i + n / 2] : v[i];
                        const float send = up ? v[i] : v[i + n / 2];
                        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
                    }
                }
            }
            const int b = hf * 4 + (lane & 3);
            if (b < B)
                gates[b][frow] = v[0] + prev[hf];
        }
        __syncthreads();
        {
            // threads (unit, b) for tid < 24*B finish the cell; whole warps stay converged for the shuffles
            const bool live = tid < LSTM_UNITS * B;
            float h = 0.f;
            if (live)
            {
                const float gi = gates[ub][uu], gf = gates[ub][LSTM_UNITS + uu], gg = gates[ub][2 * LSTM_UNITS + uu],
                            go = gates[ub][3 * LSTM_UNITS + uu];
                const float i_t = 1.0f / (1.0f + expf(-gi));
                const float f_t = 1.0f / (1.0f + expf(-gf));
                const float g_t = tanhf(gg);
                const float o_t = 1.0f / (1.0f + expf(-go));
                c = f_t * c + i_t * g_t;
                h = o_t * tanhf(c);
                out[((size_t)ub * T + t) * (2 * H) + dir * H + crank * LSTM_UNITS + uu] = h;
            }
            // four consecutive units sit in four consecutive lanes (24 % 4 == 0): gather them so the broadcast
            // into every CTA's shared memory (DSMEM) is one 16-byte store per destination instead of four
            const float h1 = __shfl_down_sync(0xffffffffu, h, 1);
            const float h2 = __shfl_down_sync(0xffffffffu, h, 2);
            const float h3 = __shfl_down_sync(0xffffffffu, h, 3);
            if (live && (uu & 3) == 0)
            {
                float *local = &hbuf[cur ^ 1][ub][crank * LSTM_UNITS + uu];
                const float4 hv = make_float4(h, h1, h2, h3);
#pragma unroll
                for (int r = 0; r < NC; ++r)
                    *reinterpret_cast<float4 *>(cluster.map_shared_rank(local, r)) = hv;
            }
        }
        cluster.sync(); // all writes of step's h landed everywhere; nobody still reads hbuf[cur]
        cur ^= 1;
    }
}

template <int H> static int lstm_cluster_launch(const float *pre, const float *whh_fwd, const float *whh_bwd, float *out,
                                                int B, int T, int chunk, cudaStream_t st)
{
    constexpr int NC = H / LSTM_UNITS;
    static bool attr_done = false; // per process; the attribute is per function, harmless to repeat per device
    if (NC > 8 && !attr_done)
    {
        DMX_CUDA(cudaFuncSetAttribute(lstm_cluster_kernel<H>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    }
    attr_done = attr_done || NC <= 8;
    if (NC > 8)
        DMX_CUDA(cudaFuncSetAttribute(lstm_cluster_kernel<H>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * NC * ((B + chunk - 1) / chunk));
    cfg.blockDim = dim3(LSTM_THREADS);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = NC;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    DMX_CUDA(cudaLaunchKernelEx(&cfg, lstm_cluster_kernel<H>, pre, whh_fwd, whh_bwd, out, B, T, chunk));
    return 0;
}

// whh_*: [4H][H] row-major (PyTorch layout); whh_t_*: the same transposed [H][4H] for the fallback kernel.
int launch_lstm_recurrent(const float *pre, const float *whh_fwd, const float *whh_bwd, const float *whh_t_fwd,
                          const float *whh_t_bwd, float *out, int B, int T, int H, cudaStream_t st)
{
    if (H == 192 || H == 384)
    {
        // a cluster keeps the hidden state of up to LSTM_BMAX items on chip; the recurrence is a latency chain (one
        // cluster barrier per step) that occupies 8 / 16 SMs per cluster, so the batch is cut into chunks of 4 items that
        // all run AT THE SAME TIME on their own clusters (16 items: 8 clusters) rather than one after the other
        static const int chunk_env = getenv("DMX_LSTM_CHUNK") ? atoi(getenv("DMX_LSTM_CHUNK")) : 0;
        int chunk = (chunk_env >= 1 && chunk_env <= LSTM_BMAX) ? chunk_env : 4;
        if (B < chunk)
            chunk = B;
        return H == 192 ? lstm_cluster_launch<192>(pre, whh_fwd, whh_bwd, out, B, T, chunk, st)
                        : lstm_cluster_launch<384>(pre, whh_fwd, whh_bwd, out, B, T, chunk, st);
    }
    if (H > 1024)
    {
        set_error("lstm: hidden size > 1024");
        return -1;
    }
    const int threads = (H + 31) / 32 * 32;
    lstm_recurrent_kernel<<<dim3(2, B), threads, 2 * H * sizeof(float), st>>>(pre, whh_t_fwd, whh_t_bwd, out, T, H);
    DMX_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// LocalState attention (reference: demucscpp_v3::local_attention, src/layers.cpp:533-721), 4 heads, 4 decay
// channels.  qkcd: [B][T][3*C + 16] = (query | key | content | query_decay) from one fused 1x1-conv GEMM.
//   dots[t, s] = <K[t], Q[s]> / sqrt(dh) - sum_n (n+1)*|t-s|/2 * 0.5*sigmoid(qd[s][h*4+n]) ; dots[s, s] = -100
//   out[s] = sum_t softmax_t(dots[:, s]) * content[t]
// One warp per query s; lanes split the head's channels; streaming softmax over the keys.
__global__ void __launch_bounds__(256) local_attention_kernel(const float *__restrict__ qkcd, float *__restrict__ out,
                                                              int T, int C)
{
    const int h = blockIdx.y, b = blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int s = blockIdx.x * 8 + warp;
    if (s >= T)
        return;
    const int dh = C / 4;
    const int ld = 3 * C + 16;
    const float *base = qkcd + (size_t)b * T * ld;
    const float *qs = base + (size_t)s * ld + h * dh;
    float q[3], acc[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
        const int c = lane + 32 * i;
        q[i] = c < dh ? __ldg(&qs[c]) : 0.f;
        acc[i] = 0.f;
    }
    float decay = 0.f; // sum_n (n+1)/2 * 0.5*sigmoid(qd[n]) : the per-|t-s| slope for this query
#pragma unroll
    for (int n = 0; n < 4; ++n)
    {
        const float qd = __ldg(&base[(size_t)s * ld + 3 * C + h * 4 + n]);
        decay += (float)(n + 1) * 0.5f * (0.5f / (1.0f + expf(-qd)));
    }
    const float inv_sqrt = 1.0f / sqrtf((float)dh);
    float m_run = -INFINITY, l_run = 0.f;
    for (int t = 0; t < T; ++t)
    {
        const float *kt = base + (size_t)t * ld + C + h * dh;
        float d = 0.f;
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
            const int c = lane + 32 * i;
            if (c < dh)
                d = fmaf(q[i], __ldg(&kt[c]), d);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            d += __shfl_xor_sync(0xffffffffu, d, o);
        float sc = d * inv_sqrt - decay * fabsf((float)(t - s));
        if (t == s)
            sc = -100.0f;
        const float m_new = fmaxf(m_run, sc);
        const float corr = (m_run == -INFINITY) ? 0.f : expf(m_run - m_new);
        const float p = expf(sc - m_new);
        l_run = l_run * corr + p;
        m_run = m_new;
        const float *ct = base + (size_t)t * ld + 2 * C + h * dh;
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
            const int c = lane + 32 * i;
            if (c < dh)
                acc[i] = acc[i] * corr + p * __ldg(&ct[c]);
        }
    }
    const float inv = 1.0f / l_run;
    float *o = out + ((size_t)b * T + s) * C + h * dh;
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
        const int c = lane + 32 * i;
        if (c < dh)
            o[c] = acc[i] * inv;
    }
}

int launch_local_attention(const float *qkcd, float *out, int B, int T, int C, cudaStream_t st)
{
    if (C % 4 || C / 4 > 96)
    {
        set_error("local_attention: unsupported channel count");
        return -1;
    }
    local_attention_kernel<<<dim3((T + 7) / 8, 4, B), 256, 0, st>>>(qkcd, out, T, C);
    DMX_LAUNCH_CHECK();
    return 0;
}

} // namespace dmx
