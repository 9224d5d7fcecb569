// Track-level orchestration on device: the reference's demucs_inference / shift_inference /
// split_inference / segment_inference (src/model_apply.cpp:60-288) with segments batched.
//   normalise by the mono mean / unbiased std           (:72-82)
//   zero-pad 0.5 s each side, shift by `offset`           (:103-121)
//   7.8 s segments at stride 0.75*7.8 s, short last chunk centre-padded with zeros (:21-43,:200-204,:260-261)
//   triangular-weighted overlap-add, divide by the weight sum (:171-179,:207-246)
//   trim to the original length, de-normalise            (:129-135,:88)
// The overlap-add is a gather (each output sample is covered by at most two segments), so no atomics
// and no zero-initialised accumulator are needed; terms are added in increasing segment order like
// the reference's loop.
#include "model.cuh"

#include <condition_variable>
#include <cstring>
#include <deque>
#include <thread>
#include <utility>

namespace dmx
{

// sum / sumsq of the mono signal (mean over the two channels), n frames
__global__ void __launch_bounds__(256) mono_stats_kernel(const float2 *__restrict__ a, int64_t n, double *stats)
{
    double ds = 0, dss = 0;
    float s = 0.f, ss = 0.f;
    int cnt = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    {
        const float2 v = __ldg(&a[i]);
        const float mono = (v.x + v.y) * 0.5f;
        s += mono;
        ss += mono * mono;
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
        double x = 0, y = 0;
        for (int i = 0; i < 8; ++i)
        {
            x += red[0][i];
            y += red[1][i];
        }
        atomicAdd(&stats[0], x);
        atomicAdd(&stats[1], y);
    }
}

__device__ __forceinline__ void track_mean_std(const double *stats, int64_t n, float &mean, float &stdv)
{
    const double s = stats[0], ss = stats[1];
    const double m = s / (double)n;
    double var = (ss - s * m) / ((double)n - 1.0);
    if (var < 0)
        var = 0;
    mean = (float)m;
    stdv = (float)sqrt(var);
}

// segs[j][i], j = blockIdx.y: normalised, shifted, centre zero padded segment k = k0 + j*kstride
__global__ void __launch_bounds__(256) extract_segments_kernel(const float2 *__restrict__ audio, int64_t L, int offset,
                                                               int64_t Ls, const double *__restrict__ stats, int k0,
                                                               int kstride, float2 *__restrict__ segs)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= SEG)
        return;
    const int k = k0 + blockIdx.y * kstride;
    float mean, stdv;
    track_mean_std(stats, L, mean, stdv);
    const int64_t off = (int64_t)k * SEG_STRIDE;
    const int64_t nk = (Ls - off) < SEG ? (Ls - off) : SEG;
    const int left = (int)((SEG - nk) / 2);
    float2 v = make_float2(0.f, 0.f);
    const int64_t j = i - left;
    if (j >= 0 && j < nk)
    {
        const int64_t src = off + j + offset - MAX_SHIFT; // index into the un-padded track
        if (src >= 0 && src < L)
        {
            const float2 a = __ldg(&audio[src]);
            v.x = (a.x - mean) / stdv;
            v.y = (a.y - mean) / stdv;
        }
    }
    segs[(size_t)blockIdx.y * SEG + i] = v;
}

// out (S, 2, L) Eigen column-major: index s + S*(ch + 2*t)
// y holds segment k in slot (k % world)*cnt + k / world (the layout a gather of round-robin shards gives;
// world = 1: slot = k).
__global__ void __launch_bounds__(256) overlap_add_kernel(const float2 *__restrict__ y, // [slots][S][SEG]
                                                          int S, int nseg, int world, int cnt, int64_t L, int offset,
                                                          int64_t Ls, const double *__restrict__ stats,
                                                          float *__restrict__ out, int64_t t_begin, int64_t t_end)
{
    const int64_t t0 = t_begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; // output samples [t_begin, t_end)
    if (t0 >= t_end)
        return;
    float mean, stdv;
    track_mean_std(stats, L, mean, stdv);
    const int64_t t = t0 + MAX_SHIFT - offset; // position in the shifted track
    const int khi = (int)(t / SEG_STRIDE);
    const float h = (float)(SEG / 2);
    for (int s = 0; s < S; ++s)
    {
        float2 num = make_float2(0.f, 0.f);
        float den = 0.f;
        for (int k = khi - 1; k <= khi; ++k)
        {
            if (k < 0 || k >= nseg)
                continue;
            const int64_t off = (int64_t)k * SEG_STRIDE;
            const int64_t nk = (Ls - off) < SEG ? (Ls - off) : SEG;
            const int64_t i = t - off;
            if (i < 0 || i >= nk)
                continue;
            const int left = (int)((SEG - nk) / 2);
            const int ii = (int)i;
            const float w = (float)(ii + 1 < SEG - ii ? ii + 1 : SEG - ii) / h;
            const size_t slot = (size_t)(k % world) * cnt + (size_t)(k / world);
            const float2 v = __ldg(&y[(slot * S + s) * SEG + left + ii]);
            num.x += w * v.x;
            num.y += w * v.y;
            den += w;
        }
        out[s + (size_t)S * (0 + 2 * t0)] = (num.x / den) * stdv + mean;
        out[s + (size_t)S * (1 + 2 * t0)] = (num.y / den) * stdv + mean;
    }
}

static int ensure(float **p, size_t *cap, size_t n)
{
    if (*cap >= n)
        return 0;
    if (*p)
        DMX_CUDA(cudaFree(*p));
    *p = nullptr;
    *cap = 0;
    DMX_CUDA(cudaMalloc(reinterpret_cast<void **>(p), n * sizeof(float)));
    *cap = n;
    return 0;
}

// The forward as the track-level entry points run it: stage tensors that only dmx_debug_tap would read (the fp32 outputs of
// the decoders' transposed convolutions, whose bf16 images are what the next layer consumes) are not written.
//
// The ~700 launches of one forward are short at small batches (4 segments per worker on 8 GPUs: ~10 us each), so the forward
// of a given (input, output, batch) is captured into a CUDA graph on its second call and replayed from then on: the
// buffers are the model's own, the fork / join of the time branch on stream2 becomes graph edges.  The first call runs
// eagerly (one-time attribute setup and lazily built tables happen there); DMX_NO_GRAPH=1, profiling and the per-layer
// taps keep the eager path.
static int track_forward(dmx_model *m, const float *d_mix, int B, float *d_out)
{
    static const bool use_graph = getenv("DMX_NO_GRAPH") == nullptr;
    struct Stages
    {
        dmx_model *m;
        explicit Stages(dmx_model *m_) : m(m_) { m->keep_stages = false; }
        ~Stages() { m->keep_stages = true; }
    } stages(m);
    if (!use_graph || m->prof.on || m->layer_taps)
        return model_forward(m, d_mix, B, d_out);
    dmx_model::FwdGraph *g = nullptr;
    for (auto &e : m->graphs)
        if (e.mix == d_mix && e.out == d_out && e.B == B)
            g = &e;
    if (!g)
    {
        if (m->graphs.size() >= 64)
        {
            // caller buffers keep changing: drop the oldest key
            if (m->graphs.front().exec)
                cudaGraphExecDestroy(m->graphs.front().exec);
            m->graphs.erase(m->graphs.begin());
        }
        m->graphs.emplace_back();
        g = &m->graphs.back();
        g->mix = d_mix;
        g->out = d_out;
        g->B = B;
    }
    if (g->exec)
    {
        DMX_CUDA(cudaGraphLaunch(g->exec, m->stream));
        m->launches += g->launches;
        m->lastB = B;
        return 0;
    }
    if (g->calls < 0 || g->calls++ == 0)
        return model_forward(m, d_mix, B, d_out);
    const int64_t before = m->launches;
    if (cudaStreamBeginCapture(m->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess)
    {
        cudaGetLastError();
        g->calls = -1;
        return model_forward(m, d_mix, B, d_out);
    }
    const int rc = model_forward(m, d_mix, B, d_out);
    cudaGraph_t graph = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(m->stream, &graph);
    if (rc || ce != cudaSuccess || !graph || cudaGraphInstantiate(&g->exec, graph, 0) != cudaSuccess)
    {
        // not capturable on this build / driver: stay eager for this key
        cudaGetLastError();
        if (graph)
            cudaGraphDestroy(graph);
        g->exec = nullptr;
        g->calls = -1;
        m->launches = before;
        return model_forward(m, d_mix, B, d_out);
    }
    cudaGraphDestroy(graph);
    g->launches = m->launches - before;
    DMX_CUDA(cudaGraphLaunch(g->exec, m->stream));
    return 0;
}

int num_segments(int64_t L, int shift)
{
    const int64_t Ls = L + MAX_SHIFT - shift;
    return (int)((Ls + SEG_STRIDE - 1) / SEG_STRIDE);
}

// Messages of the reference's per-segment progress callbacks (src/model_inference.cpp:156-344), fired
// after each batch completes (a batched device forward has no truthful mid-segment progress).
static const char *const kStageMsgs[26] = {
    "Time encoder 0", "Freq encoder 0", "Time encoder 1", "Freq encoder 1", "Time encoder 2", "Freq encoder 2",
    "Time encoder 3", "Freq encoder 3", "Freq (crosstransformer): layer 0", "Time (crosstransformer): layer 0",
    "Freq (crosstransformer): layer 1", "Time (crosstransformer): layer 1", "Freq (crosstransformer): layer 2",
    "Time (crosstransformer): layer 2", "Freq (crosstransformer): layer 3", "Time (crosstransformer): layer 3",
    "Freq (crosstransformer): layer 4", "Time (crosstransformer): layer 4", "Freq: decoder 0", "Time: decoder 0",
    "Freq: decoder 1", "Time: decoder 1", "Freq: decoder 2", "Time: decoder 2", "Freq: decoder 3", "Time: decoder 3"};

// v3: 22 callbacks per segment (src/model_inference.cpp:584-734)
static const char *const kStageMsgsV3[22] = {
    "Time encoder 0", "Freq encoder 0", "Time encoder 1", "Freq encoder 1", "Time encoder 2", "Freq encoder 2",
    "Time encoder 3", "Freq encoder 3", "Time encoder 4", "Freq encoder 4", "Shared encoder 5", "Shared decoder 0",
    "Freq decoder 1", "Time decoder 1", "Freq decoder 2", "Time decoder 2", "Freq decoder 3", "Time decoder 3",
    "Freq decoder 4", "Time decoder 4", "Freq decoder 5", "Time decoder 5"};

static int track_stats(dmx_model *m, const float *d_audio, int64_t L)
{
    cudaStream_t st = m->stream;
    DMX_CUDA(cudaMemsetAsync(m->track_stats, 0, 2 * sizeof(double), st));
    {
        ProfScope ps(m, PC_TRACK, 0.0, (double)L * 8.0);
        mono_stats_kernel<<<592, 256, 0, st>>>((const float2 *)d_audio, L, m->track_stats);
    }
    DMX_LAUNCH_CHECK();
    return 0;
}

static int check_track_args(int64_t L, int shift, int rank, int world)
{
    if (L < 2)
    {
        set_error("separate: track too short");
        return -1;
    }
    if (shift < 0 || shift >= MAX_SHIFT)
    {
        set_error("separate: shift_offset must be in [0, 22050)");
        return -1;
    }
    if (world < 1 || rank < 0 || rank >= world)
    {
        set_error("separate: bad rank/world");
        return -1;
    }
    return 0;
}

// Forward of this rank's share of the track's segments: k = rank, rank + world, ...  Results go to
// d_y[j] (j = k / world), each [S][SEG][2].  No inter-rank communication here: segments are independent
// given the track statistics, which every rank computes from the full track.
int track_forward_shard(dmx_model *m, const float *d_audio, int64_t L, int shift, int rank, int world, float *d_y,
                        dmx_progress_fn cb, void *user)
{
    if (check_track_args(L, shift, rank, world))
        return -1;
    cudaStream_t st = m->stream;
    const int S = m->S;
    const int64_t Ls = L + MAX_SHIFT - shift;
    const int nseg = num_segments(L, shift);
    DMX_CUDA(cudaEventRecord(m->ev0, st));
    if (track_stats(m, d_audio, L))
        return -1;
    const int nloc = rank < nseg ? (nseg - rank + world - 1) / world : 0;
    for (int j0 = 0; j0 < nloc; j0 += m->maxB)
    {
        const int nb = (nloc - j0) < m->maxB ? (nloc - j0) : m->maxB;
        {
            ProfScope ps(m, PC_TRACK, 0.0, (double)nb * SEG * 16.0);
            extract_segments_kernel<<<dim3((SEG + 255) / 256, nb), 256, 0, st>>>(
                (const float2 *)d_audio, L, shift, Ls, m->track_stats, rank + j0 * world, world, (float2 *)m->mix);
        }
        DMX_LAUNCH_CHECK();
        if (track_forward(m, m->mix, nb, d_y + (size_t)j0 * S * SEG * 2))
            return -1;
        if (cb)
        {
            // keep the reference's per-segment message sequence with monotone progress values
            DMX_CUDA(cudaStreamSynchronize(st));
            const bool v3 = m->arch == DMX_ARCH_HDEMUCS_V3;
            const int nmsg = v3 ? 22 : 26;
            for (int j = j0; j < j0 + nb; ++j)
                for (int sidx = 0; sidx < nmsg; ++sidx)
                    cb(((float)j + (float)(sidx + 1) / (float)nmsg) / (float)nloc, v3 ? kStageMsgsV3[sidx] : kStageMsgs[sidx],
                       user);
        }
    }
    DMX_CUDA(cudaEventRecord(m->ev1, st));
    DMX_CUDA(cudaStreamSynchronize(st));
    return 0;
}

// Weighted overlap-add of all segments (gathered in round-robin shard order), trim, de-normalise.
int track_overlap_add(dmx_model *m, const float *d_audio, int64_t L, int shift, const float *d_y_all, int world,
                      float *d_out)
{
    if (check_track_args(L, shift, 0, world))
        return -1;
    cudaStream_t st = m->stream;
    const int S = m->S;
    const int64_t Ls = L + MAX_SHIFT - shift;
    const int nseg = num_segments(L, shift);
    const int cnt = (nseg + world - 1) / world;
    DMX_CUDA(cudaEventRecord(m->ev0, st));
    if (track_stats(m, d_audio, L))
        return -1;
    {
        ProfScope ps(m, PC_TRACK, 0.0, (double)L * S * 8.0 * 2.3);
        overlap_add_kernel<<<(unsigned)((L + 255) / 256), 256, 0, st>>>((const float2 *)d_y_all, S, nseg, world, cnt, L,
                                                                       shift, Ls, m->track_stats, d_out, 0, L);
    }
    DMX_LAUNCH_CHECK();
    DMX_CUDA(cudaEventRecord(m->ev1, st));
    DMX_CUDA(cudaStreamSynchronize(st));
    return 0;
}


// ------------------------------------------------------------------------------------------------
// Contiguous-block split of a track over `world` workers (GPUs): worker r runs a run of consecutive
// segments, overlap-adds and finalises the stretch of the track that starts at its first segment and
// ends where the next worker's first segment starts.  The only data that crosses workers is the last
// SEG - SEG_STRIDE frames of a worker's last segment (the part that overlaps the next worker's first
// segment) and the two sums of the track statistics.  Replaces the chunk fan-out / merge of
// cli-apps/threaded_inference.hpp:50-190 (which re-runs whole overlapping chunks on every thread).
int plan_block(int64_t L, int shift, int rank, int world, dmx_block_plan *pl)
{
    if (check_track_args(L, shift, rank, world) || !pl)
        return -1;
    const int64_t Ls = L + MAX_SHIFT - shift;
    const int nseg = num_segments(L, shift);
    const int base = nseg / world, rem = nseg % world;
    const int k0 = rank * base + (rank < rem ? rank : rem);
    const int k1 = k0 + base + (rank < rem ? 1 : 0);
    const int64_t lead = MAX_SHIFT - shift; // shifted-track position of track frame 0
    auto clampL = [&](int64_t t) { return t < 0 ? 0 : (t > L ? L : t); };
    pl->seg_begin = k0;
    pl->seg_end = k1;
    pl->n_segments = nseg;
    pl->tail_frames = (k1 > k0 && k0 > 0) ? SEG - SEG_STRIDE : 0;
    if (k1 == k0) // more workers than segments: the idle ones are the trailing ranks
    {
        pl->in_begin = pl->in_end = pl->out_begin = pl->out_end = L;
        return 0;
    }
    int64_t last_end = (int64_t)(k1 - 1) * SEG_STRIDE + SEG;
    if (last_end > Ls)
        last_end = Ls;
    pl->in_begin = clampL((int64_t)k0 * SEG_STRIDE - lead);
    pl->in_end = clampL(last_end - lead);
    pl->out_begin = pl->in_begin;
    pl->out_end = k1 >= nseg ? L : clampL((int64_t)k1 * SEG_STRIDE - lead);
    return 0;
}

// (sum, sumsq) of the mono signal of `n` frames, ADDED to d_sums[0..1] (device doubles the caller zeroed):
// the track statistics of demucs_inference (src/model_apply.cpp:72-80) are the sums over all workers' stretches.
int track_stats_partial(dmx_model *m, const float *d_audio, int64_t n, double *d_sums)
{
    if (n <= 0)
        return 0;
    {
        ProfScope ps(m, PC_TRACK, 0.0, (double)n * 8.0);
        mono_stats_kernel<<<592, 256, 0, m->stream>>>((const float2 *)d_audio, n, d_sums);
    }
    DMX_LAUNCH_CHECK();
    return 0;
}

// As extract_segments_kernel, but the audio pointer holds only track frames [w0, w1) and the statistics are given.
__global__ void __launch_bounds__(256) extract_block_kernel(const float2 *__restrict__ win, int64_t w0, int64_t w1, int64_t L,
                                                            int offset, int64_t Ls, const double *__restrict__ stats, int k0,
                                                            float2 *__restrict__ segs)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= SEG)
        return;
    const int k = k0 + blockIdx.y;
    float mean, stdv;
    track_mean_std(stats, L, mean, stdv);
    const int64_t off = (int64_t)k * SEG_STRIDE;
    const int64_t nk = (Ls - off) < SEG ? (Ls - off) : SEG;
    const int left = (int)((SEG - nk) / 2);
    float2 v = make_float2(0.f, 0.f);
    const int64_t j = i - left;
    if (j >= 0 && j < nk)
    {
        const int64_t src = off + j + offset - MAX_SHIFT;
        if (src >= 0 && src < L && src >= w0 && src < w1)
        {
            const float2 a = __ldg(&win[src - w0]);
            v.x = (a.x - mean) / stdv;
            v.y = (a.y - mean) / stdv;
        }
    }
    segs[(size_t)blockIdx.y * SEG + i] = v;
}

int track_forward_block(dmx_model *m, const float *d_win, int64_t L, int shift, const dmx_block_plan &pl, const double *d_sums,
                        float *d_y)
{
    if (check_track_args(L, shift, 0, 1))
        return -1;
    cudaStream_t st = m->stream;
    const int S = m->S;
    const int64_t Ls = L + MAX_SHIFT - shift;
    DMX_CUDA(cudaEventRecord(m->ev0, st));
    const int nloc = pl.seg_end - pl.seg_begin;
    for (int j0 = 0; j0 < nloc; j0 += m->maxB)
    {
        const int nb = (nloc - j0) < m->maxB ? (nloc - j0) : m->maxB;
        {
            ProfScope ps(m, PC_TRACK, 0.0, (double)nb * SEG * 16.0);
            extract_block_kernel<<<dim3((SEG + 255) / 256, nb), 256, 0, st>>>((const float2 *)d_win, pl.in_begin, pl.in_end, L,
                                                                              shift, Ls, d_sums, pl.seg_begin + j0,
                                                                              (float2 *)m->mix);
        }
        DMX_LAUNCH_CHECK();
        if (track_forward(m, m->mix, nb, d_y + (size_t)j0 * S * SEG * 2))
            return -1;
    }
    DMX_CUDA(cudaEventRecord(m->ev1, st));
    return 0;
}

// Overlap-add of one worker's block: y holds its segments [k0, k1) ([k1-k0][S][SEG]); tail ([S][SEG-SEG_STRIDE]) holds the
// end of segment k0-1 (the previous worker's last one).  Writes track frames [t_begin, t_end) of sources
// [s_begin, s_end) into out_block, whose frame 0 is track frame t_begin: index s + S*(ch + 2*(t - t_begin)).
__global__ void __launch_bounds__(256)
    overlap_add_block_kernel(const float2 *__restrict__ y, const float2 *__restrict__ tail, int S, int s_begin, int s_end,
                             int nseg, int k0, int k1, int64_t L, int offset, int64_t Ls, const double *__restrict__ stats,
                             float *__restrict__ out, int64_t t_begin, int64_t t_end)
{
    const int64_t t0 = t_begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t0 >= t_end)
        return;
    float mean, stdv;
    track_mean_std(stats, L, mean, stdv);
    const int64_t t = t0 + MAX_SHIFT - offset;
    const int khi = (int)(t / SEG_STRIDE);
    const float h = (float)(SEG / 2);
    constexpr int TAIL = SEG - SEG_STRIDE;
    for (int s = s_begin; s < s_end; ++s)
    {
        float2 num = make_float2(0.f, 0.f);
        float den = 0.f;
        for (int k = khi - 1; k <= khi; ++k)
        {
            if (k < 0 || k >= nseg || k >= k1 || k < k0 - 1)
                continue;
            const int64_t off = (int64_t)k * SEG_STRIDE;
            const int64_t nk = (Ls - off) < SEG ? (Ls - off) : SEG;
            const int64_t i = t - off;
            if (i < 0 || i >= nk)
                continue;
            const int left = (int)((SEG - nk) / 2);
            const int ii = (int)i;
            const float w = (float)(ii + 1 < SEG - ii ? ii + 1 : SEG - ii) / h;
            float2 v;
            if (k >= k0)
                v = __ldg(&y[((size_t)(k - k0) * S + s) * SEG + left + ii]);
            else
                v = __ldg(&tail[(size_t)s * TAIL + (ii - SEG_STRIDE)]); // a previous segment is always a full one
            num.x += w * v.x;
            num.y += w * v.y;
            den += w;
        }
        out[s + (size_t)S * (0 + 2 * (t0 - t_begin))] = (num.x / den) * stdv + mean;
        out[s + (size_t)S * (1 + 2 * (t0 - t_begin))] = (num.y / den) * stdv + mean;
    }
}

int track_overlap_add_block(dmx_model *m, int64_t L, int shift, const dmx_block_plan &pl, const double *d_sums,
                            const float *d_y, const float *d_tail, int s_begin, int s_end, float *d_out_block)
{
    if (check_track_args(L, shift, 0, 1))
        return -1;
    if (s_begin < 0 || s_end > m->S || s_begin >= s_end)
    {
        set_error("overlap_add_block: bad source range");
        return -1;
    }
    if (pl.tail_frames > 0 && !d_tail)
    {
        set_error("overlap_add_block: this block needs the previous block's tail");
        return -1;
    }
    const int64_t n = pl.out_end - pl.out_begin;
    if (n <= 0)
        return 0;
    cudaStream_t st = m->stream;
    const int64_t Ls = L + MAX_SHIFT - shift;
    {
        ProfScope ps(m, PC_TRACK, 0.0, (double)n * (s_end - s_begin) * 8.0 * 2.3);
        overlap_add_block_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(
            (const float2 *)d_y, (const float2 *)d_tail, m->S, s_begin, s_end, pl.n_segments, pl.seg_begin, pl.seg_end, L, shift,
            Ls, d_sums, d_out_block, pl.out_begin, pl.out_end);
    }
    DMX_LAUNCH_CHECK();
    return 0;
}

// The reference's numbered per-segment progress messages for all `nseg` segments, in order (used by the entry points
// whose workers run off the calling thread: the callback may only be invoked on the caller's thread).
void track_fire_progress(dmx_model *m, int nseg, dmx_progress_fn cb, void *user)
{
    if (!cb)
        return;
    const bool v3 = m->arch == DMX_ARCH_HDEMUCS_V3;
    const int nmsg = v3 ? 22 : 26;
    for (int j = 0; j < nseg; ++j)
        for (int sidx = 0; sidx < nmsg; ++sidx)
            cb(((float)j + (float)(sidx + 1) / (float)nmsg) / (float)nseg, v3 ? kStageMsgsV3[sidx] : kStageMsgs[sidx], user);
}

void track_fire_segment_progress(const dmx_model *m, float current, float segment, dmx_progress_fn cb, void *user)
{
    if (!cb)
        return;
    const bool v3 = m->arch == DMX_ARCH_HDEMUCS_V3;
    const int nmsg = v3 ? 22 : 26;
    for (int sidx = 0; sidx < nmsg; ++sidx)
        cb(current + segment * (float)(sidx + 1) / (float)nmsg, v3 ? kStageMsgsV3[sidx] : kStageMsgs[sidx], user);
}

int track_separate(dmx_model *m, const float *d_audio, int64_t L, int shift, float *d_out, dmx_progress_fn cb,
                   void *user)
{
    if (check_track_args(L, shift, 0, 1))
        return -1;
    const int nseg = num_segments(L, shift);
    if (ensure(&m->d_y, &m->cap_y, (size_t)nseg * m->S * SEG * 2))
        return -1;
    cudaEvent_t e0 = m->ev2;
    DMX_CUDA(cudaEventRecord(e0, m->stream));
    if (track_forward_shard(m, d_audio, L, shift, 0, 1, m->d_y, cb, user))
        return -1;
    if (track_overlap_add(m, d_audio, L, shift, m->d_y, 1, d_out))
        return -1;
    // whole-call device time: from before the shard to the end of the overlap-add
    std::swap(m->ev0, m->ev2);
    return 0;
}

// Device -> pageable host memory without stalling the launching thread.  cudaMemcpyAsync into memory that is not page-locked
// blocks its caller until the data has arrived (the driver bounces it through its own pinned buffers), and the caller here is
// the thread that should be enqueuing the next batch of kernels.  So when the caller's output buffer is ordinary memory (an
// Eigen tensor, a std::vector -- what cli-apps/demucs.cpp:162-230 passes), finished stretches of the stems are handed to a
// worker thread, which moves them through a small ring of pinned buffers (4 x 8 MB per model, allocated on first use) and
// memcpy's them into place.  Measured on B200 for the 254 MB of a 3-minute track: 6 ms staged vs 12-15 ms through
// cudaMemcpy on pageable memory (and 16 ms for cudaHostRegister + copy + unregister).
struct HostStager
{
    static constexpr int SLOTS = 4;
    static constexpr size_t SLOT_BYTES = 8u << 20;
    struct Task
    {
        cudaEvent_t ready;
        const uint8_t *src;
        uint8_t *dst;
        size_t bytes;
    };
    int device = 0;
    cudaStream_t stream = nullptr;
    uint8_t *ring[SLOTS] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t slot_ev[SLOTS] = {nullptr, nullptr, nullptr, nullptr};
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    std::deque<Task> q;
    bool closing = false;
    std::string err;

    int init(int dev, cudaStream_t st)
    {
        device = dev;
        stream = st;
        for (int i = 0; i < SLOTS; ++i)
        {
            if (!ring[i])
                DMX_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&ring[i]), SLOT_BYTES, cudaHostAllocDefault));
            if (!slot_ev[i])
                DMX_CUDA(cudaEventCreateWithFlags(&slot_ev[i], cudaEventDisableTiming));
        }
        return 0;
    }
    void start()
    {
        closing = false;
        err.clear();
        th = std::thread([this] { run(); });
    }
    void post(cudaEvent_t ready, const void *src, void *dst, size_t bytes)
    {
        {
            std::lock_guard<std::mutex> lk(mu);
            q.push_back(Task{ready, static_cast<const uint8_t *>(src), static_cast<uint8_t *>(dst), bytes});
        }
        cv.notify_one();
    }
    // waits for everything posted so far; returns -1 (message in err) if a copy failed
    int finish()
    {
        {
            std::lock_guard<std::mutex> lk(mu);
            closing = true;
        }
        cv.notify_one();
        if (th.joinable())
            th.join();
        return err.empty() ? 0 : -1;
    }
    void run()
    {
        if (cudaSetDevice(device) != cudaSuccess)
            err = "stager: cudaSetDevice failed";
        for (;;)
        {
            Task t;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return closing || !q.empty(); });
                if (q.empty())
                    return;
                t = q.front();
                q.pop_front();
            }
            if (!err.empty())
            {
                cudaEventDestroy(t.ready);
                continue;
            }
            cudaError_t e = cudaStreamWaitEvent(stream, t.ready, 0);
            // chunk i travels in slot i % SLOTS; a slot is reused once its previous chunk has been memcpy'd out
            const size_t n = (t.bytes + SLOT_BYTES - 1) / SLOT_BYTES;
            for (size_t i = 0; i < n + SLOTS && e == cudaSuccess; ++i)
            {
                if (i >= SLOTS)
                {
                    const size_t j = i - SLOTS, off = j * SLOT_BYTES;
                    const size_t sz = t.bytes - off < SLOT_BYTES ? t.bytes - off : SLOT_BYTES;
                    e = cudaEventSynchronize(slot_ev[j % SLOTS]);
                    if (e == cudaSuccess)
                        memcpy(t.dst + off, ring[j % SLOTS], sz);
                }
                if (i < n && e == cudaSuccess)
                {
                    const size_t off = i * SLOT_BYTES;
                    const size_t sz = t.bytes - off < SLOT_BYTES ? t.bytes - off : SLOT_BYTES;
                    e = cudaMemcpyAsync(ring[i % SLOTS], t.src + off, sz, cudaMemcpyDeviceToHost, stream);
                    if (e == cudaSuccess)
                        e = cudaEventRecord(slot_ev[i % SLOTS], stream);
                }
            }
            if (e != cudaSuccess)
                err = std::string("stager: ") + cudaGetErrorString(e);
            cudaEventDestroy(t.ready);
        }
    }
    ~HostStager()
    {
        finish();
        for (int i = 0; i < SLOTS; ++i)
        {
            if (ring[i])
                cudaFreeHost(ring[i]);
            if (slot_ev[i])
                cudaEventDestroy(slot_ev[i]);
        }
    }
};

void stager_free(HostStager *s) { delete s; }

// acc = (first ? 0 : acc) + x * scale   (the shift average of dmx_separate_shifts)
__global__ void __launch_bounds__(256) accumulate_kernel(float4 *__restrict__ acc, const float4 *__restrict__ x, float scale,
                                                         size_t n4, int first)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
    {
        const float4 v = x[i];
        float4 a = first ? make_float4(0.f, 0.f, 0.f, 0.f) : acc[i];
        a.x += v.x * scale;
        a.y += v.y * scale;
        a.z += v.z * scale;
        a.w += v.w * scale;
        acc[i] = a;
    }
}

int track_accumulate(dmx_model *m, float *d_acc, const float *d_x, float scale, size_t n, int first)
{
    if (n % 4)
    {
        set_error("track_accumulate: element count must be a multiple of 4");
        return -1;
    }
    {
        ProfScope ps(m, PC_TRACK, 0.0, (double)n * 12.0);
        accumulate_kernel<<<1184, 256, 0, m->stream>>>(reinterpret_cast<float4 *>(d_acc), reinterpret_cast<const float4 *>(d_x), scale,
                                                       n / 4, first);
    }
    DMX_LAUNCH_CHECK();
    return 0;
}

static bool host_is_pinned(const void *p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess)
    {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

// Host -> device of a pageable buffer through the pinned ring (the memcpy of chunk i + 1 overlaps the DMA of chunk i).
static int staged_h2d(HostStager *sg, float *d_dst, const float *h_src, size_t bytes, cudaStream_t st)
{
    const uint8_t *src = reinterpret_cast<const uint8_t *>(h_src);
    uint8_t *dst = reinterpret_cast<uint8_t *>(d_dst);
    const size_t n = (bytes + HostStager::SLOT_BYTES - 1) / HostStager::SLOT_BYTES;
    for (size_t i = 0; i < n; ++i)
    {
        const int k = (int)(i % HostStager::SLOTS);
        const size_t off = i * HostStager::SLOT_BYTES;
        const size_t sz = bytes - off < HostStager::SLOT_BYTES ? bytes - off : HostStager::SLOT_BYTES;
        if (i >= HostStager::SLOTS)
            DMX_CUDA(cudaEventSynchronize(sg->slot_ev[k]));
        memcpy(sg->ring[k], src + off, sz);
        DMX_CUDA(cudaMemcpyAsync(dst + off, sg->ring[k], sz, cudaMemcpyHostToDevice, st));
        DMX_CUDA(cudaEventRecord(sg->slot_ev[k], st));
    }
    // the ring is reused by the output side of the same call: make sure the uploads have left it
    for (int k = 0; k < HostStager::SLOTS; ++k)
        DMX_CUDA(cudaEventSynchronize(sg->slot_ev[k]));
    return 0;
}

// Host-buffer form of track_separate (dmx_separate): the same work, pipelined with the transfers.  After each batch of
// segments the stretch of the output that no later segment touches (everything before the next segment's first
// sample) is overlap-added and copied to the host on a separate stream while the next batch computes, so only the
// last stretch of the 4 x stereo stems (254 MB for a 3-minute track) is exposed after the final forward.  Pinned (or
// registered) caller buffers are copied to / from directly; ordinary memory goes through the HostStager above.
int track_separate_host(dmx_model *m, const float *h_audio, int64_t L, int shift, float *h_out, dmx_progress_fn cb,
                        void *user)
{
    if (check_track_args(L, shift, 0, 1))
        return -1;
    cudaStream_t st = m->stream;
    const int S = m->S;
    const int64_t Ls = L + MAX_SHIFT - shift;
    const int nseg = num_segments(L, shift);
    if (ensure(&m->d_y, &m->cap_y, (size_t)nseg * S * SEG * 2))
        return -1;
    if (!m->stream_copy)
    {
        DMX_CUDA(cudaStreamCreateWithFlags(&m->stream_copy, cudaStreamNonBlocking));
        DMX_CUDA(cudaEventCreateWithFlags(&m->ev_copy, cudaEventDisableTiming));
    }
    const bool in_pinned = host_is_pinned(h_audio), out_pinned = host_is_pinned(h_out);
    if ((!in_pinned || !out_pinned))
    {
        if (!m->stager)
            m->stager = new HostStager();
        if (m->stager->init(m->device, m->stream_copy))
            return -1;
    }
    DMX_CUDA(cudaEventRecord(m->ev2, st));
    if (in_pinned)
        DMX_CUDA(cudaMemcpyAsync(m->d_audio, h_audio, (size_t)L * 2 * sizeof(float), cudaMemcpyHostToDevice, st));
    else if (staged_h2d(m->stager, m->d_audio, h_audio, (size_t)L * 2 * sizeof(float), st))
        return -1;
    if (track_stats(m, m->d_audio, L))
        return -1;
    if (!out_pinned)
        m->stager->start();
    int rc = 0;
    int64_t t_prev = 0;
    for (int j0 = 0; j0 < nseg && !rc; j0 += m->maxB)
    {
        const int nb = (nseg - j0) < m->maxB ? (nseg - j0) : m->maxB;
        {
            ProfScope ps(m, PC_TRACK, 0.0, (double)nb * SEG * 16.0);
            extract_segments_kernel<<<dim3((SEG + 255) / 256, nb), 256, 0, st>>>(
                (const float2 *)m->d_audio, L, shift, Ls, m->track_stats, j0, 1, (float2 *)m->mix);
        }
        if (cudaGetLastError() != cudaSuccess || track_forward(m, m->mix, nb, m->d_y + (size_t)j0 * S * SEG * 2))
        {
            rc = -1;
            break;
        }
        // samples of the (unshifted) track before segment j0 + nb begins are final now
        int64_t t_end = (int64_t)(j0 + nb) * SEG_STRIDE - (MAX_SHIFT - shift);
        if (j0 + nb >= nseg || t_end > L)
            t_end = L;
        if (t_end > t_prev)
        {
            {
                ProfScope ps(m, PC_TRACK, 0.0, (double)(t_end - t_prev) * S * 8.0 * 2.3);
                overlap_add_kernel<<<(unsigned)((t_end - t_prev + 255) / 256), 256, 0, st>>>(
                    (const float2 *)m->d_y, S, nseg, 1, nseg, L, shift, Ls, m->track_stats, m->d_out, t_prev, t_end);
            }
            const size_t off = (size_t)S * 2 * t_prev, cnt = (size_t)S * 2 * (size_t)(t_end - t_prev);
            if (out_pinned)
            {
                if (cudaGetLastError() != cudaSuccess || cudaEventRecord(m->ev_copy, st) != cudaSuccess ||
                    cudaStreamWaitEvent(m->stream_copy, m->ev_copy, 0) != cudaSuccess ||
                    cudaMemcpyAsync(h_out + off, m->d_out + off, cnt * sizeof(float), cudaMemcpyDeviceToHost, m->stream_copy) !=
                        cudaSuccess)
                    rc = -1;
            }
            else
            {
                cudaEvent_t ready = nullptr;
                if (cudaGetLastError() != cudaSuccess || cudaEventCreateWithFlags(&ready, cudaEventDisableTiming) != cudaSuccess ||
                    cudaEventRecord(ready, st) != cudaSuccess)
                    rc = -1;
                else
                    m->stager->post(ready, m->d_out + off, h_out + off, cnt * sizeof(float));
            }
            t_prev = t_end;
        }
        if (cb && !rc)
        {
            if (cudaStreamSynchronize(st) != cudaSuccess)
                rc = -1;
            const bool v3 = m->arch == DMX_ARCH_HDEMUCS_V3;
            const int nmsg = v3 ? 22 : 26;
            for (int j = j0; j < j0 + nb; ++j)
                for (int sidx = 0; sidx < nmsg; ++sidx)
                    cb(((float)j + (float)(sidx + 1) / (float)nmsg) / (float)nseg, v3 ? kStageMsgsV3[sidx] : kStageMsgs[sidx],
                       user);
        }
    }
    if (rc)
        set_error(std::string("dmx_separate: ") + cudaGetErrorString(cudaGetLastError()) + " / " + dmx_last_error());
    if (!rc && cudaEventRecord(m->ev1, st) != cudaSuccess)
        rc = -1;
    cudaStreamSynchronize(st);
    if (!out_pinned && m->stager->finish())
    {
        set_error(m->stager->err);
        rc = -1;
    }
    cudaStreamSynchronize(m->stream_copy);
    if (!rc)
        std::swap(m->ev0, m->ev2);
    return rc;
}

} // namespace dmx
