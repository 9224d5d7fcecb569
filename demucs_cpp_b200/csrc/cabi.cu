// extern "C" surface of libdemucs_b200.so (declared in include/demucs_b200.h).
#include "model.cuh"

#include <chrono>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <condition_variable>
#include <thread>
#include <vector>
#include <cuda_fp16.h>

namespace dmx
{
static thread_local std::string g_err;
void set_error(const std::string &msg) { g_err = msg; }
int num_segments(int64_t L, int shift);

static float half_bits_to_float(uint16_t h)
{
    // IEEE binary16 -> binary32, exact (src/model_load.cpp:1092-1300 widens with Eigen::half)
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu, man = h & 0x3ffu;
    uint32_t bits;
    if (exp == 0)
    {
        if (man == 0)
            bits = sign;
        else
        {
            int e = -1;
            do
            {
                ++e;
                man <<= 1;
            } while (!(man & 0x400u));
            man &= 0x3ffu;
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | (man << 13);
        }
    }
    else if (exp == 31)
        bits = sign | 0x7f800000u | (man << 13);
    else
        bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
    float f;
    memcpy(&f, &bits, 4);
    return f;
}
} // namespace dmx

using namespace dmx;

extern "C"
{

int dmx_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess)
        return 0;
    return n;
}

const char *dmx_last_error(void) { return g_err.c_str(); }

dmx_model *dmx_model_create(int arch, int device, int max_batch)
{
    if (arch != DMX_ARCH_HTDEMUCS_4S && arch != DMX_ARCH_HTDEMUCS_6S && arch != DMX_ARCH_HDEMUCS_V3)
    {
        set_error("dmx_model_create: unknown architecture");
        return nullptr;
    }
    if (max_batch < 1)
    {
        set_error("dmx_model_create: max_batch must be >= 1");
        return nullptr;
    }
    int n = dmx_device_count();
    if (n <= 0)
    {
        set_error("dmx_model_create: no CUDA device (this library has no CPU fallback)");
        return nullptr;
    }
    if (device < 0 || device >= n)
    {
        set_error("dmx_model_create: bad device index");
        return nullptr;
    }
    dmx_model *m = new dmx_model();
    m->arch = arch;
    m->device = device;
    m->maxB = max_batch;
    return m;
}

int dmx_model_set_tensor(dmx_model *m, const char *name, const uint16_t *f16, const int32_t *ne, int n_dims)
{
    if (!m || !name || !f16 || n_dims < 0 || n_dims > 4)
    {
        set_error("dmx_model_set_tensor: bad argument");
        return -1;
    }
    if (m->finalized)
    {
        set_error("dmx_model_set_tensor: model already finalized");
        return -1;
    }
    HostTensor t;
    size_t n = 1;
    for (int i = 0; i < n_dims; ++i)
    {
        if (ne[i] <= 0)
        {
            set_error("dmx_model_set_tensor: bad shape");
            return -1;
        }
        t.shape.push_back(ne[i]);
        n *= (size_t)ne[i];
    }
    t.data.resize(n);
    for (size_t i = 0; i < n; ++i)
        t.data[i] = half_bits_to_float(f16[i]);
    m->host[name] = std::move(t);
    return 0;
}

int dmx_model_finalize(dmx_model *m)
{
    if (!m)
    {
        set_error("dmx_model_finalize: null model");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    return model_finalize(m);
}

dmx_model *dmx_model_load_file(const char *path, int device, int max_batch)
{
    FILE *f = fopen(path, "rb");
    if (!f)
    {
        set_error(std::string("dmx_model_load_file: cannot open ") + path);
        return nullptr;
    }
    uint32_t magic = 0;
    if (fread(&magic, 4, 1, f) != 1)
    {
        fclose(f);
        set_error("dmx_model_load_file: short file");
        return nullptr;
    }
    int arch = 0;
    if (magic == 0x646d6334u)
        arch = DMX_ARCH_HTDEMUCS_4S;
    else if (magic == 0x646d6336u)
        arch = DMX_ARCH_HTDEMUCS_6S;
    else if (magic == 0x646d6333u)
        arch = DMX_ARCH_HDEMUCS_V3;
    else
    {
        fclose(f);
        set_error("dmx_model_load_file: invalid model data (bad magic)");
        return nullptr;
    }
    const auto t_begin = std::chrono::steady_clock::now();
    dmx_model *m = dmx_model_create(arch, device, max_batch);
    if (!m)
    {
        fclose(f);
        return nullptr;
    }
    std::vector<uint16_t> buf;
    std::vector<char> name;
    for (;;)
    {
        int32_t hdr[2];
        if (fread(hdr, 4, 2, f) != 2)
            break; // EOF (src/model_load.cpp:144-147)
        const int n_dims = hdr[0], name_len = hdr[1];
        if (n_dims < 0 || n_dims > 4 || name_len <= 0 || name_len > 4096)
        {
            set_error("dmx_model_load_file: corrupt tensor header");
            goto fail;
        }
        int32_t ne[4] = {1, 1, 1, 1};
        if (n_dims && fread(ne, 4, n_dims, f) != (size_t)n_dims)
        {
            set_error("dmx_model_load_file: truncated tensor header");
            goto fail;
        }
        name.assign(name_len + 1, 0);
        if (fread(name.data(), 1, name_len, f) != (size_t)name_len)
        {
            set_error("dmx_model_load_file: truncated tensor name");
            goto fail;
        }
        size_t n = 1;
        for (int i = 0; i < n_dims; ++i)
            n *= (size_t)(ne[i] > 0 ? ne[i] : 0);
        buf.resize(n);
        if (n == 0 || fread(buf.data(), 2, n, f) != n)
        {
            set_error(std::string("dmx_model_load_file: truncated tensor data for ") + name.data());
            goto fail;
        }
        if (dmx_model_set_tensor(m, name.data(), buf.data(), ne, n_dims))
            goto fail;
    }
    fclose(f);
    f = nullptr;
    {
        const auto t_parsed = std::chrono::steady_clock::now();
        const double pack0 = load_pack_seconds();
        if (dmx_model_finalize(m))
        {
            dmx_model_free(m);
            return nullptr;
        }
        if (getenv("DMX_LOAD_TRACE"))
        {
            const auto t_end = std::chrono::steady_clock::now();
            fprintf(stderr, "dmx_model_load_file: read + widen %.3f s, finalize %.3f s (of which repacking incl. its uploads %.3f s)\n",
                    std::chrono::duration<double>(t_parsed - t_begin).count(), std::chrono::duration<double>(t_end - t_parsed).count(),
                    load_pack_seconds() - pack0);
        }
    }
    return m;
fail:
    if (f)
        fclose(f);
    dmx_model_free(m);
    return nullptr;
}

void dmx_model_free(dmx_model *m)
{
    if (!m)
        return;
    if (m->finalized || !m->allocs.empty() || m->stream)
    {
        int prev = -1;
        cudaGetDevice(&prev);
        cudaSetDevice(m->device);
        for (cudaStream_t s : {m->stream, m->stream2, m->stream_copy})
            if (s)
                cudaStreamSynchronize(s);
        if (m->stager)
            stager_free(m->stager);
        for (auto &g : m->graphs)
            if (g.exec)
                cudaGraphExecDestroy(g.exec);
        for (void *p : m->allocs)
            cudaFree(p);
        for (float *p : {m->d_audio, m->d_y, m->d_out, m->d_tail, m->d_acc})
            if (p)
                cudaFree(p);
        for (cudaStream_t s : {m->stream, m->stream2, m->stream_copy})
            if (s)
                cudaStreamDestroy(s);
        for (cudaEvent_t e : {m->ev_fork, m->ev_join, m->ev_copy, m->ev0, m->ev1, m->ev2})
            if (e)
                cudaEventDestroy(e);
        for (cudaEvent_t e : m->prof.pool)
            cudaEventDestroy(e);
        if (prev >= 0)
            cudaSetDevice(prev);
    }
    delete m->v3;
    delete m;
}

int dmx_segment_progress(const dmx_model *m, float current_progress, float segment_progress, dmx_progress_fn cb, void *user)
{
    if (!m || !m->finalized)
    {
        set_error("dmx_segment_progress: model not loaded");
        return -1;
    }
    track_fire_segment_progress(m, current_progress, segment_progress, cb, user);
    return 0;
}

int dmx_model_arch(const dmx_model *m) { return m ? m->arch : -1; }
int dmx_model_sources(const dmx_model *m) { return m ? (m->arch == DMX_ARCH_HTDEMUCS_6S ? 6 : 4) : -1; }
int dmx_model_max_batch(const dmx_model *m) { return m ? m->maxB : -1; }
int64_t dmx_launch_count(const dmx_model *m) { return m ? m->launches : -1; }

static const char *const kProfNames[PC_COUNT] = {"stft",      "istft",      "istft_ola",        "pointwise", "gemm_conv",
                                                 "gemm_dconv", "gemm_transformer", "attention", "track",     "lstm",
                                                 "dconv_fused"};

int dmx_profile_classes(void) { return PC_COUNT; }

const char *dmx_profile_class_name(int cls) { return (cls >= 0 && cls < PC_COUNT) ? kProfNames[cls] : ""; }

int dmx_profile_begin(dmx_model *m)
{
    if (!m)
    {
        set_error("dmx_profile_begin: null model");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    m->prof.on = true;
    m->prof.used = 0;
    m->prof.recs.clear();
    for (int i = 0; i < PC_COUNT; ++i)
        m->prof.flops[i] = m->prof.bytes[i] = 0.0;
    return 0;
}

int dmx_profile_sequence(dmx_model *m, int *cls_out, int capacity)
{
    if (!m)
        return -1;
    std::lock_guard<std::mutex> lk(m->mtx);
    int n = 0; // one entry per kernel launch, in launch order
    for (const ProfRec &r : m->prof.recs)
        for (int k = 0; k < r.n; ++k, ++n)
            if (cls_out && n < capacity)
                cls_out[n] = r.cls;
    return n;
}

int dmx_profile_end(dmx_model *m, double *ms, int64_t *launches, double *flops, double *bytes)
{
    if (!m)
    {
        set_error("dmx_profile_end: null model");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    m->prof.on = false;
    DMX_CUDA(cudaSetDevice(m->device));
    DMX_CUDA(cudaStreamSynchronize(m->stream));
    for (int i = 0; i < PC_COUNT; ++i)
    {
        if (ms)
            ms[i] = 0.0;
        if (launches)
            launches[i] = 0;
        if (flops)
            flops[i] = m->prof.flops[i];
        if (bytes)
            bytes[i] = m->prof.bytes[i];
    }
    for (const ProfRec &r : m->prof.recs)
    {
        float t = 0.f;
        DMX_CUDA(cudaEventElapsedTime(&t, r.e0, r.e1));
        if (ms)
            ms[r.cls] += (double)t;
        if (launches)
            launches[r.cls] += r.n;
    }
    m->prof.recs.clear();
    m->prof.used = 0;
    return 0;
}

int dmx_model_set_precision(dmx_model *m, int precision)
{
    if (!m || (precision != DMX_PRECISION_FP32 && precision != DMX_PRECISION_BF16))
    {
        set_error("dmx_model_set_precision: bad argument");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    if (m->precision != precision && !m->graphs.empty())
    {
        // captured forwards hold the other mode's kernels
        int prev = -1;
        cudaGetDevice(&prev);
        cudaSetDevice(m->device);
        if (m->stream)
            cudaStreamSynchronize(m->stream);
        for (auto &g : m->graphs)
            if (g.exec)
                cudaGraphExecDestroy(g.exec);
        m->graphs.clear();
        if (prev >= 0)
            cudaSetDevice(prev);
    }
    m->precision = precision;
    return model_zero_attention_pads(m);
}

int dmx_model_set_layer_taps(dmx_model *m, int on)
{
    if (!m)
    {
        set_error("dmx_model_set_layer_taps: null model");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    m->layer_taps = on != 0;
    return 0;
}

int dmx_num_segments(int64_t n_frames, int shift_offset) { return num_segments(n_frames, shift_offset); }

static int grow(float **p, size_t *cap, size_t n)
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

int dmx_separate_device(dmx_model *m, const float *d_audio_lr, int64_t n_frames, int shift_offset, float *d_out)
{
    if (!m || !d_audio_lr || !d_out)
    {
        set_error("dmx_separate_device: null argument");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    if (!m->finalized)
    {
        set_error("dmx_separate_device: model not finalized");
        return -1;
    }
    DMX_CUDA(cudaSetDevice(m->device));
    return track_separate(m, d_audio_lr, n_frames, shift_offset, d_out, nullptr, nullptr);
}

int dmx_forward_shard_device(dmx_model *m, const float *d_audio_lr, int64_t n_frames, int shift_offset, int rank,
                             int world, float *d_y)
{
    if (!m || !d_audio_lr || !d_y)
    {
        set_error("dmx_forward_shard_device: null argument");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    if (!m->finalized)
    {
        set_error("dmx_forward_shard_device: model not finalized");
        return -1;
    }
    DMX_CUDA(cudaSetDevice(m->device));
    return track_forward_shard(m, d_audio_lr, n_frames, shift_offset, rank, world, d_y, nullptr, nullptr);
}

int dmx_overlap_add_device(dmx_model *m, const float *d_audio_lr, int64_t n_frames, int shift_offset,
                           const float *d_y_all, int world, float *d_out)
{
    if (!m || !d_audio_lr || !d_y_all || !d_out)
    {
        set_error("dmx_overlap_add_device: null argument");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    if (!m->finalized)
    {
        set_error("dmx_overlap_add_device: model not finalized");
        return -1;
    }
    DMX_CUDA(cudaSetDevice(m->device));
    return track_overlap_add(m, d_audio_lr, n_frames, shift_offset, d_y_all, world, d_out);
}

int dmx_plan_block(int64_t n_frames, int shift_offset, int rank, int world, dmx_block_plan *plan)
{
    if (!plan)
    {
        set_error("dmx_plan_block: null plan");
        return -1;
    }
    return plan_block(n_frames, shift_offset, rank, world, plan);
}

static int check_ready(dmx_model *m, const char *who)
{
    if (!m)
    {
        set_error(std::string(who) + ": null model");
        return -1;
    }
    if (!m->finalized)
    {
        set_error(std::string(who) + ": model not finalized");
        return -1;
    }
    return 0;
}

int dmx_track_stats_device(dmx_model *m, const float *d_audio_lr, int64_t n_frames, double *d_sums)
{
    if (check_ready(m, "dmx_track_stats_device"))
        return -1;
    if (!d_sums || (n_frames > 0 && !d_audio_lr) || n_frames < 0)
    {
        set_error("dmx_track_stats_device: bad argument");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    DMX_CUDA(cudaSetDevice(m->device));
    return track_stats_partial(m, d_audio_lr, n_frames, d_sums);
}

int dmx_forward_block_device(dmx_model *m, const float *d_audio_win, int64_t n_frames, int shift_offset,
                             const dmx_block_plan *plan, const double *d_sums, float *d_y)
{
    if (check_ready(m, "dmx_forward_block_device"))
        return -1;
    if (!plan || !d_sums || (plan->seg_end > plan->seg_begin && (!d_audio_win || !d_y)))
    {
        set_error("dmx_forward_block_device: null argument");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    DMX_CUDA(cudaSetDevice(m->device));
    return track_forward_block(m, d_audio_win, n_frames, shift_offset, *plan, d_sums, d_y);
}

int dmx_overlap_add_block_device(dmx_model *m, int64_t n_frames, int shift_offset, const dmx_block_plan *plan,
                                 const double *d_sums, const float *d_y, const float *d_prev_tail, int src_begin, int src_end,
                                 float *d_out_block)
{
    if (check_ready(m, "dmx_overlap_add_block_device"))
        return -1;
    if (!plan || !d_sums || (plan->out_end > plan->out_begin && (!d_y || !d_out_block)))
    {
        set_error("dmx_overlap_add_block_device: null argument");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    DMX_CUDA(cudaSetDevice(m->device));
    return track_overlap_add_block(m, n_frames, shift_offset, *plan, d_sums, d_y, d_prev_tail, src_begin, src_end, d_out_block);
}

int dmx_model_sync(dmx_model *m)
{
    if (check_ready(m, "dmx_model_sync"))
        return -1;
    DMX_CUDA(cudaSetDevice(m->device));
    DMX_CUDA(cudaStreamSynchronize(m->stream));
    return 0;
}

double dmx_last_device_ms(dmx_model *m)
{
    if (!m || !m->ev0 || !m->ev1)
        return -1.0;
    float t = 0.f;
    if (cudaEventElapsedTime(&t, m->ev0, m->ev1) != cudaSuccess)
        return -1.0;
    return (double)t;
}

int dmx_separate(dmx_model *m, const float *audio_lr, int64_t n_frames, int shift_offset, float *out,
                 dmx_progress_fn cb, void *user)
{
    if (!m || !audio_lr || !out)
    {
        set_error("dmx_separate: null argument");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    if (!m->finalized)
    {
        set_error("dmx_separate: model not finalized");
        return -1;
    }
    if (n_frames < 2)
    {
        set_error("dmx_separate: track too short");
        return -1;
    }
    DMX_CUDA(cudaSetDevice(m->device));
    const size_t S = (size_t)m->S;
    if (grow(&m->d_audio, &m->cap_audio, (size_t)n_frames * 2) || grow(&m->d_out, &m->cap_out, S * 2 * (size_t)n_frames))
        return -1;
    return track_separate_host(m, audio_lr, n_frames, shift_offset, out, cb, user);
}

int dmx_forward_segments(dmx_model *m, const float *segs, int n_segments, float *out)
{
    if (!m || !segs || !out)
    {
        set_error("dmx_forward_segments: null argument");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    if (!m->finalized)
    {
        set_error("dmx_forward_segments: model not finalized");
        return -1;
    }
    if (n_segments < 1 || n_segments > m->maxB)
    {
        set_error("dmx_forward_segments: n_segments must be in [1, max_batch]");
        return -1;
    }
    DMX_CUDA(cudaSetDevice(m->device));
    const size_t S = (size_t)m->S;
    if (grow(&m->d_y, &m->cap_y, (size_t)n_segments * S * SEG * 2))
        return -1;
    DMX_CUDA(cudaMemcpyAsync(m->mix, segs, (size_t)n_segments * SEG * 2 * sizeof(float), cudaMemcpyHostToDevice,
                             m->stream));
    if (model_forward(m, m->mix, n_segments, m->d_y))
        return -1;
    DMX_CUDA(cudaMemcpyAsync(out, m->d_y, (size_t)n_segments * S * SEG * 2 * sizeof(float), cudaMemcpyDeviceToHost,
                             m->stream));
    DMX_CUDA(cudaStreamSynchronize(m->stream));
    return 0;
}

int64_t dmx_debug_tap(dmx_model *m, const char *name, float *dst, int64_t capacity, int64_t dims[4])
{
    if (!m || !name)
    {
        set_error("dmx_debug_tap: null argument");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    auto it = m->taps.find(name);
    if (it == m->taps.end())
    {
        set_error(std::string("dmx_debug_tap: unknown tap ") + name);
        return -1;
    }
    const Tap &t = it->second;
    const int64_t n = (int64_t)m->lastB * t.d1 * t.d2 * t.C;
    if (dims)
    {
        dims[0] = m->lastB;
        dims[1] = t.d1;
        dims[2] = t.d2;
        dims[3] = t.C;
    }
    if (!dst)
        return n;
    if (capacity < n)
    {
        set_error("dmx_debug_tap: destination too small");
        return -1;
    }
    if (cudaSetDevice(m->device) != cudaSuccess ||
        cudaMemcpy(dst, t.ptr, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess)
    {
        set_error("dmx_debug_tap: copy failed");
        return -1;
    }
    if (t.swapped)
    {
        // stored [b][d2][d1][C]; reported [b][d1][d2][C]
        std::vector<float> tmp(dst, dst + n);
        for (int64_t b = 0; b < m->lastB; ++b)
            for (int64_t i2 = 0; i2 < t.d2; ++i2)
                for (int64_t i1 = 0; i1 < t.d1; ++i1)
                    memcpy(dst + ((b * t.d1 + i1) * t.d2 + i2) * t.C, tmp.data() + ((b * t.d2 + i2) * t.d1 + i1) * t.C,
                           (size_t)t.C * sizeof(float));
    }
    if (t.minus)
    {
        std::vector<float> sk((size_t)n);
        if (cudaMemcpy(sk.data(), t.minus, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess)
        {
            set_error("dmx_debug_tap: copy failed");
            return -1;
        }
        for (int64_t i = 0; i < n; ++i)
            dst[i] -= sk[(size_t)i];
    }
    return n;
}

int dmx_debug_linear(int device, int precision, int M, int N, int K, const float *A, const float *W, const float *bias,
                     int act, float *out, int iters, float *ms_out)
{
    if (!A || !W || !out || M < 1 || N < 1 || K < 1)
    {
        set_error("dmx_debug_linear: bad argument");
        return -1;
    }
    if (dmx_device_count() < 1)
    {
        set_error("dmx_debug_linear: no CUDA device (this library has no CPU fallback)");
        return -1;
    }
    return debug_linear(device, precision, M, N, K, A, W, bias, act, out, iters, ms_out);
}

int dmx_debug_conv(int device, int B, int U, int V, int Cin, int N, int ntaps, const int *du, const int *dv, const float *img,
                   const float *W, const float *bias, int act, int glu, int share_dv, float *out, int iters, float *ms_out)
{
    if (!img || !W || !out || !du || !dv || B < 1 || U < 1 || V < 1 || Cin < 1 || N < 1 || ntaps < 1 || (glu && (N & 1)))
    {
        set_error("dmx_debug_conv: bad argument");
        return -1;
    }
    if (dmx_device_count() < 1)
    {
        set_error("dmx_debug_conv: no CUDA device (this library has no CPU fallback)");
        return -1;
    }
    return debug_conv(device, B, U, V, Cin, N, ntaps, du, dv, img, W, bias, act, glu, share_dv, out, iters, ms_out);
}

int dmx_stft_segment(dmx_model *m, const float *seg_lr, float *spec)
{
    if (!m || !seg_lr || !spec)
    {
        set_error("dmx_stft_segment: null argument");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    if (!m->finalized)
    {
        set_error("dmx_stft_segment: model not finalized");
        return -1;
    }
    DMX_CUDA(cudaSetDevice(m->device));
    DMX_CUDA(cudaMemcpyAsync(m->mix, seg_lr, (size_t)SEG * 2 * sizeof(float), cudaMemcpyHostToDevice, m->stream));
    DMX_CUDA(cudaMemsetAsync(m->stats, 0, 2 * sizeof(double), m->stream));
    if (launch_stft(m->mix, 1, m->window, m->tw, m->spec, m->stats, m->stream))
        return -1;
    DMX_CUDA(cudaMemcpyAsync(spec, m->spec, (size_t)NBINS * LE * 4 * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
    DMX_CUDA(cudaStreamSynchronize(m->stream));
    return 0;
}

// Inverse of dmx_stft_segment: spec (2048, 336, 4) -> segment (343980, 2); identity mean/std, zero
// time branch, one "source" whose four channels are the CaC planes.
int dmx_istft_segment(dmx_model *m, const float *spec, float *seg_lr)
{
    if (!m || !spec || !seg_lr)
    {
        set_error("dmx_istft_segment: null argument");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx);
    if (!m->finalized)
    {
        set_error("dmx_istft_segment: model not finalized");
        return -1;
    }
    DMX_CUDA(cudaSetDevice(m->device));
    cudaStream_t st = m->stream;
    DMX_CUDA(cudaMemcpyAsync(m->spec, spec, (size_t)NBINS * LE * 4 * sizeof(float), cudaMemcpyHostToDevice, st));
    const float ident[2] = {0.f, 1.f};
    DMX_CUDA(cudaMemcpyAsync(m->meanstd_x, ident, sizeof ident, cudaMemcpyHostToDevice, st));
    DMX_CUDA(cudaMemcpyAsync(m->meanstd_t, ident, sizeof ident, cudaMemcpyHostToDevice, st));
    DMX_CUDA(cudaMemsetAsync(m->xt_in, 0, (size_t)SEG * 2 * sizeof(float), st));
    if (launch_istft(m->spec, 1, 1, m->meanstd_x, m->window, m->tw, m->wss, m->xt_in, m->meanstd_t, m->frames,
                     m->seg_out, st))
        return -1;
    DMX_CUDA(cudaMemcpyAsync(seg_lr, m->seg_out, (size_t)SEG * 2 * sizeof(float), cudaMemcpyDeviceToHost, st));
    DMX_CUDA(cudaStreamSynchronize(st));
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// One process, several GPUs (dmx_separate_multi) and the htdemucs_ft bag (dmx_separate_bag).
namespace
{
struct Rendezvous // reusable barrier for the worker threads (C++17: no std::barrier)
{
    std::mutex mu;
    std::condition_variable cv;
    int n, waiting = 0, gen = 0;
    explicit Rendezvous(int n_) : n(n_) {}
    void wait()
    {
        std::unique_lock<std::mutex> lk(mu);
        const int g = gen;
        if (++waiting == n)
        {
            waiting = 0;
            ++gen;
            cv.notify_all();
        }
        else
            cv.wait(lk, [&] { return gen != g; });
    }
};

struct Flag
{
    std::mutex mu;
    std::condition_variable cv;
    bool set = false;
    void raise()
    {
        {
            std::lock_guard<std::mutex> lk(mu);
            set = true;
        }
        cv.notify_all();
    }
    void wait()
    {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return set; });
    }
};
} // namespace

int dmx_separate_multi(dmx_model *const *models, int n_models, const float *audio_lr, int64_t n_frames, int shift_offset,
                       float *out, dmx_progress_fn cb, void *user)
{
    if (!models || n_models < 1 || !audio_lr || !out)
    {
        set_error("dmx_separate_multi: bad argument");
        return -1;
    }
    for (int r = 0; r < n_models; ++r)
    {
        if (check_ready(models[r], "dmx_separate_multi"))
            return -1;
        if (models[r]->arch != models[0]->arch)
        {
            set_error("dmx_separate_multi: the replicas must be the same model");
            return -1;
        }
        for (int q = 0; q < r; ++q)
            if (models[q] == models[r])
            {
                set_error("dmx_separate_multi: the same model handle was passed twice");
                return -1;
            }
    }
    if (n_models == 1)
        return dmx_separate(models[0], audio_lr, n_frames, shift_offset, out, cb, user);
    // one multi-GPU separation at a time: two calls' workers would otherwise take the replicas' locks in different orders
    static std::mutex multi_mtx;
    std::lock_guard<std::mutex> multi_lock(multi_mtx);
    const int W = n_models, S = models[0]->S;
    constexpr int TAIL = SEG - SEG_STRIDE;
    std::vector<dmx_block_plan> plans(W);
    for (int r = 0; r < W; ++r)
        if (plan_block(n_frames, shift_offset, r, W, &plans[r]))
            return -1;
    // buffers the neighbours write into must exist before the workers start
    for (int r = 0; r < W; ++r)
    {
        dmx_model *m = models[r];
        DMX_CUDA(cudaSetDevice(m->device));
        if (!m->d_tail)
            DMX_CUDA(cudaMalloc(reinterpret_cast<void **>(&m->d_tail), (size_t)S * TAIL * 2 * sizeof(float)));
        if (r + 1 < W && models[r + 1]->device != m->device)
        {
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, m->device, models[r + 1]->device) == cudaSuccess && can)
                if (cudaDeviceEnablePeerAccess(models[r + 1]->device, 0) != cudaSuccess)
                    cudaGetLastError(); // already enabled
        }
    }
    std::vector<double> part(2 * (size_t)W, 0.0);
    std::vector<std::string> errs(W);
    std::vector<Flag> tail_ready(W);
    std::atomic<bool> failed{false};
    Rendezvous rdv(W);
    auto worker = [&](int r) {
        dmx_model *m = models[r];
        const dmx_block_plan &pl = plans[r];
        std::lock_guard<std::mutex> lk(m->mtx);
        const int nloc = pl.seg_end - pl.seg_begin;
        const int64_t nin = pl.in_end - pl.in_begin, nout = pl.out_end - pl.out_begin;
        auto step = [&](int rc) {
            if (rc)
            {
                errs[r] = dmx_last_error();
                failed = true;
            }
            return rc;
        };
        auto cuda = [&](cudaError_t e, const char *what) {
            if (e != cudaSuccess)
            {
                set_error(std::string("dmx_separate_multi: ") + what + ": " + cudaGetErrorString(e));
                return step(-1);
            }
            return 0;
        };
        cudaStream_t st = m->stream;
        bool ok = !cuda(cudaSetDevice(m->device), "cudaSetDevice");
        // ---- stage 1: own window in, partial statistics out ----
        if (ok && nloc > 0)
        {
            ok = !step(grow(&m->d_audio, &m->cap_audio, (size_t)nin * 2)) && !step(grow(&m->d_y, &m->cap_y, (size_t)nloc * S * SEG * 2)) &&
                 !step(grow(&m->d_out, &m->cap_out, (size_t)S * 2 * (size_t)nout));
            ok = ok && !cuda(cudaEventRecord(m->ev2, st), "event");
            ok = ok && !cuda(cudaMemcpyAsync(m->d_audio, audio_lr + 2 * pl.in_begin, (size_t)nin * 2 * sizeof(float),
                                             cudaMemcpyHostToDevice, st), "copy in");
            ok = ok && !cuda(cudaMemsetAsync(m->track_stats, 0, 2 * sizeof(double), st), "memset");
            ok = ok && !step(track_stats_partial(m, m->d_audio + 2 * (pl.out_begin - pl.in_begin), nout, m->track_stats));
            ok = ok && !cuda(cudaMemcpyAsync(&part[2 * r], m->track_stats, 2 * sizeof(double), cudaMemcpyDeviceToHost, st), "copy");
            ok = ok && !cuda(cudaStreamSynchronize(st), "sync");
        }
        rdv.wait();
        // ---- stage 2: whole-track statistics (same summation order on every worker), forward, tail to the neighbour ----
        double tot[2] = {0.0, 0.0};
        for (int q = 0; q < W; ++q)
        {
            tot[0] += part[2 * q];
            tot[1] += part[2 * q + 1];
        }
        ok = ok && !failed;
        if (ok && nloc > 0)
        {
            ok = !cuda(cudaMemcpyAsync(m->track_stats, tot, 2 * sizeof(double), cudaMemcpyHostToDevice, st), "copy");
            ok = ok && !step(track_forward_block(m, m->d_audio, n_frames, shift_offset, pl, m->track_stats, m->d_y));
            if (ok && r + 1 < W && plans[r + 1].tail_frames > 0)
            {
                const float *src = m->d_y + ((size_t)(nloc - 1) * S * SEG + SEG_STRIDE) * 2;
                ok = !cuda(cudaMemcpy2DAsync(models[r + 1]->d_tail, (size_t)TAIL * 2 * sizeof(float), src,
                                             (size_t)SEG * 2 * sizeof(float), (size_t)TAIL * 2 * sizeof(float), S,
                                             cudaMemcpyDefault, st), "tail copy");
                ok = ok && !cuda(cudaStreamSynchronize(st), "sync");
            }
        }
        if (r + 1 < W)
            tail_ready[r + 1].raise(); // also on failure: the neighbour must not wait for ever
        // ---- stage 3: overlap-add of the own stretch, straight to the caller's buffer ----
        if (pl.tail_frames > 0)
            tail_ready[r].wait();
        ok = ok && !failed;
        if (ok && nloc > 0 && nout > 0)
        {
            ok = !step(track_overlap_add_block(m, n_frames, shift_offset, pl, m->track_stats, m->d_y, m->d_tail, 0, S, m->d_out));
            ok = ok && !cuda(cudaMemcpyAsync(out + (size_t)S * 2 * pl.out_begin, m->d_out, (size_t)S * 2 * (size_t)nout * sizeof(float),
                                             cudaMemcpyDeviceToHost, st), "copy out");
            ok = ok && !cuda(cudaEventRecord(m->ev1, st), "event");
            ok = ok && !cuda(cudaStreamSynchronize(st), "sync");
            if (ok)
                std::swap(m->ev0, m->ev2);
        }
    };
    std::vector<std::thread> th;
    for (int r = 1; r < W; ++r)
        th.emplace_back(worker, r);
    worker(0);
    for (auto &t : th)
        t.join();
    if (failed)
    {
        for (int r = 0; r < W; ++r)
            if (!errs[r].empty())
            {
                set_error("dmx_separate_multi (GPU " + std::to_string(models[r]->device) + "): " + errs[r]);
                return -1;
            }
        set_error("dmx_separate_multi: a worker failed");
        return -1;
    }
    track_fire_progress(models[0], plans[0].n_segments, cb, user);
    return 0;
}

int dmx_separate_shifts(dmx_model *m, const float *audio_lr, int64_t n_frames, const int *shift_offsets, int n_shifts, float *out,
                        dmx_progress_fn cb, void *user)
{
    if (check_ready(m, "dmx_separate_shifts"))
        return -1;
    if (!audio_lr || !out || !shift_offsets || n_shifts < 1)
    {
        set_error("dmx_separate_shifts: bad argument");
        return -1;
    }
    if (n_shifts == 1)
        return dmx_separate(m, audio_lr, n_frames, shift_offsets[0], out, cb, user);
    std::lock_guard<std::mutex> lk(m->mtx);
    DMX_CUDA(cudaSetDevice(m->device));
    const size_t n = (size_t)m->S * 2 * (size_t)n_frames;
    if (grow(&m->d_audio, &m->cap_audio, (size_t)n_frames * 2) || grow(&m->d_out, &m->cap_out, n) ||
        grow(&m->d_acc, &m->cap_acc, (n + 3) / 4 * 4))
        return -1;
    DMX_CUDA(cudaMemcpyAsync(m->d_audio, audio_lr, (size_t)n_frames * 2 * sizeof(float), cudaMemcpyHostToDevice, m->stream));
    for (int k = 0; k < n_shifts; ++k)
    {
        // each pass reports its own 0..1 progress, like successive demucs_inference calls would
        if (track_separate(m, m->d_audio, n_frames, shift_offsets[k], m->d_out, cb, user) ||
            track_accumulate(m, m->d_acc, m->d_out, 1.0f / (float)n_shifts, (n + 3) / 4 * 4, k == 0))
            return -1;
    }
    DMX_CUDA(cudaMemcpyAsync(out, m->d_acc, n * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
    DMX_CUDA(cudaStreamSynchronize(m->stream));
    return 0;
}

int dmx_separate_bag(dmx_model *const *models, int n_models, const float *audio_lr, int64_t n_frames,
                     const int *shift_offsets, float *out, dmx_progress_fn cb, void *user)
{
    if (!models || !audio_lr || !out || !shift_offsets || n_models < 1)
    {
        set_error("dmx_separate_bag: bad argument");
        return -1;
    }
    for (int i = 0; i < n_models; ++i)
    {
        if (check_ready(models[i], "dmx_separate_bag"))
            return -1;
        if (models[i]->device != models[0]->device || models[i]->S != models[0]->S)
        {
            set_error("dmx_separate_bag: the members must be on one GPU and have the same sources");
            return -1;
        }
    }
    if (n_models != models[0]->S)
    {
        set_error("dmx_separate_bag: needs one member per source (stem i is taken from member i)");
        return -1;
    }
    dmx_model *m0 = models[0];
    const int S = m0->S;
    std::lock_guard<std::mutex> lk0(m0->mtx);
    DMX_CUDA(cudaSetDevice(m0->device));
    if (grow(&m0->d_audio, &m0->cap_audio, (size_t)n_frames * 2) || grow(&m0->d_out, &m0->cap_out, (size_t)S * 2 * (size_t)n_frames))
        return -1;
    DMX_CUDA(cudaMemcpyAsync(m0->d_audio, audio_lr, (size_t)n_frames * 2 * sizeof(float), cudaMemcpyHostToDevice, m0->stream));
    DMX_CUDA(cudaStreamSynchronize(m0->stream));
    int nseg_total = 0;
    for (int i = 0; i < n_models; ++i)
    {
        dmx_model *m = models[i];
        std::unique_lock<std::mutex> lk(m->mtx, std::defer_lock);
        if (i > 0)
            lk.lock();
        dmx_block_plan pl;
        if (plan_block(n_frames, shift_offsets[i], 0, 1, &pl))
            return -1;
        nseg_total += pl.n_segments;
        if (grow(&m->d_y, &m->cap_y, (size_t)pl.n_segments * S * SEG * 2))
            return -1;
        DMX_CUDA(cudaMemsetAsync(m->track_stats, 0, 2 * sizeof(double), m->stream));
        if (track_stats_partial(m, m0->d_audio, n_frames, m->track_stats) ||
            track_forward_block(m, m0->d_audio, n_frames, shift_offsets[i], pl, m->track_stats, m->d_y) ||
            track_overlap_add_block(m, n_frames, shift_offsets[i], pl, m->track_stats, m->d_y, nullptr, i, i + 1, m0->d_out))
            return -1;
        DMX_CUDA(cudaStreamSynchronize(m->stream)); // members share m0's output buffer: one after the other
        if (cb)
        {
            // the reference runs the members one after the other, each reporting its own 0..1 progress
            // (cli-apps/demucs_ft.cpp:222-231)
            track_fire_progress(m, pl.n_segments, cb, user);
        }
    }
    (void)nseg_total;
    DMX_CUDA(cudaMemcpyAsync(out, m0->d_out, (size_t)S * 2 * (size_t)n_frames * sizeof(float), cudaMemcpyDeviceToHost, m0->stream));
    DMX_CUDA(cudaStreamSynchronize(m0->stream));
    return 0;
}

} // extern "C"
