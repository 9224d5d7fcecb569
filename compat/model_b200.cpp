// compat/model_b200.cpp -- the reference's CLI-tier API (compat/model.hpp) over the C ABI.
// Eigen::MatrixXf(2, L) and Eigen::Tensor3dXf(S, 2, L) are column-major, which is exactly the interleaved
// L,R / source-fastest layout the ABI takes: no copies or transposes on either side.
#include "model.hpp"

#include "demucs_b200.h"

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <iostream>
#include <mutex>
#include <thread>
#include <vector>

namespace
{
int env_int(const char *name, int dflt)
{
    const char *v = std::getenv(name);
    return v ? std::atoi(v) : dflt;
}

// The device-side state behind a demucs_model: the model on GPU DEMUCS_DEVICE and, with DEMUCS_NGPU = n > 1, replicas on the
// next n - 1 GPUs, handed out to the host threads that call demucs_inference on the one shared model:
//   * a lone caller (demucs.cpp / demucs_ft.cpp / demucs_v3.cpp) gets every replica and its track is split over them
//     (dmx_separate_multi);
//   * the chunk threads of the *_mt drivers (cli-apps/threaded_inference.hpp:110-123) arrive together: each takes ONE
//     free replica and runs its whole chunk there, so n chunks are in flight on n GPUs instead of queueing on one
//     context; with more threads than replicas the rest wait for the next free one.
struct Handle
{
    std::vector<dmx_model *> reps;
    std::mutex mu;
    std::condition_variable cv;
    std::vector<char> busy;
    int callers = 0;
    ~Handle()
    {
        for (dmx_model *h : reps)
            dmx_model_free(h);
    }
    std::vector<int> acquire()
    {
        std::unique_lock<std::mutex> lk(mu);
        ++callers;
        if (reps.size() > 1)
        {
            // sibling chunk threads start microseconds apart: let them register before deciding who gets what
            lk.unlock();
            std::this_thread::sleep_for(std::chrono::milliseconds(2));
            lk.lock();
        }
        cv.wait(lk, [&] { return std::find(busy.begin(), busy.end(), 0) != busy.end(); });
        std::vector<int> mine;
        const bool all_free = std::find(busy.begin(), busy.end(), 1) == busy.end();
        for (int r = 0; r < (int)reps.size(); ++r)
            if (!busy[r] && (mine.empty() || (callers == 1 && all_free)))
                mine.push_back(r);
        for (int r : mine)
            busy[r] = 1;
        return mine;
    }
    void release(const std::vector<int> &mine)
    {
        {
            std::lock_guard<std::mutex> lk(mu);
            for (int r : mine)
                busy[r] = 0;
            --callers;
        }
        cv.notify_all();
    }
};

std::shared_ptr<void> load(const std::string &path, int want_v3, bool *is4)
{
    // DEMUCS_DEVICE / DEMUCS_MAX_BATCH / DEMUCS_PRECISION (bf16 | fp32): the CLIs take positional arguments only.  Defaults
    // are the benchmarked configuration: bf16 tensor-core mode (stems >= 40 dB vs the reference), 16 segments per launch.
    dmx_model *h = dmx_model_load_file(path.c_str(), env_int("DEMUCS_DEVICE", 0), env_int("DEMUCS_MAX_BATCH", 16));
    if (!h)
    {
        std::cerr << "load_demucs_model: " << dmx_last_error() << std::endl;
        return nullptr;
    }
    const bool v3 = dmx_model_arch(h) == DMX_ARCH_HDEMUCS_V3;
    if (v3 != (want_v3 != 0))
    {
        std::cerr << "load_demucs_model: invalid model data (bad magic)" << std::endl; // src/model_load.cpp:97-102,1331-1336
        dmx_model_free(h);
        return nullptr;
    }
    const char *prec = std::getenv("DEMUCS_PRECISION");
    if (!prec || std::string(prec) != "fp32")
        dmx_model_set_precision(h, DMX_PRECISION_BF16);
    if (is4)
        *is4 = dmx_model_sources(h) == 4;
    Handle *hd = new Handle();
    hd->reps.push_back(h);
    // replicas: each GPU parses and repacks the file on a host thread of its own (0.3 s per htdemucs replica, side by side)
    const int ngpu = std::min(env_int("DEMUCS_NGPU", 1), dmx_device_count());
    std::vector<dmx_model *> extra(ngpu > 1 ? ngpu - 1 : 0, nullptr);
    std::vector<std::string> errs(extra.size());
    std::vector<std::thread> loaders;
    for (int r = 1; r < ngpu; ++r)
        loaders.emplace_back([&, r] {
            extra[r - 1] = dmx_model_load_file(path.c_str(), (env_int("DEMUCS_DEVICE", 0) + r) % dmx_device_count(),
                                               env_int("DEMUCS_MAX_BATCH", 16));
            if (!extra[r - 1])
                errs[r - 1] = dmx_last_error(); // the error string is per thread
            else if (!prec || std::string(prec) != "fp32")
                dmx_model_set_precision(extra[r - 1], DMX_PRECISION_BF16);
        });
    for (auto &t : loaders)
        t.join();
    for (size_t i = 0; i < extra.size(); ++i)
        if (extra[i])
            hd->reps.push_back(extra[i]);
    for (size_t i = 0; i < extra.size(); ++i)
        if (!extra[i])
        {
            std::cerr << "load_demucs_model (replica " << i + 1 << "): " << errs[i] << std::endl;
            delete hd;
            return nullptr;
        }
    hd->busy.assign(hd->reps.size(), 0);
    return std::shared_ptr<void>(hd, [](void *p) { delete static_cast<Handle *>(p); });
}

void trampoline(float p, const char *msg, void *user)
{
    (*static_cast<demucscpp::ProgressCallback *>(user))(p, std::string(msg));
}

Eigen::Tensor3dXf separate(void *handle, const Eigen::MatrixXf &audio, demucscpp::ProgressCallback cb)
{
    Handle *hd = static_cast<Handle *>(handle);
    dmx_model *h = hd ? hd->reps[0] : nullptr;
    if (!h || audio.rows() != 2)
    {
        std::cerr << "demucs_inference: model not loaded or audio is not (2, L)" << std::endl;
        std::exit(1);
    }
    const int S = dmx_model_sources(h);
    Eigen::Tensor3dXf out(S, 2, audio.cols());
    const int offset = rand() % 22050; // src/model_apply.cpp:114 (unseeded glibc rand, like the reference)
    std::cout << "1., apply model w/ shift, offset: " << offset << std::endl;
    const std::vector<int> mine = hd->acquire();
    std::vector<dmx_model *> reps;
    for (int r : mine)
        reps.push_back(hd->reps[r]);
    const int rc = dmx_separate_multi(reps.data(), (int)reps.size(), audio.data(), (int64_t)audio.cols(), offset, out.data(),
                                      cb ? trampoline : nullptr, &cb);
    hd->release(mine);
    if (rc != 0)
    {
        std::cerr << "demucs_inference: " << dmx_last_error() << std::endl;
        std::exit(1);
    }
    if (std::getenv("DEMUCS_VERBOSE_DEVICES"))
    {
        std::string ids;
        for (int r : mine)
            ids += (ids.empty() ? "" : ",") + std::to_string(r);
        std::cerr << "demucs_inference: " << audio.cols() << " frames on replica(s) " << ids << std::endl;
    }
    return out;
}

// ---- segment tier ----
struct SegDims
{
    int le, pad, pad_end, padded, frames, bins;
};

SegDims seg_dims(int segment_samples)
{
    // the 1/4-hop re-padding the model applies around its STFT (src/model.hpp:622-628)
    const int hop = demucscpp::FFT_HOP_SIZE;
    SegDims d;
    d.le = (segment_samples + hop - 1) / hop;
    d.pad = hop / 2 * 3;
    d.pad_end = d.pad + d.le * hop - segment_samples;
    d.padded = segment_samples + d.pad + d.pad_end;
    d.frames = segment_samples / hop + 1;
    d.bins = demucscpp::FFT_WINDOW_SIZE / 2 + 1;
    return d;
}

// A stage activation of the last forward -> the reference's host tensor.  The device keeps activations channels-last,
// (d1, d2, C); the reference's frequency-branch tensors are Tensor3dXf(C, d1, d2) column-major, i.e. [d2][d1][C] in memory, and
// its time-branch tensors are (1, C, T) column-major = [T][C] -- the tap as it is.
void fetch(dmx_model *h, const char *name, Eigen::Tensor3dXf &dst, bool freq_layout)
{
    int64_t dims[4] = {0, 0, 0, 0};
    const int64_t n = dmx_debug_tap(h, name, nullptr, 0, dims);
    if (n < 0)
        return; // this architecture does not keep that stage (e.g. the 6-source model's channel up-sampler)
    const int64_t d1 = dims[1], d2 = dims[2], C = dims[3];
    const bool ok = freq_layout ? (dst.dimension(0) == C && dst.dimension(1) == d1 && dst.dimension(2) == d2)
                                : (d1 == 1 && dst.dimension(0) == 1 && dst.dimension(1) == C && dst.dimension(2) == d2);
    if (!ok || dims[0] != 1)
    {
        std::cerr << "model_inference: stage " << name << " is (" << d1 << ", " << d2 << ", " << C
                  << ") on the device and does not fit the host tensor" << std::endl;
        std::exit(1);
    }
    if (!freq_layout || d1 == 1)
    {
        dmx_debug_tap(h, name, dst.data(), n, dims);
        return;
    }
    std::vector<float> tmp((size_t)n);
    dmx_debug_tap(h, name, tmp.data(), n, dims);
    float *out = dst.data();
    for (int64_t i = 0; i < d1; ++i)
        for (int64_t j = 0; j < d2; ++j)
            std::copy(tmp.begin() + (i * d2 + j) * C, tmp.begin() + (i * d2 + j + 1) * C, out + (j * d1 + i) * C);
}

// mix (2, L) column-major is the interleaved segment the ABI takes; the ABI returns (S, L, 2), targets_out is
// Tensor3dXf(S, 2, L) column-major = [L][2][S].
void run_segment(void *handle, const Eigen::MatrixXf &mix, Eigen::Tensor3dXf &targets_out, demucscpp::ProgressCallback &cb,
                 float current_progress, float segment_progress, dmx_model **used)
{
    Handle *hd = static_cast<Handle *>(handle);
    dmx_model *h = hd ? hd->reps[0] : nullptr;
    const int S = h ? dmx_model_sources(h) : 0;
    const int64_t L = mix.cols();
    if (!h || mix.rows() != 2 || targets_out.dimension(0) != S || targets_out.dimension(1) != 2 || targets_out.dimension(2) != L)
    {
        std::cerr << "model_inference: model not loaded, or buffers built for another segment length / source count" << std::endl;
        std::exit(1);
    }
    std::vector<float> out((size_t)S * L * 2);
    if (dmx_forward_segments(h, mix.data(), 1, out.data()) != 0)
    {
        std::cerr << "model_inference: " << dmx_last_error() << std::endl;
        std::exit(1);
    }
    float *t = targets_out.data();
    for (int s = 0; s < S; ++s)
        for (int64_t i = 0; i < L; ++i)
        {
            t[(i * 2 + 0) * S + s] = out[((size_t)s * L + i) * 2 + 0];
            t[(i * 2 + 1) * S + s] = out[((size_t)s * L + i) * 2 + 1];
        }
    if (cb)
        dmx_segment_progress(h, current_progress, segment_progress, trampoline, &cb);
    *used = h;
}
} // namespace

demucscpp::demucs_segment_buffers::demucs_segment_buffers(int nb_channels, int segment_samples_, int nb_sources)
    : segment_samples(segment_samples_)
{
    const SegDims d = seg_dims(segment_samples);
    le = d.le;
    pad = d.pad;
    pad_end = d.pad_end;
    padded_segment_samples = d.padded;
    nb_stft_frames = d.frames;
    nb_stft_bins = d.bins;
    mix.resize(nb_channels, segment_samples);
    targets_out.resize(nb_sources, nb_channels, segment_samples);
    padded_mix.resize(nb_channels, padded_segment_samples);
    z.resize(nb_channels, nb_stft_bins, nb_stft_frames);
    const int T = FREQ_BRANCH_LEN, cac = 2 * nb_channels;
    x.resize(cac, nb_stft_bins - 1, nb_stft_frames);
    x_out.resize(nb_sources * cac, nb_stft_bins - 1, nb_stft_frames);
    xt.resize(1, nb_channels, segment_samples);
    xt_out.resize(1, nb_sources * nb_channels, segment_samples);
    // level i: channels 48 << i, 2048 >> (2 i + 2) frequency rows, time-branch length ceil(previous / 4)
    Eigen::Tensor3dXf *fx[4] = {&x_0, &x_1, &x_2, &x_3}, *fs[4] = {&saved_0, &saved_1, &saved_2, &saved_3};
    Eigen::Tensor3dXf *tx[4] = {&xt_0, &xt_1, &xt_2, &xt_3}, *ts[4] = {&savedt_0, &savedt_1, &savedt_2, &savedt_3};
    const int tl[4] = {TIME_BRANCH_LEN_0, TIME_BRANCH_LEN_1, TIME_BRANCH_LEN_2, TIME_BRANCH_LEN_3};
    for (int i = 0; i < 4; ++i)
    {
        const int ch = 48 << i, rows = (nb_stft_bins - 1) >> (2 * i + 2);
        fx[i]->resize(ch, rows, T);
        fs[i]->resize(ch, rows, T);
        tx[i]->resize(1, ch, tl[i]);
        ts[i]->resize(1, ch, tl[i]);
    }
    x_3_channel_upsampled.resize(512, 8, T);
    xt_3_channel_upsampled.resize(1, 512, TIME_BRANCH_LEN_3);
}

void demucscpp::model_inference(const struct demucs_model &model, struct demucs_segment_buffers &b, struct stft_buffers &,
                                ProgressCallback cb, float current_progress, float segment_progress)
{
    dmx_model *h = nullptr;
    run_segment(model.b200.get(), b.mix, b.targets_out, cb, current_progress, segment_progress, &h);
    fetch(h, "x", b.x, true);
    fetch(h, "xt", b.xt, false);
    Eigen::Tensor3dXf *fs[4] = {&b.saved_0, &b.saved_1, &b.saved_2, &b.saved_3};
    Eigen::Tensor3dXf *ts[4] = {&b.savedt_0, &b.savedt_1, &b.savedt_2, &b.savedt_3};
    Eigen::Tensor3dXf *fx[4] = {&b.x_0, &b.x_1, &b.x_2, &b.x_3}, *tx[4] = {&b.xt_0, &b.xt_1, &b.xt_2, &b.xt_3};
    for (int i = 0; i < 4; ++i)
    {
        const std::string si = std::to_string(i);
        fetch(h, ("saved_" + si).c_str(), *fs[i], true);
        fetch(h, ("savedt_" + si).c_str(), *ts[i], false);
        if (dmx_model_sources(h) == 4 || i < 3) // 6 sources: the bottleneck stays 384 wide end to end; x_3 == saved_3's shape
        {
            fetch(h, ("x_" + si).c_str(), *fx[i], true);
            fetch(h, ("xt_" + si).c_str(), *tx[i], false);
        }
    }
    if (dmx_model_sources(h) == 4)
    {
        fetch(h, "x_3_up", b.x_3_channel_upsampled, true);
        fetch(h, "xt_3_up", b.xt_3_channel_upsampled, false);
    }
    else
    {
        fetch(h, "x_3_up", b.x_3, true);
        fetch(h, "xt_3_up", b.xt_3, false);
    }
    fetch(h, "x_out", b.x_out, true);
    fetch(h, "xt_out", b.xt_out, false);
}

demucscpp_v3::demucs_v3_segment_buffers::demucs_v3_segment_buffers(int nb_channels, int segment_samples_, int nb_sources)
    : segment_samples(segment_samples_)
{
    const SegDims d = seg_dims(segment_samples);
    le = d.le;
    pad = d.pad;
    pad_end = d.pad_end;
    padded_segment_samples = d.padded;
    nb_stft_frames = d.frames;
    nb_stft_bins = d.bins;
    mix.resize(nb_channels, segment_samples);
    targets_out.resize(nb_sources, nb_channels, segment_samples);
    const int T = demucscpp::FREQ_BRANCH_LEN, cac = 2 * nb_channels;
    x.resize(cac, nb_stft_bins - 1, nb_stft_frames);
    x_out.resize(nb_sources * cac, nb_stft_bins - 1, nb_stft_frames);
    xt.resize(1, nb_channels, segment_samples);
    xt_out.resize(1, nb_sources * nb_channels, segment_samples);
    Eigen::Tensor3dXf *fx[4] = {&x_0, &x_1, &x_2, &x_3}, *fs[4] = {&saved_0, &saved_1, &saved_2, &saved_3};
    Eigen::Tensor3dXf *tx[4] = {&xt_0, &xt_1, &xt_2, &xt_3}, *ts[4] = {&savedt_0, &savedt_1, &savedt_2, &savedt_3};
    const int tl[4] = {demucscpp::TIME_BRANCH_LEN_0, demucscpp::TIME_BRANCH_LEN_1, demucscpp::TIME_BRANCH_LEN_2,
                       demucscpp::TIME_BRANCH_LEN_3};
    for (int i = 0; i < 4; ++i)
    {
        const int ch = 48 << i, rows = (nb_stft_bins - 1) >> (2 * i + 2);
        fx[i]->resize(ch, rows, T);
        fs[i]->resize(ch, rows, T);
        tx[i]->resize(1, ch, tl[i]);
        ts[i]->resize(1, ch, tl[i]);
    }
    x_4.resize(768, 1, T);
    saved_4.resize(768, 1, T);
    x_shared_5.resize(1, 1536, T / 2);
}

void demucscpp_v3::model_v3_inference(const struct demucs_v3_model &model, struct demucs_v3_segment_buffers &b,
                                      struct demucscpp::stft_buffers &, demucscpp::ProgressCallback cb, float current_progress,
                                      float segment_progress)
{
    dmx_model *h = nullptr;
    run_segment(model.b200.get(), b.mix, b.targets_out, cb, current_progress, segment_progress, &h);
    fetch(h, "x", b.x, true);
    fetch(h, "xt", b.xt, false);
    Eigen::Tensor3dXf *fs[5] = {&b.saved_0, &b.saved_1, &b.saved_2, &b.saved_3, &b.saved_4};
    Eigen::Tensor3dXf *ts[4] = {&b.savedt_0, &b.savedt_1, &b.savedt_2, &b.savedt_3};
    Eigen::Tensor3dXf *fx[5] = {&b.x_0, &b.x_1, &b.x_2, &b.x_3, &b.x_4}, *tx[4] = {&b.xt_0, &b.xt_1, &b.xt_2, &b.xt_3};
    for (int i = 0; i < 5; ++i)
    {
        const std::string si = std::to_string(i);
        fetch(h, ("saved_" + si).c_str(), *fs[i], true);
        fetch(h, ("x_" + si).c_str(), *fx[i], true);
        if (i < 4)
        {
            fetch(h, ("savedt_" + si).c_str(), *ts[i], false);
            fetch(h, ("xt_" + si).c_str(), *tx[i], false);
        }
    }
    fetch(h, "x_shared_5", b.x_shared_5, false);
    fetch(h, "x_out", b.x_out, true);
    fetch(h, "xt_out", b.xt_out, false);
}

bool demucscpp::load_demucs_model(const std::string &model_file, struct demucs_model *model)
{
    model->b200 = load(model_file, 0, &model->is_4sources);
    return (bool)model->b200;
}

Eigen::Tensor3dXf demucscpp::demucs_inference(const struct demucs_model &model, const Eigen::MatrixXf &full_audio,
                                              ProgressCallback cb)
{
    return separate(model.b200.get(), full_audio, cb);
}

bool demucscpp_v3::load_demucs_v3_model(const std::string &model_file, struct demucs_v3_model *model)
{
    model->b200 = load(model_file, 1, nullptr);
    return (bool)model->b200;
}

Eigen::Tensor3dXf demucscpp_v3::demucs_v3_inference(const struct demucs_v3_model &model,
                                                    const Eigen::MatrixXf &full_audio, demucscpp::ProgressCallback cb)
{
    return separate(model.b200.get(), full_audio, cb);
}
