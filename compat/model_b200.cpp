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
    const int ngpu = std::min(env_int("DEMUCS_NGPU", 1), dmx_device_count());
    for (int r = 1; r < ngpu; ++r)
    {
        dmx_model *rep = dmx_model_load_file(path.c_str(), (env_int("DEMUCS_DEVICE", 0) + r) % dmx_device_count(),
                                             env_int("DEMUCS_MAX_BATCH", 16));
        if (!rep)
        {
            std::cerr << "load_demucs_model (replica " << r << "): " << dmx_last_error() << std::endl;
            delete hd;
            return nullptr;
        }
        if (!prec || std::string(prec) != "fp32")
            dmx_model_set_precision(rep, DMX_PRECISION_BF16);
        hd->reps.push_back(rep);
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
} // namespace

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
