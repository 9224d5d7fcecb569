// TEST INFRASTRUCTURE ONLY -- never linked into, called from, or shipped with the
// product library.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load oracle/_ref/libdemucs_ref.so.
//
// A thin extern "C" harness around the UNMODIFIED reference library
// (sevagh/demucs.cpp, sources compiled in place from /root/reference by
// oracle/Makefile).  It exposes, for ctypes:
//   * the track-level entry points      demucscpp::demucs_inference      (src/model_apply.cpp:60-91)
//                                       demucscpp_v3::demucs_v3_inference (src/model_apply.cpp:305-336)
//   * the per-segment forward           demucscpp::model_inference       (src/model_inference.cpp:48-475)
//                                       demucscpp_v3::model_v3_inference (src/model_inference.cpp:477-856)
//     plus named views of the per-segment scratch struct (src/model.hpp:569-647)
//     after the call, which serve as stage-level golden tensors;
//   * the DSP pair demucscpp::stft / istft (src/dsp.cpp:51-117).
// All tensors cross this boundary as raw fp32 in the reference's own Eigen
// column-major order (first index fastest), see src/tensor.hpp:24-28.
#include "dsp.hpp"
#include "model.hpp"
#include "tensor.hpp"

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>

namespace
{
struct ref_handle
{
    int arch = 0; // 4, 6 (htdemucs) or 3 (hdemucs_mmi)
    demucscpp::demucs_model m4;
    demucscpp_v3::demucs_v3_model m3;
    std::unique_ptr<demucscpp::demucs_segment_buffers> buf4;
    std::unique_ptr<demucscpp_v3::demucs_v3_segment_buffers> buf3;
    std::unique_ptr<demucscpp::stft_buffers> stft;
    std::map<std::string, const Eigen::Tensor3dXf *> views;
};

const int SEG = 343980; // int(7.8 * 44100), src/model_apply.cpp:145

void quiet_cb(float, const std::string &) {}

template <typename T> void copy_out(const T &t, float *dst)
{
    std::memcpy(dst, t.data(), sizeof(float) * (size_t)t.size());
}
} // namespace

extern "C"
{

// Loads a ggml-f16 file; the magic selects the architecture
// (src/model_load.cpp:77-102, :1331).  Returns NULL on failure.
void *ref_load(const char *path)
{
    uint32_t magic = 0;
    FILE *f = fopen(path, "rb");
    if (!f)
        return nullptr;
    if (fread(&magic, 4, 1, f) != 1)
    {
        fclose(f);
        return nullptr;
    }
    fclose(f);
    auto *h = new ref_handle();
    bool ok = false;
    if (magic == 0x646d6333)
    {
        h->arch = 3;
        ok = demucscpp_v3::load_demucs_v3_model(path, &h->m3);
    }
    else
    {
        ok = demucscpp::load_demucs_model(path, &h->m4);
        h->arch = h->m4.is_4sources ? 4 : 6;
    }
    if (!ok)
    {
        delete h;
        return nullptr;
    }
    return h;
}

void ref_free(void *hv) { delete static_cast<ref_handle *>(hv); }

int ref_arch(void *hv) { return static_cast<ref_handle *>(hv)->arch; }

int ref_nb_sources(void *hv)
{
    return static_cast<ref_handle *>(hv)->arch == 6 ? 6 : 4;
}

// Whole-track separation.  audio: (2,L) column-major = L,R interleaved.
// out: (S,2,L) column-major.  `seed` re-seeds glibc rand() so the random shift
// (src/model_apply.cpp:114) is reproducible: srand(1) gives offset 4033, the
// value an unseeded fresh process draws.  *shift_out receives rand()%22050 that
// the call will draw (computed by peeking with the same seed).
int ref_separate(void *hv, const float *audio, int64_t L, float *out,
                 unsigned seed, int *shift_out)
{
    auto *h = static_cast<ref_handle *>(hv);
    Eigen::MatrixXf a =
        Eigen::Map<const Eigen::MatrixXf>(audio, 2, (Eigen::Index)L);
    srand(seed);
    if (shift_out)
    {
        *shift_out = rand() % 22050;
        srand(seed);
    }
    Eigen::Tensor3dXf r =
        h->arch == 3 ? demucscpp_v3::demucs_v3_inference(h->m3, a, quiet_cb)
                     : demucscpp::demucs_inference(h->m4, a, quiet_cb);
    copy_out(r, out);
    return 0;
}

// One full 7.8 s segment through model_inference.  mix: (2,343980) column-major.
// targets: (S,2,343980) column-major.  Afterwards ref_view() exposes the scratch
// tensors the forward left behind.
int ref_segment(void *hv, const float *mix, float *targets)
{
    auto *h = static_cast<ref_handle *>(hv);
    int S = ref_nb_sources(hv);
    h->views.clear();
    if (h->arch == 3)
    {
        if (!h->buf3)
        {
            h->buf3 =
                std::make_unique<demucscpp_v3::demucs_v3_segment_buffers>(2, SEG,
                                                                         S);
            h->stft = std::make_unique<demucscpp::stft_buffers>(
                h->buf3->padded_segment_samples);
        }
        auto &b = *h->buf3;
        b.mix = Eigen::Map<const Eigen::MatrixXf>(mix, 2, SEG);
        demucscpp_v3::model_v3_inference(h->m3, b, *h->stft, quiet_cb, 0.f, 1.f);
        copy_out(b.targets_out, targets);
        h->views = {{"x", &b.x},
                    {"x_out", &b.x_out},
                    {"xt_out", &b.xt_out},
                    {"saved_0", &b.saved_0},
                    {"saved_1", &b.saved_1},
                    {"saved_2", &b.saved_2},
                    {"saved_3", &b.saved_3},
                    {"saved_4", &b.saved_4},
                    {"savedt_0", &b.savedt_0},
                    {"savedt_1", &b.savedt_1},
                    {"savedt_2", &b.savedt_2},
                    {"savedt_3", &b.savedt_3},
                    {"x_shared_5", &b.x_shared_5}};
    }
    else
    {
        if (!h->buf4)
        {
            h->buf4 =
                std::make_unique<demucscpp::demucs_segment_buffers>(2, SEG, S);
            h->stft = std::make_unique<demucscpp::stft_buffers>(
                h->buf4->padded_segment_samples);
        }
        auto &b = *h->buf4;
        b.mix = Eigen::Map<const Eigen::MatrixXf>(mix, 2, SEG);
        demucscpp::model_inference(h->m4, b, *h->stft, quiet_cb, 0.f, 1.f);
        copy_out(b.targets_out, targets);
        // After a full forward: saved_i/savedt_i = encoder outputs; x_3/xt_3 =
        // post-transformer bottleneck; x_2,x_1,x_0 / xt_2,xt_1,xt_0 = decoder
        // 0,1,2 outputs (the decoders reuse the encoder slots,
        // src/model_inference.cpp:314-343); x_out/xt_out = decoder 3 outputs.
        h->views = {{"x", &b.x},
                    {"xt", &b.xt},
                    {"saved_0", &b.saved_0},
                    {"saved_1", &b.saved_1},
                    {"saved_2", &b.saved_2},
                    {"saved_3", &b.saved_3},
                    {"savedt_0", &b.savedt_0},
                    {"savedt_1", &b.savedt_1},
                    {"savedt_2", &b.savedt_2},
                    {"savedt_3", &b.savedt_3},
                    {"x_3", &b.x_3},
                    {"xt_3", &b.xt_3},
                    {"x_2", &b.x_2},
                    {"xt_2", &b.xt_2},
                    {"x_1", &b.x_1},
                    {"xt_1", &b.xt_1},
                    {"x_0", &b.x_0},
                    {"xt_0", &b.xt_0},
                    {"x_out", &b.x_out},
                    {"xt_out", &b.xt_out}};
        if (h->arch == 4)
        {
            h->views["x_3_up"] = &b.x_3_channel_upsampled;
            h->views["xt_3_up"] = &b.xt_3_channel_upsampled;
        }
    }
    return 0;
}

// Shape (3 dims) of a named scratch tensor from the last ref_segment; 0 if known.
int ref_view_dims(void *hv, const char *name, int64_t *dims)
{
    auto *h = static_cast<ref_handle *>(hv);
    auto it = h->views.find(name);
    if (it == h->views.end())
        return -1;
    for (int i = 0; i < 3; ++i)
        dims[i] = it->second->dimension(i);
    return 0;
}

int ref_view_copy(void *hv, const char *name, float *dst)
{
    auto *h = static_cast<ref_handle *>(hv);
    auto it = h->views.find(name);
    if (it == h->views.end())
        return -1;
    copy_out(*it->second, dst);
    return 0;
}

// demucscpp::stft (src/dsp.cpp:51-86).  wave: (2,n) column-major; spec_out:
// interleaved re,im of the (2, 2049, n/1024+1) column-major complex tensor.
int ref_stft(const float *wave, int n, float *spec_out)
{
    demucscpp::stft_buffers sb(n);
    sb.waveform = Eigen::Map<const Eigen::MatrixXf>(wave, 2, n);
    demucscpp::stft(sb);
    std::memcpy(spec_out, sb.spec.data(),
                sizeof(std::complex<float>) * (size_t)sb.spec.size());
    return 0;
}

// demucscpp::istft (src/dsp.cpp:88-117); inverse layout of ref_stft.
int ref_istft(const float *spec, int n, float *wave_out)
{
    demucscpp::stft_buffers sb(n);
    std::memcpy(sb.spec.data(), spec,
                sizeof(std::complex<float>) * (size_t)sb.spec.size());
    demucscpp::istft(sb);
    std::memcpy(wave_out, sb.waveform.data(), sizeof(float) * 2 * (size_t)n);
    return 0;
}

} // extern "C"
