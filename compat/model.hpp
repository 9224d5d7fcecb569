// compat/model.hpp -- drop-in for the reference's src/model.hpp at the CLI tier:
// the declarations cli-apps/*.cpp and cli-apps/threaded_inference.hpp use, implemented in
// compat/model_b200.cpp on top of the C ABI of include/demucs_b200.h.  The model structs carry only
// `is_4sources` (read by cli-apps/demucs.cpp:145) and a shared handle to the device-resident model; they
// stay default-constructible, movable and usable as `const&` from several host threads
// (cli-apps/threaded_inference.hpp:110-117): calls on one device-side model serialise on its mutex inside the library;
// with DEMUCS_NGPU replicas concurrent callers are spread over the GPUs (compat/model_b200.cpp, Handle).
//
// Replaces: demucscpp::load_demucs_model        src/model.hpp:649-650
//           demucscpp::demucs_inference         src/model.hpp:658-660
//           demucscpp_v3::load_demucs_v3_model  src/model.hpp:1397-1398
//           demucscpp_v3::demucs_v3_inference   src/model.hpp:1406-1409
//           demucscpp::ProgressCallback         src/model.hpp:17
//           demucscpp::demucs_segment_buffers, model_inference            src/model.hpp:569-647,662-666
//           demucscpp_v3::demucs_v3_segment_buffers, model_v3_inference   src/model.hpp:1213-1395,1411-1415
#ifndef DEMUCS_B200_COMPAT_MODEL_HPP
#define DEMUCS_B200_COMPAT_MODEL_HPP

#include "dsp.hpp"
#include "tensor.hpp"
#include <Eigen/Dense>
#include <functional>
#include <memory>
#include <string>

namespace demucscpp
{
using ProgressCallback = std::function<void(float, const std::string &)>;

struct demucs_model
{
    bool is_4sources = true;
    std::shared_ptr<void> b200; // dmx_model*, freed with dmx_model_free
};

bool load_demucs_model(const std::string &model_file, struct demucs_model *model);

const float SEGMENT_LEN_SECS = 7.8;
const float SEGMENT_OVERLAP_SECS = 0.25;
const float MAX_SHIFT_SECS = 0.5;
const float OVERLAP = 0.25;
const float TRANSITION_POWER = 1.0;

Eigen::Tensor3dXf demucs_inference(const struct demucs_model &model, const Eigen::MatrixXf &full_audio,
                                   ProgressCallback cb);

// ---- segment tier (src/model.hpp:569-647,662-666): the public per-segment scratch struct and model_inference ----
// Same member names, shapes and (column-major) layouts as the reference's struct, so code that fills `mix`, calls
// model_inference and reads `targets_out` or a stage tensor keeps working.  The tensors are HOST copies: model_inference
// runs the segment on the GPU (dmx_forward_segments) and copies the stage activations back (dmx_debug_tap).  Not mirrored:
// `z` and `padded_mix` (the padded signal and the complex spectrogram never exist as separate tensors on the device: the
// STFT kernel pads, transforms and packs in one pass) -- they are allocated with the reference's shapes and left untouched.
// State of the stage tensors after the call = what the reference leaves behind: saved_* / savedt_* are the encoder outputs,
// x_3 / xt_3 the bottleneck after the transformer, x_3_channel_upsampled / xt_3_channel_upsampled the transformer's
// output before the down-sampler (4-source models; for 6 sources they equal x_3 / xt_3), x_2 .. x_0, x_out and
// xt_2 .. xt_0, xt_out the decoder outputs.
const int FREQ_BRANCH_LEN = 336;
const int TIME_BRANCH_LEN_IN = 343980;
const int TIME_BRANCH_LEN_0 = 85995;
const int TIME_BRANCH_LEN_1 = 21499;
const int TIME_BRANCH_LEN_2 = 5375;
const int TIME_BRANCH_LEN_3 = 1344;

struct demucs_segment_buffers
{
    int segment_samples;
    int le;
    int pad;
    int pad_end;
    int padded_segment_samples;
    int nb_stft_frames;
    int nb_stft_bins;

    Eigen::MatrixXf mix;
    Eigen::Tensor3dXf targets_out;
    Eigen::MatrixXf padded_mix;
    Eigen::Tensor3dXcf z;

    Eigen::Tensor3dXf x, x_out, x_0, x_1, x_2, x_3, x_3_channel_upsampled;
    Eigen::Tensor3dXf xt, xt_out, xt_0, xt_1, xt_2, xt_3, xt_3_channel_upsampled;
    Eigen::Tensor3dXf saved_0, saved_1, saved_2, saved_3;
    Eigen::Tensor3dXf savedt_0, savedt_1, savedt_2, savedt_3;

    demucs_segment_buffers(int nb_channels, int segment_samples, int nb_sources);
};

void model_inference(const struct demucs_model &model, struct demucs_segment_buffers &buffers, struct stft_buffers &stft_buf,
                     ProgressCallback cb, float current_progress, float segment_progress);
} // namespace demucscpp

namespace demucscpp_v3
{
struct demucs_v3_model
{
    std::shared_ptr<void> b200;
};

bool load_demucs_v3_model(const std::string &model_file, struct demucs_v3_model *model);

Eigen::Tensor3dXf demucs_v3_inference(const struct demucs_v3_model &model, const Eigen::MatrixXf &full_audio,
                                      demucscpp::ProgressCallback cb);

// ---- segment tier of the v3 model (src/model.hpp:1213-1395,1411-1415).  As above; the LSTM state and the local-attention
// index / decay tables of the reference's struct are device-internal here and have no host members. ----
struct demucs_v3_segment_buffers
{
    int segment_samples;
    int le;
    int pad;
    int pad_end;
    int padded_segment_samples;
    int nb_stft_frames;
    int nb_stft_bins;

    Eigen::MatrixXf mix;
    Eigen::Tensor3dXf targets_out;

    Eigen::Tensor3dXf x, x_out, x_0, x_1, x_2, x_3, x_4, x_shared_5;
    Eigen::Tensor3dXf xt, xt_out, xt_0, xt_1, xt_2, xt_3;
    Eigen::Tensor3dXf saved_0, saved_1, saved_2, saved_3, saved_4;
    Eigen::Tensor3dXf savedt_0, savedt_1, savedt_2, savedt_3;

    demucs_v3_segment_buffers(int nb_channels, int segment_samples, int nb_sources);
};

void model_v3_inference(const struct demucs_v3_model &model, struct demucs_v3_segment_buffers &buffers,
                        struct demucscpp::stft_buffers &stft_buf, demucscpp::ProgressCallback cb, float current_progress,
                        float segment_progress);
} // namespace demucscpp_v3

#endif
