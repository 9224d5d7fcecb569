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
} // namespace demucscpp_v3

#endif
