// compat/segment_tier_main.cpp -- drives the SEGMENT tier of the reference API (compat/model.hpp:
// demucs_segment_buffers + model_inference, demucs_v3_segment_buffers + model_v3_inference) the way the reference's own
// tests use it (test/test_model.cpp / test_layers_v3.cpp construct the buffers, fill `mix`, call model_inference and read
// the stage tensors): one segment in, every public tensor dumped for tests/test_gpu_cli.py.
//   segment_tier.main <model file> <v3: 0|1> <segment.f32: interleaved L,R floats> <out dir>
// Each tensor goes to <out dir>/<name>.f32 in its own (column-major) memory order; shapes.txt lists "name d0 d1 d2" and
// the progress callback's call count / last value.
#include "model.hpp"

#include <cstdio>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

namespace
{
std::ofstream shapes;
std::string outdir;

void dump(const char *name, const Eigen::Tensor3dXf &t)
{
    std::ofstream f(outdir + "/" + name + ".f32", std::ios::binary);
    f.write(reinterpret_cast<const char *>(t.data()), (std::streamsize)(t.size() * sizeof(float)));
    shapes << name << " " << t.dimension(0) << " " << t.dimension(1) << " " << t.dimension(2) << "\n";
}

template <class Buffers> bool fill_mix(Buffers &b, const char *path)
{
    std::ifstream f(path, std::ios::binary);
    f.read(reinterpret_cast<char *>(b.mix.data()), (std::streamsize)(b.mix.size() * sizeof(float)));
    return (size_t)f.gcount() == b.mix.size() * sizeof(float);
}
} // namespace

#define DUMP(b, field) dump(#field, (b).field)

int main(int argc, char **argv)
{
    if (argc != 5)
    {
        std::cerr << "usage: " << argv[0] << " <model> <v3: 0|1> <segment.f32> <out dir>" << std::endl;
        return 2;
    }
    const bool v3 = std::string(argv[2]) == "1";
    outdir = argv[4];
    shapes.open(outdir + "/shapes.txt");
    const int segment_samples = (int)(demucscpp::SEGMENT_LEN_SECS * demucscpp::SUPPORTED_SAMPLE_RATE);
    int calls = 0;
    float last = -1.f;
    std::string first_msg, last_msg;
    demucscpp::ProgressCallback cb = [&](float p, const std::string &msg) {
        if (calls++ == 0)
            first_msg = msg;
        last = p;
        last_msg = msg;
    };
    if (!v3)
    {
        demucscpp::demucs_model model;
        if (!demucscpp::load_demucs_model(argv[1], &model))
            return 1;
        demucscpp::demucs_segment_buffers buffers(2, segment_samples, model.is_4sources ? 4 : 6);
        demucscpp::stft_buffers stft_buf(buffers.padded_segment_samples);
        if (!fill_mix(buffers, argv[3]))
            return 3;
        demucscpp::model_inference(model, buffers, stft_buf, cb, 0.25f, 0.5f);
        DUMP(buffers, targets_out);
        DUMP(buffers, x);
        DUMP(buffers, xt);
        DUMP(buffers, saved_0);
        DUMP(buffers, saved_1);
        DUMP(buffers, saved_2);
        DUMP(buffers, saved_3);
        DUMP(buffers, savedt_0);
        DUMP(buffers, savedt_1);
        DUMP(buffers, savedt_2);
        DUMP(buffers, savedt_3);
        DUMP(buffers, x_3_channel_upsampled);
        DUMP(buffers, xt_3_channel_upsampled);
        DUMP(buffers, x_3);
        DUMP(buffers, x_2);
        DUMP(buffers, x_1);
        DUMP(buffers, x_0);
        DUMP(buffers, x_out);
        DUMP(buffers, xt_3);
        DUMP(buffers, xt_2);
        DUMP(buffers, xt_1);
        DUMP(buffers, xt_0);
        DUMP(buffers, xt_out);
        shapes << "# dims " << buffers.le << " " << buffers.pad << " " << buffers.pad_end << " " << buffers.padded_segment_samples << " "
               << buffers.nb_stft_frames << " " << buffers.nb_stft_bins << "\n";
    }
    else
    {
        demucscpp_v3::demucs_v3_model model;
        if (!demucscpp_v3::load_demucs_v3_model(argv[1], &model))
            return 1;
        demucscpp_v3::demucs_v3_segment_buffers buffers(2, segment_samples, 4);
        demucscpp::stft_buffers stft_buf(buffers.padded_segment_samples);
        if (!fill_mix(buffers, argv[3]))
            return 3;
        demucscpp_v3::model_v3_inference(model, buffers, stft_buf, cb, 0.25f, 0.5f);
        DUMP(buffers, targets_out);
        DUMP(buffers, x);
        DUMP(buffers, xt);
        DUMP(buffers, saved_0);
        DUMP(buffers, saved_1);
        DUMP(buffers, saved_2);
        DUMP(buffers, saved_3);
        DUMP(buffers, saved_4);
        DUMP(buffers, savedt_0);
        DUMP(buffers, savedt_1);
        DUMP(buffers, savedt_2);
        DUMP(buffers, savedt_3);
        DUMP(buffers, x_shared_5);
        DUMP(buffers, x_4);
        DUMP(buffers, x_3);
        DUMP(buffers, x_2);
        DUMP(buffers, x_1);
        DUMP(buffers, x_0);
        DUMP(buffers, x_out);
        DUMP(buffers, xt_3);
        DUMP(buffers, xt_2);
        DUMP(buffers, xt_1);
        DUMP(buffers, xt_0);
        DUMP(buffers, xt_out);
        shapes << "# dims " << buffers.le << " " << buffers.pad << " " << buffers.pad_end << " " << buffers.padded_segment_samples << " "
               << buffers.nb_stft_frames << " " << buffers.nb_stft_bins << "\n";
    }
    shapes << "# progress " << calls << " " << last << " | " << first_msg << " | " << last_msg << "\n";
    return 0;
}
