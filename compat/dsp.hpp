// compat/dsp.hpp -- constants of the reference's dsp.hpp that callers use (src/dsp.hpp:15-18) and the stft_buffers type name.
#ifndef DEMUCS_B200_COMPAT_DSP_HPP
#define DEMUCS_B200_COMPAT_DSP_HPP

#include "tensor.hpp"

namespace demucscpp
{
const int SUPPORTED_SAMPLE_RATE = 44100;
const int FFT_WINDOW_SIZE = 4096;
const int FFT_HOP_SIZE = 1024;

// Placeholder for the reference's host STFT scratch (src/dsp.hpp:20-101) so that model_inference keeps its signature
// (src/model.hpp:662-666).  The device path owns its STFT workspace; demucscpp::stft / istft on this struct are not
// mirrored (the C ABI's dmx_stft_segment / dmx_istft_segment are the stand-alone DSP pair, include/demucs_b200.h).
struct stft_buffers
{
    int nb_frames, nb_bins, pad;
    Eigen::MatrixXf waveform;
    explicit stft_buffers(int n_samples)
        : nb_frames(n_samples / FFT_HOP_SIZE + 1), nb_bins(FFT_WINDOW_SIZE / 2 + 1), pad(FFT_WINDOW_SIZE / 2), waveform(2, n_samples)
    {
    }
};
} // namespace demucscpp

#endif
