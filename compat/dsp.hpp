// compat/dsp.hpp -- constants of the reference's dsp.hpp that callers use (src/dsp.hpp:15-18).
#ifndef DEMUCS_B200_COMPAT_DSP_HPP
#define DEMUCS_B200_COMPAT_DSP_HPP

#include "tensor.hpp"

namespace demucscpp
{
const int SUPPORTED_SAMPLE_RATE = 44100;
const int FFT_WINDOW_SIZE = 4096;
const int FFT_HOP_SIZE = 1024;
} // namespace demucscpp

#endif
