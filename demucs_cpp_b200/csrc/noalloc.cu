// dmx_stft_segments_noalloc: the batched segment STFT for callers that must not allocate.
//
// Who needs this: code that runs where cudaMalloc / cudaFree are off limits -- a real-time audio
// callback or another latency-critical thread (cudaMalloc may take locks and synchronise the
// device), a host that hands this library a fixed pre-reserved pool and forbids any further device
// allocation, or code that has already started a region in which allocation calls are not
// permitted.  Such callers own their input and output device buffers; the only thing the STFT
// kernel still needs is its small statistics workspace (sum and sum of squares per segment, which
// the normalisation after the STFT consumes).  Everywhere else in the library that workspace comes
// from cudaMalloc at model-load time; here it is a fixed array that is part of the library image.
//
// The workspace is g_noalloc_stats below: DMX_NOALLOC_MAX_SEGMENTS x 2 doubles = 1 KiB of device
// memory on every GPU that loads this library, held until the process exits.
//
// Rules that follow from sharing one static array:
//   * every call on a device uses the same array and finds the previous call's numbers in it, so each
//     call zeroes the part it is going to use (cudaMemsetAsync on the stream, no allocation);
//   * two calls whose GPU work overlaps (different streams, or different host threads) race on it;
//     serialise them;
//   * the array's device address is looked up once per device (cudaGetSymbolAddress, with that device
//     current).  dmx_model_finalize does this for the model's device -- and touches the helper kernel so
//     that the CUDA module is already loaded -- which keeps module loading (it may allocate and
//     synchronise) out of the caller's no-allocation region.  Should a call still be the first one on its
//     device, it does the look-up itself: make that call before starting any stream capture.
#include "model.cuh"

#include <mutex>

namespace dmx
{

// In host code this name is only a handle for the symbol, NOT a device pointer: passing it to a kernel
// or to cudaMemset would fault (or quietly touch the wrong memory where the GPU can reach host
// memory).  Host code must go through cudaGetSymbolAddress; device code may name it directly.
static __device__ __align__(256) double g_noalloc_stats[2 * DMX_NOALLOC_MAX_SEGMENTS];

constexpr int kMaxDevices = 64;
static double *g_stats_addr[kMaxDevices] = {nullptr};
static std::mutex g_setup_mtx;

// Set-up, once per device, with that device current: resolve the device address of the static array.
static int noalloc_setup(int device, double **out)
{
    if (device < 0 || device >= kMaxDevices)
    {
        set_error("dmx_stft_segments_noalloc: device index out of range");
        return -1;
    }
    std::lock_guard<std::mutex> lk(g_setup_mtx);
    if (!g_stats_addr[device])
    {
        double *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_noalloc_stats);
        if (e != cudaSuccess || !p)
        {
            set_error(std::string("dmx_stft_segments_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
            return -1;
        }
        g_stats_addr[device] = p;
    }
    *out = g_stats_addr[device];
    return 0;
}

// (mean, std) per segment from the accumulated statistics, unbiased variance like the reference's
// input normalisation (src/model_inference.cpp:115-118).  Device code reads the array by name.
__global__ void noalloc_meanstd_kernel(int n, float2 *__restrict__ meanstd)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n)
        return;
    const double cnt = (double)NBINS * LE * 4;
    const double s = g_noalloc_stats[2 * b], ss = g_noalloc_stats[2 * b + 1];
    const double m = s / cnt;
    double var = (ss - s * m) / (cnt - 1.0);
    if (var < 0)
        var = 0;
    meanstd[b] = make_float2((float)m, (float)sqrt(var));
}

// Called by model_finalize with the model's device current: resolves the workspace address and forces the
// module that holds the helper kernel to load now, so that no later call has to.
int noalloc_warm_up(int device)
{
    double *p = nullptr;
    if (noalloc_setup(device, &p))
        return -1;
    cudaFuncAttributes attr;
    DMX_CUDA(cudaFuncGetAttributes(&attr, noalloc_meanstd_kernel));
    return 0;
}

} // namespace dmx

using namespace dmx;

extern "C" int dmx_stft_segments_noalloc(dmx_model *m, const float *d_segs, int n_segments, float *d_spec,
                                         float *d_meanstd, void *stream)
{
    if (!m || !d_segs || !d_spec)
    {
        set_error("dmx_stft_segments_noalloc: null argument");
        return -1;
    }
    if (n_segments < 1 || n_segments > DMX_NOALLOC_MAX_SEGMENTS)
    {
        set_error("dmx_stft_segments_noalloc: n_segments must be between 1 and " +
                  std::to_string(DMX_NOALLOC_MAX_SEGMENTS) +
                  " (the static workspace is fixed at that size); split the batch");
        return -1;
    }
    std::lock_guard<std::mutex> lk(m->mtx); // the model's tables and stream are shared with its other entry points
    if (!m->finalized)
    {
        set_error("dmx_stft_segments_noalloc: model not finalized");
        return -1;
    }
    // the caller's current device is put back on every way out
    struct DeviceGuard
    {
        int prev = -1;
        DeviceGuard() { cudaGetDevice(&prev); }
        ~DeviceGuard()
        {
            if (prev >= 0)
                cudaSetDevice(prev);
        }
    } guard;
    DMX_CUDA(cudaSetDevice(m->device));
    double *stats = nullptr;
    if (noalloc_setup(m->device, &stats))
        return -1;
    cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : m->stream;
    // the array still holds the previous call's sums: the kernel accumulates with atomics, so clear first
    DMX_CUDA(cudaMemsetAsync(stats, 0, 2 * sizeof(double) * (size_t)n_segments, st));
    if (launch_stft(d_segs, n_segments, m->window, m->tw, d_spec, stats, st))
        return -1;
    if (d_meanstd)
    {
        noalloc_meanstd_kernel<<<(n_segments + 63) / 64, 64, 0, st>>>(n_segments, reinterpret_cast<float2 *>(d_meanstd));
        DMX_LAUNCH_CHECK();
    }
    if (!stream)
        DMX_CUDA(cudaStreamSynchronize(st));
    return 0;
}
