// Device-side model: packed weights, workspace and the htdemucs forward orchestration.
#pragma once
#include "kernels.cuh"
#include "demucs_b200.h"

#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace dmx
{

struct HostTensor
{
    std::vector<float> data;
    std::vector<int> shape;
    size_t numel() const { return data.size(); }
};

// A GEMM B-operand: fp32 [Kpad][ldw] zero padded (+ padded bias).
struct Packed
{
    float *W = nullptr;
    float *bias = nullptr;
    void *Wb = nullptr; // bf16 tensor-core image
    void *Wt = nullptr; // the same in the K order of the TMA-fed kernel (layers that run on it)
    void *Wt32 = nullptr; // fp32 image in the K order of that kernel's tf32 variant
    int K = 0, N = 0, ldw = 0;
};

struct DConvW
{
    Packed c1[2], c2[2];
    float *g1w[2], *g1b[2], *g2w[2], *g2b[2], *scale[2];
    int C = 0, hid = 0, hidp = 0;
    float *fused = nullptr; // parameter block of the fused shared-memory kernel (dconv_fused.cu), if the shape has one
    float *fused_pk = nullptr; // the same pre-packed as bf16 shared-memory images for dconv_mma.cu
};

struct EncW
{
    Packed conv, rewrite;
    DConvW dc;
    float *cin_pk = nullptr; // input conv (4 CaC channels, k8) pre-packed for the same kernel (encoder 0)
    float *rw_pk = nullptr; // rewrite conv pre-packed for the fused DConv + rewrite kernel (dconv_mma.cu), if the shape has one
    int Cin = 0, C = 0;
};

struct DecW
{
    Packed rewrite, convtr;
    DConvW dc;
    int C = 0, Cout = 0;
};

struct TLayerW
{
    bool cross = false;
    Packed qkv, q, kv, out, ff1, ff2;
    float *n1w, *n1b, *n2w, *n2b, *n3w, *n3b, *now_, *nob, *g1, *g2;
};

// Per-kernel-class CUDA-event timing (bench.py's roofline): events are recorded on the model's own
// stream around every launch while profiling is on.
enum ProfClass
{
    PC_STFT = 0,
    PC_ISTFT_FRAMES,
    PC_ISTFT_OLA,
    PC_POINTWISE,
    PC_GEMM_CONV,
    PC_GEMM_DCONV,
    PC_GEMM_XFMR,
    PC_ATTENTION,
    PC_TRACK,
    PC_LSTM,
    PC_DCONV_FUSED,
    PC_COUNT
};

struct ProfRec
{
    int cls;
    int n = 1; // kernel launches inside the scope
    cudaEvent_t e0, e1;
};

struct Profiler
{
    bool on = false;
    std::vector<cudaEvent_t> pool;
    size_t used = 0;
    std::vector<ProfRec> recs;
    double flops[PC_COUNT] = {0};
    double bytes[PC_COUNT] = {0};
    cudaEvent_t get();
};

// hdemucs_mmi (v3) DConv with BiLSTM + LocalState attention (src/layers.cpp:896-1113), one sub-block
struct DConvLaW
{
    Packed c1, ih0, ih1, lin, qkcd, proj, c2;
    float *whh_t[2][2]; // [layer][direction]: W_hh transposed to [H][4H] (fallback kernel)
    float *whh[2][2];   // W_hh as stored, [4H][H] (cluster kernel keeps its rows in registers)
    float *g1w, *g1b, *g2w, *g2b, *scale;
};

struct V3Model
{
    EncW enc[4], tenc[4];
    Packed tenc4, enc4_conv, enc4_rw, enc5_conv, enc5_rw;
    float *e4n1w, *e4n1b, *e4n2w, *e4n2b, *e5n1w, *e5n1b, *e5n2w, *e5n2b;
    DConvLaW la[2][2]; // [encoder 4|5][sub-block]
    Packed d0_rw, d0_ct, d1_rw, d1_ct, td0_ct;
    float *d0n1w, *d0n1b, *d0n2w, *d0n2b, *d1n1w, *d1n1b, *d1n2w, *d1n2b, *td0n2w, *td0n2b;
    DecW cdec[4], ctdec[4];
    float *ones = nullptr;
    // workspace
    float *xt4, *g4, *r4, *x4, *y5, *g5, *r5, *x5, *d0r, *d0g, *d0ct, *xd0, *d1r, *pre, *d1ct, *x3, *td0ct, *xt3;
    float *hraw, *seq, *lpre, *l0, *l1, *lo, *qk, *att, *y2la;
};

struct Tap
{
    float *ptr = nullptr;
    int64_t d1 = 0, d2 = 0, C = 0; // per batch item dims (channels-last)
    const float *minus = nullptr;   // the buffer holds value + minus (a pre-added decoder skip): the tap reports value
    bool swapped = false;           // the buffer is stored [b][d2][d1][C] (frame-major x_out): the tap reports [b][d1][d2][C]
};

} // namespace dmx

namespace dmx
{
struct HostStager;              // track.cu: pinned ring + worker thread for pageable caller buffers
void stager_free(HostStager *s);
} // namespace dmx

struct dmx_model
{
    int arch = 0, S = 4, d_model = 512, device = 0, maxB = 1, precision = 0;
    bool finalized = false;
    std::map<std::string, dmx::HostTensor> host;
    std::vector<void *> allocs;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;          // time branch of the encoders / decoders (htdemucs), forked from `stream`
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaStream_t stream_copy = nullptr;      // device -> host copies of finished output stretches (dmx_separate)
    cudaEvent_t ev_copy = nullptr;
    std::mutex mtx;
    int64_t launches = 0;
    dmx::Profiler prof;

    // weights
    dmx::EncW enc[4], tenc[4];
    dmx::DecW dec[4], tdec[4];
    float *freq_emb2 = nullptr;
    dmx::V3Model *v3 = nullptr;
    dmx::Packed up, up_t, down, down_t;
    float *nin_w, *nin_b, *nint_w, *nint_b, *pe2d, *pe1d;
    dmx::TLayerW lay[5], lay_t[5];
    float *window = nullptr, *tw = nullptr, *wss = nullptr;

    // workspace (all sized for maxB)
    float *mix, *spec, *xt_in, *cA_f, *cA_t, *e[4], *te[4], *dx[4], *dxt[4];
    // bf16 channels-last images of the decoder inputs (x + skip), written by the producing epilogue and read by the
    // TMA-fed rewrite convolutions: [0] = bottleneck + skip, [i] = output of transposed conv i-1 + skip
    void *dimg_f[4] = {nullptr, nullptr, nullptr, nullptr}, *dimg_t[4] = {nullptr, nullptr, nullptr, nullptr};
    void *cimg_f = nullptr, *cimg_t = nullptr; // ... of the DConv results that feed the transposed convolutions
    bool use_tma = false;
    // test tier (dmx_model_set_layer_taps): copies of the token streams after norm_in and after every transformer layer
    bool keep_stages = true; // false in the track-level entry points: stage tensors nobody reads again are not written
    bool layer_taps = false;
    float *lt_f[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr}, *lt_t[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    float *h_f, *y2_f, *h_t, *y2_t;
    float *up_f, *up_tt, *xq, *tq, *ln_f, *ln_t, *kn_f, *kn_t, *qkv_f, *qkv_t, *kv_f, *kv_t, *ao_f, *ao_t, *ff_f,
        *ff_t, *x3d, *xt3d;
    float *frames, *seg_out;
    float *meanstd_x, *meanstd_t;
    double *stats = nullptr;
    size_t stats_cap = 0, stats_used = 0;
    int lastB = 0;
    std::map<std::string, dmx::Tap> taps;

    // track-level buffers (grow only)
    float *d_audio = nullptr, *d_y = nullptr, *d_out = nullptr, *d_acc = nullptr;
    size_t cap_audio = 0, cap_y = 0, cap_out = 0, cap_acc = 0;
    double *track_stats = nullptr;
    float *d_tail = nullptr; // [S][SEG - SEG_STRIDE][2]: the previous worker's tail (dmx_separate_multi)
    dmx::HostStager *stager = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr; // device time of the last track-level call
    // the forward of the track-level entry points, captured once per (input, output, batch) and replayed (track.cu)
    struct FwdGraph
    {
        const float *mix = nullptr;
        float *out = nullptr;
        int B = 0, calls = 0;
        int64_t launches = 0;
        cudaGraphExec_t exec = nullptr;
    };
    std::vector<FwdGraph> graphs;
};

namespace dmx
{
double load_pack_seconds(); // host time spent repacking weights so far (DMX_LOAD_TRACE)
int model_finalize(dmx_model *m);
int model_v3_finalize(dmx_model *m);
int model_zero_attention_pads(dmx_model *m);
int noalloc_warm_up(int device); // noalloc.cu
int model_v3_forward(dmx_model *m, const float *d_mix, int B, float *d_out);
int debug_linear(int device, int precision, int M, int N, int K, const float *A, const float *W, const float *bias,
                 int act, float *out, int iters, float *ms_out);
int debug_conv(int device, int B, int U, int V, int Cin, int N, int ntaps, const int *du, const int *dv, const float *img,
               const float *W, const float *bias, int act, int glu, int share_dv, float *out, int iters, float *ms_out);
int model_forward(dmx_model *m, const float *d_mix, int B, float *d_out); // d_out: [B][S][SEG][2]
// RAII: brackets the launches issued inside its scope with events when profiling is on
struct ProfScope
{
    dmx_model *m;
    cudaEvent_t e1 = nullptr;
    ProfScope(dmx_model *m_, int cls, double flops = 0.0, double bytes = 0.0, int nlaunch = 1);
    ~ProfScope();
};
int track_forward_shard(dmx_model *m, const float *d_audio, int64_t L, int shift, int rank, int world, float *d_y,
                        dmx_progress_fn cb, void *user);
int track_overlap_add(dmx_model *m, const float *d_audio, int64_t L, int shift, const float *d_y_all, int world,
                      float *d_out);
int track_separate(dmx_model *m, const float *d_audio, int64_t L, int shift, float *d_out, dmx_progress_fn cb,
                   void *user);
void track_fire_segment_progress(const dmx_model *m, float current, float segment, dmx_progress_fn cb, void *user);
int plan_block(int64_t L, int shift, int rank, int world, dmx_block_plan *pl);
int track_stats_partial(dmx_model *m, const float *d_audio, int64_t n, double *d_sums);
int track_forward_block(dmx_model *m, const float *d_win, int64_t L, int shift, const dmx_block_plan &pl, const double *d_sums,
                        float *d_y);
int track_overlap_add_block(dmx_model *m, int64_t L, int shift, const dmx_block_plan &pl, const double *d_sums,
                            const float *d_y, const float *d_tail, int s_begin, int s_end, float *d_out_block);
void track_fire_progress(dmx_model *m, int nseg, dmx_progress_fn cb, void *user);
int track_accumulate(dmx_model *m, float *d_acc, const float *d_x, float scale, size_t n, int first);
int track_separate_host(dmx_model *m, const float *h_audio, int64_t L, int shift, float *h_out, dmx_progress_fn cb,
                        void *user);
} // namespace dmx
