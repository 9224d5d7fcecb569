/* demucs_b200 -- C ABI of the B200-native (sm_100a) inference hot path of sevagh/demucs.cpp.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types.  The host side of the
 * reference's C++17 API (src/model.hpp: load_demucs_model / demucs_inference and their v3 twins) is
 * a thin shim over these entry points (see INTEGRATION.md and compat/model.hpp).  Every function
 * returns 0 on success and non-zero on failure (message via dmx_last_error()); nothing throws and
 * there is NO CPU fallback: without a CUDA device every compute entry point fails.
 *
 * Data conventions (identical to the reference):
 *   audio in   : float32, L stereo frames interleaved L,R,L,R = Eigen::MatrixXf(2, L) column-major
 *   stems out  : float32 (S, 2, L) Eigen column-major, i.e. index s + S*(ch + 2*t); source order
 *                drums, bass, other, vocals(, guitar, piano)        (cli-apps/demucs.cpp:185-204)
 *   weights    : the reference's ggml-style f16 file (scripts/convert-pth-to-ggml.py:111-140)
 */
#ifndef DEMUCS_B200_H
#define DEMUCS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C"
{
#endif

    typedef struct dmx_model dmx_model;

    /* Progress callback, same meaning as demucscpp::ProgressCallback (src/model.hpp:17):
     * fraction in [0,1] and a message.  Invoked on the calling thread only.
     * Only the reference's numbered stage milestones are reproduced ("Time encoder 0" ... "Time: decoder 3": 26 per segment
     * for htdemucs, 22 for hdemucs_mmi, src/model_inference.cpp), with the reference's progress values, and they are
     * fired after the launch that contains the segment has finished -- the reference's other per-segment messages (shape
     * strings, "Freq branch: normalized", "Crosstransformer finished", "Mask + istft", ...) are not emitted. */
    typedef void (*dmx_progress_fn)(float progress, const char *message, void *user);

    /* Precision modes (BASELINE.json north_star: "bf16 mode" and "fp32 mode"). */
    enum
    {
        DMX_PRECISION_FP32 = 0, /* CUDA-core fp32 everywhere; per-layer tolerance 1e-4 */
        DMX_PRECISION_BF16 = 1  /* bf16 tensor-core GEMM operands, fp32 accumulate/residual/statistics */
    };

    /* Architectures, selected by the file magic (src/model_load.cpp:77-102, :1331). */
    enum
    {
        DMX_ARCH_HTDEMUCS_4S = 4, /* 'dmc4' 0x646d6334: htdemucs and each htdemucs_ft member */
        DMX_ARCH_HTDEMUCS_6S = 6, /* 'dmc6' 0x646d6336 */
        DMX_ARCH_HDEMUCS_V3 = 3   /* 'dmc3' 0x646d6333 (hdemucs_mmi) */
    };

    /* Number of visible CUDA devices (0 if none); the library never falls back to the CPU. */
    int dmx_device_count(void);

    /* Last error message of the calling thread ("" if none). */
    const char *dmx_last_error(void);

    /* ---- model construction: replaces demucscpp::load_demucs_model (src/model.hpp:649-650,
     *      src/model_load.cpp:50-1090).  Either load a whole ggml-f16 file, or create an empty
     *      model and feed tensors one by one (what the reference's in-memory loader,
     *      src_wasm/model_load.hpp:73, does), then finalize.  `max_batch` = number of 7.8 s
     *      segments processed per launch (workspace is sized for it). ---- */
    dmx_model *dmx_model_load_file(const char *ggml_path, int device, int max_batch);
    dmx_model *dmx_model_create(int arch, int device, int max_batch);
    /* name = PyTorch state_dict key; data = f16 bits, row-major, `ne` = squeezed shape
     * (record layout of src/model_load.cpp:120-147). */
    int dmx_model_set_tensor(dmx_model *m, const char *name, const uint16_t *f16, const int32_t *ne, int n_dims);
    /* Checks that every tensor of the architecture is present with the right element count
     * (src/model_load.cpp:1065-1070,1096-1105 reject unknown/mis-sized tensors), repacks to the
     * GEMM layouts and uploads to the device. */
    int dmx_model_finalize(dmx_model *m);
    void dmx_model_free(dmx_model *m);

    int dmx_model_arch(const dmx_model *m);    /* DMX_ARCH_* */
    int dmx_model_sources(const dmx_model *m); /* 4 or 6 (demucs_model::is_4sources, src/model.hpp:287) */
    int dmx_model_max_batch(const dmx_model *m);
    int dmx_model_set_precision(dmx_model *m, int precision);

    /* ---- whole-track separation: replaces demucscpp::demucs_inference (src/model.hpp:658-660,
     *      src/model_apply.cpp:60-288): normalise by the mono mean/std, zero-pad 0.5 s, shift by
     *      `shift_offset` samples (the reference draws rand() % 22050, src/model_apply.cpp:114; the
     *      caller passes it so results are reproducible), split into 7.8 s segments with 25 %
     *      overlap, run the per-segment forward, triangular-weighted overlap-add, de-normalise.
     *      audio/out are HOST buffers; copies are inside the call. ---- */
    int dmx_separate(dmx_model *m, const float *audio_lr, int64_t n_frames, int shift_offset, float *out,
                     dmx_progress_fn cb, void *user);

    /* Same with DEVICE-resident audio and output (cudaMalloc'd by the caller, e.g. a torch tensor's
     * data_ptr).  Runs on the model's own stream and returns after it has drained. */
    int dmx_separate_device(dmx_model *m, const float *d_audio_lr, int64_t n_frames, int shift_offset,
                            float *d_out);

    /* ---- multi-GPU building blocks (one process per GPU; the reference's only parallel strategy is the
     *      chunk-level std::thread fan-out of cli-apps/threaded_inference.hpp:110-123, which this
     *      replaces with segment-level sharding): rank r of `world` runs segments r, r+world, ... of the
     *      track into d_y (device, [ceil(nseg/world)][S][343980][2]); after the shards are gathered
     *      rank-major into d_y_all ([world][ceil(nseg/world)][S][343980][2]) one rank overlap-adds. ---- */
    int dmx_forward_shard_device(dmx_model *m, const float *d_audio_lr, int64_t n_frames, int shift_offset, int rank,
                                 int world, float *d_y);
    int dmx_overlap_add_device(dmx_model *m, const float *d_audio_lr, int64_t n_frames, int shift_offset,
                               const float *d_y_all, int world, float *d_out);
    /* ---- contiguous-block split (the multi-GPU data path; replaces the chunk fan-out and merge of
     *      cli-apps/threaded_inference.hpp:50-190).  Worker `rank` of `world` runs a RUN OF CONSECUTIVE
     *      segments and finalises the stretch of the track from its first segment's start to the next
     *      worker's first segment's start, so every worker reads only its own window of the input and writes
     *      only its own stretch of the output.  Between workers travel (1) two doubles, the partial sums of
     *      the track statistics (src/model_apply.cpp:72-80), and (2) the last 85995 frames of a worker's last
     *      segment, which overlap the next worker's first segment (2.75 MB for 4 sources).
     *      All pointers are DEVICE memory of the model's GPU; the calls enqueue on the model's stream and
     *      return without waiting (synchronise with dmx_model_sync). ---- */
    typedef struct dmx_block_plan
    {
        int32_t seg_begin, seg_end; /* segments [seg_begin, seg_end) of the track belong to this worker */
        int32_t n_segments;         /* segments of the whole track (src/model_apply.cpp:186,200) */
        int32_t tail_frames;        /* frames of the previous worker's last segment this worker needs (0 or 85995) */
        int64_t in_begin, in_end;   /* track frames this worker reads */
        int64_t out_begin, out_end; /* track frames this worker finalises (a partition of [0, n_frames)) */
    } dmx_block_plan;
#define DMX_TAIL_FRAMES 85995 /* 343980 - 257985: overlap of consecutive segments (src/model_apply.cpp:145,162) */
    int dmx_plan_block(int64_t n_frames, int shift_offset, int rank, int world, dmx_block_plan *plan);
    /* Adds the (sum, sum of squares) of the mono mix of `n_frames` interleaved frames to d_sums[0..1] (device
     * doubles, zeroed by the caller).  Summed over all workers' out-stretches they are the track statistics. */
    int dmx_track_stats_device(dmx_model *m, const float *d_audio_lr, int64_t n_frames, double *d_sums);
    /* Forward of the plan's segments.  d_audio_win holds track frames [in_begin, in_end); d_sums the statistics of
     * the WHOLE track; d_y receives [seg_end - seg_begin][S][343980][2]. */
    int dmx_forward_block_device(dmx_model *m, const float *d_audio_win, int64_t n_frames, int shift_offset,
                                 const dmx_block_plan *plan, const double *d_sums, float *d_y);
    /* Triangular-weighted overlap-add, trim and de-normalise (src/model_apply.cpp:171-246,129-135,88) of the
     * worker's stretch: d_out_block (out_end - out_begin, 2, S) gets track frames [out_begin, out_end).
     * d_prev_tail ([S][85995][2], the last tail_frames frames of segment seg_begin - 1, every source) is
     * required when plan->tail_frames > 0.  Only sources [src_begin, src_end) are written -- the htdemucs_ft bag
     * takes stem i from model i (cli-apps/demucs_ft.cpp:221-241), so member i calls this with (i, i + 1). */
    int dmx_overlap_add_block_device(dmx_model *m, int64_t n_frames, int shift_offset, const dmx_block_plan *plan,
                                     const double *d_sums, const float *d_y, const float *d_prev_tail, int src_begin,
                                     int src_end, float *d_out_block);
    int dmx_model_sync(dmx_model *m);

    /* ---- one process, several GPUs: whole-track separation with HOST buffers over `n_models` replicas of the
     *      same model, one per GPU (load the same file once per device).  One host thread per GPU runs the
     *      block split above; tails cross GPUs by peer copies.  Same result as dmx_separate.  The callback is
     *      invoked on the calling thread after the workers have finished. ---- */
    int dmx_separate_multi(dmx_model *const *models, int n_models, const float *audio_lr, int64_t n_frames,
                           int shift_offset, float *out, dmx_progress_fn cb, void *user);

    /* ---- htdemucs_ft bag: replaces the four demucs_inference calls + stem pick of cli-apps/demucs_ft.cpp:
     *      221-241.  models[i] (all on one GPU) is the member fine-tuned for source i; shift_offsets[i] its
     *      own random shift (the reference draws rand() % 22050 once per member); out (S, 2, L) receives
     *      stem i of member i. ---- */
    int dmx_separate_bag(dmx_model *const *models, int n_models, const float *audio_lr, int64_t n_frames,
                         const int *shift_offsets, float *out, dmx_progress_fn cb, void *user);

    /* ---- shift averaging ("shifts > 1"): the reference's shift_inference draws ONE random offset
     *      (src/model_apply.cpp:93-138); upstream's apply_model averages several
     *      (vendor/demucs/demucs/apply.py:187-199: out = mean over k of shift-k, separate, un-shift).  This entry
     *      point is that opt-in mode: the mean over `n_shifts` passes of dmx_separate with the given offsets
     *      (n_shifts = 1 is dmx_separate).  It also multiplies the independent work units of a track by n_shifts. ---- */
    int dmx_separate_shifts(dmx_model *m, const float *audio_lr, int64_t n_frames, const int *shift_offsets, int n_shifts,
                            float *out, dmx_progress_fn cb, void *user);

    /* Device milliseconds (CUDA events on the model's stream) of the last separate / shard / overlap-add call. */
    double dmx_last_device_ms(dmx_model *m);

    /* Number of segments demucs_inference cuts a track of n_frames into for this shift
     * (src/model_apply.cpp:186,200). */
    int dmx_num_segments(int64_t n_frames, int shift_offset);

    /* ---- per-segment forward: replaces demucscpp::model_inference (src/model.hpp:662-666,
     *      src/model_inference.cpp:48-475) on a batch of full 343980-frame segments.
     *      segs: host (B, 343980, 2) interleaved; out: host (B, S, 343980, 2). ---- */
    int dmx_forward_segments(dmx_model *m, const float *segs, int n_segments, float *out);

    /* The numbered stage milestones of ONE segment as the reference's model_inference reports them (progress =
     * current_progress + segment_progress * k / 26, or k / 22 for hdemucs_mmi; src/model_inference.cpp:156-470,584-734),
     * fired on the calling thread.  For callers of dmx_forward_segments that keep the reference's ProgressCallback
     * contract at segment level (compat/model_b200.cpp: demucscpp::model_inference). */
    int dmx_segment_progress(const dmx_model *m, float current_progress, float segment_progress, dmx_progress_fn cb, void *user);

    /* Test hook standing in for the reference's public per-segment scratch struct
     * (demucs_segment_buffers, src/model.hpp:569-647): copies a named activation left by the last
     * dmx_forward_segments call to the host.  dims[4] receives (B, d1, d2, C) of the channels-last
     * device layout; returns the element count, or -1.  With dst == NULL only dims are filled. */
    int64_t dmx_debug_tap(dmx_model *m, const char *name, float *dst, int64_t capacity, int64_t dims[4]);

    /* Per-layer test tier (the reference tests one transformer layer at a time, common_encoder_layer src/layers.cpp:377-531
     * in test/test_layers.cpp:1973-2157): when switched on, every forward also keeps the token streams after norm_in +
     * positional embedding (taps "x_ct_in", "xt_ct_in") and after each of the five layers ("x_ct_0" .. "x_ct_4",
     * "xt_ct_0" .. "xt_ct_4") for dmx_debug_tap.  htdemucs only; costs device copies, so it is off by default. */
    int dmx_model_set_layer_taps(dmx_model *m, int on);

    /* Test hook for the GEMM kernels in isolation (stands in for the reference's GemmConv* tests,
     * test/test_layers.cpp:856-1387): out (M,N) = act(A (M,K) x W (N,K)^T + bias) through the production
     * fp32 CUDA-core kernel (precision 0) or the bf16 tcgen05 kernel (precision 1); act 1 = GELU.
     * All pointers are host buffers.  With iters > 0 the kernel is additionally launched `iters` times
     * back to back and its average device time (CUDA events) is returned in *ms_out. */
    int dmx_debug_linear(int device, int precision, int M, int N, int K, const float *A, const float *W,
                         const float *bias, int act, float *out, int iters, float *ms_out);

    /* Test hook for the TMA-fed convolution kernel in isolation (bf16 mode; stands in for the reference's conv tests,
     * test/test_layers.cpp:856-1387, on conv2d / conv1d src/conv.hpp:13-262): a stride-1 zero-padded convolution with
     * `ntaps` taps (du[j], dv[j]) over a channels-last image, out (B,U,V,N or N/2) = act / GLU(conv(img (B,U,V,Cin),
     * W (N, ntaps*Cin), k = tap*Cin + c) + bias).  Host buffers; operands are rounded to bf16 as the producing layer's
     * epilogue does.  share_dv = 1: taps of one plane share one staged box (production); 0: one box per tap. */
    int dmx_debug_conv(int device, int B, int U, int V, int Cin, int N, int ntaps, const int *du, const int *dv,
                       const float *img, const float *W, const float *bias, int act, int glu, int share_dv, float *out,
                       int iters, float *ms_out);

    /* Kernel launches issued by this model's forward passes so far (for bench.py's gpu_launches). */
    int64_t dmx_launch_count(const dmx_model *m);

    /* ---- measurement hook (no reference counterpart; the reference has no tracing):
     *      between begin and end every kernel launch of this model is bracketed by CUDA events on the
     *      model's stream and accumulated per kernel class.  Arrays have dmx_profile_classes() entries:
     *      device milliseconds, launch counts, algorithmic FLOPs and algorithmic bytes per class. ---- */
    int dmx_profile_classes(void);
    const char *dmx_profile_class_name(int cls);
    int dmx_profile_begin(dmx_model *m);
    int dmx_profile_end(dmx_model *m, double *ms, int64_t *launches, double *flops, double *bytes);
    /* Class index of every bracketed launch scope since dmx_profile_begin, in launch order (call before
     * dmx_profile_end); returns the number of scopes.  Used to attribute an ncu launch list to classes. */
    int dmx_profile_sequence(dmx_model *m, int *cls_out, int capacity);

    /* ---- stand-alone DSP pair: replaces demucscpp::stft / demucscpp::istft (src/dsp.hpp:103-104,
     *      src/dsp.cpp:51-117) on one 343980-frame stereo segment, through the same kernels the
     *      forward uses.  spec: host (2048, 336, 4) = [bin][frame](ReL, ImL, ReR, ImR), un-normalised. */
    int dmx_stft_segment(dmx_model *m, const float *seg_lr, float *spec);
    int dmx_istft_segment(dmx_model *m, const float *spec, float *seg_lr);

    /* ---- allocation-free batched segment STFT.  No counterpart in the reference (its stft,
     *      src/dsp.cpp:51-86, works on host vectors it owns); this is an extra for callers that are not
     *      allowed to allocate device memory when they call in -- real-time audio threads, hosts that
     *      run from a fixed pre-reserved pool, code inside a no-allocation region.  Same computation
     *      and same kernel as dmx_stft_segment, but: input and output are DEVICE buffers the caller
     *      owns (d_segs (n, 343980, 2) interleaved -> d_spec (n, 2048, 336, 4); optional d_meanstd
     *      (n, 2) receives each segment's spectrogram mean and unbiased std, may be NULL), and the
     *      kernel's statistics workspace is a fixed array inside the library image instead of
     *      cudaMalloc'd memory.  The call performs no cudaMalloc/cudaFree and calls no allocating
     *      routine.  Limits and caveats:
     *        - at most DMX_NOALLOC_MAX_SEGMENTS segments per call, larger n is rejected with an error;
     *        - the workspace (1 KiB) exists on every GPU that loads the library, until process exit;
     *        - all calls on a device share it: it is cleared at the start of each call, and calls
     *          whose GPU work overlaps race with each other -- serialise them;
     *        - the workspace's device address is resolved once per device; dmx_model_finalize already does
     *          it (and loads the CUDA module) for the model's device, so nothing of the kind is left for
     *          the call itself.  If a call nevertheless is the first on its device it resolves the address
     *          then: let that call happen before beginning any stream capture;
     *        - the caller's current CUDA device is restored before returning;
     *        - `stream` is a cudaStream_t to enqueue on (the call then returns without waiting), or
     *          NULL to use the model's own stream and wait for completion. ---- */
#define DMX_NOALLOC_MAX_SEGMENTS 64
    int dmx_stft_segments_noalloc(dmx_model *m, const float *d_segs, int n_segments, float *d_spec, float *d_meanstd,
                                  void *stream);

#ifdef __cplusplus
}
#endif
#endif /* DEMUCS_B200_H */
