// Launcher declarations for the demucs_b200 kernels (implemented in fft.cu, gemm.cu, pointwise.cu,
// attention.cu).  All launchers return 0 on success, -1 after set_error().
#pragma once
#include "common.cuh"

namespace dmx
{

// ---------------------------------------------------------------------------------------------
// "tap-GEMM": every convolution / transposed convolution / linear layer of the model expressed as
//   C[m, n] = epilogue( sum_{tap, c} prologue(A[row(m, tap), c]) * W[tap*Cin + c, n] + bias[n] )
// over channels-last activations.  An output row m decodes to (b, u_o, v_o) on a per-item grid
// Uo x Vo (time branch: U = 1, V = samples; freq branch: U = frequency bins, V = frames); tap j
// reads input row (u_o*su + du[j], v_o*sv + dv[j]) of the Ui x Vi input grid, zero outside it.
// Output column n = r*Nco + co is stored at grid position (u_o*ous + r*our + ouo, v_o*ovs + r*ovr + ovo)
// of the Uout x Vout output grid, channel co, dropped outside it -- which lets ConvTranspose
// k=8,s=4 run as ONE GEMM with two taps and N = 4*Cout, the crop folded into (ouo/ovo).
//   reference ops covered: conv1d/conv2d(+_fused_gelu) src/conv.hpp:13-262, conv*_tr src/conv.hpp:264-524,
//   linear layers of common_encoder_layer src/layers.cpp:426-514.
struct GemmParams
{
    // A operand
    const float *A = nullptr;
    const float *A2 = nullptr; // optional second tensor added element-wise (decoder skip, src/encdec.cpp:172)
    int lda = 0, Cin = 0, K = 0;
    int ntaps = 1;
    int du[9] = {0}, dv[9] = {0};
    int Ui = 1, Vi = 1, Uo = 1, Vo = 1, su = 1, sv = 1;
    int M = 0;
    // A prologue: GroupNorm(1 group) + GELU of the producer's raw output (src/layers.cpp:51-94)
    int pro = 0;
    const double *pro_stats = nullptr; // [items][2] (sum, sumsq)
    double pro_count = 1.0;
    const float *pro_w = nullptr, *pro_b = nullptr;
    int pro_per_u = 0; // item = b*Ui + u_i (freq rows) or b (time)
    // B operand: packed fp32 [Kpad][ldw], zero padded
    const float *W = nullptr;
    int ldw = 0;
    const void *Wb = nullptr; // bf16 image for the tensor-core kernel (gemm_tc.cu), same (k, n) contents
    const float *bias = nullptr;
    int N = 0;
    // epilogue
    int act = 0;                   // 1 = exact-erf GELU
    int glu = 0;                   // 1 = columns packed as (a, gate) pairs -> a*sigmoid(gate)
    const float *rowadd = nullptr; // out += rowadd[u_o*rowadd_ld + co]   (frequency embedding)
    int rowadd_ld = 0;
    const float *gamma = nullptr;  // out = C_old + gamma[n]*val           (LayerScale residual)
    const float *addC = nullptr;   // out += addC[same element as C]  (the NEXT decoder's skip, src/encdec.cpp:172, added
                                   // once here instead of in every tap of that decoder's 3x3 / k3 gather)
    double *stats = nullptr;       // accumulate (sum, sumsq) of the stored values per item
    int stat_per_u = 0;
    int out_bf16 = 0;              // tensor-core kernel only, identity mapping only: C holds bf16 (ldc in elements)
    int out_pack = 0;              // ... C holds TILE-PACKED bf16 [row/128][ldc/8][row%128][8]: the next GEMM's A image
    const void *Apk = nullptr;     // tensor-core kernel only: A operand given as tile-packed bf16 (1x1 / linear only)
    int pack_T = 0;                // out_pack: rows per batch item in the packed space (multiple of 128), 0 = flat rows
    // output mapping
    float *C = nullptr;
    int ldc = 0, Nco = 0;
    int Uout = 1, Vout = 1, ous = 1, our = 0, ouo = 0, ovs = 1, ovr = 0, ovo = 0;
    int out_vu = 0; // output stored [b][v][u][c] instead of [b][u][v][c] (the last frequency decoder: frame-major for the iSTFT)
    // filled in by launch_gemm_tc: reciprocals for fast_div by Uo*Vo, Vo (row -> (b, u_o, v_o)) and Nco
    unsigned long long div_per = 0, div_vo = 0, div_nco = 0;
    int no_bulk = 0; // development switch: keep the transposing epilogue
    int dbg_epi = 0; // development switch (DMX_DBG_EPI): 4 = transposed convolutions keep the generic transposing epilogue
    // halo mode of the tensor-core kernel (stride-1 multi-tap convolutions; see gemm_tc.cu), also set by the launcher
    int halo_nu = 0, halo_dumin = 0, halo_dvmin = 0, halo_rows = 0, halo_tpr = 0; // planes, first du / dv, patch rows, tiles per (b,u)
    unsigned long long div_tpr = 0, div_uo = 0, div_c8 = 0, div_rows = 0;
    // optional bf16 copy of the output, same element indexing as C (identity or scatter mapping, not with bulk / packed
    // outputs): the channels-last IMAGE the next convolution's TMA loads read
    void *Cbf = nullptr;
    int skip_c = 0; // scatter mapping with Cbf: do not store the fp32 output at all (production: only the image is read again)
    int Cbf_ld = 0; // channels per image row (>= Nco; 48-channel images are stored with a 64-channel pitch, see TmaPlan)
    // TMA-fed persistent mode (gemm_tc.cu: launch_gemm_tma): A is a bf16 channels-last image [B][Ui][Vi][Cin] read through a
    // tensor map; Wt = the weights in that kernel's K order (TmaPlan); the t_* fields are filled in by the launcher
    const void *Aimg = nullptr;
    const void *Wt = nullptr;
    int t_nbox = 0, t_box_du[9] = {0}, t_box_dv[9] = {0}, t_rows = 0;
    int t_tpb = 0, t_shift16 = 0; // taps per box; descriptor offset (16-byte units) between consecutive taps of a box
    int t_cw = 0, t_kw = 0, t_nchunk = 0, t_cps = 1, t_pitch = 0, t_box_bytes = 0, t_layout = 0;
    int Aimg_ld = 0; // channels per image row of Aimg (0: Cin)
    // tf32 variant: A is the ordinary fp32 channels-last tensor p.A (lda channels per row), read through a tensor map in
    // 32-channel chunks (128-byte rows) and multiplied with kind::tf32 MMAs; Wt32 = fp32 weights in that kernel's K order
    const void *Wt32 = nullptr;
    int t_tf32 = 0;
    int t_noshare = 0; // set by the caller: one box per tap instead of one per plane (test hook)
};

// K order and operand boxes of the TMA-fed kernel for a stride-1 tap set: a BOX is a run of input rows of one plane
// (one du) that TMA lands in shared memory once per tile; with share_dv all taps of a plane read the same box at
// row offsets dv - dvmin (equally spaced), otherwise every tap has its own box.  cw = 64 / 32 / 16 channels = one
// 128 / 64 / 32-byte swizzle row; K' runs over (chunk group, box, chunk in group, tap in box, channel in chunk).
struct TmaPlan
{
    int nbox = 0, box_du[9], box_dv[9], rows = 0;
    int ntap = 0, tap_id[9]; // taps in box order: box b holds taps tap_id[b * tpb .. + tpb)
    int tpb = 0, shift = 0;  // taps per box (the same for every box) and the row shift between consecutive taps of a box
    int cw = 0, nchunk = 0, cps = 1, pitch = 0, box_bytes = 0, layout = 0; // cps: chunks that share one ring slot / barrier
    int kw = 0;     // channels of a chunk that enter the GEMM (= cw, except 48-channel images: cw 64, kw 48)
    int cpitch = 0; // channels per image row the producer must write (Cin, or 64 for 48-channel images: a tensor-map copy
                    // costs per ROW whatever its width, so one 128-byte row beats three 32-byte ones)
};
int tc_tma_plan(int ntaps, const int *du, const int *dv, int Cin, bool share_dv, TmaPlan *pl);
int tc_tma_cpitch(int C); // image row pitch in channels for a C-channel tensor
// the same for the tf32 variant (fp32 activations, 32-channel chunks; channels beyond Cin are zero-filled by the copy and
// meet zero weight rows): plan, K' size, original K index of K' index kp (-1: padding), packed fp32 image size / index
int tc_tma_plan32(int ntaps, const int *du, const int *dv, int Cin, bool share_dv, TmaPlan *pl);
int tc_tma_k32(const TmaPlan &pl);
int tc_tma_korig32(const TmaPlan &pl, int Cin, int kp);
size_t tc_packed_elems32(int K, int N);
size_t tc_packed_index32(int K, int N, int k, int n);
// original K index (tap * Cin + c) of K' index kp
int tc_tma_korig(const TmaPlan &pl, int Cin, int kp);
int launch_gemm_tma(const GemmParams &p, cudaStream_t st);
bool tc_tma_available();

int launch_gemm(const GemmParams &p, cudaStream_t st);

// gemm_tc.cu: the same operator on tcgen05 tensor cores (bf16 operands, fp32 accumulate in TMEM)
int tc_init_device();
int tc_tile_n(int N);
size_t tc_packed_elems(int K, int N);
size_t tc_packed_index(int K, int N, int k, int n);
int launch_gemm_tc(const GemmParams &p, const void *Wb, cudaStream_t st);

// ---------------------------------------------------------------------------------------------
// fft.cu
int fft_init_device();
int launch_stft(const float *mix, int B, const float *window, const float *tw, float *spec, double *stats,
                cudaStream_t st);
// xout element (b, bin k, frame t, channel c) sits at ((b * 2048 * 336) * C) + k * sk + t * st_ + c: bin-major
// (sk = 336 * C, st_ = C) or frame-major (sk = C, st_ = 2048 * C; what the last decoder writes in production)
int launch_istft(const float *xout, int B, int S, const float *meanstd, const float *window, const float *tw,
                 const float *wss, const float *xt_out, const float *meanstd_t, float *frames_scratch, float *out,
                 cudaStream_t st, int frame_major = 0);

// ---------------------------------------------------------------------------------------------
// pointwise.cu
// sum / sumsq of n_per_item floats per batch item -> stats[b][2] (double, must be zeroed)
int launch_sum_stats(const float *x, int B, size_t n_per_item, double *stats, cudaStream_t st);
// x = (x - mean)/(std + 1e-5) with unbiased std; writes (mean, std) per item  (src/model_inference.cpp:115-144)
int launch_normalize_input(float *x, int B, size_t n_per_item, const double *stats, float *meanstd,
                           cudaStream_t st);
// DConv tail: x[r][c] += scale[c] * GN(y)[r][c] * sigmoid(GN(y)[r][c+C])  (src/layers.cpp:240-253)
int launch_gn_glu_residual(const float *y, float *x, int B, int U, int V, int C, const double *stats,
                           int per_u, const float *gw, const float *gb, const float *scale, cudaStream_t st,
                           void *img = nullptr, int img_ld = 0);
// LayerNorm writing TILE-PACKED bf16 (the tcgen05 GEMM's A image, see GemmParams::Apk) instead of fp32 rows
int launch_layer_norm_pack(const float *x, void *ypk, int rows, int C, const float *w, const float *b, cudaStream_t st);
// LayerNorm over the last dim (unbiased var), optional additive table add[row % add_rows][c]
int launch_layer_norm(const float *x, float *y, int rows, int C, const float *w, const float *b,
                      const float *add, int add_rows, cudaStream_t st);
// MyGroupNorm apply: x = (x-mean)*rstd*w[c]+b[c] per batch item over (rows_per_item, C)
int launch_group_norm_apply(float *x, int B, int rows_per_item, int C, const double *stats, const float *w,
                            const float *b, cudaStream_t st);

// ---------------------------------------------------------------------------------------------
// dconv_fused.cu: whole DConv branch of a frequency row resident in shared memory (C = 48 / 96 levels)
bool dconv_fused_supported(int C, int hid, int V);
int dconv_fused_param_floats(int C, int hid);
int dconv_fused_init_device();
int launch_dconv_fused(float *x, const float *params, int items, int C, int hid, cudaStream_t st);
// dconv_mma.cu: the same with both convolutions on tensor cores (mma.sync, bf16 operands); bf16 mode only
bool dconv_mma_supported(int C, int hid, int V);
int dconv_mma_init_device();
int dconv_mma_pack_floats(int C, int hid);
void dconv_mma_pack(int C, int hid, const float *blk, float *out);
int launch_dconv_mma(float *x, const void *packed, int items, int C, int hid, cudaStream_t st);
bool dconv_mma_rw_supported(int C, int hid, int V);
int dconv_mma_rw_pack_floats(int C);
void dconv_mma_rw_pack(int C, const float *W, const float *bias, float *out);
int dconv_mma_cin_pack_floats(int C);
void dconv_mma_cin_pack(int C, const float *W, const float *bias, float *out);
int launch_dconv_mma_rw(float *x, const void *packed, const void *rw, float *out, const float *rowadd, int F, int items,
                        int C, int hid, const float *spec, const void *cin, cudaStream_t st);
bool dconv_time_mma_supported(int C, int hid, int hs);
int launch_dconv_time_mma(float *x, float *hbuf, const void *packed, int B, int L, int C, int hid, int hs,
                          double *const *stats, cudaStream_t st);

// ---------------------------------------------------------------------------------------------
// v3_kernels.cu (hdemucs_mmi)
int launch_group_norm_g(const float *x, float *y, int B, int rows_in, int rows_out, int row0, int C, int G,
                        double *stats_zeroed, const float *w, const float *b, int act, int glu, cudaStream_t st);
int launch_lstm_recurrent(const float *pre, const float *whh_fwd, const float *whh_bwd, const float *whh_t_fwd,
                          const float *whh_t_bwd, float *out, int B, int T, int H, cudaStream_t st);
int launch_local_attention(const float *qkcd, float *out, int B, int T, int C, cudaStream_t st);

// ---------------------------------------------------------------------------------------------
// attention.cu: softmax(Q K^T / sqrt(dh)) V per head, heads = contiguous dh-wide column blocks
// (src/layers.cpp:442-482).  Q: [B*T][ldq], K/V: [B*S][ldk|ldv]; O: [B*T][ldo].
int launch_attention(const float *Q, int ldq, const float *K, int ldk, const float *V, int ldv, float *O, int ldo,
                     int B, int T, int S, int heads, int dh, cudaStream_t st);

// attention_tc.cu: the same attention on tcgen05 tensor cores; Q/K/V bf16, O fp32
int attention_tc_init_device();
// Q/K/V are TILE-PACKED bf16 ([item*Tpad/128 + tile][ld/8][128][8], per-item rows padded to Tpad / Spad); q_col,
// k_col, v_col = first column of the tensor inside its packed matrix; output tile-packed bf16 (flat rows), d = heads*dh.
int launch_attention_tc(const void *Qp, int ldq, int q_col, const void *Kp, int ldk, int k_col, const void *Vp, int ldv,
                        int v_col, void *Opk, int B, int T, int Tpad, int S, int Spad, int heads, int dh, cudaStream_t st);

} // namespace dmx
