// htdemucs (4s / 6s / ft member) on device: weight repacking, workspace and the per-segment forward
// (reference: demucscpp::model_inference, src/model_inference.cpp:48-475) over a batch of segments.
//
// Device layout: channels-last fp32.  Time branch [B][L][C]; frequency branch [B][F][T=336][C]
// (C fastest, then frames, then bins).  With it every conv of the model is a tap-GEMM on strided
// views (kernels.cuh) and no transpose/shuffle of the reference is ever
// materialised; transformer tokens are simply the rows of the [8][336][C] bottleneck (row f*336+t),
// with the 2-D positional table built in that order.
#include "model_common.cuh"

namespace dmx
{

// ------------------------------------------------------------------------------------------------
// profiling
cudaEvent_t Profiler::get()
{
    if (used == pool.size())
    {
        cudaEvent_t e;
        cudaEventCreate(&e);
        pool.push_back(e);
    }
    return pool[used++];
}

ProfScope::ProfScope(dmx_model *m_, int cls, double flops, double bytes, int nlaunch) : m(m_)
{
    m->launches += nlaunch;
    if (!m->prof.on)
        return;
    ProfRec r;
    r.cls = cls;
    r.n = nlaunch;
    r.e0 = m->prof.get();
    r.e1 = m->prof.get();
    e1 = r.e1;
    m->prof.recs.push_back(r);
    m->prof.flops[cls] += flops;
    m->prof.bytes[cls] += bytes;
    cudaEventRecord(r.e0, m->stream);
}

ProfScope::~ProfScope()
{
    if (e1)
        cudaEventRecord(e1, m->stream);
}

static int pack_tlayer(dmx_model *m, TLayerW &L, const std::string &p, bool cross, int d)
{
    L.cross = cross;
    const std::string attn = p + (cross ? "cross_attn." : "self_attn.");
    if (cross)
    {
        if (pack_linear(m, L.q, attn + "in_proj_weight", attn + "in_proj_bias", 3 * d, d, 0, d) ||
            pack_linear(m, L.kv, attn + "in_proj_weight", attn + "in_proj_bias", 3 * d, d, d, 2 * d))
            return -1;
    }
    else if (pack_linear(m, L.qkv, attn + "in_proj_weight", attn + "in_proj_bias", 3 * d, d, 0, 3 * d))
        return -1;
    if (pack_linear(m, L.out, attn + "out_proj.weight", attn + "out_proj.bias", d, d, 0, d) ||
        pack_linear(m, L.ff1, p + "linear1.weight", p + "linear1.bias", 4 * d, d, 0, 4 * d) ||
        pack_linear(m, L.ff2, p + "linear2.weight", p + "linear2.bias", d, 4 * d, 0, d))
        return -1;
    if (upload_vec(m, &L.n1w, p + "norm1.weight", d) || upload_vec(m, &L.n1b, p + "norm1.bias", d) ||
        upload_vec(m, &L.n2w, p + "norm2.weight", d) || upload_vec(m, &L.n2b, p + "norm2.bias", d) ||
        upload_vec(m, &L.now_, p + "norm_out.weight", d) || upload_vec(m, &L.nob, p + "norm_out.bias", d) ||
        upload_vec(m, &L.g1, p + "gamma_1.scale", d) || upload_vec(m, &L.g2, p + "gamma_2.scale", d))
        return -1;
    L.n3w = L.n3b = nullptr;
    if (cross && (upload_vec(m, &L.n3w, p + "norm3.weight", d) || upload_vec(m, &L.n3b, p + "norm3.bias", d)))
        return -1;
    return 0;
}

// 2-D sinusoidal embedding (src/crosstransformer.cpp:7-53) in token order row = f*T + t.
static std::vector<float> make_pe2d(int d, int height, int width)
{
    std::vector<float> pe((size_t)height * width * d);
    const int half = d / 2;
    std::vector<float> div(half / 2);
    for (int j = 0; j < half / 2; ++j)
        div[j] = expf((float)(2 * j) * (-logf(10000.0f) / (float)half));
    for (int f = 0; f < height; ++f)
        for (int t = 0; t < width; ++t)
        {
            float *row = &pe[((size_t)f * width + t) * d];
            for (int j = 0; j < half / 2; ++j)
            {
                const float vw = (float)t * div[j], vh = (float)f * div[j];
                row[2 * j] = sinf(vw);
                row[2 * j + 1] = cosf(vw);
                row[half + 2 * j] = sinf(vh);
                row[half + 2 * j + 1] = cosf(vh);
            }
        }
    return pe;
}

// 1-D sinusoidal embedding (src/crosstransformer.cpp:55-77): cos first half, sin second half.
static std::vector<float> make_pe1d(int length, int d)
{
    std::vector<float> pe((size_t)length * d);
    const int half = d / 2;
    for (int t = 0; t < length; ++t)
        for (int i = 0; i < half; ++i)
        {
            const float div = (float)i / (float)(half - 1);
            const float phase = (float)t / powf(10000.0f, div);
            pe[(size_t)t * d + i] = cosf(phase);
            pe[(size_t)t * d + half + i] = sinf(phase);
        }
    return pe;
}

// ------------------------------------------------------------------------------------------------
int model_finalize(dmx_model *m)
{
    if (m->finalized)
        return 0;
    if (m->arch == DMX_ARCH_HDEMUCS_V3)
        return model_v3_finalize(m);
    DMX_CUDA(cudaSetDevice(m->device));
    DMX_CUDA(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking));
    DMX_CUDA(cudaStreamCreateWithFlags(&m->stream2, cudaStreamNonBlocking));
    DMX_CUDA(cudaEventCreateWithFlags(&m->ev_fork, cudaEventDisableTiming));
    DMX_CUDA(cudaEventCreateWithFlags(&m->ev_join, cudaEventDisableTiming));
    DMX_CUDA(cudaEventCreate(&m->ev0));
    DMX_CUDA(cudaEventCreate(&m->ev1));
    DMX_CUDA(cudaEventCreate(&m->ev2));
    if (fft_init_device() || tc_init_device() || attention_tc_init_device() || dconv_fused_init_device() ||
        dconv_mma_init_device())
        return -1;
    const bool is4 = m->arch == DMX_ARCH_HTDEMUCS_4S;
    m->S = is4 ? 4 : 6;
    m->d_model = is4 ? 512 : 384;
    const int S = m->S, d = m->d_model;

    for (int i = 0; i < 4; ++i)
    {
        const int cin_f = i == 0 ? 4 : CH[i], cin_t = i == 0 ? 2 : CH[i], C = CH[i + 1];
        const std::string e = "encoder." + std::to_string(i), te = "tencoder." + std::to_string(i);
        m->enc[i].Cin = cin_f;
        m->enc[i].C = C;
        m->tenc[i].Cin = cin_t;
        m->tenc[i].C = C;
        // tf32 ring-kernel images: the 1x1 rewrites of both branches and the strided convolutions of frequency encoders 1-3
        const TmaPlan p_rw = enc_plan32(C, false), p_cv = enc_plan32(cin_f >= 48 ? cin_f : 48, true);
        if (pack_conv(m, m->enc[i].conv, e + ".conv.weight", e + ".conv.bias", C, cin_f, 8, false, nullptr, i > 0 ? &p_cv : nullptr) ||
            pack_conv(m, m->enc[i].rewrite, e + ".rewrite.weight", e + ".rewrite.bias", 2 * C, C, 1, true, nullptr, &p_rw) ||
            pack_dconv(m, m->enc[i].dc, e, C) ||
            pack_conv(m, m->tenc[i].conv, te + ".conv.weight", te + ".conv.bias", C, cin_t, 8, false) ||
            pack_conv(m, m->tenc[i].rewrite, te + ".rewrite.weight", te + ".rewrite.bias", 2 * C, C, 1, true, nullptr, &p_rw) ||
            pack_dconv(m, m->tenc[i].dc, te, C))
            return -1;
        if (dconv_mma_rw_supported(C, m->enc[i].dc.hid, LE))
        {
            const HostTensor *rw = get(m, e + ".rewrite.weight", (size_t)2 * C * C), *rb = get(m, e + ".rewrite.bias", 2 * C);
            if (!rw || !rb)
                return -1;
            std::vector<float> pk((size_t)dconv_mma_rw_pack_floats(C));
            dconv_mma_rw_pack(C, rw->data.data(), rb->data.data(), pk.data());
            if (upload(m, &m->enc[i].rw_pk, pk))
                return -1;
            if (cin_f == 4)
            {
                const HostTensor *cw = get(m, e + ".conv.weight", (size_t)C * 4 * 8), *cb = get(m, e + ".conv.bias", C);
                if (!cw || !cb)
                    return -1;
                std::vector<float> ck((size_t)dconv_mma_cin_pack_floats(C));
                dconv_mma_cin_pack(C, cw->data.data(), cb->data.data(), ck.data());
                if (upload(m, &m->enc[i].cin_pk, ck))
                    return -1;
            }
        }
    }
    for (int i = 0; i < 4; ++i)
    {
        const int C = CH[4 - i];
        const int cout_f = i == 3 ? 4 * S : CH[3 - i], cout_t = i == 3 ? 2 * S : CH[3 - i];
        const std::string dn = "decoder." + std::to_string(i), tn = "tdecoder." + std::to_string(i);
        m->dec[i].C = C;
        m->dec[i].Cout = cout_f;
        m->tdec[i].C = C;
        m->tdec[i].Cout = cout_t;
        // 3x3 rewrite weight (2C, C, 3, 3): tap = kh*3 + kw with kh along frequency, kw along frames
        const TmaPlan plf = rewrite_plan(C, true), plt = rewrite_plan(C, false);
        const TmaPlan pcf = convtr_plan(C, true), pct = convtr_plan(C, false);
        if (pack_conv(m, m->dec[i].rewrite, dn + ".rewrite.weight", dn + ".rewrite.bias", 2 * C, C, 9, true, &plf) ||
            pack_dconv(m, m->dec[i].dc, dn, C) ||
            pack_convtr(m, m->dec[i].convtr, dn + ".conv_tr.weight", dn + ".conv_tr.bias", C, cout_f, 8, &pcf) ||
            pack_conv(m, m->tdec[i].rewrite, tn + ".rewrite.weight", tn + ".rewrite.bias", 2 * C, C, 3, true, &plt) ||
            pack_dconv(m, m->tdec[i].dc, tn, C) ||
            pack_convtr(m, m->tdec[i].convtr, tn + ".conv_tr.weight", tn + ".conv_tr.bias", C, cout_t, 8, &pct))
            return -1;
    }
    {
        // x_0(c, f, t) += 2.0 * emb(f, c)   (src/model_inference.cpp:163-179; 10 * 0.2 folded)
        const HostTensor *t = get(m, "freq_emb.embedding.weight", 512 * 48);
        if (!t)
            return -1;
        std::vector<float> e2(t->data);
        for (auto &v : e2)
            v *= 2.0f;
        if (upload(m, &m->freq_emb2, e2))
            return -1;
    }
    if (is4)
    {
        if (pack_conv(m, m->up, "channel_upsampler.weight", "channel_upsampler.bias", 512, 384, 1, false) ||
            pack_conv(m, m->up_t, "channel_upsampler_t.weight", "channel_upsampler_t.bias", 512, 384, 1, false) ||
            pack_conv(m, m->down, "channel_downsampler.weight", "channel_downsampler.bias", 384, 512, 1, false) ||
            pack_conv(m, m->down_t, "channel_downsampler_t.weight", "channel_downsampler_t.bias", 384, 512, 1, false))
            return -1;
    }
    if (upload_vec(m, &m->nin_w, "crosstransformer.norm_in.weight", d) ||
        upload_vec(m, &m->nin_b, "crosstransformer.norm_in.bias", d) ||
        upload_vec(m, &m->nint_w, "crosstransformer.norm_in_t.weight", d) ||
        upload_vec(m, &m->nint_b, "crosstransformer.norm_in_t.bias", d))
        return -1;
    for (int i = 0; i < 5; ++i)
    {
        if (pack_tlayer(m, m->lay[i], "crosstransformer.layers." + std::to_string(i) + ".", i % 2 == 1, d) ||
            pack_tlayer(m, m->lay_t[i], "crosstransformer.layers_t." + std::to_string(i) + ".", i % 2 == 1, d))
            return -1;
    }
    // the reference rejects files with unknown tensors (src/model_load.cpp:1065-1070): so do we
    {
        const size_t expect = is4 ? 533 : 525;
        if (m->host.size() != expect)
        {
            set_error("model_finalize: expected " + std::to_string(expect) + " tensors, got " +
                      std::to_string(m->host.size()));
            return -1;
        }
    }
    {
        const std::vector<float> w = make_window();
        if (upload(m, &m->window, w) || upload(m, &m->wss, make_wss(w)) || upload(m, &m->tw, make_twiddles()) ||
            upload(m, &m->pe2d, make_pe2d(d, 8, T)) || upload(m, &m->pe1d, make_pe1d(TLEN3, d)))
            return -1;
    }
    m->host.clear(); // host fp32 copies are no longer needed

    // ---------------- workspace ----------------
    const size_t B = (size_t)m->maxB;
    auto A = [&](float **p, size_t n) { return dev_alloc_t(m, p, B * n); };
    int rc = 0;
    rc |= A(&m->mix, (size_t)SEG * 2);
    rc |= A(&m->spec, (size_t)NBINS * T * 4);
    rc |= A(&m->xt_in, (size_t)SEG * 2);
    rc |= A(&m->cA_f, (size_t)512 * T * 48);
    rc |= A(&m->cA_t, (size_t)TLEN0 * 48);
    for (int i = 0; i < 4; ++i)
    {
        rc |= A(&m->e[i], (size_t)FREQS[i + 1] * T * CH[i + 1]);
        rc |= A(&m->te[i], (size_t)TLENS[i + 1] * CH[i + 1]);
        rc |= A(&m->dx[i], (size_t)FREQS[3 - i] * T * m->dec[i].Cout);
        rc |= A(&m->dxt[i], (size_t)TLENS[3 - i] * m->tdec[i].Cout);
    }
    rc |= A(&m->h_f, (size_t)512 * T * 8);
    rc |= A(&m->y2_f, (size_t)512 * T * 96);
    rc |= A(&m->h_t, (size_t)TLEN0 * 8);
    rc |= A(&m->y2_t, (size_t)TLEN0 * 96);
    const size_t TF = 8 * T, TT = TLEN3;
    rc |= A(&m->up_f, TF * d);
    rc |= A(&m->up_tt, TT * d);
    rc |= A(&m->xq, TF * d);
    rc |= A(&m->tq, TT * d);
    rc |= A(&m->ln_f, TF * d);
    rc |= A(&m->ln_t, TT * d);
    rc |= A(&m->kn_f, TT * d);
    rc |= A(&m->kn_t, TF * d);
    rc |= A(&m->qkv_f, TF * 3 * d);
    rc |= A(&m->qkv_t, TT * 3 * d);
    rc |= A(&m->kv_f, TT * 2 * d);
    rc |= A(&m->kv_t, TF * 2 * d);
    rc |= A(&m->ao_f, TF * d);
    rc |= A(&m->ao_t, TT * d);
    rc |= A(&m->ff_f, TF * 4 * d);
    rc |= A(&m->ff_t, TT * 4 * d);
    rc |= A(&m->x3d, TF * 384);
    rc |= A(&m->xt3d, TT * 384);
    rc |= A(&m->frames, (size_t)S * T * NFFT * 2);
    rc |= A(&m->seg_out, (size_t)S * SEG * 2);
    rc |= A(&m->meanstd_x, 2);
    rc |= A(&m->meanstd_t, 2);
    m->use_tma = tc_tma_available();
    if (m->use_tma)
    {
        rc |= dev_alloc(m, &m->cimg_f, B * (size_t)512 * T * 64 * 2);
        rc |= dev_alloc(m, &m->cimg_t, B * (size_t)TLEN0 * 64 * 2);
    }
    if (m->use_tma)
        for (int i = 0; i < 4 && !rc; ++i)
        {
            // decoder i reads [B][FREQS[4-i]][T][CH[4-i]] / [B][TLENS[4-i]][CH[4-i]] as bf16
            rc |= dev_alloc(m, &m->dimg_f[i], B * (size_t)FREQS[4 - i] * T * tc_tma_cpitch(CH[4 - i]) * 2);
            rc |= dev_alloc(m, &m->dimg_t[i], B * (size_t)TLENS[4 - i] * tc_tma_cpitch(CH[4 - i]) * 2);
        }
    if (rc)
        return -1;
    // pad channels of the DConv hidden tensors are never written: they must hold finite values
    DMX_CUDA(cudaMemset(m->h_f, 0, B * (size_t)512 * T * 8 * sizeof(float)));
    DMX_CUDA(cudaMemset(m->h_t, 0, B * (size_t)TLEN0 * 8 * sizeof(float)));
    // statistics arena: (sum, sumsq) doubles; DConv needs up to 512 items per batch element and layer
    m->stats_cap = B * 2 * (size_t)(2 * 2 * 2 * (512 + 128 + 32 + 8) * 2 + 16 * 2 * 2 + 64);
    if (dev_alloc_t(m, &m->stats, m->stats_cap) || dev_alloc_t(m, &m->track_stats, 2))
        return -1;
    if (noalloc_warm_up(m->device))
        return -1;
    m->finalized = true;
    return model_zero_attention_pads(m);
}

// The tensor-core attention reads Q/K/V as tile-packed bf16 with every item's rows padded to whole 128-row
// tiles; pad rows are never written by the projections, so they must hold zeros (0 x garbage could be NaN in
// the P.V product).  Called at finalize and whenever the precision mode changes (fp32 mode reuses the buffers).
int model_zero_attention_pads(dmx_model *m)
{
    if (!m->finalized || m->arch == DMX_ARCH_HDEMUCS_V3)
        return 0;
    const size_t B = (size_t)m->maxB, d = (size_t)m->d_model, TF = 8 * T, TT = TLEN3;
    DMX_CUDA(cudaSetDevice(m->device));
    DMX_CUDA(cudaMemset(m->qkv_f, 0, B * TF * 3 * d * sizeof(float)));
    DMX_CUDA(cudaMemset(m->qkv_t, 0, B * TT * 3 * d * sizeof(float)));
    DMX_CUDA(cudaMemset(m->kv_f, 0, B * TT * 2 * d * sizeof(float)));
    DMX_CUDA(cudaMemset(m->kv_t, 0, B * TF * 2 * d * sizeof(float)));
    return 0;
}

double load_pack_seconds() { return LoadTimer::total(); }

int model_forward(dmx_model *m, const float *d_mix, int B, float *d_out)
{
    if (m->arch == DMX_ARCH_HDEMUCS_V3)
        return model_v3_forward(m, d_mix, B, d_out);
    if (!m->finalized)
    {
        set_error("model_forward: model not finalized");
        return -1;
    }
    if (B < 1 || B > m->maxB)
    {
        set_error("model_forward: batch exceeds max_batch");
        return -1;
    }
    cudaStream_t st = m->stream;
    Fwd f{m, B, st};
    const int S = m->S, d = m->d_model;
    m->stats_used = 0;
    m->lastB = B;
    DMX_CUDA(cudaMemsetAsync(m->stats, 0, m->stats_cap * sizeof(double), st));

    // ---- STFT, CaC packing and the two input normalisations (src/model_inference.cpp:64-144) ----
    double *sx = f.stat_slots(B), *sxt = f.stat_slots(B);
    {
        // algorithmic bytes: read the segment, write the CaC spectrogram (13.8 MB / segment)
        ProfScope ps(m, PC_STFT, 0.0, (double)B * ((double)PADDED * 2 * 4 + (double)NBINS * T * 4 * 4));
        if (launch_stft(d_mix, B, m->window, m->tw, m->spec, sx, st))
            return -1;
    }
    {
        ProfScope ps(m, PC_POINTWISE, 0.0, (double)B * NBINS * T * 4 * 8.0);
        if (launch_normalize_input(m->spec, B, (size_t)NBINS * T * 4, sx, m->meanstd_x, st))
            return -1;
    }
    DMX_CUDA(cudaMemcpyAsync(m->xt_in, d_mix, (size_t)B * SEG * 2 * sizeof(float), cudaMemcpyDeviceToDevice, st));
    {
        ProfScope ps(m, PC_POINTWISE, 0.0, (double)B * SEG * 2 * 4.0);
        if (launch_sum_stats(m->xt_in, B, (size_t)SEG * 2, sxt, st))
            return -1;
    }
    {
        ProfScope ps(m, PC_POINTWISE, 0.0, (double)B * SEG * 2 * 8.0);
        if (launch_normalize_input(m->xt_in, B, (size_t)SEG * 2, sxt, m->meanstd_t, st))
            return -1;
    }

    // ---- encoders (src/model_inference.cpp:155-212, src/encdec.cpp:8-164) ----
    // The time and frequency branches are independent through the encoders and through the decoders: the time
    // branch runs on a second stream so its small deep-level launches fill SMs the frequency branch leaves idle.
    // (Profiling mode keeps one stream: its per-class CUDA events bracket launches on m->stream.)
    cudaStream_t st_t = m->prof.on ? st : m->stream2;
    auto fork = [&]() -> int {
        if (st_t == st)
            return 0;
        DMX_CUDA(cudaEventRecord(m->ev_fork, st));
        DMX_CUDA(cudaStreamWaitEvent(st_t, m->ev_fork, 0));
        return 0;
    };
    auto join = [&]() -> int {
        if (st_t == st)
            return 0;
        DMX_CUDA(cudaEventRecord(m->ev_join, st_t));
        DMX_CUDA(cudaStreamWaitEvent(st, m->ev_join, 0));
        return 0;
    };
    if (fork())
        return -1;
    // bf16 mode: the 48- / 96-channel 1x1 rewrites of the encoders run on the ring kernel's tf32 variant, straight from the
    // fp32 activations (per 16 segments: 212 -> 142, 114 -> 92, 217 -> 185 us).  The deeper rewrites and the strided
    // convolutions of frequency encoders 1-3 were slower there (fp32 operands double the L2 -> SM traffic; 374 -> 440,
    // 167 -> 231, 164 -> 204 us) and stay on the one-tile gather kernel (DMX_TF32_CONV=1 / DMX_NO_TF32_ENC=1 to switch)
    static const bool tf32_env = getenv("DMX_NO_TF32_ENC") == nullptr;
    static const bool tf32_conv = getenv("DMX_TF32_CONV") != nullptr;
    const bool tf32e = tf32_env && m->use_tma && m->precision == DMX_PRECISION_BF16;
    const float *xin_f = m->spec, *xin_t = m->xt_in;
    for (int i = 0; i < 4; ++i)
    {
        const EncW &te = m->tenc[i];
        const EncW &fe = m->enc[i];
        const int C = fe.C;
        // time: conv k8 s4 p2 (+GELU), ceil-mode length
        f.st = st_t;
        {
            GemmParams p = f.base(xin_t, te.Cin, te.Cin, te.conv, m->cA_t, C, 1, TLENS[i], 1, TLENS[i + 1]);
            p.ntaps = 8;
            p.sv = 4;
            for (int t = 0; t < 8; ++t)
            {
                p.du[t] = 0;
                p.dv[t] = t - 2;
            }
            if (te.Cin == 2 && TLENS[i] % 2 == 0)
            {
                // stereo input: the same convolution over PAIRS of samples (4 values per row, k4 s2 p1) -- identical
                // K ordering (tap*2 + ch), but rows are 16 bytes, so the gather runs on the vector path
                p.lda = p.Cin = 4;
                p.Vi = TLENS[i] / 2;
                p.ntaps = 4;
                p.sv = 2;
                for (int t = 0; t < 4; ++t)
                    p.dv[t] = t - 1;
            }
            p.act = 1;
            f.gemm(p, PC_GEMM_CONV);
            f.dconv(m->cA_t, 1, TLENS[i + 1], te.dc, m->h_t, m->y2_t, 0);
            GemmParams r = f.base(m->cA_t, C, C, te.rewrite, m->te[i], C, 1, TLENS[i + 1], 1, TLENS[i + 1]);
            r.glu = 1;
            r.Nco = C;
            if (tf32e && C <= 96)
                r.Wt32 = te.rewrite.Wt32;
            f.gemm(r, PC_GEMM_CONV);
        }
        // freq: conv k8 s4 p2 along frequency (+GELU), frames are independent
        f.st = st;
        {
            const bool fused_rw =
                m->precision == DMX_PRECISION_BF16 && fe.rw_pk && fe.dc.fused_pk && dconv_mma_rw_supported(C, fe.dc.hid, T);
            const bool fused_cin = fused_rw && fe.cin_pk && fe.Cin == 4 && FREQS[i] == 4 * FREQS[i + 1];
            if (!fused_cin)
            {
                GemmParams p = f.base(xin_f, fe.Cin, fe.Cin, fe.conv, m->cA_f, C, FREQS[i], T, FREQS[i + 1], T);
                p.ntaps = 8;
                p.su = 4;
                for (int t = 0; t < 8; ++t)
                {
                    p.du[t] = t - 2;
                    p.dv[t] = 0;
                }
                p.act = 1;
                if (tf32e && i > 0 && tf32_conv)
                    p.Wt32 = fe.conv.Wt32;
                f.gemm(p, PC_GEMM_CONV);
            }
            if (fused_rw)
            {
                // (input conv +) DConv + rewrite / GLU / frequency embedding of this level in one launch (row stays on the SM)
                if (!f.rc)
                {
                    const int F = FREQS[i + 1], hid = fe.dc.hid;
                    ProfScope ps(m, PC_DCONV_FUSED,
                                 2.0 * (double)B * F * T * (2.0 * (3.0 * C * hid + 2.0 * 2.0 * C * hid) + 2.0 * C * C),
                                 (double)B * F * T * C * 8.0);
                    f.rc = launch_dconv_mma_rw(m->cA_f, fe.dc.fused_pk, fe.rw_pk, m->e[i], i == 0 ? m->freq_emb2 : nullptr, F,
                                               B * F, C, hid, fused_cin ? xin_f : nullptr, fe.cin_pk, st);
                }
            }
            else
            {
                f.dconv(m->cA_f, FREQS[i + 1], T, fe.dc, m->h_f, m->y2_f, 1);
                GemmParams r = f.base(m->cA_f, C, C, fe.rewrite, m->e[i], C, FREQS[i + 1], T, FREQS[i + 1], T);
                r.glu = 1;
                r.Nco = C;
                if (i == 0)
                {
                    r.rowadd = m->freq_emb2;
                    r.rowadd_ld = 48;
                }
                else if (tf32e && C <= 96)
                    r.Wt32 = fe.rewrite.Wt32;
                f.gemm(r, PC_GEMM_CONV);
            }
        }
        xin_t = m->te[i];
        xin_f = m->e[i];
    }

    f.st = st;
    if (join())
        return -1;
    // ---- cross-transformer (src/model_inference.cpp:214-309, src/crosstransformer.cpp:205-339) ----
    const int TF = 8 * T, TT = TLEN3;
    const float *tok_f = m->e[3], *tok_t = m->te[3];
    if (m->arch == DMX_ARCH_HTDEMUCS_4S)
    {
        f.gemm(f.linear(m->e[3], 384, m->up, m->up_f, d, TF), PC_GEMM_XFMR);
        f.gemm(f.linear(m->te[3], 384, m->up_t, m->up_tt, d, TT), PC_GEMM_XFMR);
        tok_f = m->up_f;
        tok_t = m->up_tt;
    }
    f.layer_norm(tok_f, m->xq, B * TF, d, m->nin_w, m->nin_b, m->pe2d, TF);
    f.layer_norm(tok_t, m->tq, B * TT, d, m->nint_w, m->nint_b, m->pe1d, TT);
    // per-layer test tier (the reference tests common_encoder_layer in isolation, test/test_layers.cpp:1973-2157): keep
    // the token streams after norm_in and after every layer
    auto keep_layer = [&](int slot) -> int {
        if (!m->layer_taps)
            return 0;
        const size_t nf = (size_t)m->maxB * TF * d, nt = (size_t)m->maxB * TT * d;
        if (!m->lt_f[slot] && (dev_alloc_t(m, &m->lt_f[slot], nf) || dev_alloc_t(m, &m->lt_t[slot], nt)))
            return -1;
        DMX_CUDA(cudaMemcpyAsync(m->lt_f[slot], m->xq, (size_t)B * TF * d * sizeof(float), cudaMemcpyDeviceToDevice, st));
        DMX_CUDA(cudaMemcpyAsync(m->lt_t[slot], m->tq, (size_t)B * TT * d * sizeof(float), cudaMemcpyDeviceToDevice, st));
        return 0;
    };
    if (keep_layer(0))
        return -1;
    for (int i = 0; i < 5; ++i)
    {
        if (i % 2 == 0)
        {
            f.tlayer(m->lay[i], m->xq, TF, nullptr, 0, m->ln_f, m->qkv_f, m->kv_f, m->ao_f, m->ff_f);
            f.tlayer(m->lay_t[i], m->tq, TT, nullptr, 0, m->ln_t, m->qkv_t, m->kv_t, m->ao_t, m->ff_t);
        }
        else
        {
            // freq queries attend the current time tokens; time queries attend the freq tokens from BEFORE
            // this layer's update (old_x, src/crosstransformer.cpp:286-295): normalise both key sets first.
            f.layer_norm_a(m->tq, m->kn_f, B * TT, d, m->lay[i].n2w, m->lay[i].n2b);
            f.layer_norm_a(m->xq, m->kn_t, B * TF, d, m->lay_t[i].n2w, m->lay_t[i].n2b);
            f.tlayer(m->lay[i], m->xq, TF, m->kn_f, TT, m->ln_f, m->qkv_f, m->kv_f, m->ao_f, m->ff_f);
            f.tlayer(m->lay_t[i], m->tq, TT, m->kn_t, TF, m->ln_t, m->qkv_t, m->kv_t, m->ao_t, m->ff_t);
        }
        if (f.rc || keep_layer(i + 1))
            return -1;
    }
    const float *dec_in_f = m->xq, *dec_in_t = m->tq;
    // bf16 mode: the decoders' rewrite convolutions run on the TMA-fed kernel whenever the (x + skip) image exists
    const bool tma = m->use_tma && m->precision == DMX_PRECISION_BF16;
    const void *img_f = nullptr, *img_t = nullptr;
    static const bool tma_ct_env = getenv("DMX_TMA_CONVTR") != nullptr;
    const bool tma_ct = tma && tma_ct_env;
    if (m->arch == DMX_ARCH_HTDEMUCS_4S)
    {
        GemmParams pd = f.linear(m->xq, d, m->down, m->x3d, 384, TF), pdt = f.linear(m->tq, d, m->down_t, m->xt3d, 384, TT);
        if (tma)
        {
            // decoder 0's skip is added here and the sum also leaves as the bf16 image its rewrite conv reads
            pd.addC = m->e[3];
            pd.Cbf = m->dimg_f[0];
            pdt.addC = m->te[3];
            pdt.Cbf = m->dimg_t[0];
            pd.Cbf_ld = pdt.Cbf_ld = tc_tma_cpitch(384);
            img_f = m->dimg_f[0];
            img_t = m->dimg_t[0];
        }
        f.gemm(pd, PC_GEMM_XFMR);
        f.gemm(pdt, PC_GEMM_XFMR);
        dec_in_f = m->x3d;
        dec_in_t = m->xt3d;
    }

    // ---- decoders (src/model_inference.cpp:314-344, src/encdec.cpp:166-361) ----
    if (fork())
        return -1;
    for (int i = 0; i < 4; ++i)
    {
        const DecW &fd = m->dec[i];
        const DecW &td = m->tdec[i];
        const int C = fd.C, F = FREQS[4 - i], Lin = TLENS[4 - i], Lout = TLENS[3 - i];
        f.st = st;
        {
            // (x + skip) -> 3x3 conv -> GLU
            GemmParams p = f.base(dec_in_f, C, C, fd.rewrite, m->cA_f, C, F, T, F, T);
            if (i == 0 && !img_f)
                p.A2 = m->e[3]; // later decoders find their skip already added by the previous convT's epilogue
            if (img_f && fd.rewrite.Wt)
            {
                p.Aimg = img_f;
                p.Aimg_ld = tc_tma_cpitch(C);
                p.Wt = fd.rewrite.Wt;
            }
            p.ntaps = 9;
            for (int kh = 0; kh < 3; ++kh)
                for (int kw = 0; kw < 3; ++kw)
                {
                    p.du[kh * 3 + kw] = kh - 1;
                    p.dv[kh * 3 + kw] = kw - 1;
                }
            p.glu = 1;
            p.Nco = C;
            f.gemm(p, PC_GEMM_CONV);
            // (the transposed convolutions can run on the TMA-fed kernel too -- DMX_TMA_CONVTR, 384-channel level only -- but they
            // are bound by their scatter epilogue, and bf16 image stores in the fused DConv kernels cost more than the operand
            // path saved: 41.3 -> 42.4 ms/step with all four levels on it)
            const bool cimg = f.dconv(m->cA_f, F, T, fd.dc, m->h_f, m->y2_f, 1, tma_ct ? m->cimg_f : nullptr, tc_tma_cpitch(C));
            // ConvTranspose k(8,1) s(4,1) along frequency as a 2-tap GEMM, crop [2, 2+4F)
            GemmParams q = f.base(m->cA_f, C, C, fd.convtr, m->dx[i], fd.Cout, F, T, F + 1, T);
            if (cimg && fd.convtr.Wt)
            {
                q.Aimg = m->cimg_f;
                q.Aimg_ld = tc_tma_cpitch(C);
                q.Wt = fd.convtr.Wt;
            }
            q.ntaps = 2;
            q.du[0] = -1;
            q.du[1] = 0;
            q.dv[0] = q.dv[1] = 0;
            q.Nco = fd.Cout;
            q.Uout = 4 * F;
            q.Vout = T;
            q.ous = 4;
            q.our = 1;
            q.ouo = -2;
            q.out_vu = i == 3; // the spectrogram heads leave frame-major ([b][t][bin][c]): what the iSTFT reads in one piece
            q.act = i < 3 ? 1 : 0;
            if (i < 3)
                q.addC = m->e[2 - i]; // the next decoder's skip (src/encdec.cpp:172), added once here
            img_f = nullptr;
            if (tma && i < 3)
            {
                q.Cbf = m->dimg_f[i + 1];
                q.Cbf_ld = tc_tma_cpitch(fd.Cout);
                q.skip_c = !m->keep_stages && m->dec[i + 1].rewrite.Wt != nullptr; // only the image is read again
                img_f = m->dimg_f[i + 1];
            }
            f.gemm(q, PC_GEMM_CONV);
        }
        f.st = st_t;
        {
            GemmParams p = f.base(dec_in_t, C, C, td.rewrite, m->cA_t, C, 1, Lin, 1, Lin);
            if (i == 0 && !img_t)
                p.A2 = m->te[3];
            if (img_t && td.rewrite.Wt)
            {
                p.Aimg = img_t;
                p.Aimg_ld = tc_tma_cpitch(C);
                p.Wt = td.rewrite.Wt;
            }
            p.ntaps = 3;
            for (int t = 0; t < 3; ++t)
            {
                p.du[t] = 0;
                p.dv[t] = t - 1;
            }
            p.glu = 1;
            p.Nco = C;
            f.gemm(p, PC_GEMM_CONV);
            const bool cimg = f.dconv(m->cA_t, 1, Lin, td.dc, m->h_t, m->y2_t, 0, tma_ct ? m->cimg_t : nullptr, tc_tma_cpitch(C));
            GemmParams q = f.base(m->cA_t, C, C, td.convtr, m->dxt[i], td.Cout, 1, Lin, 1, Lin + 1);
            if (cimg && td.convtr.Wt)
            {
                q.Aimg = m->cimg_t;
                q.Aimg_ld = tc_tma_cpitch(C);
                q.Wt = td.convtr.Wt;
            }
            q.ntaps = 2;
            q.dv[0] = -1;
            q.dv[1] = 0;
            q.du[0] = q.du[1] = 0;
            q.Nco = td.Cout;
            q.Uout = 1;
            q.Vout = Lout;
            q.ous = 0;
            q.ovs = 4;
            q.ovr = 1;
            q.ovo = -2;
            q.act = i < 3 ? 1 : 0;
            if (i < 3)
                q.addC = m->te[2 - i];
            img_t = nullptr;
            if (tma && i < 3)
            {
                q.Cbf = m->dimg_t[i + 1];
                q.Cbf_ld = tc_tma_cpitch(td.Cout);
                q.skip_c = !m->keep_stages && m->tdec[i + 1].rewrite.Wt != nullptr;
                img_t = m->dimg_t[i + 1];
            }
            f.gemm(q, PC_GEMM_CONV);
        }
        dec_in_f = m->dx[i];
        dec_in_t = m->dxt[i];
    }
    f.st = st;
    if (f.rc || join())
        return -1;

    // ---- de-normalise, per-source iSTFT, add time branch (src/model_inference.cpp:352-474) ----
    {
        // algorithmic bytes of iSTFT + mix (66.0 MB / segment for 4 sources)
        ProfScope ps(m, PC_ISTFT_FRAMES, 0.0,
                     (double)B * ((double)4 * S * NBINS * T * 4 + (double)2 * S * SEG * 4 + (double)S * 2 * SEG * 4), 2);
        if (launch_istft(m->dx[3], B, S, m->meanstd_x, m->window, m->tw, m->wss, m->dxt[3], m->meanstd_t, m->frames,
                         d_out, st, 1))
            return -1;
    }

    m->taps.clear();
    set_tap(m, "x", m->spec, NBINS, T, 4);
    set_tap(m, "xt", m->xt_in, 1, SEG, 2);
    for (int i = 0; i < 4; ++i)
    {
        const std::string si = std::to_string(i);
        set_tap(m, ("saved_" + si).c_str(), m->e[i], FREQS[i + 1], T, CH[i + 1]);
        set_tap(m, ("savedt_" + si).c_str(), m->te[i], 1, TLENS[i + 1], CH[i + 1]);
    }
    if (m->layer_taps)
        for (int i = 0; i < 6; ++i)
        {
            const std::string tag = i == 0 ? "in" : std::to_string(i - 1);
            set_tap(m, ("x_ct_" + tag).c_str(), m->lt_f[i], 8, T, d);
            set_tap(m, ("xt_ct_" + tag).c_str(), m->lt_t[i], 1, TLEN3, d);
        }
    set_tap(m, "x_3_up", m->xq, 8, T, d);
    set_tap(m, "xt_3_up", m->tq, 1, TLEN3, d);
    const bool pre3 = tma && m->arch == DMX_ARCH_HTDEMUCS_4S; // x3d / xt3d hold value + skip (see above)
    set_tap(m, "x_3", const_cast<float *>(m->arch == DMX_ARCH_HTDEMUCS_4S ? m->x3d : m->xq), 8, T,
            m->arch == DMX_ARCH_HTDEMUCS_4S ? 384 : d, pre3 ? m->e[3] : nullptr);
    set_tap(m, "xt_3", const_cast<float *>(m->arch == DMX_ARCH_HTDEMUCS_4S ? m->xt3d : m->tq), 1, TLEN3,
            m->arch == DMX_ARCH_HTDEMUCS_4S ? 384 : d, pre3 ? m->te[3] : nullptr);
    const char *fn[4] = {"x_2", "x_1", "x_0", "x_out"};
    const char *tn[4] = {"xt_2", "xt_1", "xt_0", "xt_out"};
    for (int i = 0; i < 4; ++i)
    {
        set_tap(m, fn[i], m->dx[i], FREQS[3 - i], T, m->dec[i].Cout, i < 3 ? m->e[2 - i] : nullptr, i == 3);
        set_tap(m, tn[i], m->dxt[i], 1, TLENS[3 - i], m->tdec[i].Cout, i < 3 ? m->te[2 - i] : nullptr);
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------
// Test hook: out[M][N] = act(A[M][K] x W[N][K]^T + bias) through the production GEMM kernels (fp32 CUDA-core
// or bf16 tcgen05 path), with temporary device buffers.  Lets the tests check the tensor-core kernel's
// operand layouts and descriptors in isolation against numpy.
int debug_linear(int device, int precision, int M, int N, int K, const float *A, const float *W, const float *bias,
                 int act, float *out, int iters, float *ms_out)
{
    DMX_CUDA(cudaSetDevice(device));
    dmx_model tmp;
    tmp.device = device;
    int rc = 0;
    float *dA = nullptr, *dC = nullptr;
    Packed pk;
    do
    {
        if ((rc = tc_init_device()))
            break;
        rc = pack(
            &tmp, pk, K, N, false, [=](int k, int n) { return W[(size_t)n * K + k]; },
            [=](int n) { return bias ? bias[n] : 0.f; });
        if (rc)
            break;
        if ((rc = dev_alloc_t(&tmp, &dA, (size_t)M * K)) || (rc = dev_alloc_t(&tmp, &dC, (size_t)M * N)))
            break;
        if (cudaMemcpy(dA, A, (size_t)M * K * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess ||
            cudaMemset(dC, 0, (size_t)M * N * sizeof(float)) != cudaSuccess)
        {
            set_error("debug_linear: copy failed");
            rc = -1;
            break;
        }
        GemmParams p;
        p.A = dA;
        p.lda = K;
        p.Cin = K;
        p.K = K;
        p.W = pk.W;
        p.Wb = pk.Wb;
        p.ldw = pk.ldw;
        p.bias = pk.bias;
        p.N = N;
        p.Ui = p.Uo = 1;
        p.Vi = p.Vo = M;
        p.M = M;
        p.C = dC;
        p.ldc = N;
        p.Nco = N;
        p.Uout = 1;
        p.Vout = M;
        p.act = act;
        rc = precision == DMX_PRECISION_BF16 ? launch_gemm_tc(p, pk.Wb, 0) : launch_gemm(p, 0);
        if (rc)
            break;
        if (iters > 0 && ms_out)
        {
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0);
            cudaEventCreate(&e1);
            cudaDeviceSynchronize();
            cudaEventRecord(e0, 0);
            for (int it = 0; it < iters && !rc; ++it)
                rc = precision == DMX_PRECISION_BF16 ? launch_gemm_tc(p, pk.Wb, 0) : launch_gemm(p, 0);
            cudaEventRecord(e1, 0);
            cudaEventSynchronize(e1);
            float t = 0.f;
            cudaEventElapsedTime(&t, e0, e1);
            *ms_out = t / (float)iters;
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
            if (rc)
                break;
        }
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess)
        {
            set_error(std::string("debug_linear: kernel failed: ") + cudaGetErrorString(e));
            rc = -1;
            break;
        }
        if (cudaMemcpy(out, dC, (size_t)M * N * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess)
        {
            set_error("debug_linear: copy back failed");
            rc = -1;
        }
    } while (0);
    for (void *q : tmp.allocs)
        cudaFree(q);
    tmp.allocs.clear();
    return rc;
}

// Test hook for the TMA-fed convolution kernel in isolation: a stride-1 "same" convolution with an arbitrary tap set over a
// channels-last image, out = act / GLU(conv(img, W) + bias).  img (B, U, V, Cin) and W (N, ntaps * Cin) (k = tap * Cin + c)
// are host fp32 (rounded to bf16 here, as the producing epilogue / the weight packer would); out (B, U, V, N or N / 2).
int debug_conv(int device, int B, int U, int V, int Cin, int N, int ntaps, const int *du, const int *dv, const float *img,
               const float *W, const float *bias, int act, int glu, int share_dv, float *out, int iters, float *ms_out)
{
    DMX_CUDA(cudaSetDevice(device));
    if (!tc_tma_available())
    {
        set_error("debug_conv: the TMA path is not available");
        return -1;
    }
    dmx_model tmp;
    tmp.device = device;
    int rc = 0;
    void *dimg = nullptr;
    float *dC = nullptr;
    Packed pk;
    const int K = ntaps * Cin, nout = glu ? N / 2 : N;
    const size_t npos = (size_t)B * U * V;
    do
    {
        TmaPlan pl;
        if ((rc = tc_init_device()) || (rc = tc_tma_plan(ntaps, du, dv, Cin, share_dv != 0, &pl)))
        {
            if (rc)
                set_error("debug_conv: bad tap set / channel count");
            break;
        }
        rc = pack(
            &tmp, pk, K, N, glu != 0, [=](int k, int n) { return W[(size_t)n * K + k]; }, [=](int n) { return bias ? bias[n] : 0.f; },
            &pl, Cin);
        if (rc)
            break;
        const int cp = tc_tma_cpitch(Cin);
        std::vector<uint16_t> hb(npos * cp, 0x7fc0); // pad channels hold NaN: they must never reach the GEMM
        for (size_t i = 0; i < npos; ++i)
            for (int c = 0; c < Cin; ++c)
                hb[i * cp + c] = f32_to_bf16_bits(img[i * Cin + c]);
        if ((rc = dev_alloc(&tmp, &dimg, hb.size() * 2)) || (rc = dev_alloc_t(&tmp, &dC, npos * nout)))
            break;
        if (cudaMemcpy(dimg, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice) != cudaSuccess ||
            cudaMemset(dC, 0, npos * nout * sizeof(float)) != cudaSuccess)
        {
            set_error("debug_conv: copy failed");
            rc = -1;
            break;
        }
        GemmParams p;
        p.Aimg = dimg;
        p.Aimg_ld = cp;
        p.Wt = pk.Wt;
        p.lda = Cin;
        p.Cin = Cin;
        p.K = K;
        p.ntaps = ntaps;
        for (int t = 0; t < ntaps; ++t)
        {
            p.du[t] = du[t];
            p.dv[t] = dv[t];
        }
        p.bias = pk.bias;
        p.N = N;
        p.Ui = p.Uo = U;
        p.Vi = p.Vo = V;
        p.M = B * U * V;
        p.C = dC;
        p.ldc = nout;
        p.Nco = nout;
        p.Uout = U;
        p.Vout = V;
        p.act = act;
        p.glu = glu;
        p.t_noshare = share_dv ? 0 : 1;
        if ((rc = launch_gemm_tma(p, 0)))
            break;
        if (iters > 0 && ms_out)
        {
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0);
            cudaEventCreate(&e1);
            cudaDeviceSynchronize();
            cudaEventRecord(e0, 0);
            for (int it = 0; it < iters && !rc; ++it)
                rc = launch_gemm_tma(p, 0);
            cudaEventRecord(e1, 0);
            cudaEventSynchronize(e1);
            float t = 0.f;
            cudaEventElapsedTime(&t, e0, e1);
            *ms_out = t / (float)iters;
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
            if (rc)
                break;
        }
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess)
        {
            set_error(std::string("debug_conv: kernel failed: ") + cudaGetErrorString(e));
            rc = -1;
            break;
        }
        if (cudaMemcpy(out, dC, npos * nout * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess)
        {
            set_error("debug_conv: copy back failed");
            rc = -1;
        }
    } while (0);
    for (void *q : tmp.allocs)
        cudaFree(q);
    tmp.allocs.clear();
    return rc;
}

} // namespace dmx
