// hdemucs_mmi (v3 hybrid Demucs) on device: weight repacking, workspace and the per-segment forward
// (reference: demucscpp_v3::model_v3_inference, src/model_inference.cpp:477-856) over a batch of segments.
// Same channels-last layout and tap-GEMM machinery as the htdemucs model (model.cu); what is specific here:
//   * DConv with compress 4 in encoders 0-3 and no DConv in the decoders              (src/layers.cpp:723-894)
//   * encoder 4 collapses the 8 remaining frequency bins into 1 and merges the time branch ("inject"),
//     encoder 5 / decoder 0 are shared, GroupNorm with 4 groups around them           (src/encdec.cpp:519-726)
//   * the DConv of encoders 4/5 carries a 2-layer BiLSTM and LocalState attention     (src/layers.cpp:896-1113)
//   * the time decoder chain restarts from frequency decoder 1's pre-ConvT activation (src/model_inference.cpp:686-692)
#include "model_common.cuh"

namespace dmx
{

static const int T5 = 168; // shared-branch length after encoder 5 (src/model.hpp:679-681)

static int pack_la(dmx_model *m, DConvLaW &w, const std::string &p, int C, int hid)
{
    if (pack_conv(m, w.c1, p + "0.weight", p + "0.bias", hid, C, 3, false))
        return -1;
    if (upload_vec(m, &w.g1w, p + "1.weight", hid) || upload_vec(m, &w.g1b, p + "1.bias", hid))
        return -1;
    const std::string L = p + "3.lstm.";
    for (int layer = 0; layer < 2; ++layer)
    {
        const int in = layer == 0 ? hid : 2 * hid;
        const std::string sl = "_l" + std::to_string(layer);
        const HostTensor *wf = get(m, L + "weight_ih" + sl, (size_t)4 * hid * in);
        const HostTensor *wb = get(m, L + "weight_ih" + sl + "_reverse", (size_t)4 * hid * in);
        const HostTensor *bif = get(m, L + "bias_ih" + sl, 4 * hid), *bhf = get(m, L + "bias_hh" + sl, 4 * hid);
        const HostTensor *bib = get(m, L + "bias_ih" + sl + "_reverse", 4 * hid),
                         *bhb = get(m, L + "bias_hh" + sl + "_reverse", 4 * hid);
        if (!wf || !wb || !bif || !bhf || !bib || !bhb)
            return -1;
        const float *wfd = wf->data.data(), *wbd = wb->data.data();
        const float *b1 = bif->data.data(), *b2 = bhf->data.data(), *b3 = bib->data.data(), *b4 = bhb->data.data();
        const int G4 = 4 * hid;
        // input projections of both directions in one GEMM: columns [0,4h) forward, [4h,8h) backward;
        // the two bias vectors of each direction are summed (src/lstm.cpp:89-103)
        if (pack(
                m, layer == 0 ? w.ih0 : w.ih1, in, 2 * G4, false,
                [=](int k, int n) { return n < G4 ? wfd[(size_t)n * in + k] : wbd[(size_t)(n - G4) * in + k]; },
                [=](int n) { return n < G4 ? b1[n] + b2[n] : b3[n - G4] + b4[n - G4]; }))
            return -1;
        for (int dir = 0; dir < 2; ++dir)
        {
            const HostTensor *hh = get(m, L + "weight_hh" + sl + (dir ? "_reverse" : ""), (size_t)G4 * hid);
            if (!hh)
                return -1;
            std::vector<float> t((size_t)hid * G4);
            for (int j = 0; j < G4; ++j)
                for (int k = 0; k < hid; ++k)
                    t[(size_t)k * G4 + j] = hh->data[(size_t)j * hid + k];
            if (upload(m, &w.whh_t[layer][dir], t) || upload(m, &w.whh[layer][dir], hh->data))
                return -1;
        }
    }
    if (pack_linear(m, w.lin, p + "3.linear.weight", p + "3.linear.bias", hid, 2 * hid, 0, hid))
        return -1;
    {
        // fused 1x1 convs of LocalState: (query | key | content | query_decay)
        const HostTensor *q = get(m, p + "4.query.weight", (size_t)hid * hid), *k = get(m, p + "4.key.weight", (size_t)hid * hid),
                         *c = get(m, p + "4.content.weight", (size_t)hid * hid),
                         *d = get(m, p + "4.query_decay.weight", (size_t)16 * hid);
        const HostTensor *qb = get(m, p + "4.query.bias", hid), *kb = get(m, p + "4.key.bias", hid),
                         *cb = get(m, p + "4.content.bias", hid), *db = get(m, p + "4.query_decay.bias", 16);
        if (!q || !k || !c || !d || !qb || !kb || !cb || !db)
            return -1;
        const float *qd = q->data.data(), *kd = k->data.data(), *cd = c->data.data(), *dd = d->data.data();
        const float *qbd = qb->data.data(), *kbd = kb->data.data(), *cbd = cb->data.data(), *dbd = db->data.data();
        if (pack(
                m, w.qkcd, hid, 3 * hid + 16, false,
                [=](int kk, int n) {
                    if (n < hid)
                        return qd[(size_t)n * hid + kk];
                    if (n < 2 * hid)
                        return kd[(size_t)(n - hid) * hid + kk];
                    if (n < 3 * hid)
                        return cd[(size_t)(n - 2 * hid) * hid + kk];
                    return dd[(size_t)(n - 3 * hid) * hid + kk];
                },
                [=](int n) {
                    if (n < hid)
                        return qbd[n];
                    if (n < 2 * hid)
                        return kbd[n - hid];
                    if (n < 3 * hid)
                        return cbd[n - 2 * hid];
                    return dbd[n - 3 * hid];
                }))
            return -1;
    }
    if (pack_linear(m, w.proj, p + "4.proj.weight", p + "4.proj.bias", hid, hid, 0, hid))
        return -1;
    if (pack_conv(m, w.c2, p + "5.weight", p + "5.bias", 2 * C, hid, 1, false))
        return -1;
    if (upload_vec(m, &w.g2w, p + "6.weight", 2 * C) || upload_vec(m, &w.g2b, p + "6.bias", 2 * C) ||
        upload_vec(m, &w.scale, p + "8.scale", C))
        return -1;
    return 0;
}

int model_v3_finalize(dmx_model *m)
{
    if (m->finalized)
        return 0;
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
    m->S = 4;
    m->d_model = 0;
    if (m->host.size() != 395) // the reference rejects files with unknown tensors (src/model_load.cpp:2142-2147)
    {
        set_error("model_v3_finalize: expected 395 tensors, got " + std::to_string(m->host.size()));
        return -1;
    }
    V3Model *v = new V3Model();
    m->v3 = v;
    for (int i = 0; i < 4; ++i)
    {
        const int cin_f = i == 0 ? 4 : CH[i], cin_t = i == 0 ? 2 : CH[i], C = CH[i + 1];
        const std::string e = "encoder." + std::to_string(i), te = "tencoder." + std::to_string(i);
        v->enc[i].Cin = cin_f;
        v->enc[i].C = C;
        v->tenc[i].Cin = cin_t;
        v->tenc[i].C = C;
        if (pack_conv(m, v->enc[i].conv, e + ".conv.weight", e + ".conv.bias", C, cin_f, 8, false) ||
            pack_conv(m, v->enc[i].rewrite, e + ".rewrite.weight", e + ".rewrite.bias", 2 * C, C, 1, true) ||
            pack_dconv(m, v->enc[i].dc, e, C, 4) ||
            pack_conv(m, v->tenc[i].conv, te + ".conv.weight", te + ".conv.bias", C, cin_t, 8, false) ||
            pack_conv(m, v->tenc[i].rewrite, te + ".rewrite.weight", te + ".rewrite.bias", 2 * C, C, 1, true) ||
            pack_dconv(m, v->tenc[i].dc, te, C, 4))
            return -1;
    }
    if (pack_conv(m, v->tenc4, "tencoder.4.conv.weight", "tencoder.4.conv.bias", 768, 384, 8, false) ||
        pack_conv(m, v->enc4_conv, "encoder.4.conv.weight", "encoder.4.conv.bias", 768, 384, 8, false) ||
        pack_conv(m, v->enc4_rw, "encoder.4.rewrite.weight", "encoder.4.rewrite.bias", 1536, 768, 1, false) ||
        pack_conv(m, v->enc5_conv, "encoder.5.conv.weight", "encoder.5.conv.bias", 1536, 768, 4, false) ||
        pack_conv(m, v->enc5_rw, "encoder.5.rewrite.weight", "encoder.5.rewrite.bias", 3072, 1536, 1, false))
        return -1;
    if (upload_vec(m, &v->e4n1w, "encoder.4.norm1.weight", 768) || upload_vec(m, &v->e4n1b, "encoder.4.norm1.bias", 768) ||
        upload_vec(m, &v->e4n2w, "encoder.4.norm2.weight", 1536) || upload_vec(m, &v->e4n2b, "encoder.4.norm2.bias", 1536) ||
        upload_vec(m, &v->e5n1w, "encoder.5.norm1.weight", 1536) || upload_vec(m, &v->e5n1b, "encoder.5.norm1.bias", 1536) ||
        upload_vec(m, &v->e5n2w, "encoder.5.norm2.weight", 3072) || upload_vec(m, &v->e5n2b, "encoder.5.norm2.bias", 3072))
        return -1;
    for (int e = 0; e < 2; ++e)
        for (int j = 0; j < 2; ++j)
            if (pack_la(m, v->la[e][j], "encoder." + std::to_string(4 + e) + ".dconv.layers." + std::to_string(j) + ".",
                        e == 0 ? 768 : 1536, e == 0 ? 192 : 384))
                return -1;
    if (pack_conv(m, v->d0_rw, "decoder.0.rewrite.weight", "decoder.0.rewrite.bias", 3072, 1536, 3, false) ||
        pack_convtr(m, v->d0_ct, "decoder.0.conv_tr.weight", "decoder.0.conv_tr.bias", 1536, 768, 4) ||
        pack_conv(m, v->d1_rw, "decoder.1.rewrite.weight", "decoder.1.rewrite.bias", 1536, 768, 9, false) ||
        pack_convtr(m, v->d1_ct, "decoder.1.conv_tr.weight", "decoder.1.conv_tr.bias", 768, 384, 8) ||
        pack_convtr(m, v->td0_ct, "tdecoder.0.conv_tr.weight", "tdecoder.0.conv_tr.bias", 768, 384, 8))
        return -1;
    if (upload_vec(m, &v->d0n1w, "decoder.0.norm1.weight", 3072) || upload_vec(m, &v->d0n1b, "decoder.0.norm1.bias", 3072) ||
        upload_vec(m, &v->d0n2w, "decoder.0.norm2.weight", 768) || upload_vec(m, &v->d0n2b, "decoder.0.norm2.bias", 768) ||
        upload_vec(m, &v->d1n1w, "decoder.1.norm1.weight", 1536) || upload_vec(m, &v->d1n1b, "decoder.1.norm1.bias", 1536) ||
        upload_vec(m, &v->d1n2w, "decoder.1.norm2.weight", 384) || upload_vec(m, &v->d1n2b, "decoder.1.norm2.bias", 384) ||
        upload_vec(m, &v->td0n2w, "tdecoder.0.norm2.weight", 384) || upload_vec(m, &v->td0n2b, "tdecoder.0.norm2.bias", 384))
        return -1;
    for (int l = 0; l < 4; ++l)
    {
        const int C = CH[4 - l];
        const int cout_f = l == 3 ? 16 : CH[3 - l], cout_t = l == 3 ? 8 : CH[3 - l];
        const std::string dn = "decoder." + std::to_string(l + 2), tn = "tdecoder." + std::to_string(l + 1);
        v->cdec[l].C = C;
        v->cdec[l].Cout = cout_f;
        v->ctdec[l].C = C;
        v->ctdec[l].Cout = cout_t;
        const TmaPlan plf = rewrite_plan(C, true), plt = rewrite_plan(C, false);
        if (pack_conv(m, v->cdec[l].rewrite, dn + ".rewrite.weight", dn + ".rewrite.bias", 2 * C, C, 9, true, &plf) ||
            pack_convtr(m, v->cdec[l].convtr, dn + ".conv_tr.weight", dn + ".conv_tr.bias", C, cout_f) ||
            pack_conv(m, v->ctdec[l].rewrite, tn + ".rewrite.weight", tn + ".rewrite.bias", 2 * C, C, 3, true, &plt) ||
            pack_convtr(m, v->ctdec[l].convtr, tn + ".conv_tr.weight", tn + ".conv_tr.bias", C, cout_t))
            return -1;
    }
    {
        const HostTensor *t = get(m, "freq_emb.embedding.weight", 512 * 48);
        if (!t)
            return -1;
        std::vector<float> e2(t->data);
        for (auto &x : e2)
            x *= 2.0f;
        if (upload(m, &m->freq_emb2, e2))
            return -1;
        if (upload(m, &v->ones, std::vector<float>(3072, 1.0f)))
            return -1;
    }
    {
        const std::vector<float> w = make_window();
        if (upload(m, &m->window, w) || upload(m, &m->wss, make_wss(w)) || upload(m, &m->tw, make_twiddles()))
            return -1;
    }
    m->host.clear();

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
        rc |= A(&m->dx[i], (size_t)FREQS[3 - i] * T * v->cdec[i].Cout);
        rc |= A(&m->dxt[i], (size_t)TLENS[3 - i] * v->ctdec[i].Cout);
        if (i > 0 && tc_tma_available())
        {
            // bf16 images of the common decoders' inputs (transposed conv i-1 + skip), read by the TMA-fed rewrite convolutions
            m->use_tma = true;
            rc |= dev_alloc(m, &m->dimg_f[i], B * (size_t)FREQS[4 - i] * T * tc_tma_cpitch(CH[4 - i]) * 2);
            rc |= dev_alloc(m, &m->dimg_t[i], B * (size_t)TLENS[4 - i] * tc_tma_cpitch(CH[4 - i]) * 2);
        }
    }
    rc |= A(&m->h_f, (size_t)512 * T * 12);
    rc |= A(&m->y2_f, (size_t)512 * T * 96);
    rc |= A(&m->h_t, (size_t)TLEN0 * 12);
    rc |= A(&m->y2_t, (size_t)TLEN0 * 96);
    rc |= A(&v->xt4, (size_t)T * 768);
    rc |= A(&v->g4, (size_t)T * 768);
    rc |= A(&v->r4, (size_t)T * 1536);
    rc |= A(&v->x4, (size_t)T * 768);
    rc |= A(&v->y5, (size_t)T5 * 1536);
    rc |= A(&v->g5, (size_t)T5 * 1536);
    rc |= A(&v->r5, (size_t)T5 * 3072);
    rc |= A(&v->x5, (size_t)T5 * 1536);
    rc |= A(&v->d0r, (size_t)T5 * 3072);
    rc |= A(&v->d0g, (size_t)T5 * 1536);
    rc |= A(&v->d0ct, (size_t)(2 * T5 + 2) * 768);
    rc |= A(&v->xd0, (size_t)T * 768);
    rc |= A(&v->d1r, (size_t)T * 1536);
    rc |= A(&v->pre, (size_t)T * 768);
    rc |= A(&v->d1ct, (size_t)8 * T * 384);
    rc |= A(&v->x3, (size_t)8 * T * 384);
    rc |= A(&v->td0ct, (size_t)(4 * T + 4) * 384);
    rc |= A(&v->xt3, (size_t)TLEN3 * 384);
    const size_t TH = (size_t)T * 192; // == T5 * 384
    rc |= A(&v->hraw, TH);
    rc |= A(&v->seq, TH);
    rc |= A(&v->lpre, TH * 8);
    rc |= A(&v->l0, TH * 2);
    rc |= A(&v->l1, TH * 2);
    rc |= A(&v->lo, TH);
    rc |= A(&v->qk, TH * 3 + (size_t)T * 16);
    rc |= A(&v->att, TH);
    rc |= A(&v->y2la, (size_t)T * 1536);
    rc |= A(&m->frames, (size_t)4 * T * NFFT * 2);
    rc |= A(&m->seg_out, (size_t)4 * SEG * 2);
    rc |= A(&m->meanstd_x, 2);
    rc |= A(&m->meanstd_t, 2);
    if (rc)
        return -1;
    DMX_CUDA(cudaMemset(m->h_f, 0, B * (size_t)512 * T * 12 * sizeof(float)));
    DMX_CUDA(cudaMemset(m->h_t, 0, B * (size_t)TLEN0 * 12 * sizeof(float)));
    m->stats_cap = B * 2 * (size_t)(2 * 2 * (512 + 128 + 32 + 8) * 2 + 512);
    if (dev_alloc_t(m, &m->stats, m->stats_cap) || dev_alloc_t(m, &m->track_stats, 2))
        return -1;
    if (noalloc_warm_up(m->device))
        return -1;
    m->finalized = true;
    return 0;
}

// DConv with BiLSTM + LocalState attention, in place on x [B][Tn][C] (src/layers.cpp:896-1113)
static void dconv_la(Fwd &f, float *x, int Tn, int C, int hid, const DConvLaW *w2)
{
    dmx_model *m = f.m;
    V3Model *v = m->v3;
    const int B = f.B;
    for (int j = 0; j < 2; ++j)
    {
        const DConvLaW &w = w2[j];
        const int dil = 1 << j;
        GemmParams p = f.base(x, C, C, w.c1, v->hraw, hid, 1, Tn, 1, Tn);
        p.ntaps = 3;
        for (int t = 0; t < 3; ++t)
        {
            p.du[t] = 0;
            p.dv[t] = (t - 1) * dil;
        }
        f.gemm(p, PC_GEMM_DCONV);
        f.group_norm_g(v->hraw, v->seq, Tn, Tn, 0, hid, 1, w.g1w, w.g1b, 1, 0);          // GroupNorm(1) + GELU
        // 2-layer BiLSTM: hoisted input projections (both directions at once) + sequential recurrence
        f.gemm(f.linear(v->seq, hid, w.ih0, v->lpre, 8 * hid, Tn), PC_GEMM_DCONV);
        if (!f.rc)
        {
            ProfScope ps(m, PC_LSTM, 2.0 * B * Tn * 2.0 * 4.0 * hid * hid);
            f.rc = launch_lstm_recurrent(v->lpre, w.whh[0][0], w.whh[0][1], w.whh_t[0][0], w.whh_t[0][1], v->l0, B, Tn, hid, f.st);
        }
        f.gemm(f.linear(v->l0, 2 * hid, w.ih1, v->lpre, 8 * hid, Tn), PC_GEMM_DCONV);
        if (!f.rc)
        {
            ProfScope ps(m, PC_LSTM, 2.0 * B * Tn * 2.0 * 4.0 * hid * hid);
            f.rc = launch_lstm_recurrent(v->lpre, w.whh[1][0], w.whh[1][1], w.whh_t[1][0], w.whh_t[1][1], v->l1, B, Tn, hid, f.st);
        }
        // Linear(2h -> h) + skip: lo = seq + W l1 + b
        if (!f.rc && cudaMemcpyAsync(v->lo, v->seq, (size_t)B * Tn * hid * sizeof(float), cudaMemcpyDeviceToDevice, f.st) !=
                         cudaSuccess)
        {
            set_error("dconv_la: copy failed");
            f.rc = -1;
        }
        GemmParams pl = f.linear(v->l1, 2 * hid, w.lin, v->lo, hid, Tn);
        pl.gamma = v->ones;
        f.gemm(pl, PC_GEMM_DCONV);
        // LocalState attention: fused q/k/content/decay projections, attention, output projection + residual
        f.gemm(f.linear(v->lo, hid, w.qkcd, v->qk, 3 * hid + 16, Tn), PC_GEMM_DCONV);
        if (!f.rc)
        {
            ProfScope ps(m, PC_ATTENTION, 4.0 * B * (double)Tn * Tn * hid);
            f.rc = launch_local_attention(v->qk, v->att, B, Tn, hid, f.st);
        }
        GemmParams pp = f.linear(v->att, hid, w.proj, v->lo, hid, Tn);
        pp.gamma = v->ones;
        f.gemm(pp, PC_GEMM_DCONV);
        // 1x1 -> GroupNorm(1) -> GLU -> LayerScale -> residual
        double *s2 = f.stat_slots(B);
        GemmParams q = f.base(v->lo, hid, hid, w.c2, v->y2la, 2 * C, 1, Tn, 1, Tn);
        q.stats = s2;
        q.stat_per_u = 0;
        f.gemm(q, PC_GEMM_DCONV);
        if (!f.rc)
        {
            ProfScope ps(m, PC_POINTWISE, 0.0, (double)B * Tn * C * 16.0);
            f.rc = launch_gn_glu_residual(v->y2la, x, B, 1, Tn, C, s2, 0, w.g2w, w.g2b, w.scale, f.st);
        }
    }
}

int model_v3_forward(dmx_model *m, const float *d_mix, int B, float *d_out)
{
    if (!m->finalized || !m->v3)
    {
        set_error("model_v3_forward: model not finalized");
        return -1;
    }
    if (B < 1 || B > m->maxB)
    {
        set_error("model_forward: batch exceeds max_batch");
        return -1;
    }
    V3Model *v = m->v3;
    cudaStream_t st = m->stream;
    Fwd f{m, B, st};
    m->stats_used = 0;
    m->lastB = B;
    DMX_CUDA(cudaMemsetAsync(m->stats, 0, m->stats_cap * sizeof(double), st));

    // ---- STFT, CaC, input normalisation (src/model_inference.cpp:489-582) ----
    double *sx = f.stat_slots(B), *sxt = f.stat_slots(B);
    {
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

    // time branch on a second stream through the branch-parallel parts (as in model.cu; one stream while profiling)
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
    // ---- encoders 0-3 (src/encdec.cpp:363-517) ----
    const float *xin_f = m->spec, *xin_t = m->xt_in;
    for (int i = 0; i < 4; ++i)
    {
        const EncW &te = v->tenc[i];
        const EncW &fe = v->enc[i];
        const int C = fe.C;
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
            f.gemm(r, PC_GEMM_CONV);
        }
        f.st = st;
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
            f.gemm(p, PC_GEMM_CONV);
            f.dconv(m->cA_f, FREQS[i + 1], T, fe.dc, m->h_f, m->y2_f, 1);
            GemmParams r = f.base(m->cA_f, C, C, fe.rewrite, m->e[i], C, FREQS[i + 1], T, FREQS[i + 1], T);
            r.glu = 1;
            r.Nco = C;
            if (i == 0)
            {
                r.rowadd = m->freq_emb2;
                r.rowadd_ld = 48;
            }
            f.gemm(r, PC_GEMM_CONV);
        }
        xin_t = m->te[i];
        xin_f = m->e[i];
    }
    f.st = st;
    if (join())
        return -1;

    // ---- time encoder 4: plain conv 384 -> 768, k8 s4 p2 (src/encdec.cpp:519-530) ----
    {
        GemmParams p = f.base(m->te[3], 384, 384, v->tenc4, v->xt4, 768, 1, TLEN3, 1, T);
        p.ntaps = 8;
        p.sv = 4;
        for (int t = 0; t < 8; ++t)
        {
            p.du[t] = 0;
            p.dv[t] = t - 2;
        }
        f.gemm(p, PC_GEMM_CONV);
    }
    // ---- encoder 4 (src/encdec.cpp:532-573): k(8,1) s(4,1) p0 conv over the 8 bins, added onto the time branch ----
    {
        GemmParams p = f.base(m->e[3], 384, 384, v->enc4_conv, v->xt4, 768, 8, T, 1, T);
        p.ntaps = 8;
        p.su = 4;
        for (int t = 0; t < 8; ++t)
        {
            p.du[t] = t;
            p.dv[t] = 0;
        }
        p.gamma = v->ones; // xt4 += conv(x_3): the "inject" merge
        f.gemm(p, PC_GEMM_CONV);
        f.group_norm_g(v->xt4, v->g4, T, T, 0, 768, 4, v->e4n1w, v->e4n1b, 1, 0);
        dconv_la(f, v->g4, T, 768, 192, v->la[0]);
        f.gemm(f.linear(v->g4, 768, v->enc4_rw, v->r4, 1536, T), PC_GEMM_CONV);
        f.group_norm_g(v->r4, v->x4, T, T, 0, 1536, 4, v->e4n2w, v->e4n2b, 0, 1);
    }
    // ---- shared encoder 5 (src/encdec.cpp:575-615): conv1d k4 s2 p1 -> 168 steps ----
    {
        GemmParams p = f.base(v->x4, 768, 768, v->enc5_conv, v->y5, 1536, 1, T, 1, T5);
        p.ntaps = 4;
        p.sv = 2;
        for (int t = 0; t < 4; ++t)
        {
            p.du[t] = 0;
            p.dv[t] = t - 1;
        }
        f.gemm(p, PC_GEMM_CONV);
        f.group_norm_g(v->y5, v->g5, T5, T5, 0, 1536, 4, v->e5n1w, v->e5n1b, 1, 0);
        dconv_la(f, v->g5, T5, 1536, 384, v->la[1]);
        f.gemm(f.linear(v->g5, 1536, v->enc5_rw, v->r5, 3072, T5), PC_GEMM_CONV);
        f.group_norm_g(v->r5, v->x5, T5, T5, 0, 3072, 4, v->e5n2w, v->e5n2b, 0, 1);
    }
    // ---- shared decoder 0 (src/encdec.cpp:617-656): input is the skip alone ----
    {
        GemmParams p = f.base(v->x5, 1536, 1536, v->d0_rw, v->d0r, 3072, 1, T5, 1, T5);
        p.ntaps = 3;
        for (int t = 0; t < 3; ++t)
        {
            p.du[t] = 0;
            p.dv[t] = t - 1;
        }
        f.gemm(p, PC_GEMM_CONV);
        f.group_norm_g(v->d0r, v->d0g, T5, T5, 0, 3072, 4, v->d0n1w, v->d0n1b, 0, 1);
        // ConvTranspose1d k4 s2: 168 -> 338 (kept whole: GroupNorm statistics span it), then crop [1, 337)
        GemmParams q = f.base(v->d0g, 1536, 1536, v->d0_ct, v->d0ct, 768, 1, T5, 1, T5 + 1);
        q.ntaps = 2;
        q.dv[0] = -1;
        q.dv[1] = 0;
        q.Nco = 768;
        q.Uout = 1;
        q.Vout = 2 * T5 + 2;
        q.ous = 0;
        q.ovs = 2;
        q.ovr = 1;
        q.ovo = 0;
        f.gemm(q, PC_GEMM_CONV);
        f.group_norm_g(v->d0ct, v->xd0, 2 * T5 + 2, T, 1, 768, 4, v->d0n2w, v->d0n2b, 1, 0);
    }
    // ---- frequency decoder 1 (src/encdec.cpp:658-698) ----
    {
        GemmParams p = f.base(v->xd0, 768, 768, v->d1_rw, v->d1r, 1536, 1, T, 1, T);
        p.A2 = v->x4; // skip = saved_4
        p.ntaps = 9;
        for (int kh = 0; kh < 3; ++kh)
            for (int kw = 0; kw < 3; ++kw)
            {
                p.du[kh * 3 + kw] = kh - 1;
                p.dv[kh * 3 + kw] = kw - 1;
            }
        f.gemm(p, PC_GEMM_CONV);
        f.group_norm_g(v->d1r, v->pre, T, T, 0, 1536, 4, v->d1n1w, v->d1n1b, 0, 1);       // -> `pre` (768 ch)
        // ConvTranspose2d k(8,1) s(4,1): 1 bin -> 8 bins, no crop
        GemmParams q = f.base(v->pre, 768, 768, v->d1_ct, v->d1ct, 384, 1, T, 2, T);
        q.ntaps = 2;
        q.du[0] = -1;
        q.du[1] = 0;
        q.Nco = 384;
        q.Uout = 8;
        q.Vout = T;
        q.ous = 4;
        q.our = 1;
        q.ouo = 0;
        f.gemm(q, PC_GEMM_CONV);
        f.group_norm_g(v->d1ct, v->x3, 8 * T, 8 * T, 0, 384, 4, v->d1n2w, v->d1n2b, 1, 0);
    }
    // ---- time decoder 0 (src/encdec.cpp:700-726): restarts the time branch from `pre` ----
    {
        GemmParams q = f.base(v->pre, 768, 768, v->td0_ct, v->td0ct, 384, 1, T, 1, T + 1);
        q.ntaps = 2;
        q.dv[0] = -1;
        q.dv[1] = 0;
        q.Nco = 384;
        q.Uout = 1;
        q.Vout = 4 * T + 4;
        q.ous = 0;
        q.ovs = 4;
        q.ovr = 1;
        q.ovo = 0;
        f.gemm(q, PC_GEMM_CONV);
        f.group_norm_g(v->td0ct, v->xt3, 4 * T + 4, TLEN3, 2, 384, 4, v->td0n2w, v->td0n2b, 1, 0);
    }
    // ---- common decoders, no DConv (src/encdec.cpp:728-863) ----
    const float *dec_in_f = v->x3, *dec_in_t = v->xt3;
    const bool tma = m->use_tma && m->precision == DMX_PRECISION_BF16;
    const void *img_f = nullptr, *img_t = nullptr;
    if (fork())
        return -1;
    for (int i = 0; i < 4; ++i)
    {
        const DecW &fd = v->cdec[i];
        const DecW &td = v->ctdec[i];
        const int C = fd.C, F = FREQS[4 - i], Lin = TLENS[4 - i], Lout = TLENS[3 - i];
        f.st = st;
        {
            GemmParams p = f.base(dec_in_f, C, C, fd.rewrite, m->cA_f, C, F, T, F, T);
            if (i == 0)
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
            GemmParams q = f.base(m->cA_f, C, C, fd.convtr, m->dx[i], fd.Cout, F, T, F + 1, T);
            q.ntaps = 2;
            q.du[0] = -1;
            q.du[1] = 0;
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
            if (tma && i < 3 && m->dimg_f[i + 1])
            {
                q.Cbf = m->dimg_f[i + 1];
                q.Cbf_ld = tc_tma_cpitch(fd.Cout);
                q.skip_c = !m->keep_stages && v->cdec[i + 1].rewrite.Wt != nullptr; // only the image is read again
                img_f = m->dimg_f[i + 1];
            }
            f.gemm(q, PC_GEMM_CONV);
        }
        f.st = st_t;
        {
            GemmParams p = f.base(dec_in_t, C, C, td.rewrite, m->cA_t, C, 1, Lin, 1, Lin);
            if (i == 0)
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
            GemmParams q = f.base(m->cA_t, C, C, td.convtr, m->dxt[i], td.Cout, 1, Lin, 1, Lin + 1);
            q.ntaps = 2;
            q.dv[0] = -1;
            q.dv[1] = 0;
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
            if (tma && i < 3 && m->dimg_t[i + 1])
            {
                q.Cbf = m->dimg_t[i + 1];
                q.Cbf_ld = tc_tma_cpitch(td.Cout);
                q.skip_c = !m->keep_stages && v->ctdec[i + 1].rewrite.Wt != nullptr;
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
    {
        ProfScope ps(m, PC_ISTFT_FRAMES, 0.0,
                     (double)B * ((double)16 * NBINS * T * 4 + (double)8 * SEG * 4 + (double)4 * 2 * SEG * 4), 2);
        if (launch_istft(m->dx[3], B, 4, m->meanstd_x, m->window, m->tw, m->wss, m->dxt[3], m->meanstd_t, m->frames, d_out,
                         st, 1))
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
    set_tap(m, "saved_4", v->x4, 1, T, 768);
    set_tap(m, "x_shared_5", v->x5, 1, T5, 1536);
    set_tap(m, "x_4", v->xd0, 1, T, 768);
    set_tap(m, "x_3", v->x3, 8, T, 384);
    set_tap(m, "xt_3", v->xt3, 1, TLEN3, 384);
    const char *fn[4] = {"x_2", "x_1", "x_0", "x_out"};
    const char *tn[4] = {"xt_2", "xt_1", "xt_0", "xt_out"};
    for (int i = 0; i < 4; ++i)
    {
        set_tap(m, fn[i], m->dx[i], FREQS[3 - i], T, v->cdec[i].Cout, i < 3 ? m->e[2 - i] : nullptr, i == 3);
        set_tap(m, tn[i], m->dxt[i], 1, TLENS[3 - i], v->ctdec[i].Cout, i < 3 ? m->te[2 - i] : nullptr);
    }
    return 0;
}

} // namespace dmx
