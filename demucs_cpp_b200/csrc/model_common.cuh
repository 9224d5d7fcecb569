// Helpers shared by the htdemucs (model.cu) and hdemucs_mmi (model_v3.cu) device models: allocation,
// weight repacking into GEMM B-operands, constant DSP tables and the forward-pass launch helpers.
#pragma once
#include "model.cuh"

#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace dmx
{

static const int FREQS[5] = {2048, 512, 128, 32, 8};
static const int TLENS[5] = {SEG, TLEN0, TLEN1, TLEN2, TLEN3};
static const int CH[5] = {0, 48, 96, 192, 384}; // CH[i+1] = channels after encoder i
static const int T = LE;

// ------------------------------------------------------------------------------------------------
// allocation helpers
static int dev_alloc(dmx_model *m, void **p, size_t bytes)
{
    DMX_CUDA(cudaMalloc(p, bytes ? bytes : 16));
    m->allocs.push_back(*p);
    return 0;
}

template <typename Tp> static int dev_alloc_t(dmx_model *m, Tp **p, size_t n)
{
    return dev_alloc(m, reinterpret_cast<void **>(p), n * sizeof(Tp));
}

static int upload(dmx_model *m, float **dst, const std::vector<float> &v)
{
    if (dev_alloc_t(m, dst, v.size()))
        return -1;
    DMX_CUDA(cudaMemcpy(*dst, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice));
    return 0;
}

static const HostTensor *get(dmx_model *m, const std::string &name, size_t numel)
{
    auto it = m->host.find(name);
    if (it == m->host.end())
    {
        set_error("model: missing tensor " + name);
        return nullptr;
    }
    if (it->second.numel() != numel)
    {
        set_error("model: tensor " + name + " has " + std::to_string(it->second.numel()) + " elements, expected " +
                  std::to_string(numel));
        return nullptr;
    }
    return &it->second;
}

static int upload_vec(dmx_model *m, float **dst, const std::string &name, size_t n)
{
    const HostTensor *t = get(m, name, n);
    if (!t)
        return -1;
    return upload(m, dst, t->data);
}

static inline int round_up(int x, int k) { return (x + k - 1) / k * k; }

// Host-side repacking at load time runs over ~40 M weights several times with strided index arithmetic: split the column
// range over a few threads when the layer is big enough to pay for them (load time).
template <typename F> static void par_for(int n, size_t work, F fn)
{
    static const int maxt = [] {
        const char *e = getenv("DMX_LOAD_THREADS");
        const int hw = (int)std::thread::hardware_concurrency();
        const int t = e ? atoi(e) : (hw > 0 ? (hw < 8 ? hw : 8) : 4);
        return t < 1 ? 1 : t;
    }();
    const int nt = work < (size_t)1 << 18 ? 1 : (n < maxt ? n : maxt);
    if (nt <= 1)
    {
        fn(0, n);
        return;
    }
    std::vector<std::thread> th;
    for (int t = 0; t < nt; ++t)
        th.emplace_back([=] { fn((int)((long long)n * t / nt), (int)((long long)n * (t + 1) / nt)); });
    for (auto &x : th)
        x.join();
}

struct LoadTimer
{
    static double &total()
    {
        static double t = 0.0;
        return t;
    }
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    ~LoadTimer()
    {
        static std::mutex mu; // replicas may be loaded from several host threads at once
        std::lock_guard<std::mutex> lk(mu);
        total() += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
};

static inline uint16_t f32_to_bf16_bits(float f)
{
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u)
        return (uint16_t)((u >> 16) | 0x40); // NaN
    u += 0x7fffu + ((u >> 16) & 1u);          // round to nearest even
    return (uint16_t)(u >> 16);
}

// Generic packer: element (k, n) of the GEMM B operand given by f(k, n); bias by fb(n).
// If glu: output columns are interleaved so that packed 2j = original j, packed 2j+1 = original j + N/2.
template <typename FW, typename FB>
static int pack(dmx_model *m, Packed &p, int K, int N, bool glu, FW fw, FB fb, const TmaPlan *tp = nullptr, int tpCin = 0,
                const TmaPlan *tp32 = nullptr)
{
    p.K = K;
    p.N = N;
    p.ldw = round_up(N, 128);
    const int Kpad = round_up(K, 16);
    LoadTimer load_timer;
    const size_t work = (size_t)K * N;
    std::vector<float> w((size_t)Kpad * p.ldw, 0.f), b(p.ldw, 0.f);
    const int ldw = p.ldw;
    par_for(N, work, [&, ldw](int n0, int n1) {
        for (int n = n0; n < n1; ++n)
        {
            const int src = glu ? ((n & 1) ? (n >> 1) + N / 2 : (n >> 1)) : n;
            for (int k = 0; k < K; ++k)
                w[(size_t)k * ldw + n] = fw(k, src);
            b[n] = fb(src);
        }
    });
    if (upload(m, &p.W, w) || upload(m, &p.bias, b))
        return -1;
    // bf16 image for the tcgen05 kernel: same (k, n) contents, round-to-nearest-even, zero padded
    {
        std::vector<uint16_t> wb(tc_packed_elems(K, N), 0);
        par_for(N, work, [&, ldw](int n0, int n1) {
            for (int n = n0; n < n1; ++n)
                for (int k = 0; k < K; ++k)
                    wb[tc_packed_index(K, N, k, n)] = f32_to_bf16_bits(w[(size_t)k * ldw + n]);
        });
        if (dev_alloc(m, &p.Wb, wb.size() * sizeof(uint16_t)))
            return -1;
        DMX_CUDA(cudaMemcpy(p.Wb, wb.data(), wb.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
    }
    if (tp)
    {
        // the same weights in the K order of the TMA-fed kernel (chunk, box, tap, channel): row k' of that image is
        // row tc_tma_korig(k') of this one
        std::vector<uint16_t> wt(tc_packed_elems(K, N), 0);
        par_for(K, work, [&, ldw](int k0, int k1) {
            for (int kp = k0; kp < k1; ++kp)
            {
                const int k = tc_tma_korig(*tp, tpCin, kp);
                for (int n = 0; n < N; ++n)
                    wt[tc_packed_index(K, N, kp, n)] = f32_to_bf16_bits(w[(size_t)k * ldw + n]);
            }
        });
        if (dev_alloc(m, &p.Wt, wt.size() * sizeof(uint16_t)))
            return -1;
        DMX_CUDA(cudaMemcpy(p.Wt, wt.data(), wt.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
    }
    if (tp32)
    {
        // fp32 image for the tf32 variant: K' covers whole 32-channel chunks, rows of channels past Cin are zero
        const int Kp = tc_tma_k32(*tp32);
        std::vector<float> wt(tc_packed_elems32(Kp, N), 0.f);
        for (int kp = 0; kp < Kp; ++kp)
        {
            const int k = tc_tma_korig32(*tp32, tpCin, kp);
            if (k < 0)
                continue;
            for (int n = 0; n < N; ++n)
                wt[tc_packed_index32(Kp, N, kp, n)] = w[(size_t)k * p.ldw + n];
        }
        if (dev_alloc(m, &p.Wt32, wt.size() * sizeof(float)))
            return -1;
        DMX_CUDA(cudaMemcpy(p.Wt32, wt.data(), wt.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    return 0;
}

// Conv weight (Cout, Cin, ntaps) row-major -> B[tap*Cin + c][cout]   (src/conv.hpp:97-110 equivalent)
static int pack_conv(dmx_model *m, Packed &p, const std::string &wname, const std::string &bname, int Cout, int Cin,
                     int ntaps, bool glu, const TmaPlan *tp = nullptr, const TmaPlan *tp32 = nullptr)
{
    const HostTensor *w = get(m, wname, (size_t)Cout * Cin * ntaps);
    const HostTensor *b = get(m, bname, Cout);
    if (!w || !b)
        return -1;
    const float *wd = w->data.data();
    const float *bd = b->data.data();
    return pack(
        m, p, Cin * ntaps, Cout, glu,
        [=](int k, int n) {
            const int tap = k / Cin, c = k % Cin;
            return wd[((size_t)n * Cin + c) * ntaps + tap];
        },
        [=](int n) { return bd[n]; }, tp, Cin, tp32);
}

// tf32 plans of the encoders: the 1x1 rewrite (one tap) and the k8 s4 convolution along frequency (eight planes
// 4 u - 2 ... 4 u + 5, src/encdec.cpp:24-50); both read the fp32 activations as they are
static TmaPlan enc_plan32(int C, bool conv8)
{
    int du[9] = {0}, dv[9] = {0};
    const int nt = conv8 ? 8 : 1;
    for (int t = 0; t < nt; ++t)
        du[t] = conv8 ? t - 2 : 0;
    TmaPlan pl;
    tc_tma_plan32(nt, du, dv, C, true, &pl);
    return pl;
}

// The K order of the TMA-fed kernel for the decoders' "rewrite" convolutions: 3 x 3 over (frequency, frames), tap =
// kh * 3 + kw (kh along frequency), or k3 along time (src/encdec.cpp:178-196,270-288)
static TmaPlan rewrite_plan(int C, bool freq)
{
    int du[9], dv[9];
    const int nt = freq ? 9 : 3;
    for (int t = 0; t < nt; ++t)
    {
        du[t] = freq ? t / 3 - 1 : 0;
        dv[t] = freq ? t % 3 - 1 : t - 1;
    }
    TmaPlan pl;
    static const bool share = getenv("DMX_TMA_NO_SHARE") == nullptr;
    tc_tma_plan(nt, du, dv, C, share, &pl);
    return pl;
}

// ConvTranspose with kernel k = 2*stride (k8 s4, or v3's k4 s2): weight (Cin, Cout, k) -> two taps
// (x[i-1] with kernel index r+stride, x[i] with kernel index r), N = stride*Cout columns n = r*Cout + co
// (src/conv.hpp:264-474 equivalent)
static int pack_convtr(dmx_model *m, Packed &p, const std::string &wname, const std::string &bname, int Cin, int Cout,
                       int ksize = 8, const TmaPlan *tp = nullptr)
{
    const HostTensor *w = get(m, wname, (size_t)Cin * Cout * ksize);
    const HostTensor *b = get(m, bname, Cout);
    if (!w || !b)
        return -1;
    const float *wd = w->data.data();
    const float *bd = b->data.data();
    const int stride = ksize / 2;
    return pack(
        m, p, 2 * Cin, stride * Cout, false,
        [=](int k, int n) {
            const int tap = k / Cin, c = k % Cin;
            const int r = n / Cout, co = n % Cout;
            const int kk = tap == 0 ? r + stride : r;
            return wd[((size_t)c * Cout + co) * ksize + kk];
        },
        [=](int n) { return bd[n % Cout]; }, tp, Cin);
}

// The K order of the TMA-fed kernel for a transposed convolution as a two-tap GEMM: x[i - 1], x[i] along frequency
// (two planes) or along time (one box, rows shifted by one)
static TmaPlan convtr_plan(int C, bool freq)
{
    const int du[2] = {freq ? -1 : 0, 0}, dv[2] = {freq ? 0 : -1, 0};
    TmaPlan pl;
    static const bool share = getenv("DMX_TMA_NO_SHARE") == nullptr;
    tc_tma_plan(2, du, dv, C, share, &pl);
    return pl;
}

// Linear weight rows [row0, row0+N) of an (out, in) matrix -> B[in][n]
static int pack_linear(dmx_model *m, Packed &p, const std::string &wname, const std::string &bname, int out_total,
                       int in, int row0, int N)
{
    const HostTensor *w = get(m, wname, (size_t)out_total * in);
    const HostTensor *b = get(m, bname, out_total);
    if (!w || !b)
        return -1;
    const float *wd = w->data.data();
    const float *bd = b->data.data();
    return pack(
        m, p, in, N, false, [=](int k, int n) { return wd[(size_t)(row0 + n) * in + k]; },
        [=](int n) { return bd[row0 + n]; });
}

static int upload_vec_padded(dmx_model *m, float **dst, const std::string &name, size_t n, size_t npad)
{
    const HostTensor *t = get(m, name, n);
    if (!t)
        return -1;
    std::vector<float> v(t->data);
    v.resize(npad, 0.f);
    return upload(m, dst, v);
}

static int pack_dconv(dmx_model *m, DConvW &d, const std::string &prefix, int C, int compress = 8)
{
    d.C = C;
    d.hid = C / compress;
    // The hidden tensor is stored with its channel count rounded up to a multiple of 4 (6 -> 8) so that its
    // rows stay 16-byte aligned for vector loads; the pad channels carry zero GroupNorm / conv weights.
    d.hidp = round_up(d.hid, 4);
    for (int j = 0; j < 2; ++j)
    {
        const std::string p = prefix + ".dconv.layers." + std::to_string(j) + ".";
        if (pack_conv(m, d.c1[j], p + "0.weight", p + "0.bias", d.hid, C, 3, false))
            return -1;
        {
            const HostTensor *w = get(m, p + "3.weight", (size_t)2 * C * d.hid);
            const HostTensor *b = get(m, p + "3.bias", 2 * C);
            if (!w || !b)
                return -1;
            const float *wd = w->data.data();
            const float *bd = b->data.data();
            const int hid = d.hid;
            if (pack(
                    m, d.c2[j], d.hidp, 2 * C, false, [=](int k, int n) { return k < hid ? wd[(size_t)n * hid + k] : 0.f; },
                    [=](int n) { return bd[n]; }))
                return -1;
        }
        if (upload_vec_padded(m, &d.g1w[j], p + "1.weight", d.hid, d.hidp) ||
            upload_vec_padded(m, &d.g1b[j], p + "1.bias", d.hid, d.hidp) ||
            upload_vec(m, &d.g2w[j], p + "4.weight", 2 * C) || upload_vec(m, &d.g2b[j], p + "4.bias", 2 * C) ||
            upload_vec(m, &d.scale[j], p + "6.scale", C))
            return -1;
    }
    if (dconv_fused_supported(C, d.hid, LE) || dconv_time_mma_supported(C, d.hid, d.hidp))
    {
        // parameter block of the fused kernel: per sub-block W1[3][C][HID] | b1 | g1w | g1b | W2[2C][HID] | b2 | g2w | g2b | scale
        std::vector<float> blk;
        blk.reserve(dconv_fused_param_floats(C, d.hid));
        const int hid = d.hid;
        for (int j = 0; j < 2; ++j)
        {
            const std::string p = prefix + ".dconv.layers." + std::to_string(j) + ".";
            const HostTensor *w1 = get(m, p + "0.weight", (size_t)hid * C * 3), *b1 = get(m, p + "0.bias", hid);
            const HostTensor *g1w = get(m, p + "1.weight", hid), *g1b = get(m, p + "1.bias", hid);
            const HostTensor *w2 = get(m, p + "3.weight", (size_t)2 * C * hid), *b2 = get(m, p + "3.bias", 2 * C);
            const HostTensor *g2w = get(m, p + "4.weight", 2 * C), *g2b = get(m, p + "4.bias", 2 * C);
            const HostTensor *sc = get(m, p + "6.scale", C);
            if (!w1 || !b1 || !g1w || !g1b || !w2 || !b2 || !g2w || !g2b || !sc)
                return -1;
            for (int tap = 0; tap < 3; ++tap)
                for (int c = 0; c < C; ++c)
                    for (int jj = 0; jj < hid; ++jj)
                        blk.push_back(w1->data[((size_t)jj * C + c) * 3 + tap]);
            blk.insert(blk.end(), b1->data.begin(), b1->data.end());
            blk.insert(blk.end(), g1w->data.begin(), g1w->data.end());
            blk.insert(blk.end(), g1b->data.begin(), g1b->data.end());
            blk.insert(blk.end(), w2->data.begin(), w2->data.end());
            blk.insert(blk.end(), b2->data.begin(), b2->data.end());
            blk.insert(blk.end(), g2w->data.begin(), g2w->data.end());
            blk.insert(blk.end(), g2b->data.begin(), g2b->data.end());
            blk.insert(blk.end(), sc->data.begin(), sc->data.end());
        }
        if ((int)blk.size() != dconv_fused_param_floats(C, hid))
        {
            set_error("pack_dconv: fused parameter block size mismatch");
            return -1;
        }
        if (upload(m, &d.fused, blk))
            return -1;
        if (const int nf = dconv_mma_pack_floats(C, hid))
        {
            std::vector<float> pk((size_t)nf);
            dconv_mma_pack(C, hid, blk.data(), pk.data());
            if (upload(m, &d.fused_pk, pk))
                return -1;
        }
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------
// constant tables, computed exactly like the reference does (float libm calls)
static std::vector<float> make_window()
{
    // src/dsp.hpp:59-75
    const float PI = 3.14159265359F;
    std::vector<float> w(NFFT);
    const float floatN = (float)(NFFT + 1);
    for (int n = 0; n < NFFT; ++n)
        w[n] = 0.5F * (1.0F - cosf(2.0F * PI * (float)n / (floatN - 1)));
    return w;
}

static std::vector<float> make_wss(const std::vector<float> &w)
{
    // src/dsp.hpp:77-100 with nb_frames = 340
    const int nb_frames = PADDED / HOP + 1;
    std::vector<float> out(OLA_LEN, 0.f);
    for (int i = 0; i < nb_frames; ++i)
        for (int j = 0; j < NFFT; ++j)
            out[i * HOP + j] += w[j] * w[j];
    return out;
}

static std::vector<float> make_twiddles()
{
    std::vector<float> t(2 * NFFT);
    for (int k = 0; k < NFFT; ++k)
    {
        const double a = -2.0 * M_PI * (double)k / (double)NFFT;
        t[2 * k] = (float)cos(a);
        t[2 * k + 1] = (float)sin(a);
    }
    return t;
}

// ------------------------------------------------------------------------------------------------
// forward helpers
struct Fwd
{
    dmx_model *m;
    int B;
    cudaStream_t st;
    int rc = 0;

    double *stat_slots(size_t items)
    {
        double *p = m->stats + m->stats_used;
        m->stats_used += 2 * items;
        if (m->stats_used > m->stats_cap)
        {
            set_error("statistics arena overflow");
            rc = -1;
            return m->stats;
        }
        return p;
    }
    void gemm(const GemmParams &p, int cls)
    {
        if (rc)
            return;
        // algorithmic bytes: the input grid read once (+ skip tensor), the output written once (+ residual read)
        const double nout = p.glu ? p.N / 2 : p.N;
        const double esz_out = p.out_bf16 ? 2.0 : 4.0;
        double bytes = (double)B * p.Ui * p.Vi * p.Cin * 4.0 * (p.A2 ? 2.0 : 1.0) + (double)p.M * nout * esz_out;
        if (p.Aimg && p.Wt && m->precision == DMX_PRECISION_BF16)
            bytes = (double)B * p.Ui * p.Vi * p.Cin * 2.0 + (double)p.M * nout * esz_out;
        if (p.Cbf)
            bytes += (double)B * p.Uout * p.Vout * p.Nco * 2.0;
        if (p.gamma)
            bytes += (double)p.M * nout * 4.0;
        ProfScope ps(m, cls, 2.0 * (double)p.M * (double)p.N * (double)p.K, bytes);
        if (m->precision == DMX_PRECISION_BF16 && ((p.Aimg && p.Wt) || p.Wt32))
            rc = launch_gemm_tma(p, st);
        else if (m->precision == DMX_PRECISION_BF16 && p.Wb)
            rc = launch_gemm_tc(p, p.Wb, st);
        else
            rc = launch_gemm(p, st);
    }
    // common fields: single batch grid (U, V) in and out, plain output mapping
    GemmParams base(const float *Ain, int lda, int Cin, const Packed &w, float *C, int ldc, int Ui, int Vi, int Uo,
                    int Vo)
    {
        GemmParams p;
        p.A = Ain;
        p.lda = lda;
        p.Cin = Cin;
        p.K = w.K;
        p.W = w.W;
        p.Wb = w.Wb;
        p.ldw = w.ldw;
        p.bias = w.bias;
        p.N = w.N;
        p.Ui = Ui;
        p.Vi = Vi;
        p.Uo = Uo;
        p.Vo = Vo;
        p.M = B * Uo * Vo;
        p.C = C;
        p.ldc = ldc;
        p.Nco = w.N;
        p.Uout = Uo;
        p.Vout = Vo;
        return p;
    }

    // DConv residual branch (src/layers.cpp:152-375) in place on x [B][U][V][C]; GroupNorm items are
    // frequency rows (per_u = 1) or whole batch elements (time branch, U = 1).
    // img (bf16 mode): the result is also written as a bf16 channels-last image with img_ld channels per row (what a
    // TMA-fed convolution that follows would read); returns whether the path taken wrote it -- only the GEMM +
    // gn_glu_residual form (the 384-channel level) does: the image stores measurably slowed the fused mma.sync kernels.
    bool dconv(float *x, int U, int V, const DConvW &w, float *h, float *y2, int per_u, void *img = nullptr, int img_ld = 0)
    {
        const int C = w.C, hid = w.hid, hidp = w.hidp;
        if (per_u && w.fused && dconv_fused_supported(C, hid, V))
        {
            if (rc)
                return false;
            // one HBM read + one write of x; MACs: 2 sub-blocks x (k3 conv + the 1x1 conv evaluated twice)
            ProfScope ps(m, PC_DCONV_FUSED, 2.0 * 2.0 * (double)B * U * V * (3.0 * C * hid + 2.0 * 2.0 * C * hid),
                         (double)B * U * V * C * 8.0);
            if (m->precision == DMX_PRECISION_BF16 && w.fused_pk && dconv_mma_supported(C, hid, V))
                rc = launch_dconv_mma(x, w.fused_pk, B * U, C, hid, st);
            else
                rc = launch_dconv_fused(x, w.fused, B * U, C, hid, st);
            return false;
        }
        if ((per_u || U == 1) && w.fused_pk && m->precision == DMX_PRECISION_BF16 && dconv_time_mma_supported(C, hid, hidp))
        {
            if (rc)
                return false;
            // bf16 mode, tiled form: three tensor-core launches per sub-block, y never stored.  GroupNorm items are
            // the batch elements (time branch) or the B*U frequency rows (the 192-channel level, too wide for the
            // row-resident kernel above)
            const int items = per_u ? B * U : B;
            double *sl[4] = {stat_slots(items), stat_slots(items), stat_slots(items), stat_slots(items)};
            ProfScope ps(m, PC_DCONV_FUSED, 2.0 * 2.0 * (double)B * U * V * (3.0 * C * hid + 2.0 * 2.0 * C * hid),
                         2.0 * (double)B * U * V * (3.0 * C + 3.0 * hidp) * 4.0, 6);
            rc = launch_dconv_time_mma(x, h, w.fused_pk, items, V, C, hid, hidp, sl, st);
            return false;
        }
        const size_t items = (size_t)B * (per_u ? U : 1);
        for (int j = 0; j < 2; ++j)
        {
            const int dil = 1 << j;
            double *s1 = stat_slots(items), *s2 = stat_slots(items);
            GemmParams p = base(x, C, C, w.c1[j], h, hidp, U, V, U, V);
            p.ntaps = 3;
            for (int t = 0; t < 3; ++t)
            {
                p.du[t] = 0;
                p.dv[t] = (t - 1) * dil;
            }
            p.stats = s1;
            p.stat_per_u = per_u;
            gemm(p, PC_GEMM_DCONV);
            GemmParams q = base(h, hidp, hidp, w.c2[j], y2, 2 * C, U, V, U, V);
            q.pro = 1;
            q.pro_stats = s1;
            q.pro_count = per_u ? (double)V * hid : (double)U * V * hid;
            q.pro_w = w.g1w[j];
            q.pro_b = w.g1b[j];
            q.pro_per_u = per_u;
            q.stats = s2;
            q.stat_per_u = per_u;
            gemm(q, PC_GEMM_DCONV);
            if (!rc)
            {
                ProfScope ps(m, PC_POINTWISE, 0.0, (double)B * U * V * C * 4.0 * 4.0);
                rc = launch_gn_glu_residual(y2, x, B, U, V, C, s2, per_u, w.g2w[j], w.g2b[j], w.scale[j], st, j == 1 ? img : nullptr,
                                            img_ld);
            }
        }
        return img != nullptr && m->precision == DMX_PRECISION_BF16;
    }

    void layer_norm(const float *x, float *y, int rows, int C, const float *w, const float *b, const float *add = nullptr,
                    int add_rows = 0)
    {
        if (rc)
            return;
        ProfScope ps(m, PC_POINTWISE, 0.0, (double)rows * C * 8.0);
        rc = launch_layer_norm(x, y, rows, C, w, b, add, add_rows, st);
    }

    // GroupNorm(G) over (rows, C/G) per batch item, optional GELU / GLU / row crop (v3, src/layers.hpp:125-227)
    void group_norm_g(const float *x, float *y, int rows_in, int rows_out, int row0, int C, int G, const float *w,
                      const float *b, int act, int glu)
    {
        if (rc)
            return;
        double *s = stat_slots((size_t)B * G);
        ProfScope ps(m, PC_POINTWISE, 0.0, (double)B * rows_in * C * 8.0 + (double)B * rows_out * (glu ? C / 2 : C) * 4.0, 2);
        rc = launch_group_norm_g(x, y, B, rows_in, rows_out, row0, C, G, s, w, b, act, glu, st);
    }

    // tokens-as-rows linear layer: [B*Tn][Cin] x W -> [B*Tn][N]
    GemmParams linear(const float *x, int Cin, const Packed &w, float *y, int ldc, int Tn)
    {
        return base(x, Cin, Cin, w, y, ldc, 1, Tn, 1, Tn);
    }

    // LayerNorm feeding a GEMM: bf16 mode writes the tcgen05 kernel's tile-packed A image, fp32 mode plain rows
    void layer_norm_a(const float *x, float *y, int rows, int C, const float *w, const float *b)
    {
        if (m->precision != DMX_PRECISION_BF16)
        {
            layer_norm(x, y, rows, C, w, b);
            return;
        }
        if (rc)
            return;
        ProfScope ps(m, PC_POINTWISE, 0.0, (double)rows * C * 6.0);
        rc = launch_layer_norm_pack(x, y, rows, C, w, b, st);
    }

    // linear layer whose A operand is tile-packed bf16 (bf16 mode) or fp32 rows (fp32 mode)
    GemmParams linear_a(const float *x, int Cin, const Packed &w, float *y, int ldc, int Tn)
    {
        GemmParams p = linear(x, Cin, w, y, ldc, Tn);
        if (m->precision == DMX_PRECISION_BF16)
            p.Apk = x;
        return p;
    }

    // one transformer layer (src/layers.cpp:377-531) on q [B*Tq][d]; kn = already-normalised key source
    // for cross layers (nullptr for self attention; produced by layer_norm_a).
    // bf16 mode: every GEMM operand inside the layer travels as bf16 -- LayerNorm, attention and FFN1 write the
    // next GEMM's tile-packed A image directly, the projections write Q/K/V as bf16 rows -- while the residual
    // stream q, the statistics and the softmax stay fp32.
    void tlayer(const TLayerW &L, float *q, int Tq, const float *kn, int Tk, float *ln, float *qkv, float *kv, float *ao,
                float *ff)
    {
        const int d = m->d_model, dh = d / 8;
        const bool tcm = m->precision == DMX_PRECISION_BF16;
        layer_norm_a(q, ln, B * Tq, d, L.n1w, L.n1b);
        if (!L.cross)
        {
            GemmParams pq = linear_a(ln, d, L.qkv, qkv, 3 * d, Tq);
            pq.out_pack = tcm; // Q | K | V tile-packed, per-item rows padded to whole tiles: attention operand images
            pq.pack_T = tcm ? round_up(Tq, 128) : 0;
            gemm(pq, PC_GEMM_XFMR);
            if (!rc)
            {
                ProfScope ps(m, PC_ATTENTION, 4.0 * B * (double)Tq * Tq * d);
                if (tcm)
                {
                    rc = launch_attention_tc(qkv, 3 * d, 0, qkv, 3 * d, d, qkv, 3 * d, 2 * d, ao, B, Tq, round_up(Tq, 128), Tq,
                                             round_up(Tq, 128), 8, dh, st);
                }
                else
                    rc = launch_attention(qkv, 3 * d, qkv + d, 3 * d, qkv + 2 * d, 3 * d, ao, d, B, Tq, Tq, 8, dh, st);
            }
        }
        else
        {
            GemmParams pq = linear_a(ln, d, L.q, qkv, d, Tq);
            pq.out_pack = tcm;
            pq.pack_T = tcm ? round_up(Tq, 128) : 0;
            gemm(pq, PC_GEMM_XFMR);
            GemmParams pkv = linear_a(kn, d, L.kv, kv, 2 * d, Tk);
            pkv.out_pack = tcm;
            pkv.pack_T = tcm ? round_up(Tk, 128) : 0;
            gemm(pkv, PC_GEMM_XFMR);
            if (!rc)
            {
                ProfScope ps(m, PC_ATTENTION, 4.0 * B * (double)Tq * Tk * d);
                if (tcm)
                {
                    rc = launch_attention_tc(qkv, d, 0, kv, 2 * d, 0, kv, 2 * d, d, ao, B, Tq, round_up(Tq, 128), Tk,
                                             round_up(Tk, 128), 8, dh, st);
                }
                else
                    rc = launch_attention(qkv, d, kv, 2 * d, kv + d, 2 * d, ao, d, B, Tq, Tk, 8, dh, st);
            }
        }
        GemmParams po = linear_a(ao, d, L.out, q, d, Tq);
        po.gamma = L.g1;
        gemm(po, PC_GEMM_XFMR);
        // pre-FFN norm: norm3 for cross layers, norm2 for self layers (src/crosstransformer.cpp:111-113)
        layer_norm_a(q, ln, B * Tq, d, L.cross ? L.n3w : L.n2w, L.cross ? L.n3b : L.n2b);
        GemmParams p1 = linear_a(ln, d, L.ff1, ff, 4 * d, Tq);
        p1.act = 1;
        p1.out_pack = tcm;
        gemm(p1, PC_GEMM_XFMR);
        GemmParams p2 = linear_a(ff, 4 * d, L.ff2, q, d, Tq);
        p2.gamma = L.g2;
        double *s = stat_slots(B);
        p2.stats = s;
        p2.stat_per_u = 0;
        gemm(p2, PC_GEMM_XFMR);
        if (!rc)
        {
            ProfScope ps(m, PC_POINTWISE, 0.0, (double)B * Tq * d * 8.0);
            rc = launch_group_norm_apply(q, B, Tq, d, s, L.now_, L.nob, st);
        }
    }
};

static void set_tap(dmx_model *m, const char *name, float *p, int64_t d1, int64_t d2, int64_t C,
                    const float *minus = nullptr, bool swapped = false)
{
    Tap t;
    t.ptr = p;
    t.minus = minus;
    t.swapped = swapped;
    t.d1 = d1;
    t.d2 = d2;
    t.C = C;
    m->taps[name] = t;
}

} // namespace dmx
