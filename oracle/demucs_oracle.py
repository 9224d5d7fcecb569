"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product; the product path is
the CUDA library behind include/demucs_b200.h and fails loudly without it.  Only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.

A CPU restatement (numpy + torch CPU ops, fp32 or fp64) of the reference's inference
hot path for htdemucs 4s/6s/ft: the per-segment forward `demucscpp::model_inference`
(src/model_inference.cpp:48-475) and the track-level normalise / shift / split /
overlap-add around it (src/model_apply.cpp:60-288).  It follows the REFERENCE's
semantics, not upstream PyTorch's, wherever they differ: unbiased
(N-1) variance in every norm, edge-including mirror padding, ceil-mode conv lengths,
exact-erf GELU, `std + eps` input normalisation.

Pinning: tests/test_oracle_vs_ref.py checks this file against the compiled reference
(oracle/_ref, built by oracle/Makefile from /root/reference/src) on the same weights
and inputs, stage by stage and end to end, and tests/golden/*.npz holds committed
vectors produced by the reference itself (tests/golden/make_golden.py).

Tensors here use the reference's logical index order (e.g. freq activations are
(C, F, T), time activations (C, L)), as ordinary C-order arrays.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

SEG = 343980          # int(7.8 * 44100)                      src/model_apply.cpp:145
STRIDE = 257985       # int((1 - 0.25) * SEG)                 src/model_apply.cpp:162
MAX_SHIFT = 22050     # int(0.5 * 44100)                      src/model_apply.cpp:103-104
NFFT, HOP = 4096, 1024  #                                     src/dsp.hpp:16-18
LE = 336              # ceil(SEG / HOP)                       src/model.hpp:619
PAD = 1536            # floor(HOP/2)*3                        src/model.hpp:620
PAD_END = PAD + LE * HOP - SEG  # 1620                        src/model.hpp:621
TLEN = [343980, 85995, 21499, 5375, 1344]  #                  src/model.hpp:19-24


# ----------------------------------------------------------------------------- primitives
def hann_window():
    """Periodic Hann in float32 with the reference's truncated PI (src/dsp.hpp:59-75)."""
    PI = np.float32(3.14159265359)
    n = np.arange(NFFT, dtype=np.float32)
    arg = (np.float32(2.0) * PI * n / np.float32(NFFT)).astype(np.float32)
    return (np.float32(0.5) * (np.float32(1.0) - np.cos(arg).astype(np.float32))).astype(np.float32)


def window_sumsquare(nb_frames):
    """src/dsp.hpp:77-100 -- float32 accumulation in frame order."""
    w = hann_window()
    out = np.zeros(NFFT + HOP * (nb_frames - 1), dtype=np.float32)
    w2 = (w * w).astype(np.float32)
    for i in range(nb_frames):
        out[i * HOP:i * HOP + NFFT] += w2
    return out


def mirror_pad_edge(x, left, right):
    """Mirror padding that INCLUDES the boundary sample: padded[left-1-i] = x[i]
    (src/model_inference.cpp:22-46 and src/dsp.cpp:19-38)."""
    n = x.shape[-1]
    return torch.cat([x[..., :left].flip(-1), x, x[..., n - right:].flip(-1)], dim=-1)


def gelu(x):
    """Exact erf GELU, src/conv.hpp:203-204, src/layers.cpp:83-85."""
    return 0.5 * x * (1.0 + torch.erf(x / math.sqrt(2.0)))


def var_unbiased(x, dims):
    """demucscpp::calculate_variance: sum((x-mean)^2)/(n-1), src/layers.hpp:76-95."""
    m = x.mean(dim=dims, keepdim=True)
    n = 1
    for d in dims:
        n *= x.shape[d]
    return m, ((x - m) ** 2).sum(dim=dims, keepdim=True) / (n - 1)


def group_norm1(x, w, b, eps=1e-5):
    """GroupNorm with ONE group over (C, W) per leading item, N-1 variance, eps inside
    the sqrt (src/layers.cpp:9-49).  x: (items, C, W)."""
    m, v = var_unbiased(x, (1, 2))
    return (x - m) / torch.sqrt(v + eps) * w[None, :, None] + b[None, :, None]


def layer_norm(x, w, b, eps=1e-5):
    """LayerNorm over the last dim with N-1 variance (src/layers.cpp:121-150)."""
    m, v = var_unbiased(x, (x.dim() - 1,))
    return (x - m) / torch.sqrt(v + eps) * w + b


def glu(x, dim):
    """a * sigmoid(b), src/layers.cpp:96-119."""
    a, b = x.chunk(2, dim=dim)
    return a * (1.0 / (1.0 + torch.exp(-b)))


def conv1d_ceil(x, w, b, stride, pad, dilation=1):
    """Cross-correlation whose output length is ceil((L+2p-d(k-1)-1)/s)+1 with zero
    fill past the end (src/conv.hpp:25-34,54-61).  x: (B, Cin, L); w: (Cout, Cin, k)."""
    L = x.shape[-1]
    k = w.shape[-1]
    out_len = int(math.ceil((L + 2 * pad - dilation * (k - 1) - 1) / stride)) + 1
    need = (out_len - 1) * stride + dilation * (k - 1) + 1 - 2 * pad
    if need > L:
        x = F.pad(x, (0, need - L))
    return F.conv1d(x, w, b, stride=stride, padding=pad, dilation=dilation)[..., :out_len]


# ----------------------------------------------------------------------------- STFT pair
def stft(wave):
    """demucscpp::stft (src/dsp.cpp:51-86,119-149): centre pad 2048 (edge-including
    mirror), hop 1024, periodic Hann, rFFT-4096, times 1/sqrt(4096).
    wave: (2, n) -> complex (2, 2049, n//1024 + 1)."""
    dt = wave.dtype
    win = torch.from_numpy(hann_window()).to(dt)
    p = mirror_pad_edge(wave, NFFT // 2, NFFT // 2)
    frames = p.unfold(-1, NFFT, HOP)                       # (2, nb_frames, 4096)
    spec = torch.fft.rfft(frames * win, dim=-1) * (1.0 / math.sqrt(float(NFFT)))
    return spec.transpose(1, 2)                            # (2, 2049, frames)


def istft(spec, n):
    """demucscpp::istft (src/dsp.cpp:88-117,151-185): times sqrt(4096), UNSCALED inverse
    real FFT (Eigen FFT 'Speedy' sets Unscaled, vendor/eigen/unsupported/Eigen/FFT:186-197),
    times window / 4096 / (wss + 1e-8), overlap-add, drop 2048 at both ends.
    spec: complex (2, 2049, frames) -> (2, n)."""
    cdt = spec.dtype
    dt = torch.float64 if cdt == torch.complex128 else torch.float32
    nb_frames = spec.shape[-1]
    win = torch.from_numpy(hann_window()).to(dt)
    wss = torch.from_numpy(window_sumsquare(nb_frames)).to(dt)
    fr = torch.fft.irfft(spec.transpose(1, 2) * math.sqrt(float(NFFT)), n=NFFT, dim=-1) * float(NFFT)
    fr = fr * win / float(NFFT)                            # (2, frames, 4096)
    out = torch.zeros(2, NFFT + HOP * (nb_frames - 1), dtype=dt)
    for j in range(nb_frames):
        out[:, j * HOP:j * HOP + NFFT] += fr[:, j] / (wss[j * HOP:j * HOP + NFFT] + 1e-8)
    return out[:, NFFT // 2:NFFT // 2 + n]


# ----------------------------------------------------------------------------- blocks
class Weights:
    """name -> torch tensor view of a ggml file read by demucs_cpp_b200.weights.read_ggml."""

    def __init__(self, tensors, dtype=torch.float32):
        self.t = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dtype) for k, v in tensors.items()}
        self.dtype = dtype
        self.is_v3 = "tencoder.4.conv.weight" in self.t
        last = "decoder.5.conv_tr.weight" if self.is_v3 else "decoder.3.conv_tr.weight"
        self.S = self.t[last].shape[1] // 4
        self.is_4s = "channel_upsampler.weight" in self.t
        self.d_model = 0 if self.is_v3 else self.t["crosstransformer.norm_in.weight"].shape[0]

    def __getitem__(self, k):
        return self.t[k]


def dconv(W, prefix, y):
    """demucscpp::apply_dconv (src/layers.cpp:152-375).  y: (items, C, T); for the freq
    branch every frequency row is an item.  Two residual sub-blocks, dilation 1 and 2:
    conv k3 -> GroupNorm(1)+GELU -> 1x1 -> GroupNorm(1) -> GLU -> LayerScale -> +=."""
    for d in range(2):
        p = "%s.dconv.layers.%d." % (prefix, d)
        dil = 2 ** d
        h = F.conv1d(y, W[p + "0.weight"], W[p + "0.bias"], padding=dil, dilation=dil)
        h = gelu(group_norm1(h, W[p + "1.weight"], W[p + "1.bias"]))
        h = F.conv1d(h, W[p + "3.weight"][:, :, None], W[p + "3.bias"])
        h = glu(group_norm1(h, W[p + "4.weight"], W[p + "4.bias"]), 1)
        y = y + h * W[p + "6.scale"][None, :, None]
    return y


def freq_encoder(W, i, x):
    """demucscpp::apply_freq_encoder (src/encdec.cpp:8-80).  x: (Cin, F, T) -> (Cout, F/4, T)."""
    p = "encoder.%d" % i
    y = F.conv2d(x[None], W[p + ".conv.weight"][..., None], W[p + ".conv.bias"],
                 stride=(4, 1), padding=(2, 0))[0]
    y = gelu(y)                                            # (C, F', T)
    y = dconv(W, p, y.permute(1, 0, 2)).permute(1, 0, 2)   # items = frequency rows
    y = torch.einsum("oc,cft->oft", W[p + ".rewrite.weight"], y) + W[p + ".rewrite.bias"][:, None, None]
    return glu(y, 0)


def time_encoder(W, i, xt):
    """demucscpp::apply_time_encoder (src/encdec.cpp:82-164).  xt: (Cin, L) -> (Cout, L')."""
    p = "tencoder.%d" % i
    y = gelu(conv1d_ceil(xt[None], W[p + ".conv.weight"], W[p + ".conv.bias"], 4, 2))
    y = dconv(W, p, y)
    y = F.conv1d(y, W[p + ".rewrite.weight"][:, :, None], W[p + ".rewrite.bias"])
    return glu(y, 1)[0]


def freq_decoder(W, i, x, skip, last):
    """demucscpp::apply_freq_decoder (src/encdec.cpp:166-256)."""
    p = "decoder.%d" % i
    y = F.conv2d((x + skip)[None], W[p + ".rewrite.weight"], W[p + ".rewrite.bias"], padding=1)[0]
    y = glu(y, 0)
    y = dconv(W, p, y.permute(1, 0, 2)).permute(1, 0, 2)
    z = F.conv_transpose2d(y[None], W[p + ".conv_tr.weight"][..., None], W[p + ".conv_tr.bias"],
                           stride=(4, 1))[0]
    if not last:
        z = gelu(z)
    return z[:, 2:z.shape[1] - 2, :]


def time_decoder(W, i, xt, skip, out_len, last):
    """demucscpp::apply_time_decoder (src/encdec.cpp:258-361)."""
    p = "tdecoder.%d" % i
    y = F.conv1d((xt + skip)[None], W[p + ".rewrite.weight"], W[p + ".rewrite.bias"], padding=1)
    y = glu(y, 1)
    y = dconv(W, p, y)
    z = F.conv_transpose1d(y, W[p + ".conv_tr.weight"], W[p + ".conv_tr.bias"], stride=4)[0]
    if not last:
        z = gelu(z)
    return z[:, 2:2 + out_len]


def sin_embedding_2d(d_model, height, width, dtype):
    """create_2d_sin_embedding (src/crosstransformer.cpp:7-53): first half of the channels
    encodes the WIDTH (time-frame) index, second half the HEIGHT (frequency) index,
    sin/cos interleaved.  Returns (d_model, height, width)."""
    pe = torch.zeros(d_model, height, width, dtype=dtype)
    half = d_model // 2
    div = np.exp(np.arange(0, half, 2, dtype=np.float32) * np.float32(-math.log(10000.0) / half)).astype(np.float32)
    div = torch.from_numpy(div)
    wpos = torch.arange(width, dtype=torch.float32)[None, :] * div[:, None]     # float32 like the ref
    hpos = torch.arange(height, dtype=torch.float32)[None, :] * div[:, None]
    pe[0:half:2] = torch.sin(wpos).to(dtype)[:, None, :]
    pe[1:half:2] = torch.cos(wpos).to(dtype)[:, None, :]
    pe[half::2] = torch.sin(hpos).to(dtype)[:, :, None]
    pe[half + 1::2] = torch.cos(hpos).to(dtype)[:, :, None]
    return pe


def sin_embedding_1d(length, dim, dtype):
    """create_sin_embedding (src/crosstransformer.cpp:55-77): cos first half, sin second
    half, phase = t / 10000^(i/(half-1)).  Returns (length, dim)."""
    half = dim // 2
    div = (np.arange(half, dtype=np.float32) / np.float32(half - 1)).astype(np.float32)
    denom = np.power(np.float32(10000.0), div).astype(np.float32)
    phase = (np.arange(length, dtype=np.float32)[:, None] / denom[None, :]).astype(np.float32)
    ph = torch.from_numpy(phase)
    return torch.cat([torch.cos(ph), torch.sin(ph)], dim=1).to(dtype)


def transformer_layer(W, p, q, k, cross):
    """demucscpp::common_encoder_layer (src/layers.cpp:377-531).  q: (T, C); k: (S, C).
    Self layers reuse norm1 for keys and their norm2 acts as the pre-FFN norm
    (src/crosstransformer.cpp:79-136); cross layers have norm1/norm2/norm3."""
    C = q.shape[1]
    nh = 8
    dh = C // nh
    attn = "cross_attn" if cross else "self_attn"
    qn = layer_norm(q, W[p + "norm1.weight"], W[p + "norm1.bias"])
    kn = layer_norm(k, W[p + "norm2.weight"], W[p + "norm2.bias"]) if cross else qn
    ipw, ipb = W[p + attn + ".in_proj_weight"], W[p + attn + ".in_proj_bias"]
    Q = qn @ ipw[:C].T + ipb[:C]
    K = kn @ ipw[C:2 * C].T + ipb[C:2 * C]
    V = kn @ ipw[2 * C:].T + ipb[2 * C:]
    out = torch.empty_like(Q)
    for h in range(nh):                                    # contiguous dh-channel blocks per head
        sl = slice(h * dh, (h + 1) * dh)
        dots = Q[:, sl] @ K[:, sl].T / math.sqrt(float(dh))
        dots = dots - dots.max(dim=1, keepdim=True).values
        e = torch.exp(dots)
        out[:, sl] = (e / e.sum(dim=1, keepdim=True)) @ V[:, sl]
    o = out @ W[p + attn + ".out_proj.weight"].T + W[p + attn + ".out_proj.bias"]
    q = q + o * W[p + "gamma_1.scale"]
    n3 = "norm3" if cross else "norm2"
    f = layer_norm(q, W[p + n3 + ".weight"], W[p + n3 + ".bias"])
    f = gelu(f @ W[p + "linear1.weight"].T + W[p + "linear1.bias"])
    f = f @ W[p + "linear2.weight"].T + W[p + "linear2.bias"]
    q = q + f * W[p + "gamma_2.scale"]
    # MyGroupNorm = one group over (T, C), N-1 variance, per-channel affine (:517-530)
    m, v = var_unbiased(q, (0, 1))
    return (q - m) / torch.sqrt(v + 1e-5) * W[p + "norm_out.weight"] + W[p + "norm_out.bias"]


def crosstransformer(W, x, xt, taps=None):
    """demucscpp::apply_crosstransformer (src/crosstransformer.cpp:205-339).
    x: (C, Fr, T1) freq branch, xt: (C, T2) time branch; returns same shapes.
    Freq tokens are ordered t*Fr + f (:227-238).  With `taps`, the token streams after norm_in + positional
    embedding ("x_ct_in", "xt_ct_in") and after each of the five layers ("x_ct_<i>", "xt_ct_<i>": the output of
    common_encoder_layer, src/layers.cpp:377-531) are kept, freq as (C, Fr, T1), time as (1, C, T2)."""
    C, Fr, T1 = x.shape
    dt = x.dtype
    pe2 = sin_embedding_2d(C, Fr, T1, dt).permute(2, 1, 0).reshape(T1 * Fr, C)
    xq = x.permute(2, 1, 0).reshape(T1 * Fr, C)
    xq = layer_norm(xq, W["crosstransformer.norm_in.weight"], W["crosstransformer.norm_in.bias"]) + pe2
    tq = xt.T
    tq = layer_norm(tq, W["crosstransformer.norm_in_t.weight"], W["crosstransformer.norm_in_t.bias"]) \
        + sin_embedding_1d(tq.shape[0], C, dt)

    def keep(tag):
        if taps is not None:
            taps["x_ct_" + tag] = xq.reshape(T1, Fr, C).permute(2, 1, 0).clone()
            taps["xt_ct_" + tag] = tq.T[None].clone()

    keep("in")
    for idx in range(5):
        pf = "crosstransformer.layers.%d." % idx
        pt = "crosstransformer.layers_t.%d." % idx
        if idx % 2 == 0:
            xq = transformer_layer(W, pf, xq, xq, False)
            tq = transformer_layer(W, pt, tq, tq, False)
        else:
            old = xq
            xq = transformer_layer(W, pf, xq, tq, True)
            tq = transformer_layer(W, pt, tq, old, True)
        keep(str(idx))
    return xq.reshape(T1, Fr, C).permute(2, 1, 0), tq.T


# ----------------------------------------------------------------------------- per-segment forward
def segment_forward(W, mix, taps=None):
    """demucscpp::model_inference (src/model_inference.cpp:48-475).
    mix: (2, 343980) tensor -> (S, 2, 343980).  `taps`, if a dict, receives the
    intermediate tensors under the names oracle/ref_harness.cpp exposes."""
    dt = W.dtype
    mix = mix.to(dt)
    S = W.S
    tp = taps if taps is not None else {}

    padded = mirror_pad_edge(mix, PAD, PAD_END)                          # :64  (2, 347136)
    spec = stft(padded)                                                  # :71  (2, 2049, 340)
    z = spec[:, :NFFT // 2, 2:2 + LE]                                    # :75-99
    x = torch.stack([z[0].real, z[0].imag, z[1].real, z[1].imag]).to(dt)  # CaC (4, 2048, 336)
    mean, var = var_unbiased(x, (0, 1, 2))
    std = torch.sqrt(var)
    x = (x - mean) / (std + 1e-5)                                        # :115-124
    meant, vart = var_unbiased(mix, (0, 1))
    stdt = torch.sqrt(vart)
    xt = (mix - meant) / (stdt + 1e-5)                                   # :127-144
    tp["x"], tp["xt"] = x, xt[None]

    saved, savedt = [], []
    for i in range(4):                                                   # :155-212
        xt = time_encoder(W, i, xt)
        x = freq_encoder(W, i, x)
        if i == 0:
            x = x + 2.0 * W["freq_emb.embedding.weight"].T[:, :, None]   # :163-179 (10 * 0.2)
        saved.append(x)
        savedt.append(xt)
        tp["saved_%d" % i], tp["savedt_%d" % i] = x, xt[None]

    if W.is_4s:                                                          # :214-287
        x = torch.einsum("oc,cft->oft", W["channel_upsampler.weight"], x) + W["channel_upsampler.bias"][:, None, None]
        xt = W["channel_upsampler_t.weight"] @ xt + W["channel_upsampler_t.bias"][:, None]
    x, xt = crosstransformer(W, x, xt, taps)
    if W.is_4s:
        tp["x_3_up"], tp["xt_3_up"] = x, xt[None]
        x = torch.einsum("oc,cft->oft", W["channel_downsampler.weight"], x) + W["channel_downsampler.bias"][:, None, None]
        xt = W["channel_downsampler_t.weight"] @ xt + W["channel_downsampler_t.bias"][:, None]
    tp["x_3"], tp["xt_3"] = x, xt[None]

    names = ["2", "1", "0", "out"]
    for i in range(4):                                                   # :314-344
        x = freq_decoder(W, i, x, saved[3 - i], last=(i == 3))
        xt = time_decoder(W, i, xt, savedt[3 - i], TLEN[3 - i], last=(i == 3))
        tp["x_" + names[i]], tp["xt_" + names[i]] = x, xt[None]

    x = x * std + mean                                                   # :380-393 (no eps)
    xt = xt * stdt + meant                                               # :396-405
    out = torch.empty(S, 2, SEG, dtype=dt)
    for s in range(S):                                                   # :409-474
        xs = x[4 * s:4 * s + 4]
        zc = torch.complex(torch.stack([xs[0], xs[2]]), torch.stack([xs[1], xs[3]]))
        zc = F.pad(zc, (2, 2, 0, 1))                                     # 2 zero frames each side, 1 zero bin
        w = istft(zc, SEG + PAD + PAD_END)
        out[s] = w[:, PAD:PAD + SEG] + xt[2 * s:2 * s + 2]
    return out


# ----------------------------------------------------------------------------- hdemucs_mmi (v3)
def group_norm_g(x, w, b, groups, eps=1e-5):
    """demucscpp_v3::groupnorm::generalized_group_norm (src/layers.hpp:130-170): statistics per group over
    (dim0, C/G, W) JOINTLY, N-1 variance.  x: (dim0, C, W)."""
    d0, C, Wd = x.shape
    xg = x.reshape(d0, groups, C // groups, Wd)
    m, v = var_unbiased(xg, (0, 2, 3))
    y = ((xg - m) / torch.sqrt(v + eps)).reshape(d0, C, Wd)
    return y * w[None, :, None] + b[None, :, None]


def lstm_bidir2(W, p, x):
    """demucscpp_v3::lstm_forward (src/lstm.cpp:68-147): 2-layer bidirectional LSTM over the WHOLE sequence
    (no 200-step framing), PyTorch gate order i,f,g,o, zero initial state.  x: (T, in) -> (T, 2h)."""
    inp = x
    for layer in range(2):
        outs = []
        for rev, suf in ((False, ""), (True, "_reverse")):
            wih, whh = W[p + "weight_ih_l%d%s" % (layer, suf)], W[p + "weight_hh_l%d%s" % (layer, suf)]
            bias = W[p + "bias_ih_l%d%s" % (layer, suf)] + W[p + "bias_hh_l%d%s" % (layer, suf)]
            hsz = whh.shape[1]
            pre = inp @ wih.T + bias                      # input projections of all steps
            h = torch.zeros(hsz, dtype=x.dtype)
            c = torch.zeros(hsz, dtype=x.dtype)
            out = torch.empty(inp.shape[0], hsz, dtype=x.dtype)
            steps = range(inp.shape[0] - 1, -1, -1) if rev else range(inp.shape[0])
            for t in steps:
                g = pre[t] + whh @ h
                i_t = 1.0 / (1.0 + torch.exp(-g[:hsz]))
                f_t = 1.0 / (1.0 + torch.exp(-g[hsz:2 * hsz]))
                g_t = torch.tanh(g[2 * hsz:3 * hsz])
                o_t = 1.0 / (1.0 + torch.exp(-g[3 * hsz:]))
                c = f_t * c + i_t * g_t
                h = o_t * torch.tanh(c)
                out[t] = h
            outs.append(out)
        inp = torch.cat(outs, dim=1)
    return inp


def local_attention(W, p, x):
    """demucscpp_v3::local_attention (src/layers.cpp:533-721).  x: (C, T) -> x + proj(attn).  4 heads, 4 decay
    channels; dots[t, s] = <K[:, t], Q[:, s]>/sqrt(dh) + sum_n kernel[n, |t-s|] * 0.5*sigmoid(qd[h*4+n, s]);
    diagonal := -100; softmax over t (keys) for each query s."""
    C, T = x.shape
    nh, nd = 4, 4
    dh = C // nh
    q = W[p + "query.weight"] @ x + W[p + "query.bias"][:, None]
    k = W[p + "key.weight"] @ x + W[p + "key.bias"][:, None]
    cont = W[p + "content.weight"] @ x + W[p + "content.bias"][:, None]
    qd = W[p + "query_decay.weight"] @ x + W[p + "query_decay.bias"][:, None]
    qd = 0.5 / (1.0 + torch.exp(-qd))                                     # :595-596
    idx = torch.arange(T)
    delta = (idx[:, None] - idx[None, :]).abs().to(x.dtype)               # |t - s|
    out = torch.empty_like(x)
    for h in range(nh):
        sl = slice(h * dh, (h + 1) * dh)
        dots = k[sl].T @ q[sl] / math.sqrt(float(dh))                     # [t, s]
        for n in range(nd):
            dots = dots + (-(n + 1) * delta / math.sqrt(float(nd))) * qd[h * nd + n][None, :]
        dots = torch.where(torch.eye(T, dtype=torch.bool), torch.full_like(dots, -100.0), dots)
        w = torch.softmax(dots, dim=0)                                    # over keys t
        out[sl] = cont[sl] @ w                                            # out[c, s] = sum_t w[t, s] content[c, t]
    return x + W[p + "proj.weight"] @ out + W[p + "proj.bias"][:, None]


def dconv_v3(W, prefix, y):
    """apply_dconv_v3 (src/layers.cpp:723-894): like dconv() with hidden = C/4."""
    return dconv(W, prefix, y)


def dconv_v3_lstm_attn(W, prefix, y):
    """apply_dconv_v3_encoder_4_5 (src/layers.cpp:896-1113).  y: (1, C, T).  Per sub-block:
    conv k3 dil -> GN(1)+GELU -> BiLSTM(2 layers)+Linear+skip -> LocalAttention -> 1x1 -> GN(1) -> GLU -> scale -> +=."""
    for d in range(2):
        p = "%s.dconv.layers.%d." % (prefix, d)
        dil = 2 ** d
        h = F.conv1d(y, W[p + "0.weight"], W[p + "0.bias"], padding=dil, dilation=dil)
        h = gelu(group_norm1(h, W[p + "1.weight"], W[p + "1.bias"]))[0]            # (hid, T)
        seq = h.T                                                                   # (T, hid)
        lo = lstm_bidir2(W, p + "3.lstm.", seq) @ W[p + "3.linear.weight"].T + W[p + "3.linear.bias"] + seq
        h = local_attention(W, p + "4.", lo.T)[None]                                # (1, hid, T)
        h = F.conv1d(h, W[p + "5.weight"][:, :, None], W[p + "5.bias"])
        h = glu(group_norm1(h, W[p + "6.weight"], W[p + "6.bias"]), 1)
        y = y + h * W[p + "8.scale"][None, :, None]
    return y


def segment_forward_v3(W, mix, taps=None):
    """demucscpp_v3::model_v3_inference (src/model_inference.cpp:477-856).  mix: (2, 343980) -> (4, 2, 343980)."""
    dt = W.dtype
    mix = mix.to(dt)
    tp = taps if taps is not None else {}
    padded = mirror_pad_edge(mix, PAD, PAD_END)
    z = stft(padded)[:, :NFFT // 2, 2:2 + LE]
    x = torch.stack([z[0].real, z[0].imag, z[1].real, z[1].imag]).to(dt)
    mean, var = var_unbiased(x, (0, 1, 2))
    std = torch.sqrt(var)
    x = (x - mean) / (std + 1e-5)
    meant, vart = var_unbiased(mix, (0, 1))
    stdt = torch.sqrt(vart)
    xt = (mix - meant) / (stdt + 1e-5)
    tp["x"], tp["xt"] = x, xt[None]
    saved, savedt = [], []
    for i in range(4):                                                   # :586-655 (encdec.cpp:363-517)
        xt = time_encoder(W, i, xt)
        x = freq_encoder(W, i, x)
        if i == 0:
            x = x + 2.0 * W["freq_emb.embedding.weight"].T[:, :, None]
        saved.append(x)
        savedt.append(xt)
        tp["saved_%d" % i], tp["savedt_%d" % i] = x, xt[None]
    # tencoder 4: plain conv, no activation (encdec.cpp:519-530)
    xt4 = conv1d_ceil(xt[None], W["tencoder.4.conv.weight"], W["tencoder.4.conv.bias"], 4, 2)   # (1, 768, 336)
    tp["xt_4"] = xt4
    # encoder 4 (encdec.cpp:532-573): k(8,1) s(4,1) conv collapses 8 bins to 1, + inject, GN(4)+GELU, DConv, 1x1, GN(4), GLU
    y = F.conv2d(x[None], W["encoder.4.conv.weight"][..., None], W["encoder.4.conv.bias"], stride=(4, 1))[0]  # (768,1,336)
    y = y.permute(1, 0, 2) + xt4                                          # (1, 768, 336)
    y = gelu(group_norm_g(y, W["encoder.4.norm1.weight"], W["encoder.4.norm1.bias"], 4))
    y = dconv_v3_lstm_attn(W, "encoder.4", y)
    y = F.conv1d(y, W["encoder.4.rewrite.weight"][:, :, None], W["encoder.4.rewrite.bias"])
    y = group_norm_g(y, W["encoder.4.norm2.weight"], W["encoder.4.norm2.bias"], 4)
    x4 = glu(y, 1)                                                        # (1, 768, 336)
    tp["saved_4"] = x4.permute(1, 0, 2)                                   # reference keeps it as (768, 1, 336)
    # shared encoder 5 (encdec.cpp:575-615)
    y = conv1d_ceil(x4, W["encoder.5.conv.weight"], W["encoder.5.conv.bias"], 2, 1)               # (1, 1536, 168)
    y = gelu(group_norm_g(y, W["encoder.5.norm1.weight"], W["encoder.5.norm1.bias"], 4))
    y = dconv_v3_lstm_attn(W, "encoder.5", y)
    y = F.conv1d(y, W["encoder.5.rewrite.weight"][:, :, None], W["encoder.5.rewrite.bias"])
    y = group_norm_g(y, W["encoder.5.norm2.weight"], W["encoder.5.norm2.bias"], 4)
    x5 = glu(y, 1)                                                        # (1, 1536, 168)
    tp["x_shared_5"] = x5
    # shared decoder 0 (encdec.cpp:617-656): input = skip only
    y = F.conv1d(x5, W["decoder.0.rewrite.weight"], W["decoder.0.rewrite.bias"], padding=1)
    y = glu(group_norm_g(y, W["decoder.0.norm1.weight"], W["decoder.0.norm1.bias"], 4), 1)
    y = F.conv_transpose1d(y, W["decoder.0.conv_tr.weight"], W["decoder.0.conv_tr.bias"], stride=2)
    y = gelu(group_norm_g(y, W["decoder.0.norm2.weight"], W["decoder.0.norm2.bias"], 4))
    xd0 = y[:, :, 1:1 + LE]                                               # (1, 768, 336)
    tp["x_4"] = xd0
    # freq decoder 1 (encdec.cpp:658-698): returns `pre`, seeds the time decoder chain
    y = (xd0.permute(1, 0, 2) + x4.permute(1, 0, 2))[None]                # (1, 768, 1, 336)
    y = F.conv2d(y, W["decoder.1.rewrite.weight"], W["decoder.1.rewrite.bias"], padding=1)[0]     # (1536, 1, 336)
    y = group_norm_g(y.permute(1, 0, 2), W["decoder.1.norm1.weight"], W["decoder.1.norm1.bias"], 4).permute(1, 0, 2)
    pre = glu(y, 0)                                                       # (768, 1, 336)
    y = F.conv_transpose2d(pre[None], W["decoder.1.conv_tr.weight"][..., None], W["decoder.1.conv_tr.bias"],
                           stride=(4, 1))[0]                              # (384, 8, 336)
    x = gelu(group_norm_g(y.permute(1, 0, 2), W["decoder.1.norm2.weight"], W["decoder.1.norm2.bias"], 4)).permute(1, 0, 2)
    tp["x_3"] = x
    # time decoder 0 (encdec.cpp:700-726)
    y = F.conv_transpose1d(pre.permute(1, 0, 2), W["tdecoder.0.conv_tr.weight"], W["tdecoder.0.conv_tr.bias"], stride=4)
    y = gelu(group_norm_g(y, W["tdecoder.0.norm2.weight"], W["tdecoder.0.norm2.bias"], 4))
    xt = y[0, :, 2:2 + TLEN[4]]                                           # (384, 1344)
    tp["xt_3"] = xt[None]
    # common decoders, no DConv (encdec.cpp:728-863)
    names = ["2", "1", "0", "out"]
    for l in range(4):
        pf, pt = "decoder.%d" % (l + 2), "tdecoder.%d" % (l + 1)
        y = F.conv2d((x + saved[3 - l])[None], W[pf + ".rewrite.weight"], W[pf + ".rewrite.bias"], padding=1)[0]
        y = glu(y, 0)
        y = F.conv_transpose2d(y[None], W[pf + ".conv_tr.weight"][..., None], W[pf + ".conv_tr.bias"], stride=(4, 1))[0]
        if l < 3:
            y = gelu(y)
        x = y[:, 2:y.shape[1] - 2, :]
        yt = F.conv1d((xt + savedt[3 - l])[None], W[pt + ".rewrite.weight"], W[pt + ".rewrite.bias"], padding=1)
        yt = glu(yt, 1)
        yt = F.conv_transpose1d(yt, W[pt + ".conv_tr.weight"], W[pt + ".conv_tr.bias"], stride=4)[0]
        if l < 3:
            yt = gelu(yt)
        xt = yt[:, 2:2 + TLEN[3 - l]]
        tp["x_" + names[l]], tp["xt_" + names[l]] = x, xt[None]
    x = x * std + mean
    xt = xt * stdt + meant
    out = torch.empty(4, 2, SEG, dtype=dt)
    for s_ in range(4):
        xs = x[4 * s_:4 * s_ + 4]
        zc = torch.complex(torch.stack([xs[0], xs[2]]), torch.stack([xs[1], xs[3]]))
        zc = F.pad(zc, (2, 2, 0, 1))
        w = istft(zc, SEG + PAD + PAD_END)
        out[s_] = w[:, PAD:PAD + SEG] + xt[2 * s_:2 * s_ + 2]
    return out


# ----------------------------------------------------------------------------- track level
def segment_plan(length):
    """Chunk (offset, length) list of split_inference (src/model_apply.cpp:200-204)."""
    return [(o, min(SEG, length - o)) for o in range(0, length, STRIDE)]


def triangle_weight():
    """src/model_apply.cpp:171-179 (TRANSITION_POWER = 1)."""
    h = SEG // 2
    w = np.zeros(SEG, dtype=np.float32)
    w[:h] = np.linspace(1, h, h, dtype=np.float32)
    w[SEG - h:] = w[:h][::-1]
    return (w / w.max()).astype(np.float32)


def separate(W, audio, shift_offset, forward=None):
    """demucscpp::demucs_inference (src/model_apply.cpp:60-288) with the random shift
    (`rand() % 22050`, :114) passed in explicitly.  audio: (2, L) -> (S, 2, L)."""
    dt = W.dtype
    forward = forward or ((lambda m: segment_forward_v3(W, m)) if getattr(W, "is_v3", False)
                          else (lambda m: segment_forward(W, m)))
    a = torch.as_tensor(audio).to(dt)
    L = a.shape[1]
    ref = a.mean(dim=0)                                                  # :72
    rmean = ref.mean()
    rstd = torch.sqrt(((ref - rmean) ** 2).sum() / (L - 1))              # :75-77 (N-1)
    a = (a - rmean) / rstd
    padded = F.pad(a, (MAX_SHIFT, MAX_SHIFT))                            # :108-112
    shifted = padded[:, shift_offset:shift_offset + L + MAX_SHIFT - shift_offset]  # :119-120
    Ls = shifted.shape[1]
    out = torch.zeros(W.S, 2, Ls, dtype=dt)
    sumw = torch.zeros(Ls, dtype=dt)
    w = torch.from_numpy(triangle_weight()).to(dt)
    for off, n in segment_plan(Ls):                                      # :200-235
        left = (SEG - n) // 2                                            # centre zero pad, :21-43,:260-261
        seg = torch.zeros(2, SEG, dtype=dt)
        seg[:, left:left + n] = shifted[:, off:off + n]
        y = forward(seg)[:, :, left:left + n]                            # :274-285
        out[:, :, off:off + n] += w[:n] * y
        sumw[off:off + n] += w[:n]
    out = out / sumw                                                     # :237-246
    out = out[:, :, MAX_SHIFT - shift_offset:MAX_SHIFT - shift_offset + L]  # :129-135
    return out * rstd + rmean                                            # :88
