"""Seeded synthetic weights, config and batches for the benchmark/parity harness.

The reference ships no FastSpeech2 checkpoint, its HiFi-GAN weights are stripped from the mount and
its speakers.json / stats.json are 0-byte files, so every measurement uses
seeded random-init weights with the reference's state_dict layout and a
fabricated config (schemas: gather.py:28-30, 56-69).  Pure torch-CPU generation, deterministic for a
given torch build, so the build container and the GPU box produce identical tensors.
"""
import math

import numpy as np
import torch

# --- model constants of config/JVS-VCTK_langemb_configs/JVS-VCTK_1/{model,preprocess}.yaml -------
MODEL_CONFIG = {
    "transformer": {
        "encoder_layer": 4, "encoder_head": 2, "encoder_hidden": 256,
        "decoder_layer": 6, "decoder_head": 2, "decoder_hidden": 256,
        "conv_filter_size": 1024, "conv_kernel_size": [9, 1],
        "encoder_dropout": 0.2, "decoder_dropout": 0.2,
    },
    "variance_predictor": {"filter_size": 256, "kernel_size": 3, "dropout": 0.5},
    "variance_embedding": {"pitch_quantization": "linear", "energy_quantization": "linear", "n_bins": 256},
    "jdit": {"use_jdit": False, "prenet_hidden": 256},
    "multi_speaker": True,
    "max_seq_len": 1000,
    "vocoder": {"model": "HiFi-GAN", "speaker": "universal"},
    "speaker_generation": {"GMM_mixtures": 3},
    "discriminator": {"hidden": 256},
}
PREPROCESS_CONFIG = {
    "audio": {"sampling_rate": 22050, "max_wav_value": 32768.0},
    "stft": {"filter_length": 1024, "hop_length": 256, "win_length": 1024},
    "mel": {"n_mel_channels": 80, "mel_fmin": 0, "mel_fmax": 8000},
    "pitch": {"feature": "phoneme_level", "normalization": True},
    "energy": {"feature": "phoneme_level", "normalization": True},
    "speaker_generation": {"metadata": {"gender": {"M": 0, "F": 1}, "language": {"ja": 0, "en": 1}}},
}
HIFIGAN_CONFIG = {
    "resblock": "1", "upsample_rates": [8, 8, 2, 2], "upsample_kernel_sizes": [16, 16, 4, 4],
    "upsample_initial_channel": 512, "resblock_kernel_sizes": [3, 7, 11],
    "resblock_dilation_sizes": [[1, 3, 5], [1, 3, 5], [1, 3, 5]],
    "num_mels": 80, "hop_size": 256, "sampling_rate": 22050,
}
N_VOCAB = 429          # len(text.symbols.symbols) + 1  (transformer/Models.py:40)
N_ACCENT = 5           # transformer/Models.py:58-60
D_MODEL = 256
N_MEL = 80
GMM_K = 3
META_DIM = 4
SAMPLING_RATE = 22050
HOP = 256
STATS_JSON = {"pitch": [-2.5, 8.0, 0.0, 1.0], "energy": [-1.5, 7.0, 0.0, 1.0]}


def speakers_json(n_speakers=209):
    """{name: [id, gender, language]} (gather.py:28-30): 100 JVS + 109 VCTK style entries."""
    out = {}
    for i in range(n_speakers):
        ja = i < 100
        out[("jvs%03d" % (i + 1)) if ja else ("p%03d" % (125 + i))] = [i, "MF"[i % 2], "ja" if ja else "en"]
    return out


def sinusoid_table(n_position, d_hid=D_MODEL):
    """transformer/Models.py:10-30 — float64 numpy angles pos / 10000^(2*(j//2)/d), sin on even and cos
    on odd columns, cast to fp32.  The 256 denominators use scalar np.power exactly as the reference's
    per-element call does; the rest is elementwise IEEE arithmetic."""
    denom = np.array([np.power(10000, 2 * (j // 2) / d_hid) for j in range(d_hid)], dtype=np.float64)
    pos = np.arange(n_position, dtype=np.float64)[:, None]
    table = pos / denom[None, :]
    table[:, 0::2] = np.sin(table[:, 0::2])
    table[:, 1::2] = np.cos(table[:, 1::2])
    return torch.FloatTensor(table)


def _normal(g, shape, std):
    return torch.randn(shape, generator=g, dtype=torch.float32) * std


def make_fs2_state_dict(seed=1234, n_speakers=209, duration_mode="fixed3", stats=STATS_JSON):
    """Random-init FastSpeech2 weights keyed exactly like the reference state_dict (Appendix A).

    duration_mode: "fixed3" → duration predictor linear weight 0, bias ln 4 (every valid phoneme gets
    exactly 3 frames: the deterministic benchmark shape); "spread" → bias ln 4 only.
    PostNet BatchNorm statistics are randomised so BN-folding bugs cannot hide.
    """
    g = torch.Generator().manual_seed(seed)
    sd = {}
    d, dff = D_MODEL, 1024

    def lin(name, out_f, in_f, k=None, gain=1.0):
        shape = (out_f, in_f) if k is None else (out_f, in_f, k)
        fan_in = in_f * (k or 1)
        sd[name + ".weight"] = _normal(g, shape, gain / math.sqrt(fan_in))
        sd[name + ".bias"] = _normal(g, (out_f,), 0.05)

    def ln(name, n):
        sd[name + ".weight"] = 1.0 + _normal(g, (n,), 0.1)
        sd[name + ".bias"] = _normal(g, (n,), 0.1)

    pe = sinusoid_table(MODEL_CONFIG["max_seq_len"] + 1).unsqueeze(0)
    sd["encoder.position_enc"] = pe.clone()
    sd["decoder.position_enc"] = pe.clone()
    w = _normal(g, (N_VOCAB, d), 1.0); w[0] = 0
    sd["encoder.src_word_emb.weight"] = w
    w = _normal(g, (N_ACCENT, d), 1.0); w[0] = 0
    sd["encoder.src_accent_emb.weight"] = w
    for part, n_layers in (("encoder", 4), ("decoder", 6)):
        for n in range(n_layers):
            p = "%s.layer_stack.%d." % (part, n)
            for nm in ("w_qs", "w_ks", "w_vs", "fc"):
                lin(p + "slf_attn." + nm, d, d)
            ln(p + "slf_attn.layer_norm", d)
            lin(p + "pos_ffn.w_1", dff, d, 9, gain=math.sqrt(2.0))
            lin(p + "pos_ffn.w_2", d, dff, 1)
            ln(p + "pos_ffn.layer_norm", d)
    sd["variance_adaptor.pitch_bins"] = torch.linspace(stats["pitch"][0], stats["pitch"][1], 255)
    sd["variance_adaptor.energy_bins"] = torch.linspace(stats["energy"][0], stats["energy"][1], 255)
    for nm in ("duration", "pitch", "energy"):
        p = "variance_adaptor.%s_predictor." % nm
        lin(p + "conv_layer.conv1d_1.conv", d, d, 3, gain=math.sqrt(2.0))
        ln(p + "conv_layer.layer_norm_1", d)
        lin(p + "conv_layer.conv1d_2.conv", d, d, 3, gain=math.sqrt(2.0))
        ln(p + "conv_layer.layer_norm_2", d)
        lin(p + "linear_layer", 1, d, gain=2.0 if nm != "duration" else 0.5)
    if duration_mode == "fixed3":
        sd["variance_adaptor.duration_predictor.linear_layer.weight"].zero_()
    sd["variance_adaptor.duration_predictor.linear_layer.bias"].fill_(math.log(4.0))
    sd["variance_adaptor.pitch_embedding.weight"] = _normal(g, (256, d), 0.5)
    sd["variance_adaptor.energy_embedding.weight"] = _normal(g, (256, d), 0.5)
    lin("mel_linear", N_MEL, d)
    chans = [N_MEL, 512, 512, 512, 512, N_MEL]
    for i in range(5):
        p = "postnet.convolutions.%d." % i
        lin(p + "0.conv", chans[i + 1], chans[i], 5, gain=1.0)
        c = chans[i + 1]
        sd[p + "1.weight"] = 1.0 + _normal(g, (c,), 0.2)
        sd[p + "1.bias"] = _normal(g, (c,), 0.1)
        sd[p + "1.running_mean"] = _normal(g, (c,), 0.2)
        sd[p + "1.running_var"] = 0.5 + torch.rand((c,), generator=g)
        sd[p + "1.num_batches_tracked"] = torch.tensor(1000, dtype=torch.int64)
    sd["speaker_emb.weight"] = _normal(g, (n_speakers, d), 0.5)
    lin("speaker_enc.pi_linear.0", GMM_K, META_DIM)
    lin("speaker_enc.sigma_linear.0", GMM_K * d, META_DIM)
    lin("speaker_enc.mu_linear", GMM_K * d, META_DIM)
    sd["speaker_enc.sigma_linear.0.bias"] -= 1.0      # softplus(.) ~ 0.3: compact mixture components
    return sd


def hifigan_layers(cfg=HIFIGAN_CONFIG):
    """(name, kind, c_in, c_out, k, stride, dilation) for every conv of the generator in forward
    order (hifigan/models.py:113-147)."""
    ch0 = cfg["upsample_initial_channel"]
    out = [("conv_pre", "conv", cfg["num_mels"], ch0, 7, 1, 1)]
    ch = ch0
    for i, (u, k) in enumerate(zip(cfg["upsample_rates"], cfg["upsample_kernel_sizes"])):
        out.append(("ups.%d" % i, "convT", ch0 // 2 ** i, ch0 // 2 ** (i + 1), k, u, 1))
        ch = ch0 // 2 ** (i + 1)
        for j, (kk, dd) in enumerate(zip(cfg["resblock_kernel_sizes"], cfg["resblock_dilation_sizes"])):
            rb = i * len(cfg["resblock_kernel_sizes"]) + j
            for m, dil in enumerate(dd):
                out.append(("resblocks.%d.convs1.%d" % (rb, m), "conv", ch, ch, kk, 1, dil))
            for m in range(len(dd)):
                out.append(("resblocks.%d.convs2.%d" % (rb, m), "conv", ch, ch, kk, 1, 1))
    out.append(("conv_post", "conv", ch, 1, 7, 1, 1))
    return out


def make_hifigan_state_dict(seed=4321, weight_norm=True, cfg=HIFIGAN_CONFIG):
    """Random-init generator weights.  Fan-in-scaled normals (the stock N(0, 0.01) init of
    hifigan/models.py:10-13 collapses the signal to DC and makes an SNR bar meaningless, SURVEY.md
    §8(c)).  weight_norm=True emits the checkpoint layout (weight_g/weight_v/bias, 234 entries) that
    utils/model.py:62-68 loads before remove_weight_norm(); False emits folded weight/bias."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, kind, cin, cout, k, stride, dil in hifigan_layers(cfg):
        if kind == "conv":
            shape = (cout, cin, k)
            gain = 1.0
            if ".convs1." in name:
                gain = 1.4                   # leaky-relu halves the variance of its input
            if ".convs2." in name:
                gain = 0.6                   # keep the residual branch modest so x stays O(1)
            if name == "conv_post":
                gain = 0.8
            std = gain / math.sqrt(cin * k)
        else:
            shape = (cin, cout, k)           # ConvTranspose1d layout
            std = math.sqrt(2.0 * stride / (cin * k))
        v = _normal(g, shape, std)
        b = _normal(g, (cout,), 0.02)
        if weight_norm:
            norm = v.flatten(1).norm(dim=1).view(-1, 1, 1)       # dim=0: per first-dim slice
            gscale = 1.0 + 0.2 * (torch.rand(norm.shape, generator=g) - 0.5)
            sd[name + ".bias"] = b
            sd[name + ".weight_g"] = norm * gscale
            sd[name + ".weight_v"] = v
        else:
            sd[name + ".weight"] = v
            sd[name + ".bias"] = b
    return sd


def fold_weight_norm(sd):
    """w = g * v / ||v|| over all dims but 0 (torch.nn.utils.weight_norm default dim=0;
    hifigan/models.py:167-174 strips it once at load)."""
    out = {}
    for k, v in sd.items():
        if k.endswith(".weight_v"):
            base = k[: -len(".weight_v")]
            gg = sd[base + ".weight_g"]
            norm = v.flatten(1).norm(dim=1).view(-1, *([1] * (v.dim() - 1)))
            out[base + ".weight"] = v * (gg / norm)
        elif k.endswith(".weight_g"):
            continue
        else:
            out[k] = v
    return out


META_COMBOS = ((1, 0, 1, 0), (1, 0, 0, 1), (0, 1, 1, 0), (0, 1, 0, 1))   # (M,F | ja,en) one-hots


def make_batch(B, L, seed, lens=None, accents_const=None):
    """Synthetic phoneme batch: ids ~ U{1..428}, 0 only as pad; accents U{0..3} or a
    constant; speaker metadata one-hots cycling the four (gender, language) combos."""
    g = torch.Generator().manual_seed(seed)
    if lens is None:
        lens = torch.full((B,), L, dtype=torch.int64)
    else:
        lens = torch.as_tensor(lens, dtype=torch.int64)
    Lmax = int(lens.max())
    texts = torch.randint(1, N_VOCAB, (B, Lmax), generator=g, dtype=torch.int64)
    if accents_const is None:
        accents = torch.randint(0, 4, (B, Lmax), generator=g, dtype=torch.int64)
    else:
        accents = torch.full((B, Lmax), accents_const, dtype=torch.int64)
    pad = torch.arange(Lmax)[None, :] >= lens[:, None]
    texts[pad] = 0
    accents[pad] = 0
    meta = torch.tensor([META_COMBOS[i % 4] for i in range(B)], dtype=torch.float32)
    return {"texts": texts, "accents": accents, "src_lens": lens, "max_src_len": Lmax, "speaker_meta": meta}


def c4_lengths(n=4096):
    """Config 4: numpy default_rng(0).integers(20, 301, 4096)."""
    return np.random.default_rng(0).integers(20, 301, n)
