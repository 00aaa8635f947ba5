"""Drop-in for the reference ``model.FastSpeech2`` (model/fastspeech2.py:15-303) and
``SpeakerMetaEncoder`` (:306-341), inference only.

Same constructor, ``forward`` / ``synthesize_from_speaker_emb`` / ``speaker_gen`` /
``speaker_distribution`` signatures, the same ``state_dict`` keys and shapes
and the same 10/12-tuple outputs, so synthesize.py, examples_gen.py and speaker_gen.py run unmodified
against it.  The sub-modules below are *parameter containers only*: all arithmetic happens in
libmasg.so (hand-written sm_100a CUDA) through the C ABI in include/masg.h.  No CPU fallback: calling
with CPU tensors, in training mode, or with an unsupported configuration raises.
"""
import ctypes
import json
import os

import torch
import torch.nn as nn
from torch.distributions import Categorical, Independent, MixtureSameFamily, Normal

from . import _lib
from .synthetic import N_VOCAB, N_ACCENT, sinusoid_table


class _MHA(nn.Module):
    def __init__(self, d, n_head):
        super().__init__()
        self.w_qs = nn.Linear(d, d)
        self.w_ks = nn.Linear(d, d)
        self.w_vs = nn.Linear(d, d)
        self.layer_norm = nn.LayerNorm(d)
        self.fc = nn.Linear(d, d)


class _FFN(nn.Module):
    def __init__(self, d, d_hid, ks):
        super().__init__()
        self.w_1 = nn.Conv1d(d, d_hid, kernel_size=ks[0], padding=(ks[0] - 1) // 2)
        self.w_2 = nn.Conv1d(d_hid, d, kernel_size=ks[1], padding=(ks[1] - 1) // 2)
        self.layer_norm = nn.LayerNorm(d)


class _FFTBlock(nn.Module):
    def __init__(self, d, n_head, d_hid, ks):
        super().__init__()
        self.slf_attn = _MHA(d, n_head)
        self.pos_ffn = _FFN(d, d_hid, ks)


class _Coder(nn.Module):
    def __init__(self, cfg, part, with_emb):
        super().__init__()
        tr = cfg["transformer"]
        d = tr[part + "_hidden"]
        if with_emb:
            self.src_word_emb = nn.Embedding(N_VOCAB, d, padding_idx=0)
            self.src_accent_emb = nn.Embedding(N_ACCENT, d, padding_idx=0)
        self.position_enc = nn.Parameter(sinusoid_table(cfg["max_seq_len"] + 1, d).unsqueeze(0), requires_grad=False)
        self.layer_stack = nn.ModuleList(
            [_FFTBlock(d, tr[part + "_head"], tr["conv_filter_size"], tr["conv_kernel_size"]) for _ in range(tr[part + "_layer"])])


class _ConvWrap(nn.Module):
    def __init__(self, cin, cout, k):
        super().__init__()
        self.conv = nn.Conv1d(cin, cout, kernel_size=k, padding=(k - 1) // 2)


class _VariancePredictor(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        d = cfg["transformer"]["encoder_hidden"]
        f = cfg["variance_predictor"]["filter_size"]
        k = cfg["variance_predictor"]["kernel_size"]
        self.conv_layer = nn.Module()
        self.conv_layer.conv1d_1 = _ConvWrap(d, f, k)
        self.conv_layer.layer_norm_1 = nn.LayerNorm(f)
        self.conv_layer.conv1d_2 = _ConvWrap(f, f, k)
        self.conv_layer.layer_norm_2 = nn.LayerNorm(f)
        self.linear_layer = nn.Linear(f, 1)


class _VarianceAdaptor(nn.Module):
    def __init__(self, pre, cfg, config_path):
        super().__init__()
        self.duration_predictor = _VariancePredictor(cfg)
        self.pitch_predictor = _VariancePredictor(cfg)
        self.energy_predictor = _VariancePredictor(cfg)
        ve = cfg["variance_embedding"]
        n_bins = ve["n_bins"]
        with open(os.path.join(config_path, "stats.json")) as f:      # model/modules.py:41-46
            stats = json.load(f)
        import numpy as np
        for nm in ("pitch", "energy"):
            lo, hi = stats[nm][:2]
            if ve[nm + "_quantization"] == "log":
                bins = torch.exp(torch.linspace(np.log(lo), np.log(hi), n_bins - 1))
            else:
                bins = torch.linspace(lo, hi, n_bins - 1)
            setattr(self, nm + "_bins", nn.Parameter(bins, requires_grad=False))
        d = cfg["transformer"]["encoder_hidden"]
        self.pitch_embedding = nn.Embedding(n_bins, d)
        self.energy_embedding = nn.Embedding(n_bins, d)


class _PostNet(nn.Module):
    def __init__(self, n_mel=80, dim=512, k=5, n=5):
        super().__init__()
        self.convolutions = nn.ModuleList()
        chans = [n_mel] + [dim] * (n - 1) + [n_mel]
        for i in range(n):
            self.convolutions.append(nn.Sequential(_ConvWrap(chans[i], chans[i + 1], k), nn.BatchNorm1d(chans[i + 1])))


def _owner_ctx(module):
    return module._masg_owner()._context()


class SpeakerMetaEncoder(nn.Module):
    """TacoSpawn metadata → GMM (model/fastspeech2.py:306-341).  ``forward`` returns the same
    ``MixtureSameFamily(Categorical(pi), Independent(Normal(mu, sigma), 1))`` object; the three tiny
    linears + softmax/softplus run in one CUDA kernel (masg_gmm_params)."""

    def __init__(self, preprocess_config, model_config):
        super().__init__()
        self.model_config = model_config
        self.metadata_list = preprocess_config["speaker_generation"]["metadata"]
        self.input_dim = sum(len(x) for x in self.metadata_list.values())
        self.K = model_config["speaker_generation"]["GMM_mixtures"]
        self.D = model_config["transformer"]["encoder_hidden"]
        self.pi_linear = nn.Sequential(nn.Linear(self.input_dim, self.K), nn.Softmax(dim=1))
        self.sigma_linear = nn.Sequential(nn.Linear(self.input_dim, self.K * self.D), nn.Softplus())
        self.mu_linear = nn.Linear(self.input_dim, self.K * self.D)
        self._masg_owner = None

    def params(self, meta):
        ctx = _owner_ctx(self)
        meta = meta.to(device=ctx.device, dtype=torch.float32).reshape(-1, self.input_dim).contiguous()
        n = meta.shape[0]
        pi = torch.empty(n, self.K, device=ctx.device)
        mu = torch.empty(n, self.K, self.D, device=ctx.device)
        sigma = torch.empty(n, self.K, self.D, device=ctx.device)
        ctx.check(ctx.lib.masg_gmm_params(ctx.h, ctx.stream(), _lib.ptr(meta), n, _lib.ptr(pi), _lib.ptr(mu), _lib.ptr(sigma)),
                  "masg_gmm_params")
        return pi, mu, sigma

    def forward(self, input):
        pi, mu, sigma = self.params(input)
        return MixtureSameFamily(Categorical(pi), Independent(Normal(loc=mu, scale=sigma), 1))


class FastSpeech2(nn.Module):
    """FastSpeech2 — B200-native inference drop-in (reference: model/fastspeech2.py:15)."""

    def __init__(self, preprocess_config, model_config, config_path, precision=None):
        super().__init__()
        self.model_config = model_config
        self.preprocess_config = preprocess_config
        tr = model_config["transformer"]
        unsupported = []
        if tr["encoder_hidden"] != 256 or tr["decoder_hidden"] != 256: unsupported.append("hidden != 256")
        if tr["encoder_head"] != 2 or tr["decoder_head"] != 2: unsupported.append("heads != 2")
        if tr["conv_filter_size"] != 1024: unsupported.append("conv_filter_size != 1024")
        if model_config["variance_predictor"]["filter_size"] != 256 or model_config["variance_predictor"]["kernel_size"] != 3:
            unsupported.append("variance predictor shape")
        if model_config["variance_embedding"]["n_bins"] != 256: unsupported.append("n_bins != 256")
        if model_config["jdit"]["use_jdit"]: unsupported.append("use_jdit")
        if preprocess_config["pitch"]["feature"] != "phoneme_level" or preprocess_config["energy"]["feature"] != "phoneme_level":
            unsupported.append("frame-level pitch/energy")
        if preprocess_config["mel"]["n_mel_channels"] != 80: unsupported.append("n_mel != 80")
        if unsupported:
            raise NotImplementedError("configuration not supported by the sm_100a kernels (no fallback): " + ", ".join(unsupported))
        self.use_jdit = False
        self.encoder = _Coder(model_config, "encoder", True)
        self.variance_adaptor = _VarianceAdaptor(preprocess_config, model_config, config_path)
        self.decoder = _Coder(model_config, "decoder", False)
        self.mel_linear = nn.Linear(tr["decoder_hidden"], preprocess_config["mel"]["n_mel_channels"])
        self.postnet = _PostNet()
        self.speaker_enc = None
        self.speaker_emb = None
        if model_config["multi_speaker"]:
            with open(os.path.join(config_path, "speakers.json"), "r") as f:       # fastspeech2.py:39-45
                n_speaker = len(json.load(f))
            self.speaker_emb = nn.Embedding(n_speaker, tr["encoder_hidden"])
            self.speaker_enc = SpeakerMetaEncoder(preprocess_config, model_config)
            owner = self
            self.speaker_enc._masg_owner = lambda: owner
        self._precision = precision or os.environ.get("MASG_PRECISION", "bf16")
        self._ctx = None
        self._dirty = True

    # ---- weight lifecycle: any change of storage/values marks the repacked copies stale ----
    def _apply(self, fn, *a, **k):
        self._dirty = True
        return super()._apply(fn, *a, **k)

    def load_state_dict(self, *a, **k):
        self._dirty = True
        return super().load_state_dict(*a, **k)

    def set_precision(self, precision):
        if precision != self._precision:
            self._precision = precision
            self._ctx = None
            self._dirty = True

    def _context(self):
        dev = self.mel_linear.weight.device
        if dev.type != "cuda":
            raise RuntimeError("FastSpeech2 (B200-native) requires its parameters on a CUDA device; got %s. "
                               "There is no CPU fallback." % dev)
        if self._ctx is None or self._ctx.device != dev:
            self._ctx = _lib.Context(dev, self._precision)
            self._dirty = True
        if self._dirty:
            if self.training:
                raise RuntimeError("training mode is not supported (inference-only path); call .eval()")
            self._ctx.load_state("fs2", self.state_dict())
            self._dirty = False
        return self._ctx

    # ---- the hot path -----------------------------------------------------------------------------
    def _synth(self, texts, src_lens, max_src_len, speaker_emb, accents, p_control, e_control, d_control, d_targets,
               max_mel_len, mels, p_targets, e_targets):
        if mels is not None:
            raise NotImplementedError("ground-truth mels are a training-path input (out of scope)")
        if self.training:
            raise RuntimeError("training mode is not supported (inference-only path); call .eval()")
        ctx = self._context()
        dev = ctx.device
        if texts.device != dev:
            raise RuntimeError("texts is on %s but the model is on %s (no CPU fallback)" % (texts.device, dev))
        if accents is None:
            raise TypeError("accents is required (transformer/Models.py:101 embeds it unconditionally)")
        B, L = texts.shape
        if int(max_src_len) != L:
            raise ValueError("max_src_len (%d) must equal texts.shape[1] (%d)" % (int(max_src_len), L))
        texts = texts.to(torch.int64).contiguous()
        accents = accents.to(device=dev, dtype=torch.int64).contiguous()
        acc1d = 1 if accents.dim() == 1 else 0
        src_lens_in = src_lens
        lens = src_lens.to(device=dev, dtype=torch.int64).contiguous()
        spk = speaker_emb.to(device=dev, dtype=torch.float32).contiguous()
        if spk.shape != (B, 256):
            raise ValueError("speaker embedding must be [B,256], got %s" % (tuple(spk.shape),))
        f32 = dict(dtype=torch.float32, device=dev)
        pitch = torch.empty(B, L, **f32); energy = torch.empty(B, L, **f32)
        log_d = torch.empty(B, L, **f32); d_rounded = torch.empty(B, L, **f32)
        mel_lens = torch.empty(B, dtype=torch.int64, device=dev)
        src_mask = torch.empty(B, L, dtype=torch.uint8, device=dev)
        d_t = p_t = e_t = None
        if d_targets is not None:
            d_t = d_targets.to(**f32).contiguous()
        if p_targets is not None:                     # teacher-forced pitch / energy (model/modules.py:80-100)
            p_t = p_targets.to(**f32).contiguous()
        if e_targets is not None:
            e_t = e_targets.to(**f32).contiguous()
        for t_ in (d_t, p_t, e_t):
            if t_ is not None and tuple(t_.shape) != (B, L):
                raise ValueError("duration / pitch / energy targets must be [B, L] (phoneme level)")
        tmax = ctypes.c_int64(0)
        ctx.check(ctx.lib.masg_fs2_phase_a(
            ctx.h, ctx.stream(), _lib.ptr(texts), _lib.ptr(accents), acc1d, _lib.ptr(lens), B, L, _lib.ptr(spk),
            float(p_control), float(e_control), float(d_control), _lib.ptr(d_t), _lib.ptr(p_t), _lib.ptr(e_t), _lib.ptr(pitch), _lib.ptr(energy),
            _lib.ptr(log_d), _lib.ptr(d_rounded), _lib.ptr(mel_lens), _lib.ptr(src_mask), ctypes.byref(tmax)), "masg_fs2_phase_a")
        T = int(max_mel_len) if max_mel_len is not None else int(tmax.value)
        if T <= 0:
            raise RuntimeError("all predicted durations are zero: nothing to synthesise")
        mel = torch.empty(B, T, 80, **f32); post = torch.empty(B, T, 80, **f32)
        mel_mask = torch.empty(B, T, dtype=torch.uint8, device=dev)
        ctx.check(ctx.lib.masg_fs2_phase_b(ctx.h, ctx.stream(), T, _lib.ptr(mel), _lib.ptr(post), _lib.ptr(mel_mask)), "masg_fs2_phase_b")
        if d_targets is not None:
            d_rounded = d_targets
        return (mel, post, pitch, energy, log_d, d_rounded, src_mask.view(torch.bool), mel_mask.view(torch.bool),
                src_lens_in, mel_lens)

    def forward(self, speakers, texts, src_lens, max_src_len, mels=None, mel_lens=None, max_mel_len=None, p_targets=None,
                e_targets=None, d_targets=None, p_control=1.0, e_control=1.0, d_control=1.0, accents=None, speaker_meta=None):
        """model/fastspeech2.py:52-174 (inference branch).  Returns the 10-tuple, plus
        (speaker_emb_p, speaker_emb_s) when multi_speaker."""
        if self.speaker_emb is None:
            raise NotImplementedError("multi_speaker=False is not supported by this path")
        with torch.no_grad():
            speaker_emb_s = self.speaker_emb(speakers)
            out = self._synth(texts, src_lens, max_src_len, speaker_emb_s, accents, p_control, e_control, d_control,
                              d_targets, max_mel_len, mels, p_targets, e_targets)
            speaker_emb_p = self.speaker_enc(speaker_meta) if speaker_meta is not None else None
        return out + (speaker_emb_p, speaker_emb_s)

    def synthesize_from_speaker_emb(self, speakers, texts, src_lens, max_src_len, mels=None, mel_lens=None, max_mel_len=None,
                                    p_targets=None, e_targets=None, d_targets=None, p_control=1.0, e_control=1.0,
                                    d_control=1.0, accents=None, speaker_emb=None):
        """model/fastspeech2.py:187-303.  ``speakers`` is ignored (callers pass raw_texts there)."""
        with torch.no_grad():
            return self._synth(texts, src_lens, max_src_len, speaker_emb, accents, p_control, e_control, d_control,
                               d_targets, max_mel_len, mels, p_targets, e_targets)

    def speaker_distribution(self, speaker_meta):
        """model/fastspeech2.py:182-185."""
        with torch.no_grad():
            return self.speaker_enc(speaker_meta)

    def speaker_gen(self, speaker_meta, seed=None):
        """model/fastspeech2.py:176-180: one sample per metadata row, [N,256].  Fused Philox
        categorical+normal kernel; the seed is drawn from torch's default generator unless given, so
        torch.manual_seed() makes it reproducible."""
        with torch.no_grad():
            pi, mu, sigma = self.speaker_enc.params(speaker_meta)
            return gmm_sample(self._context(), pi, mu, sigma, pi.shape[0], seed=seed)


def gmm_sample(ctx, pi, mu, scale, n, seed=None, offset=0, k_inject=None, eps_inject=None, return_k=False):
    """MixtureSameFamily.sample() on the device (masg_gmm_sample).  pi [Np,K], mu/scale [Np,K,D], Np ∈ {1, n}."""
    Np, K, D = mu.shape
    if seed is None:
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())
    out = torch.empty(n, D, dtype=torch.float32, device=ctx.device)
    k_out = torch.empty(n, dtype=torch.int32, device=ctx.device)
    pi = pi.to(device=ctx.device, dtype=torch.float32).contiguous()
    mu = mu.to(device=ctx.device, dtype=torch.float32).contiguous()
    scale = scale.to(device=ctx.device, dtype=torch.float32).contiguous()
    if k_inject is not None:
        k_inject = k_inject.to(device=ctx.device, dtype=torch.int32).contiguous()
    if eps_inject is not None:
        eps_inject = eps_inject.to(device=ctx.device, dtype=torch.float32).contiguous()
    ctx.check(ctx.lib.masg_gmm_sample(ctx.h, ctx.stream(), _lib.ptr(pi), _lib.ptr(mu), _lib.ptr(scale), n, Np, K, D,
                                      ctypes.c_uint64(seed), ctypes.c_uint64(offset), _lib.ptr(k_inject), _lib.ptr(eps_inject),
                                      _lib.ptr(out), _lib.ptr(k_out)), "masg_gmm_sample")
    return (out, k_out) if return_k else out
