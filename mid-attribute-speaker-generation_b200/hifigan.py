"""Drop-in for the reference ``hifigan.Generator`` / ``hifigan.AttrDict`` (hifigan/models.py:112-174,
hifigan/__init__.py:4-6) and ``utils.model.vocoder_infer`` (utils/model.py:74-92).

The call order of utils/model.py:61-69 works unchanged: ``Generator(h)`` → ``load_state_dict(ckpt["generator"])``
(weight-normed keys ``weight_g``/``weight_v``/``bias``, 234 entries) → ``eval()`` →
``remove_weight_norm()`` (156 entries) → ``to(device)`` → ``forward(mel[B,80,T]) → wav[B,1,256T]``.
The sub-modules are parameter containers; the arithmetic runs in libmasg.so (include/masg.h).
"""
import os

import torch
import torch.nn as nn

from . import _lib
from .synthetic import HIFIGAN_CONFIG, hifigan_layers


class AttrDict(dict):
    def __init__(self, *args, **kwargs):
        super(AttrDict, self).__init__(*args, **kwargs)
        self.__dict__ = self


class _WNConv(nn.Module):
    """Parameter container with torch weight_norm naming (dim=0): weight_g [n0,1,1], weight_v [n0,n1,k]."""

    def __init__(self, shape, n_bias):
        super().__init__()
        self.bias = nn.Parameter(torch.zeros(n_bias))
        v = torch.randn(shape) * 0.01
        self.weight_g = nn.Parameter(v.flatten(1).norm(dim=1).view(-1, 1, 1))
        self.weight_v = nn.Parameter(v)

    def remove_weight_norm(self):
        if hasattr(self, "weight_v"):
            v, g = self.weight_v.data, self.weight_g.data
            w = v * (g / v.flatten(1).norm(dim=1).view(-1, 1, 1))
            del self.weight_g
            del self.weight_v
            self.weight = nn.Parameter(w)


class _ResBlock(nn.Module):
    def __init__(self, ch, k):
        super().__init__()
        self.convs1 = nn.ModuleList([_WNConv((ch, ch, k), ch) for _ in range(3)])
        self.convs2 = nn.ModuleList([_WNConv((ch, ch, k), ch) for _ in range(3)])


class Generator(nn.Module):
    """HiFi-GAN V1 generator — B200-native inference drop-in (reference: hifigan/models.py:112)."""

    def __init__(self, h, precision=None):
        super().__init__()
        self.h = h
        want = HIFIGAN_CONFIG
        for key in ("upsample_rates", "upsample_kernel_sizes", "upsample_initial_channel", "resblock_kernel_sizes",
                    "resblock_dilation_sizes"):
            got = h[key] if isinstance(h, dict) else getattr(h, key)
            if list(got) != list(want[key]) if isinstance(want[key], list) else got != want[key]:
                raise NotImplementedError("HiFi-GAN config %s=%r is not supported by the sm_100a kernels (V1 only, no fallback)" % (key, got))
        self.num_kernels = 3
        self.num_upsamples = 4
        self.conv_pre = _WNConv((512, 80, 7), 512)
        self.ups = nn.ModuleList()
        self.resblocks = nn.ModuleList()
        for name, kind, cin, cout, k, stride, dil in hifigan_layers(want):
            if kind == "convT":
                self.ups.append(_WNConv((cin, cout, k), cout))
                for kk in want["resblock_kernel_sizes"]:
                    self.resblocks.append(_ResBlock(cout, kk))
        self.conv_post = _WNConv((1, 32, 7), 1)
        self._precision = precision or os.environ.get("MASG_PRECISION", "bf16")
        self._ctx = None
        self._dirty = True

    def _apply(self, fn, *a, **k):
        self._dirty = True
        return super()._apply(fn, *a, **k)

    def load_state_dict(self, *a, **k):
        self._dirty = True
        return super().load_state_dict(*a, **k)

    def remove_weight_norm(self):
        print("Removing weight norm...")                       # hifigan/models.py:168
        for m in self.modules():
            if isinstance(m, _WNConv):
                m.remove_weight_norm()
        self._dirty = True

    def set_precision(self, precision):
        if precision != self._precision:
            self._precision, self._ctx, self._dirty = precision, None, True

    def _context(self):
        dev = self.conv_pre.bias.device
        if dev.type != "cuda":
            raise RuntimeError("Generator (B200-native) requires its parameters on a CUDA device; got %s. "
                               "There is no CPU fallback." % dev)
        if self._ctx is None or self._ctx.device != dev:
            self._ctx = _lib.Context(dev, self._precision)
            self._dirty = True
        if self._dirty:
            self._ctx.load_state("hifigan", self.state_dict())
            self._dirty = False
        return self._ctx

    def forward(self, x, channels_last=False, return_int16=False):
        """x: mel [B,80,T] (or [B,T,80] with channels_last=True) → wav fp32 [B,1,256·T]."""
        ctx = self._context()
        if x.device != ctx.device:
            raise RuntimeError("mel is on %s but the vocoder is on %s (no CPU fallback)" % (x.device, ctx.device))
        x = x.detach().to(torch.float32).contiguous()
        if channels_last:
            B, T, C = x.shape
        else:
            B, C, T = x.shape
        if C != 80:
            raise ValueError("expected 80 mel channels, got %d" % C)
        wav = torch.empty(B, 1, T * 256, dtype=torch.float32, device=ctx.device)
        w16 = torch.empty(B, T * 256, dtype=torch.int16, device=ctx.device) if return_int16 else None
        ctx.check(ctx.lib.masg_hifigan_forward(ctx.h, ctx.stream(), _lib.ptr(x), 1 if channels_last else 0, B, T,
                                               _lib.ptr(wav), _lib.ptr(w16)), "masg_hifigan_forward")
        return (wav, w16) if return_int16 else wav


def vocoder_infer(mels, vocoder, model_config, preprocess_config, lengths=None):
    """utils/model.py:74-92 for the HiFi-GAN branch: mel [B,80,T] → list of int16 numpy arrays cut to
    ``lengths``.  The ×32768 and the C-cast int16 conversion are fused into the conv_post kernel, so the
    D2H copy moves 2 bytes per sample instead of 4."""
    name = model_config["vocoder"]["model"]
    if name != "HiFi-GAN":
        raise NotImplementedError("only the HiFi-GAN vocoder is on the B200 path")
    if float(preprocess_config["audio"]["max_wav_value"]) != 32768.0:
        raise NotImplementedError("max_wav_value must be 32768.0")
    with torch.no_grad():
        _, w16 = vocoder(mels, return_int16=True)
    wavs = [w for w in w16.cpu().numpy()]
    for i in range(len(mels)):
        if lengths is not None:
            wavs[i] = wavs[i][: lengths[i]]
    return wavs
