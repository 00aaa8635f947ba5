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

    def _prep(self, x, channels_last):
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
        return ctx, x, B, T

    def forward(self, x, channels_last=False, return_int16=False, int16_only=False):
        """x: mel [B,80,T] (or [B,T,80] with channels_last=True) → wav fp32 [B,1,256·T].
        return_int16: also the int16 waveform [B, 256·T] (utils/model.py:82-85, fused into the last kernel);
        int16_only: skip the fp32 waveform altogether (4 of the last kernel's 70 bytes per sample) and return only int16."""
        ctx, x, B, T = self._prep(x, channels_last)
        want16 = return_int16 or int16_only
        wav = None if int16_only else torch.empty(B, 1, T * 256, dtype=torch.float32, device=ctx.device)
        w16 = torch.empty(B, T * 256, dtype=torch.int16, device=ctx.device) if want16 else None
        ctx.check(ctx.lib.masg_hifigan_forward(ctx.h, ctx.stream(), _lib.ptr(x), 1 if channels_last else 0, B, T,
                                               _lib.ptr(wav), _lib.ptr(w16)), "masg_hifigan_forward")
        if int16_only:
            return w16
        return (wav, w16) if return_int16 else wav

    def forward_packed(self, x, mel_lens, channels_last=False, hop=256):
        """Vocoder + int16 cast + per-utterance trim in one pass (utils/model.py:74-92 without its host work): returns
        (packed int16 [sum(mel_lens)·hop] on the device, offsets int64 [B+1] on the device); utterance b is
        packed[offsets[b]:offsets[b+1]].  ``mel_lens``: int64 [B] tensor (device or host)."""
        ctx, x, B, T = self._prep(x, channels_last)
        lens = torch.as_tensor(mel_lens).to(device=ctx.device, dtype=torch.int64).contiguous()
        if lens.shape != (B,):
            raise ValueError("mel_lens must be [B]")
        packed = torch.empty(B * T * hop, dtype=torch.int16, device=ctx.device)          # capacity: the padded size bounds the total
        offsets = torch.empty(B + 1, dtype=torch.int64, device=ctx.device)
        ctx.check(ctx.lib.masg_hifigan_forward_packed(ctx.h, ctx.stream(), _lib.ptr(x), 1 if channels_last else 0, B, T, _lib.ptr(lens),
                                                      int(hop), _lib.ptr(packed), _lib.ptr(offsets)), "masg_hifigan_forward_packed")
        return packed, offsets


def vocoder_infer(mels, vocoder, model_config, preprocess_config, lengths=None):
    """utils/model.py:74-92 for the HiFi-GAN branch: mel [B,80,T] → list of int16 numpy arrays cut to
    ``lengths``.  The ×32768, the C-cast int16 conversion AND the trim are fused into the conv_post kernel (packed
    output), so the D2H copy moves 2 bytes per VALID sample instead of 4 per padded one."""
    name = model_config["vocoder"]["model"]
    if name != "HiFi-GAN":
        raise NotImplementedError("only the HiFi-GAN vocoder is on the B200 path")
    if float(preprocess_config["audio"]["max_wav_value"]) != 32768.0:
        raise NotImplementedError("max_wav_value must be 32768.0")
    hop = int(preprocess_config["stft"]["hop_length"]) if "stft" in preprocess_config else 256
    with torch.no_grad():
        if lengths is None:
            return [w for w in vocoder(mels, int16_only=True).cpu().numpy()]
        # trimmed on the device: one D2H copy of exactly the valid samples, then zero-copy views per utterance.  The reference
        # slices to `lengths` samples (utils/model.py:88-90); callers pass mel_len * hop, which is what the packed kernel keeps.
        lens = torch.as_tensor(lengths).to(torch.int64).cpu()
        T = mels.shape[2]
        if bool((lens % hop != 0).any()) or bool((lens > T * hop).any()):
            w16 = vocoder(mels, int16_only=True).cpu().numpy()           # arbitrary sample counts: slice on the host
            return [w16[i][: int(lens[i])] for i in range(len(w16))]
        packed, offsets = vocoder.forward_packed(mels, lens // hop, hop=hop)
        off = offsets.cpu().numpy()
        flat = packed[: int(off[-1])].cpu().numpy()
    return [flat[off[i]: off[i + 1]] for i in range(len(off) - 1)]
