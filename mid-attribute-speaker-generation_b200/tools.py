"""Drop-ins for the host-side helpers around the hot path (utils/tools.py), de-hosted.

``synth_samples`` keeps the reference's signature (utils/tools.py:228) and file outputs, but where the reference does, PER
UTTERANCE, two ``.item()`` syncs (:233-234), three device-to-host slices (:235-246), a ``json.load(stats.json)`` (:248-252), a
matplotlib figure (:254-262) and a synchronous ``wavfile.write`` (:272-274), this version does per CALL: one batched D2H of the
lengths (and of pitch / energy / durations / mel only when plotting is asked for), one read of stats.json, one vocoder pass whose
last kernel already emits trimmed int16 audio (packed), one D2H of exactly the valid samples, and hands the files to a background
writer.  Plotting is opt-in (``plot=True`` with a ``plot_fn`` such as the reference's own ``plot_mel``): matplotlib is not part of
the synthesis path (and is not installed in this image).
"""
import json
import os

import numpy as np
import torch

from .wavio import WavWriter


def expand(values, durations):
    """utils/tools.py:166-170 — repeat values[i] max(0, int(durations[i])) times (int() truncates toward zero) — vectorised."""
    values = np.asarray(values)
    reps = np.maximum(np.trunc(np.asarray(durations, dtype=np.float64)).astype(np.int64), 0)
    n = min(len(values), len(reps))                      # zip() semantics
    return np.repeat(values[:n], reps[:n])


def get_mask_from_lengths(lengths, max_len=None):
    """utils/tools.py:155-163 (True marks PAD) on the tensor's own device (the reference moves it to a module-level global)."""
    if max_len is None:
        max_len = int(torch.max(lengths).item())
    ids = torch.arange(0, int(max_len), device=lengths.device)
    return ids[None, :] >= lengths[:, None]


def synth_samples(args, targets, predictions, vocoder, model_config, preprocess_config, path, basename=None, plot=False,
                  plot_fn=None, writer=None):
    """utils/tools.py:228-274 (``basename=None``: one file pair per ``targets[0][i]``) and examples_gen.py:68-109
    (``basename`` given: that name for every utterance, as the script does for its batch-1 calls).

    predictions: the 10/12-tuple of ``FastSpeech2.forward`` / ``synthesize_from_speaker_emb``.  vocoder: the drop-in
    ``hifigan.Generator`` (anything with ``forward_packed(mel[B,T,80], mel_lens, channels_last=True)``).
    Writes ``<path>/<basename>.wav`` (int16, byte-identical container to scipy's) and, with ``plot=True``,
    ``<path>/<basename>.png`` through ``plot_fn(data, stats, titles)`` — called with exactly the arrays the reference hands to
    ``plot_mel``.  Returns the list of wav paths (written by the time the call returns, unless the caller owns ``writer``)."""
    if model_config["vocoder"]["model"] != "HiFi-GAN":
        raise NotImplementedError("only the HiFi-GAN vocoder is on the B200 path")
    if float(preprocess_config["audio"]["max_wav_value"]) != 32768.0:
        raise NotImplementedError("max_wav_value must be 32768.0")
    n = len(predictions[0])
    names = [basename] * n if basename is not None else list(targets[0])
    hop = int(preprocess_config["stft"]["hop_length"])
    rate = int(preprocess_config["audio"]["sampling_rate"])
    mel_lens_dev = predictions[9]
    # ---- one vocoder pass; the trim to mel_len * hop (utils/model.py:88-90) happens in its last kernel ----
    with torch.no_grad():
        packed, offsets = vocoder.forward_packed(predictions[1], mel_lens_dev, channels_last=True, hop=hop)
    # ---- one batched D2H of everything the host needs ----
    off = offsets.cpu().numpy()
    flat = packed[: int(off[-1])].cpu().numpy()
    if plot:
        if plot_fn is None:
            raise ValueError("plot=True needs plot_fn (e.g. the reference's utils.tools.plot_mel): plotting is opt-in")
        src_lens = torch.as_tensor(predictions[8]).cpu().numpy()
        mel_lens = mel_lens_dev.cpu().numpy()
        mel = predictions[1].detach().cpu().numpy()
        pitch_all, energy_all, dur_all = (predictions[i].detach().cpu().numpy() for i in (2, 3, 5))
        with open(os.path.join(args.config, "stats.json")) as f:          # once per call (the reference re-reads it per utterance)
            stats = json.load(f)
            stats = stats["pitch"] + stats["energy"][:2]
        for i in range(n):
            sl, ml = int(src_lens[i]), int(mel_lens[i])
            duration = dur_all[i, :sl]
            if preprocess_config["pitch"]["feature"] == "phoneme_level":
                pitch = expand(pitch_all[i, :sl], duration)
            else:
                pitch = pitch_all[i, :ml]
            if preprocess_config["energy"]["feature"] == "phoneme_level":
                energy = expand(energy_all[i, :sl], duration)
            else:
                energy = energy_all[i, :ml]
            fig = plot_fn([(mel[i, :ml].T, pitch, energy)], stats, ["Synthetized Spectrogram"])
            if fig is not None and hasattr(fig, "savefig"):
                fig.savefig(os.path.join(path, "{}.png".format(names[i])))
    # ---- wav emission off this thread ----
    own = writer is None
    w = WavWriter() if own else writer
    paths = []
    try:
        for i in range(n):
            p = os.path.join(path, "{}.wav".format(names[i]))
            w.submit(p, rate, flat[off[i]: off[i + 1]], copy=False)       # `flat` is ours: the views stay valid
            paths.append(p)
    finally:
        if own:
            w.close()
    return paths
