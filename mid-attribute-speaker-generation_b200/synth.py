"""Batched synthesis driver: the vectorised replacement for the per-speaker batch-1 loops of
examples_gen.py:247-273 / examples_gen_distri.py:209-218, and the batch scheduler that shards
utterances over ranks.  Host logic only; all arithmetic is in libmasg.so.
"""
import ctypes
import heapq

import numpy as np
import torch

from . import _lib
from .synthetic import HOP, SAMPLING_RATE, META_COMBOS


class Synthesizer:
    """FastSpeech2 + HiFi-GAN sharing one libmasg context (one per GPU / process)."""

    def __init__(self, fs2, vocoder, precision="bf16"):
        dev = next(fs2.parameters()).device
        self.device = dev
        self.ctx = _lib.Context(dev, precision)
        for m in (fs2, vocoder):
            m._precision = precision
            m._ctx = self.ctx
            m._dirty = True
        self.fs2, self.vocoder = fs2, vocoder
        fs2._context()
        vocoder._context()
        self._pin = {}

    def synthesize(self, texts, accents, src_lens, speaker_emb, p_control=1.0, e_control=1.0, d_control=1.0):
        """Device tensors in → (wav int16 [B, 256·T] on device, mel_lens int64 [B], postnet mel [B,T,80])."""
        out = self.fs2.synthesize_from_speaker_emb(None, texts, src_lens, texts.shape[1], p_control=p_control,
                                                   e_control=e_control, d_control=d_control, accents=accents,
                                                   speaker_emb=speaker_emb)
        _, w16 = self.vocoder(out[1], channels_last=True, return_int16=True)
        return w16, out[9], out[1]

    def _pinned(self, name, shape, dtype):
        t = self._pin.get(name)
        if t is None or t.shape != tuple(shape) or t.dtype != dtype:
            t = torch.empty(shape, dtype=dtype).pin_memory()
            self._pin[name] = t
        return t

    def synthesize_host(self, texts, accents, src_lens, speaker_emb, t_cap, p_control=1.0, e_control=1.0, d_control=1.0):
        """End-to-end through the C ABI with HOST buffers (masg_synthesize_host): pinned ids/lengths/embeddings are
        copied H2D, FastSpeech2 + HiFi-GAN run, int16 waveforms and mel lengths come back D2H — all inside the call.
        Returns (wav int16 [B, 256·t_cap] pinned host tensor, mel_lens int64 [B] host, T_max, h2d_bytes, d2h_bytes)."""
        B, L = texts.shape
        h_t = self._pinned("texts", (B, L), torch.int64); h_t.copy_(texts)
        h_a = self._pinned("accents", (B, L), torch.int64); h_a.copy_(accents)
        h_l = self._pinned("lens", (B,), torch.int64); h_l.copy_(src_lens)
        h_s = self._pinned("spk", (B, 256), torch.float32); h_s.copy_(speaker_emb)
        h_w = self._pinned("wav", (B, 256 * t_cap), torch.int16)
        h_m = self._pinned("mel_lens", (B,), torch.int64)
        tmax = ctypes.c_int64(0)
        ctx = self.ctx
        ctx.check(ctx.lib.masg_synthesize_host(ctx.h, ctx.stream(), _lib.ptr(h_t), _lib.ptr(h_a), _lib.ptr(h_l), B, L,
                                               _lib.ptr(h_s), float(p_control), float(e_control), float(d_control), int(t_cap),
                                               _lib.ptr(h_w), _lib.ptr(h_m), ctypes.byref(tmax)), "masg_synthesize_host")
        T = int(tmax.value)
        h2d = h_t.numel() * 8 + h_a.numel() * 8 + h_l.numel() * 8 + h_s.numel() * 4
        d2h = B * T * 256 * 2 + B * 8
        return h_w, h_m, T, h2d, d2h


def sample_speakers(fs2, n_sampled, n_interp, seed=0):
    """Config-3 speaker set: n_sampled embeddings from speaker_gen (cycling the four
    metadata combos) + n_interp from InterpolateGMM(a, b) over the 6 unordered combo pairs with
    t ∈ {0.25, 0.5, 0.75}.  Returns [n_sampled + n_interp, 256] on the model's device."""
    from .distributions import InterpolateGMM
    dev = next(fs2.parameters()).device
    g = torch.Generator().manual_seed(seed)
    metas = torch.tensor([META_COMBOS[i % 4] for i in range(n_sampled)], dtype=torch.float32, device=dev)
    embs = [fs2.speaker_gen(metas, seed=int(torch.randint(0, 2 ** 62, (1,), generator=g)))] if n_sampled else []
    if n_interp:
        combos = torch.tensor(META_COMBOS, dtype=torch.float32, device=dev)
        dists = [fs2.speaker_distribution(combos[i:i + 1]) for i in range(4)]
        pairs = [(a, b) for a in range(4) for b in range(a + 1, 4)]
        cache = {}
        rates = (0.25, 0.5, 0.75)
        ctx = fs2._context()
        from .fastspeech2 import gmm_sample
        for i in range(n_interp):
            a, b = pairs[i % len(pairs)]
            t = rates[(i // len(pairs)) % len(rates)]
            if (a, b) not in cache:
                cache[(a, b)] = InterpolateGMM(dists[a], dists[b])
            ig = cache[(a, b)]
            ig.interpolate_rate(t)
            comp = ig.component_distribution.base_dist
            embs.append(gmm_sample(ctx, ig.mixture_distribution.probs.reshape(1, -1), comp.loc.reshape(1, -1, 256),
                                   comp.scale.reshape(1, -1, 256), 1, seed=int(torch.randint(0, 2 ** 62, (1,), generator=g))))
    return torch.cat(embs, 0)


# ------------------------------------------------------------------------------------------------
# on-disk speaker artefacts, interchangeable with the reference's
# ------------------------------------------------------------------------------------------------
def speaker_names(gender, language, n, start=1):
    """The reference's naming of generated speakers: '<gender>_<language>gen<NNN>' (examples_gen.py:267,
    examples_gen_distri.py:214)."""
    return ["%s_%sgen%03d" % (gender, language, i) for i in range(start, start + n)]


def save_speaker_embeddings(out_dir, names, embs):
    """One `<name>.npy` per speaker holding a float32 [1, 256] array — exactly what `np.save(output_path / speaker_name,
    speaker_emb)` leaves behind (examples_gen.py:272, examples_gen_distri.py:218), written from one batched [N, 256] tensor."""
    import os
    embs = embs.detach().to("cpu", torch.float32).numpy() if torch.is_tensor(embs) else np.asarray(embs, dtype=np.float32)
    assert embs.ndim == 2 and embs.shape[0] == len(names), "embs must be [N, D] with one name per row"
    os.makedirs(out_dir, exist_ok=True)
    paths = []
    for name, e in zip(names, embs):
        path = os.path.join(str(out_dir), name + ".npy")
        np.save(path, e[None, :].copy())
        paths.append(path)
    return paths


def load_speaker_embeddings(paths, device=None):
    """Stack reference-format `.npy` speaker embeddings ([1, D] or [D] each) into one [N, D] float32 tensor, ready for
    `synthesize_from_speaker_emb(..., speaker_emb=...)` (the reference feeds them one by one, synthesize.py:105-131)."""
    rows = []
    for p in paths:
        a = np.load(str(p)).astype(np.float32)
        assert a.ndim in (1, 2) and (a.ndim == 1 or a.shape[0] == 1), "%s: expected a [1, D] or [D] array, got %s" % (p, a.shape)
        rows.append(a.reshape(-1))
    assert rows and all(r.shape == rows[0].shape for r in rows), "speaker embeddings differ in size"
    t = torch.from_numpy(np.stack(rows, 0))
    return t.to(device) if device is not None else t


# ------------------------------------------------------------------------------------------------
# batch scheduler (host logic, CPU-testable): length-bucketed batches, LPT assignment to ranks
# ------------------------------------------------------------------------------------------------
def utterance_cost(L, frames_per_phoneme=3.0):
    """Algorithmic FLOPs of one utterance: linear terms + attention."""
    T = L * frames_per_phoneme
    return L * 25.43e6 + T * (43.33e6 + 614.11e6) + 1024.0 * (4 * L * L + 6 * T * T)


def plan_batches(lengths, frame_budget=40000, max_batch=64, frames_per_phoneme=3.0):
    """Sort by length, cut into batches whose *padded* frame count (B · T_max) stays under the budget.
    Returns a list of index arrays (into ``lengths``)."""
    lengths = np.asarray(lengths)
    order = np.argsort(-lengths, kind="stable")
    batches, cur = [], []
    for i in order:
        tmax = lengths[cur[0]] * frames_per_phoneme if cur else lengths[i] * frames_per_phoneme
        if cur and ((len(cur) + 1) * tmax > frame_budget or len(cur) >= max_batch):
            batches.append(np.array(cur))
            cur = []
        cur.append(int(i))
    if cur:
        batches.append(np.array(cur))
    return batches


def assign_batches(batches, lengths, world_size, frames_per_phoneme=3.0):
    """Longest-processing-time-first assignment of batches to ranks; returns per-rank lists of batch
    indices.  Deterministic (ties → lower rank), so every rank derives the same plan with no
    communication."""
    lengths = np.asarray(lengths)
    cost = [sum(utterance_cost(lengths[i], frames_per_phoneme) for i in b) for b in batches]
    heap = [(0.0, r) for r in range(world_size)]
    heapq.heapify(heap)
    out = [[] for _ in range(world_size)]
    for bi in sorted(range(len(batches)), key=lambda k: (-cost[k], k)):
        load, r = heapq.heappop(heap)
        out[r].append(bi)
        heapq.heappush(heap, (load + cost[bi], r))
    return out


def shard_batch(B, rank, world_size):
    """Contiguous split of a batch dimension (configs 3/5: 64 → 8×8, 16 → 8×2)."""
    per = (B + world_size - 1) // world_size
    lo = min(rank * per, B)
    return lo, min(lo + per, B)


def audio_seconds(mel_lens):
    return float(np.sum(np.asarray(mel_lens, dtype=np.float64)) * HOP / SAMPLING_RATE)
