"""Batched synthesis driver: the vectorised replacement for the per-speaker batch-1 loops of
examples_gen.py:247-273 / examples_gen_distri.py:209-218, and the batch scheduler that shards
utterances over ranks.  Host logic only; all arithmetic is in libmasg.so.
"""
import ctypes
import heapq
import os

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
        w16 = self.vocoder(out[1], channels_last=True, int16_only=True)
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

    def synthesize_host_packed(self, texts, accents, src_lens, speaker_emb, cap_samples, p_control=1.0, e_control=1.0, d_control=1.0):
        """The same through masg_synthesize_host_packed: the per-utterance trim (utils/model.py:88-90) happens in the vocoder's
        last kernel and ONE D2H copy moves exactly the valid samples.  Returns (packed int16 [cap_samples] pinned host tensor,
        offsets int64 [B+1] host, mel_lens int64 [B] host, T_max, h2d_bytes, d2h_bytes); utterance b is
        packed[offsets[b]:offsets[b+1]]."""
        B, L = texts.shape
        h_t = self._pinned("texts", (B, L), torch.int64); h_t.copy_(texts)
        h_a = self._pinned("accents", (B, L), torch.int64); h_a.copy_(accents)
        h_l = self._pinned("lens", (B,), torch.int64); h_l.copy_(src_lens)
        h_s = self._pinned("spk", (B, 256), torch.float32); h_s.copy_(speaker_emb)
        h_w = self._pinned("wav_packed", (int(cap_samples),), torch.int16)
        h_o = self._pinned("offsets", (B + 1,), torch.int64)
        h_m = self._pinned("mel_lens", (B,), torch.int64)
        tmax = ctypes.c_int64(0)
        ctx = self.ctx
        ctx.check(ctx.lib.masg_synthesize_host_packed(ctx.h, ctx.stream(), _lib.ptr(h_t), _lib.ptr(h_a), _lib.ptr(h_l), B, L,
                                                      _lib.ptr(h_s), float(p_control), float(e_control), float(d_control),
                                                      ctypes.c_int64(int(cap_samples)), _lib.ptr(h_w), _lib.ptr(h_o), _lib.ptr(h_m),
                                                      ctypes.byref(tmax)), "masg_synthesize_host_packed")
        h2d = h_t.numel() * 8 + h_a.numel() * 8 + h_l.numel() * 8 + h_s.numel() * 4
        d2h = int(h_o[B]) * 2 + (B + 1) * 8
        return h_w, h_o, h_m, int(tmax.value), h2d, d2h

    def synthesize_host_packed_async(self, texts, accents, src_lens, speaker_emb, cap_samples, p_control=1.0, e_control=1.0, d_control=1.0):
        """Deferred-copy form (masg_synthesize_host_packed_async): returns a handle as soon as the batch is enqueued; ``handle.wait()``
        yields the tuple of ``synthesize_host_packed`` once the audio has landed.  Two sets of pinned buffers alternate, so keep at
        most two handles un-waited; the D2H copy of one batch overlaps the next batch's phase A."""
        B, L = texts.shape
        k = self._async_i = getattr(self, "_async_i", 0) ^ 1
        tag = "a%d_" % k
        h_t = self._pinned(tag + "texts", (B, L), torch.int64); h_t.copy_(texts)
        h_a = self._pinned(tag + "accents", (B, L), torch.int64); h_a.copy_(accents)
        h_l = self._pinned(tag + "lens", (B,), torch.int64); h_l.copy_(src_lens)
        h_s = self._pinned(tag + "spk", (B, 256), torch.float32); h_s.copy_(speaker_emb)
        h_w = self._pinned(tag + "wav_packed", (int(cap_samples),), torch.int16)
        h_o = self._pinned(tag + "offsets", (B + 1,), torch.int64)
        h_m = self._pinned(tag + "mel_lens", (B,), torch.int64)
        tmax, ticket = ctypes.c_int64(0), ctypes.c_int(-1)
        ctx = self.ctx
        ctx.check(ctx.lib.masg_synthesize_host_packed_async(ctx.h, ctx.stream(), _lib.ptr(h_t), _lib.ptr(h_a), _lib.ptr(h_l), B, L,
                                                            _lib.ptr(h_s), float(p_control), float(e_control), float(d_control),
                                                            ctypes.c_int64(int(cap_samples)), _lib.ptr(h_w), _lib.ptr(h_o), _lib.ptr(h_m),
                                                            ctypes.byref(tmax), ctypes.byref(ticket)), "masg_synthesize_host_packed_async")
        h2d = h_t.numel() * 8 + h_a.numel() * 8 + h_l.numel() * 8 + h_s.numel() * 4
        d2h = int(h_o[B]) * 2 + (B + 1) * 8
        return _PendingAudio(ctx, ticket.value, (h_w, h_o, h_m, int(tmax.value), h2d, d2h))

    def generate(self, texts, speakers, out_dir=None, speaker_names=None, text_ids=None, n_speakers=None, seed=0, p_control=1.0,
                 e_control=1.0, d_control=1.0, frame_budget=40000, max_batch=64, frames_per_phoneme=6.0, writer=None,
                 save_embeddings=True, sampling_rate=SAMPLING_RATE):
        """"N speakers x M texts" in one call — the batched form of the reference's generation scripts, which synthesise one
        (speaker, sentence) pair per Python iteration at batch size 1 (examples_gen.py:247-273, examples_gen_distri.py:209-218).

        texts:    tokenised batch {"texts", "accents" int64 [M, L] zero padded, "src_lens" int64 [M]} (frontend.Frontend.tokenize).
        speakers: float tensor [N, 256] of embeddings; or a distribution with ``.sample()`` (the drop-in InterpolateGMM /
                  BarycenterGMM, or one loaded with distributions.load_reference_distribution) together with ``n_speakers``; or a
                  metadata one-hot tensor [N, 4], sampled through ``speaker_gen`` (examples_gen.py:256-268).
        Every (speaker, text) pair becomes one utterance; utterances are sorted into length-bucketed batches under a padded-frame
        budget (plan_batches), each batch runs through masg_synthesize_host_packed (trim on the device, one D2H of the valid
        samples) and, with ``out_dir``, a background writer emits ``<speaker>_<text id>.wav`` while the next batch is synthesised;
        the embeddings are saved as ``<speaker>.npy`` ([1, 256] float32, the reference's format).
        Returns a list of dicts (speaker, text_id, path or wav, samples, mel_len, and the batch / row / padded length the utterance
        was synthesised in — the reference's outputs depend on the padding) in (speaker, text) order."""
        from .wavio import WavWriter
        from .fastspeech2 import gmm_sample
        ids, acc, lens = (torch.as_tensor(texts[k]).to("cpu", torch.int64) for k in ("texts", "accents", "src_lens"))
        M = ids.shape[0]
        dev = self.device
        # ---- speakers -> [N, 256] on the host (one or two launches, one D2H) ----
        if torch.is_tensor(speakers) and speakers.dim() == 2 and speakers.shape[1] == 256:
            emb = speakers.detach().to("cpu", torch.float32)
        elif torch.is_tensor(speakers) and speakers.dim() == 2:
            emb = self.fs2.speaker_gen(speakers.to(dev), seed=seed).cpu()
        elif hasattr(speakers, "component_distribution"):
            if not n_speakers:
                raise ValueError("n_speakers is required when `speakers` is a distribution")
            comp = speakers.component_distribution.base_dist
            emb = gmm_sample(self.ctx, speakers.mixture_distribution.probs.reshape(1, -1).to(dev), comp.loc.reshape(1, comp.loc.shape[-2], -1).to(dev),
                             comp.scale.reshape(1, comp.scale.shape[-2], -1).to(dev), int(n_speakers), seed=seed).cpu()
        else:
            raise TypeError("speakers must be an [N, 256] embedding tensor, an [N, 4] metadata tensor or a mixture distribution")
        N = emb.shape[0]
        speaker_names = list(speaker_names) if speaker_names is not None else ["spk%03d" % (i + 1) for i in range(N)]
        text_ids = list(text_ids) if text_ids is not None else ["%04d" % i for i in range(M)]
        if len(speaker_names) != N or len(text_ids) != M:
            raise ValueError("speaker_names / text_ids do not match the speakers / texts")
        # ---- N x M utterances, length-bucketed ----
        utt_text = np.tile(np.arange(M), N)
        utt_spk = np.repeat(np.arange(N), M)
        utt_len = lens.numpy()[utt_text]
        batches = plan_batches(utt_len, frame_budget=frame_budget, max_batch=max_batch, frames_per_phoneme=frames_per_phoneme)
        own_writer = writer is None and out_dir is not None
        if own_writer:
            writer = WavWriter()
        results = [None] * (N * M)
        try:
            if out_dir is not None and save_embeddings:
                save_speaker_embeddings(out_dir, speaker_names, emb)
            def collect(bi, idx, L, handle):
                h_w, h_o, h_m, T, _, _ = handle.wait()
                off = h_o.numpy()
                flat = h_w.numpy()
                for k, u in enumerate(idx):
                    wav = flat[off[k]: off[k + 1]]
                    rec = {"speaker": speaker_names[utt_spk[u]], "text_id": text_ids[utt_text[u]], "samples": int(off[k + 1] - off[k]),
                           "mel_len": int(h_m[k]), "batch": bi, "row": k, "padded_to": L}
                    if out_dir is not None:
                        rec["path"] = os.path.join(str(out_dir), "%s_%s.wav" % (rec["speaker"], rec["text_id"]))
                        writer.submit(rec["path"], sampling_rate, wav, copy=True)     # the pinned buffer is reused two batches later
                    else:
                        rec["wav"] = wav.copy()
                    results[int(u)] = rec

            # deferred-copy calls, one batch in flight behind the one being collected: the D2H copy and the file hand-over of batch
            # n overlap the synthesis of batch n + 1
            prev = None
            for bi, idx in enumerate(batches):
                t_idx = torch.from_numpy(utt_text[idx])
                L = int(utt_len[idx].max())
                b_ids, b_acc, b_len = ids[t_idx, :L].contiguous(), acc[t_idx, :L].contiguous(), lens[t_idx].contiguous()
                b_emb = emb[torch.from_numpy(utt_spk[idx])].contiguous()
                cap = int(len(idx) * L * 2 * frames_per_phoneme) * HOP
                while True:
                    try:
                        handle = self.synthesize_host_packed_async(b_ids, b_acc, b_len, b_emb, cap, p_control, e_control, d_control)
                        break
                    except RuntimeError as e:
                        if "capacity" not in str(e):
                            raise
                        cap *= 2                                      # durations longer than the estimate: retry with more room
                        if prev is not None:                          # the retry reuses the pinned set of the batch still in flight
                            collect(*prev)
                            prev = None
                if prev is not None:
                    collect(*prev)
                prev = (bi, idx, L, handle)
            if prev is not None:
                collect(*prev)
        finally:
            if own_writer:
                writer.close()
        return results


class _PendingAudio:
    """Handle of a deferred-copy synthesis call."""

    def __init__(self, ctx, ticket, result):
        self._ctx, self._ticket, self._result = ctx, ticket, result

    def wait(self):
        if self._ticket is not None:
            self._ctx.check(self._ctx.lib.masg_synthesize_host_wait(self._ctx.h, self._ticket), "masg_synthesize_host_wait")
            self._ticket = None
        return self._result


class SpeakerSampler:
    """A speaker set drawn from the TacoSpawn prior and its mid-attribute interpolations, re-sampled on every call with TWO
    kernel launches whatever its size (the reference draws one speaker per Python iteration: examples_gen.py:258-273,
    examples_gen_distri.py:209-218).  Default set = config 3 of BASELINE.json: ``n_sampled`` embeddings from
    ``speaker_gen`` cycling the four (gender, language) metadata combos + ``n_interp`` from ``InterpolateGMM(a, b)`` over the 6
    unordered combo pairs with t in {0.25, 0.5, 0.75}.  The priors are built once (the reference, too, loads a pickled
    distribution once); their K*K-component parameters are stacked so that one batched ``masg_gmm_sample`` launch serves all
    interpolated speakers."""

    def __init__(self, fs2, n_sampled, n_interp, seed=0, metas=None, rates=(0.25, 0.5, 0.75)):
        from .distributions import InterpolateGMM
        self.fs2, self.ctx, self.seed, self.calls = fs2, fs2._context(), int(seed), 0
        dev = self.ctx.device
        combos = torch.tensor(META_COMBOS, dtype=torch.float32, device=dev)
        self.meta = (torch.as_tensor(metas, dtype=torch.float32).to(dev) if metas is not None
                     else combos[torch.arange(n_sampled, device=dev) % 4]).reshape(-1, 4)[:n_sampled].contiguous()
        self.n_sampled, self.n_interp = n_sampled, n_interp
        self.interp_spec = []
        if n_interp:
            dists = [fs2.speaker_distribution(combos[i:i + 1]) for i in range(4)]
            pairs = [(a, b) for a in range(4) for b in range(a + 1, 4)]
            cache, pis, mus, scs = {}, [], [], []
            for i in range(n_interp):
                a, b = pairs[i % len(pairs)]
                t = rates[(i // len(pairs)) % len(rates)]
                if (a, b) not in cache:
                    cache[(a, b)] = InterpolateGMM(dists[a], dists[b])
                ig = cache[(a, b)]
                ig.interpolate_rate(t)
                comp = ig.component_distribution.base_dist
                pis.append(ig.mixture_distribution.probs.reshape(1, -1).clone())
                mus.append(comp.loc.reshape(1, -1, 256).clone())
                scs.append(comp.scale.reshape(1, -1, 256).clone())
                self.interp_spec.append((a, b, t))
            self.pi, self.mu, self.sc = torch.cat(pis).contiguous(), torch.cat(mus).contiguous(), torch.cat(scs).contiguous()

    def sample(self):
        """[n_sampled + n_interp, 256] on the model's device; a new draw per call (Philox seed = seed + 2 * call [+ 1])."""
        from .fastspeech2 import gmm_sample
        self.calls += 1
        embs = []
        if self.n_sampled:
            embs.append(self.fs2.speaker_gen(self.meta, seed=self.seed + 2 * self.calls))
        if self.n_interp:
            embs.append(gmm_sample(self.ctx, self.pi, self.mu, self.sc, self.n_interp, seed=self.seed + 2 * self.calls + 1))
        return torch.cat(embs, 0)


def sample_speakers(fs2, n_sampled, n_interp, seed=0):
    """One draw of the config-3 speaker set (see SpeakerSampler): [n_sampled + n_interp, 256] on the model's device."""
    return SpeakerSampler(fs2, n_sampled, n_interp, seed=seed).sample()


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


# ------------------------------------------------------------------------------------------------
# optional gather of the output waveforms over NVLink (SURVEY.md §8(e) "Collective"): the only data-path collective
# ------------------------------------------------------------------------------------------------
class WaveformGather:
    """All-gather of every rank's packed int16 audio (row b = packed[offsets[b]:offsets[b+1]]) so that rank 0 — or every rank —
    holds the whole job's waveforms.  Lengths differ per rank, so the exchange is two collectives on a side stream: an
    ``all_gather`` of the [total samples, utterances] counts, then one ``all_gather_into_tensor`` each of the offsets and of
    the sample buffers padded to the largest rank's size (C4: ~1.0 GB of int16 in total, ~1-2 ms at NVLink-5 rates; C3:
    19.7 MB per rank).  ``start()`` returns at once — the next batch's kernels run on the compute stream while NCCL moves the
    previous batch — and ``wait()`` yields [(packed, offsets)] per rank, trimmed.  Backend-agnostic through torch.distributed:
    NCCL over NVLink / NVSwitch on the GPUs, gloo in the CPU tests of the bookkeeping."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.stream = None
        self._pending = None

    def start(self, packed, offsets, n_max=None, b_max=None):
        """Enqueue the exchange.  ``n_max`` / ``b_max`` (upper bounds, identical on every rank, on the samples and utterances any
        rank holds — e.g. B * T_cap * hop and B): with them the sizes of the padded exchange are known up front and ``start`` does
        not touch the host; without them the counts are all-gathered first and read back (one host synchronisation, which stalls
        the launch of the next batch: measured +1.7 ms per C3 step on 8 GPUs)."""
        dist = self.dist
        dev = packed.device
        cuda = dev.type == "cuda"
        if cuda and self.stream is None:
            self.stream = torch.cuda.Stream(device=dev)
        B = offsets.numel() - 1
        known = n_max is not None and b_max is not None
        ready = torch.cuda.current_stream(dev).record_event() if cuda else None

        def run():
            nm, bm = (int(n_max), int(b_max)) if known else (None, None)
            if not known:
                counts = torch.stack([offsets[B].to(torch.int64), torch.tensor(B, dtype=torch.int64, device=dev)])
                all_counts = torch.empty(self.world * 2, dtype=torch.int64, device=dev)
                dist.all_gather_into_tensor(all_counts, counts, group=self.group)
                host = all_counts.view(self.world, 2).cpu()                      # sizes of the padded exchange (16 bytes per rank)
                nm, bm = int(host[:, 0].max()), int(host[:, 1].max())
            nm = max(nm, 1)
            if packed.numel() > nm or B > bm:
                raise ValueError("WaveformGather: this rank holds more than n_max / b_max")
            send_w = torch.zeros(nm, dtype=torch.int16, device=dev)
            send_w[: packed.numel()] = packed                                    # samples beyond offsets[B] are never read back
            send_o = torch.full((bm + 2,), -1, dtype=torch.int64, device=dev)    # [B | offsets (B + 1) | -1 padding]
            send_o[0] = B
            send_o[1: B + 2] = offsets
            recv_w = torch.empty(self.world * nm, dtype=torch.int16, device=dev)
            recv_o = torch.empty(self.world * (bm + 2), dtype=torch.int64, device=dev)
            # neither NCCL nor gloo has a 16-bit integer type: the samples travel as bytes
            dist.all_gather_into_tensor(recv_w.view(torch.uint8), send_w.view(torch.uint8), group=self.group)
            dist.all_gather_into_tensor(recv_o, send_o, group=self.group)
            # the bookkeeping goes to the host on THIS stream, so that wait() never has to synchronise with the compute stream
            host_o = recv_o
            if cuda:
                host_o = torch.empty(recv_o.shape, dtype=torch.int64).pin_memory()
                host_o.copy_(recv_o, non_blocking=True)
            return recv_w.view(self.world, -1), recv_o.view(self.world, -1), host_o.view(self.world, -1)

        if cuda:
            with torch.cuda.stream(self.stream):
                self.stream.wait_event(ready)
                packed.record_stream(self.stream); offsets.record_stream(self.stream)
                res = run()
                done = self.stream.record_event()
        else:
            res, done = run(), None
        self._pending = (res, done, dev)
        return self

    def wait(self):
        """[(packed int16 [n_r], offsets int64 [B_r + 1])] for r in ranks; the compute stream waits for the side stream."""
        (w, o, host), done, dev = self._pending
        self._pending = None
        if done is not None:
            done.synchronize()                                                   # the side stream only: long finished when waited a step later
            torch.cuda.current_stream(dev).wait_event(done)
        out = []
        for r in range(self.world):
            nb = int(host[r, 0])
            out.append((w[r, : int(host[r, 1 + nb])], o[r, 1: nb + 2]))
        return out


def gather_waveforms(packed, offsets, group=None):
    """Blocking form of WaveformGather: every rank returns the list of all ranks' (packed, offsets)."""
    return WaveformGather(group).start(packed, offsets).wait()
