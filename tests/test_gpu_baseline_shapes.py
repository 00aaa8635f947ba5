"""GPU parity at BASELINE.json's own shapes (run with -m gpu on a B200): the CUDA path in the precision
bench.py times (bf16 mode) against the CPU oracle — not properties — on

  C2  B = 1, L = 50, accents = 4, FastSpeech2 + HiFi-GAN (the latency configuration)
  C3  rows of the bench batch itself (make_batch(64, 200, seed=1236), T = 600)
  C4  two real scheduler batches out of plan_batches(c4_lengths()): the shortest and a mid-length one,
      the SAME padded batch replayed by the oracle (pad-leak parity)
  C5  one 1000-phoneme utterance (T = 3000: rebuilt position table + k_attention_f16p at S = 3000)

Rows of a batch are independent given the padded shape (every op on the path is per-utterance; the pad leak of
Appendix B comes from the row's OWN padding), so the oracle replays a subset of rows that includes the longest
one (same L_max and T_max as the full batch) instead of all 64.

Tolerances (north_star): durations / mel_lens / masks bit-exact; mel max-abs <= 1e-2 (bf16 mode);
waveform SNR >= 30 dB per utterance over its valid samples.
"""
import numpy as np
import pytest
import torch

import masg_b200  # noqa: F401
from masg_b200 import synthetic as syn
from masg_b200.synth import plan_batches
from oracle import restate
from tests import helpers

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
MEL_TOL = 1e-2
SNR_MIN = 30.0


def _embeddings(B, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(B, 256, generator=g) * 0.5


def _gpu(fs2, voc, b, emb, rows=None, **kw):
    gb = helpers.to_gpu(b)
    sel = (lambda t: t) if rows is None else (lambda t: t[torch.tensor(rows, device=t.device)].contiguous())
    acc = gb["accents"] if gb["accents"].dim() == 1 else sel(gb["accents"])
    out = fs2.synthesize_from_speaker_emb(None, sel(gb["texts"]), sel(gb["src_lens"]), b["max_src_len"], accents=acc,
                                          speaker_emb=sel(emb.to(DEV)), **kw)
    wav = voc(out[1], channels_last=True)
    torch.cuda.synchronize()
    return out, wav


def _oracle_rows(sd, hsd, b, emb, rows, vocoder_rows=None, **kw):
    """Oracle on a subset of rows of the padded batch (rows must contain a longest utterance)."""
    rows = list(rows)
    assert int(b["src_lens"][rows].max()) == int(b["src_lens"].max()), "subset must keep the batch's L_max"
    acc = b["accents"] if b["accents"].dim() == 1 else b["accents"][rows]
    ref = restate.fs2_forward(sd, b["texts"][rows], acc, b["src_lens"][rows], b["max_src_len"], emb[rows], **kw)
    vr = list(range(len(rows))) if vocoder_rows is None else [rows.index(r) for r in vocoder_rows]
    ref_wav = restate.hifigan_forward(hsd, ref[1][vr].transpose(1, 2))
    return ref, ref_wav, vr


def _check(fs2, voc, sd, b, emb, out, wav, ref, ref_wav, rows, vr, tag):
    """(1) the integer decisions of the un-forced run against the oracle's: durations, mel lengths, masks equal; pitch / energy
    buckets equal (energy with the pitch forced: its input contains the pitch embedding) — a flip is accepted only on a boundary
    of the reference's own value (helpers.check_unforced_*) and is counted;
    (2) rows without any flip: the UN-FORCED mel / postnet mel / waveform against the oracle at the north_star tolerances;
    (3) all checked rows again with the oracle's pitch / energy / durations teacher-forced (p_targets / e_targets / d_targets,
    model/modules.py:80-100,128-130 — 'given identical predictor outputs'), so rows with a boundary flip are compared too."""
    rows_t = torch.tensor(rows)
    pb, eb = sd["variance_adaptor.pitch_bins"], sd["variance_adaptor.energy_bins"]
    d_flips = helpers.check_unforced_durations(out[4].cpu()[rows_t], ref[4], out[5].cpu()[rows_t], ref[5], tag)
    fp = helpers.check_unforced_buckets(out[2].cpu()[rows_t], ref[2], pb, tag + " pitch")
    fe = helpers.check_unforced_buckets(out[3].cpu()[rows_t], ref[3], eb, tag + " energy (un-forced)", assert_boundary=False)
    # the energy predictor reads x + pitch_embedding: its bucket decision is checked 'given identical predictor outputs'
    # upstream, i.e. with the oracle's pitch teacher-forced (a pitch-bucket flip legitimately changes the energy near it)
    out_p, _ = _gpu(fs2, voc, b, emb, rows=rows, p_targets=ref[2].to(DEV))
    helpers.check_unforced_buckets(out_p[3], ref[3], eb, tag + " energy (pitch forced)")
    assert torch.equal(out[6].cpu()[rows_t], ref[6]), tag + ": src masks differ"
    if d_flips == 0:
        assert torch.equal(out[9].cpu()[rows_t], ref[9]), tag + ": mel_lens differ"
        assert out[0].shape[1] == ref[0].shape[1], tag + ": T_max differs"
        assert torch.equal(out[7].cpu()[rows_t], ref[7]), tag + ": mel masks differ"
    clean = [k for k in range(len(rows)) if d_flips == 0 and not fp[k].any() and not fe[k].any()]

    def compare(o, w, sel, what):
        e_mel = e_post = 0.0
        snrs = []
        for k in sel:
            r = k if o[0].shape[0] == len(rows) else rows[k]
            e_mel = max(e_mel, (o[0][r].cpu() - ref[0][k]).abs().max().item())          # valid AND padded frames of the padded batch
            e_post = max(e_post, (o[1][r].cpu() - ref[1][k]).abs().max().item())
            n = int(ref[9][k]) * syn.HOP
            if k in vr and n:
                snrs.append(helpers.snr_db(ref_wav[vr.index(k), 0, :n].numpy(), w[r, 0, :n].cpu().numpy()))
        print("%s [%s, %d rows]: mel max-abs %.3e postnet %.3e (tol %.0e); end-to-end waveform SNR %s dB" %
              (tag, what, len(sel), e_mel, e_post, MEL_TOL, ", ".join("%.1f" % s_ for s_ in snrs)))
        assert e_mel <= MEL_TOL and e_post <= MEL_TOL, tag + " " + what
        assert not snrs or min(snrs) >= SNR_MIN, tag + " " + what

    if clean:
        compare(out, wav, clean, "un-forced")
    out_f, wav_f = _gpu(fs2, voc, b, emb, rows=rows, p_targets=ref[2].to(DEV), e_targets=ref[3].to(DEV), d_targets=ref[5].to(DEV))
    assert torch.equal(out_f[9].cpu(), ref[9]) and torch.equal(out_f[7].cpu(), ref[7]) and out_f[0].shape == ref[0].shape
    compare(out_f, wav_f, list(range(len(rows))), "pitch / energy / durations teacher-forced from the oracle")
    return out_f, wav_f


def test_c2_batch1_l50_matches_oracle():
    """BASELINE config 2 exactly: B = 1, L = 50, "no accent" (constant 4), FastSpeech2 + HiFi-GAN, bf16 mode."""
    fs2, sd = helpers.fs2_on_gpu(1234, "fixed3", "bf16")
    voc, hsd = helpers.hifigan_on_gpu(4321, "bf16")
    b = syn.make_batch(1, 50, seed=1235, accents_const=4)
    emb = _embeddings(1, 2)
    out, wav = _gpu(fs2, voc, b, emb)
    assert out[0].shape == (1, 150, 80) and wav.shape == (1, 1, 150 * 256)
    ref, ref_wav, vr = _oracle_rows(sd, restate.fold_weight_norm(hsd), b, emb, [0])
    _check(fs2, voc, sd, b, emb, out, wav, ref, ref_wav, [0], vr, "C2")


def test_c3_bench_batch_rows_match_oracle():
    """Rows 0-3 of the batch bench.py times (64 x 200 phonemes, T = 600), bf16 mode, against the oracle at B = 4."""
    fs2, sd = helpers.fs2_on_gpu(1234, "fixed3", "bf16")
    voc, hsd = helpers.hifigan_on_gpu(4321, "bf16")
    b = syn.make_batch(64, 200, seed=1234 + 2)
    emb = _embeddings(64, 3)
    out, wav = _gpu(fs2, voc, b, emb)
    assert out[0].shape == (64, 600, 80) and wav.shape == (64, 1, 600 * 256)
    rows = [0, 1, 2, 3]
    ref, ref_wav, vr = _oracle_rows(sd, restate.fold_weight_norm(hsd), b, emb, rows)
    _check(fs2, voc, sd, b, emb, out, wav, ref, ref_wav, rows, vr, "C3 rows 0-3")
    # rows outside the checked subset: same integer decisions as everyone else in this fixed-duration batch
    assert torch.equal(out[9].cpu(), torch.full((64,), 600)) and not out[7].any()


def test_c3_shape_spread_durations_unforced():
    """200-phoneme utterances with the 'spread' duration head (T varies per utterance, ~4 frames per phoneme): bf16-mode
    durations, mel lengths and T_max must equal the oracle's WITHOUT teacher forcing; a flip is tolerated only where the
    oracle's own value sits on a rounding / bin boundary to within the measured prediction error; all 16 rows are compared."""
    fs2, sd = helpers.fs2_on_gpu(11, "spread", "bf16")
    voc, hsd = helpers.hifigan_on_gpu(4321, "bf16")
    B = 16
    b = syn.make_batch(B, 200, seed=99)
    emb = _embeddings(B, 4)
    out, wav = _gpu(fs2, voc, b, emb)
    rows = list(range(B))
    ref = restate.fs2_forward(sd, b["texts"], b["accents"], b["src_lens"], 200, emb)
    r = int(ref[9].argmax())
    ref_wav = restate.hifigan_forward(restate.fold_weight_norm(hsd), ref[1][r:r + 1].transpose(1, 2))
    print("spread durations: T per utterance %d..%d" % (int(ref[9].min()), int(ref[9].max())))
    _check(fs2, voc, sd, b, emb, out, wav, ref, ref_wav, rows, [r], "C3-shape spread")


def test_c5_one_utterance_t3000_matches_oracle():
    """BASELINE config 5 shape for one utterance: 1000 phonemes -> T = 3000 in bf16 mode (fp16 decoder, k_attention_f16p at
    S = 3000, the position table rebuilt for T > max_seq_len: transformer/Models.py:157-164), FastSpeech2 + HiFi-GAN."""
    fs2, sd = helpers.fs2_on_gpu(1234, "fixed3", "bf16")
    voc, hsd = helpers.hifigan_on_gpu(4321, "bf16")
    b = syn.make_batch(2, 1000, seed=1234 + 4)
    emb = _embeddings(2, 5)
    out, wav = _gpu(fs2, voc, b, emb)
    assert out[0].shape == (2, 3000, 80)
    ref, ref_wav, vr = _oracle_rows(sd, restate.fold_weight_norm(hsd), b, emb, [0])
    _check(fs2, voc, sd, b, emb, out, wav, ref, ref_wav, [0], vr, "C5 utterance 0")


def test_c4_scheduler_batches_replayed_by_oracle():
    """Two batches formed by the C4 scheduler (plan_batches over numpy default_rng(0).integers(20, 301, 4096)): the shortest
    one and a mid-length one.  The oracle replays the same padded batch (longest, a middle and the shortest row), so the last
    10 valid frames of each shorter row (PostNet pad leak, Appendix B-2) and the last ~10.3 frames of its waveform (vocoder
    pad leak, B-3) are compared too."""
    fs2, sd = helpers.fs2_on_gpu(1234, "fixed3", "bf16")
    voc, hsd = helpers.hifigan_on_gpu(4321, "bf16")
    hsd_f = restate.fold_weight_norm(hsd)
    lengths = syn.c4_lengths(4096)
    batches = plan_batches(lengths, frame_budget=40000, max_batch=64)
    mid = min(range(len(batches)), key=lambda i: abs(int(lengths[batches[i]].max()) - 150))
    for name, idx in (("shortest", batches[-1]), ("mid", batches[mid])):
        lens = lengths[idx]
        b = syn.make_batch(len(idx), 0, seed=1237 + 7919 * int(idx[0]), lens=lens)
        emb = _embeddings(len(idx), 6)
        out, wav = _gpu(fs2, voc, b, emb)
        order = np.argsort(-lens, kind="stable")
        rows = [int(order[0]), int(order[len(order) // 2]), int(order[-1])]
        voc_rows = rows if name == "shortest" else rows[1:]                 # the long rows cost the CPU oracle seconds each
        ref, ref_wav, vr = _oracle_rows(sd, hsd_f, b, emb, rows, vocoder_rows=voc_rows)
        out_f, wav_f = _check(fs2, voc, sd, b, emb, out, wav, ref, ref_wav, rows, vr,
                              "C4 %s batch (B=%d, L %d..%d)" % (name, len(idx), lens.min(), lens.max()))
        # the pad-leak regions on their own (same integer decisions as the oracle): last 10 valid frames of the shortest row,
        # last 11 frames of its waveform
        k = len(rows) - 1
        T = int(ref[9][k])
        tail = (out_f[1].cpu()[k, T - 10:T] - ref[1][k, T - 10:T]).abs().max().item()
        n = T * syn.HOP
        kk = vr.index(k)
        wtail = helpers.snr_db(ref_wav[kk, 0, n - 11 * syn.HOP:n].numpy(), wav_f[k, 0, n - 11 * syn.HOP:n].cpu().numpy())
        print("  pad-leak regions of the shortest row (T=%d of %d): postnet tail max-abs %.3e, waveform tail SNR %.1f dB" %
              (T, ref[0].shape[1], tail, wtail))
        assert tail <= MEL_TOL and wtail >= SNR_MIN - 5.0                   # 11 frames only: a shorter window, same error
