"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI
(ctypes → libmasg.so), against the CPU oracle on the same seeded inputs and weights, against the
reference-generated golden fixtures, and — at BASELINE.json's full sizes — through size-independent
properties.  Tolerances are the ones BASELINE.json's north_star states:
  durations / LengthRegulator / bucketize / masks: bit-exact
  mel: max-abs <= 1e-4 (fp32 mode), <= 1e-2 (bf16 tcgen05 mode)
  waveform: SNR >= 30 dB
"""
import ctypes
import os

import numpy as np
import pytest
import torch

import masg_b200  # noqa: F401
from masg_b200 import synthetic as syn
from oracle import restate
from tests import helpers

pytestmark = pytest.mark.gpu

# bf16x3 = the tensor-core "1e-4" mode (every FastSpeech2 contraction split-bf16, ~16 mantissa bits per operand): measured
# 1.0e-4 .. 1.1e-4 on the three golden fixtures after 10 FFT blocks + PostNet, so its bar is 2e-4 (a real TF32 path, 10 bits, would
# sit near 1e-3); the strict 1e-4 bar is met by the fp32 mode
MEL_TOL = {"fp32": 1e-4, "bf16": 1e-2, "bf16x3": 2e-4}
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ctx():
    from masg_b200 import _lib
    return _lib.default_context(DEV, "fp32")


def P(t):
    from masg_b200 import _lib
    return _lib.ptr(t)


# ------------------------------------------------------------------------------------------------
# bit-exact integer / index kernels
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("B,L,C", [(3, 7, 8), (1, 1, 256), (64, 200, 256), (16, 1000, 256), (5, 33, 64)])
def test_length_regulator_bit_exact(ctx, B, L, C):
    g = torch.Generator().manual_seed(B * 1000 + L)
    x = torch.randn(B, L, C, generator=g)
    d = torch.round(torch.rand(B, L, generator=g) * 6) * (1.3 if L != 200 else 1.0)    # fractional products, zeros
    d[:, ::5] = 0.0
    if B > 2:
        d[1] = 0.0                                                                    # an all-pad utterance
    want, want_len = restate.length_regulate(x, d)
    T = want.shape[1]
    if T == 0:
        pytest.skip("degenerate")
    xg, dg = x.to(DEV), d.to(DEV)
    out = torch.full((B, T, C), 7.0, device=DEV)
    lens = torch.empty(B, dtype=torch.int64, device=DEV)
    mask = torch.empty(B, T, dtype=torch.uint8, device=DEV)
    ctx.check(ctx.lib.masg_length_regulate(ctx.h, ctx.stream(), P(xg), P(dg), B, L, C, T, P(out), P(lens), P(mask)), "lr")
    assert torch.equal(lens.cpu(), want_len)
    assert torch.equal(out.cpu(), want)                                               # bit-exact gather + zero pad
    assert torch.equal(mask.cpu().bool(), restate.mask_from_lengths(want_len, T))
    assert int(want_len.sum()) == int(d.to(torch.float64).trunc().clamp(min=0).sum())  # conservation
    lens2 = torch.empty(B, dtype=torch.int64, device=DEV)
    ctx.check(ctx.lib.masg_mel_lengths(ctx.h, ctx.stream(), P(dg), B, L, P(lens2)), "mel_lengths")
    assert torch.equal(lens2.cpu(), want_len)
    # explicit max_len larger than the batch max: zero padded (utils/tools.py:363-381)
    out2 = torch.full((B, T + 5, C), 7.0, device=DEV)
    ctx.check(ctx.lib.masg_length_regulate(ctx.h, ctx.stream(), P(xg), P(dg), B, L, C, T + 5, P(out2), None, None), "lr")
    assert torch.equal(out2.cpu(), restate.length_regulate(x, d, T + 5)[0])


@pytest.mark.parametrize("d_ctl", [1.0, 1.3, 0.5])
def test_duration_rounding_bit_exact(ctx, d_ctl):
    g = torch.Generator().manual_seed(3)
    log_d = torch.randn(20000, generator=g) * 1.5 + 1.0
    halves = torch.log(torch.arange(1, 40, dtype=torch.float32) + 0.5)                # exp(x)-1 lands on .5 boundaries
    log_d = torch.cat([log_d, halves, torch.tensor([0.0, -10.0, -0.0, 5.0])])
    want = restate.duration_round(log_d, d_ctl)
    out = torch.empty_like(log_d, device=DEV)
    log_dg = log_d.to(DEV)                                                            # keep alive across the call
    ctx.check(ctx.lib.masg_duration_round(ctx.h, ctx.stream(), P(log_dg), d_ctl, log_d.numel(), P(out)), "dur")
    out = out.cpu()
    # The kernel evaluates exp in double and rounds it to fp32 (k_duration), i.e. it implements the duration rule with a
    # correctly rounded fp32 exp.  That is bit-exact on EVERY input against the same rule evaluated through a float64 exp
    # (numpy), including the 39 engineered inputs whose exp(x) - 1 lands within an ulp of a .5 rounding boundary ...
    e32 = np.exp(log_d.numpy().astype(np.float64)).astype(np.float32)
    want_cr = np.maximum(np.rint(e32 - np.float32(1.0)).astype(np.float32) * np.float32(d_ctl), np.float32(0.0))
    assert np.array_equal(out.numpy(), want_cr)
    # ... and against torch's own CPU fp32 exp (SLEEF, <= 1 ulp, not correctly rounded) wherever that exp is not itself
    # on a boundary: all 20000 random inputs and the special values.  On the engineered boundary inputs torch CPU, torch
    # CUDA (expf, 2 ulp) and a correctly rounded exp can legitimately disagree by one rounding step: the reference's result
    # there depends on its libm, so the correctly rounded value above is the pinned one.
    n_rand = 20000
    assert torch.equal(out[:n_rand], want[:n_rand])
    assert torch.equal(out[-4:], want[-4:])
    n_diff = int((out[n_rand:-4] != want[n_rand:-4]).sum())
    print("engineered .5-boundary inputs where torch's CPU expf rounds the other way: %d of 39" % n_diff)
    assert (out[n_rand:] - want[n_rand:]).abs().max() <= d_ctl + 1e-6


def test_bucketize_embedding_bit_exact(ctx):
    g = torch.Generator().manual_seed(4)
    bins = torch.linspace(-2.5, 8.0, 255)
    v = torch.cat([torch.randn(5000, generator=g) * 3 + 2, bins[::7], torch.tensor([-100.0, 100.0, 0.0, -2.5, 8.0])])
    table = torch.randn(256, 256, generator=g)
    x = torch.randn(v.numel(), 256, generator=g)
    want_idx = torch.bucketize(v, bins)
    assert torch.equal(want_idx, restate.bucketize(v, bins))
    want = x + table[want_idx]
    xg = x.to(DEV)
    idx = torch.empty(v.numel(), dtype=torch.int64, device=DEV)
    vg, bg, tg = v.to(DEV), bins.to(DEV), table.to(DEV)                               # keep alive across the call
    ctx.check(ctx.lib.masg_bucketize_embed_add(ctx.h, ctx.stream(), P(vg), P(bg), 255, P(tg),
                                               v.numel(), 256, P(xg), P(idx)), "bucket")
    assert torch.equal(idx.cpu(), want_idx)
    assert torch.equal(xg.cpu(), want)


# ------------------------------------------------------------------------------------------------
# FastSpeech2
# ------------------------------------------------------------------------------------------------
def _run_fs2(model, batch, emb, **kw):
    gb = helpers.to_gpu(batch)
    out = model.synthesize_from_speaker_emb(None, gb["texts"], gb["src_lens"], batch["max_src_len"], accents=gb["accents"],
                                            speaker_emb=emb.to(DEV), **kw)
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_fs2_matches_oracle_ragged_batch(precision):
    model, sd = helpers.fs2_on_gpu(11, "spread", precision)
    b = syn.make_batch(4, 40, seed=21, lens=[40, 31, 17, 5])
    g = torch.Generator().manual_seed(2)
    emb = torch.randn(4, 256, generator=g) * 0.5
    taps = {}
    ref = restate.fs2_forward(sd, b["texts"], b["accents"], b["src_lens"], 40, emb, p_control=0.9, d_control=1.2, taps=taps)
    cx = model._context()
    with pytest.raises(RuntimeError):
        cx.tap("encoder_out", (4, 40, 256))                                   # taps are opt-in (they cost two activation copies)
    cx.enable_taps(True)
    out = _run_fs2(model, b, emb, p_control=0.9, d_control=1.2)
    tol = MEL_TOL[precision]
    enc = cx.tap("encoder_out", (4, 40, 256)).cpu()
    assert (enc - taps["encoder_out"]).abs().max() <= tol * 2
    T_ = out[0].shape[1]
    if torch.equal(out[5].cpu(), ref[5]):
        assert (cx.tap("lr_out", (4, T_, 256)).cpu() - taps["lr_out"]).abs().max() <= 4.0 + tol   # a bucket flip moves one embedding row
        assert (cx.tap("decoder_out", (4, T_, 256)).cpu() - taps["decoder_out"]).abs().max() <= 1.0
    cx.enable_taps(False)
    for i, name in ((2, "pitch"), (3, "energy"), (4, "log_d")):
        assert (out[i].cpu() - ref[i]).abs().max() <= tol * 4, name
    assert torch.equal(out[6].cpu(), ref[6]) and torch.equal(out[8].cpu(), b["src_lens"])
    # un-forced integer durations: equal in BOTH precisions (a flip only on an exact rounding boundary of the reference itself)
    flips = helpers.check_unforced_durations(out[4], ref[4], out[5], ref[5], precision)
    if precision == "fp32" or flips == 0:
        assert flips == 0 and torch.equal(out[9].cpu(), ref[9])
    # mel parity on the same padded batch with the oracle's durations teacher-forced
    ref_t = restate.fs2_forward(sd, b["texts"], b["accents"], b["src_lens"], 40, emb, p_control=0.9, d_control=1.2, d_target=ref[5])
    out_t = _run_fs2(model, b, emb, p_control=0.9, d_control=1.2, d_targets=ref[5].to(DEV))
    assert torch.equal(out_t[9].cpu(), ref_t[9]) and torch.equal(out_t[7].cpu(), ref_t[7])
    assert out_t[0].shape == ref_t[0].shape and out_t[0].dtype == torch.float32 and out_t[9].dtype == torch.int64
    e0 = (out_t[0].cpu() - ref_t[0]).abs().max().item()
    e1 = (out_t[1].cpu() - ref_t[1]).abs().max().item()                        # valid AND padded frames
    print("mel max-abs err %s: mel %.3e postnet %.3e (tol %.0e)" % (precision, e0, e1, tol))
    assert e0 <= tol and e1 <= tol


@pytest.mark.parametrize("precision", ["fp32", "bf16", "bf16x3"])
@pytest.mark.parametrize("name", ["fs2_ragged", "fs2_c1", "fs2_jsut"])
def test_fs2_matches_reference_golden(golden_dir, precision, name):
    """Against outputs of the unmodified reference (oracle/gen_golden.py); fs2_jsut = real phoneme / accent sequences from the
    vendored jsut-lab labels.  Un-forced first: every integer decision (durations, mel lengths, masks, pitch / energy buckets)
    must equal the reference's — a flip is accepted only on a boundary of the reference's own value, and not at all in fp32
    mode.  Then the mel comparison 'given identical predictor outputs': the reference's pitch / energy / durations
    teacher-forced through p_targets / e_targets / d_targets (model/modules.py:80-100,128-130)."""
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    model, sd = helpers.fs2_on_gpu(int(g["weights_seed"]), str(g["duration_mode"]), precision)
    texts = torch.from_numpy(g["texts"]).to(DEV)
    args = dict(accents=torch.from_numpy(g["accents"]).to(DEV), speaker_emb=torch.from_numpy(g["speaker_emb"]).to(DEV),
                p_control=float(g["p_control"]), d_control=float(g["d_control"]))
    lens = torch.from_numpy(g["src_lens"]).to(DEV)
    out = model.synthesize_from_speaker_emb(["raw text is ignored"], texts, lens, texts.shape[1], **args)
    tag = "%s %s" % (name, precision)
    flips = helpers.check_unforced_durations(out[4], g["log_d"], out[5], g["d_rounded"], tag)
    fp = helpers.check_unforced_buckets(out[2], g["pitch"], sd["variance_adaptor.pitch_bins"], tag + " pitch")
    # energy 'given identical predictor outputs' upstream: its predictor reads x + pitch_embedding, so with the pitch forced
    out_p = model.synthesize_from_speaker_emb(None, texts, lens, texts.shape[1], p_targets=torch.from_numpy(g["pitch"]).to(DEV), **args)
    fe = helpers.check_unforced_buckets(out_p[3], g["energy"], sd["variance_adaptor.energy_bins"], tag + " energy (pitch forced)")
    n_flips = flips + int(fp.sum()) + int(fe.sum())
    if precision == "fp32":
        assert n_flips == 0
    tol = MEL_TOL[precision]
    if n_flips == 0:                                   # un-forced outputs comparable as they are
        assert torch.equal(out[9].cpu(), torch.from_numpy(g["mel_lens"]))
        assert np.abs(out[0].cpu().numpy() - g["mel"]).max() <= tol
        assert np.abs(out[1].cpu().numpy() - g["postnet_mel"]).max() <= tol
    out = model.synthesize_from_speaker_emb(["raw text is ignored"], texts, lens, texts.shape[1],
                                            d_targets=torch.from_numpy(g["d_rounded"]).to(DEV), p_targets=torch.from_numpy(g["pitch"]).to(DEV),
                                            e_targets=torch.from_numpy(g["energy"]).to(DEV), **args)
    assert torch.equal(out[9].cpu(), torch.from_numpy(g["mel_lens"]))
    assert torch.equal(out[5].cpu(), torch.from_numpy(g["d_rounded"]))
    e0, e1 = np.abs(out[0].cpu().numpy() - g["mel"]).max(), np.abs(out[1].cpu().numpy() - g["postnet_mel"]).max()
    print("golden %s: mel max-abs %.3e postnet %.3e (tol %.0e), un-forced flips %d" % (tag, e0, e1, tol, n_flips))
    assert e0 <= tol and e1 <= tol
    assert torch.equal(out[7].cpu(), torch.from_numpy(g["mel_mask"]))


def test_fs2_teacher_forced_pitch_energy_matches_oracle():
    """p_targets / e_targets (model/modules.py:80-100): the embedding comes from bucketize(target); the returned prediction is
    the raw predictor output, NOT scaled by p_control."""
    model, sd = helpers.fs2_on_gpu(11, "spread", "fp32")
    b = syn.make_batch(3, 23, seed=31, lens=[23, 14, 6])
    g = torch.Generator().manual_seed(8)
    emb = torch.randn(3, 256, generator=g) * 0.5
    pt = torch.rand(3, 23, generator=g) * 10 - 2.5
    et = torch.rand(3, 23, generator=g) * 8 - 1.5
    ref = restate.fs2_forward(sd, b["texts"], b["accents"], b["src_lens"], 23, emb, p_control=0.7, p_target=pt, e_target=et)
    out = _run_fs2(model, b, emb, p_control=0.7, p_targets=pt.to(DEV), e_targets=et.to(DEV))
    assert torch.equal(out[9].cpu(), ref[9]) and torch.equal(out[5].cpu(), ref[5])
    for i in (0, 1, 2, 3, 4):
        assert (out[i].cpu() - ref[i]).abs().max() <= 1e-4, i
    ref_u = restate.fs2_forward(sd, b["texts"], b["accents"], b["src_lens"], 23, emb, p_control=0.7)
    assert (ref[2] * 0.7 - ref_u[2]).abs().max() <= 1e-6          # the un-forced path scales, the forced one does not


def test_fs2_forward_signature_and_speaker_lookup():
    model, sd = helpers.fs2_on_gpu(11, "spread", "fp32")
    b = syn.make_batch(2, 9, seed=3, lens=[9, 6])
    gb = helpers.to_gpu(b)
    speakers = torch.tensor([3, 150], device=DEV)
    out = model(speakers, gb["texts"], gb["src_lens"], 9, accents=gb["accents"], speaker_meta=gb["speaker_meta"])
    assert len(out) == 12
    emb = sd["speaker_emb.weight"][speakers.cpu()]
    ref = restate.fs2_forward(sd, b["texts"], b["accents"], b["src_lens"], 9, emb)
    assert (out[1].cpu() - ref[1]).abs().max() <= 1e-4 and torch.equal(out[9].cpu(), ref[9])
    assert torch.equal(out[11].cpu(), emb)
    pi, mu, sigma = restate.gmm_params(sd, b["speaker_meta"])
    assert (out[10].component_distribution.base_dist.loc.cpu() - mu).abs().max() < 1e-5
    # accents given as a 1-D [L] tensor broadcast over the batch (synthesize.py:234)
    acc1 = torch.full((9,), 4, dtype=torch.int64)
    out1 = model.synthesize_from_speaker_emb(None, gb["texts"], gb["src_lens"], 9, accents=acc1.to(DEV), speaker_emb=emb.to(DEV))
    ref1 = restate.fs2_forward(sd, b["texts"], acc1, b["src_lens"], 9, emb)
    assert (out1[1].cpu() - ref1[1]).abs().max() <= 1e-4
    with pytest.raises(TypeError):
        model.synthesize_from_speaker_emb(None, gb["texts"], gb["src_lens"], 9, speaker_emb=emb.to(DEV))
    with pytest.raises(RuntimeError):
        model.synthesize_from_speaker_emb(None, b["texts"], b["src_lens"], 9, accents=b["accents"], speaker_emb=emb)   # CPU tensors


def test_fs2_long_sequence_position_table():
    """L and T beyond max_seq_len=1000: the rebuilt sinusoid table path (transformer/Models.py:86-98,157-164)."""
    model, sd = helpers.fs2_on_gpu(1234, "fixed3", "fp32")
    b = syn.make_batch(1, 1010, seed=8)
    emb = torch.zeros(1, 256)
    ref = restate.fs2_forward(sd, b["texts"], b["accents"], b["src_lens"], 1010, emb)
    out = _run_fs2(model, b, emb)
    assert out[0].shape == (1, 3030, 80) and torch.equal(out[9].cpu(), ref[9])
    assert (out[1].cpu() - ref[1]).abs().max() <= 2e-4


# ------------------------------------------------------------------------------------------------
# HiFi-GAN
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_hifigan_matches_oracle_and_golden(golden_dir, precision):
    voc, hsd = helpers.hifigan_on_gpu(4321, precision)
    g = np.load(os.path.join(golden_dir, "hifigan_small.npz"))
    mel = torch.from_numpy(g["mel"])
    wav, w16 = voc(mel.to(DEV), return_int16=True)
    torch.cuda.synchronize()
    assert wav.shape == (2, 1, 20 * 256) and wav.dtype == torch.float32
    snr = min(helpers.snr_db(g["wav"][i], wav[i].cpu().numpy()) for i in range(2))
    print("HiFi-GAN %s: SNR vs reference golden %.1f dB, max-abs %.3e" % (precision, snr, np.abs(wav.cpu().numpy() - g["wav"]).max()))
    assert snr >= (80.0 if precision == "fp32" else 30.0)
    if precision == "fp32":
        assert np.abs(wav.cpu().numpy() - g["wav"]).max() <= 1e-4
    # int16 conversion is exactly the reference's numpy cast of OUR float waveform (utils/model.py:82-85)
    assert np.array_equal(w16.cpu().numpy(), restate.wav_to_int16(wav.squeeze(1)))
    # ragged, odd T and channels-last input
    g2 = torch.Generator().manual_seed(5)
    mel2 = torch.randn(3, 80, 37, generator=g2)
    ref2 = restate.hifigan_forward(hsd, mel2)
    out2 = voc(mel2.transpose(1, 2).contiguous().to(DEV), channels_last=True)
    assert out2.shape == (3, 1, 37 * 256)
    snr2 = min(helpers.snr_db(ref2[i].numpy(), out2[i].cpu().numpy()) for i in range(3))
    assert snr2 >= (80.0 if precision == "fp32" else 30.0)


def test_vocoder_infer_dropin():
    from masg_b200.hifigan import vocoder_infer
    voc, hsd = helpers.hifigan_on_gpu(4321, "fp32")
    g2 = torch.Generator().manual_seed(6)
    mel = torch.randn(2, 80, 16, generator=g2)
    wavs = vocoder_infer(mel.to(DEV), voc, syn.MODEL_CONFIG, syn.PREPROCESS_CONFIG, lengths=[16 * 256, 9 * 256])
    ref = restate.wav_to_int16(restate.hifigan_forward(hsd, mel).squeeze(1))
    assert wavs[0].dtype == np.int16 and len(wavs[0]) == 16 * 256 and len(wavs[1]) == 9 * 256
    assert np.abs(wavs[0].astype(np.int32) - ref[0].astype(np.int32)).max() <= 4
    assert np.abs(wavs[1].astype(np.int32) - ref[1][: 9 * 256].astype(np.int32)).max() <= 4


# ------------------------------------------------------------------------------------------------
# speaker prior
# ------------------------------------------------------------------------------------------------
def test_gmm_params_sampling_and_interpolation(golden_dir):
    from masg_b200.fastspeech2 import gmm_sample
    from masg_b200.distributions import InterpolateGMM, BarycenterGMM
    model, sd = helpers.fs2_on_gpu(1234, "fixed3", "fp32")
    gold = np.load(os.path.join(golden_dir, "gmm_prior.npz"))
    metas = torch.tensor(syn.META_COMBOS, dtype=torch.float32, device=DEV)
    pi, mu, sigma = model.speaker_enc.params(metas)
    assert np.abs(pi.cpu().numpy() - gold["pi"]).max() < 1e-6
    assert np.abs(mu.cpu().numpy() - gold["mu"]).max() < 1e-6 and np.abs(sigma.cpu().numpy() - gold["sigma"]).max() < 1e-6
    dist = model.speaker_distribution(metas)
    assert dist.batch_shape == (4,) and dist.event_shape == (256,)
    # exact sampling with injected draws
    gen = torch.Generator().manual_seed(1)
    N = 64
    meta_n = metas[torch.arange(N) % 4]
    pn, mn, sn = model.speaker_enc.params(meta_n)
    k = torch.randint(0, 3, (N,), generator=gen).to(torch.int32)
    eps = torch.randn(N, 256, generator=gen)
    out = gmm_sample(model._context(), pn, mn, sn, N, k_inject=k, eps_inject=eps)
    want = restate.gmm_sample_injected(mn.cpu(), sn.cpu(), k.long(), eps)
    assert torch.equal(out.cpu(), want)
    # internal Philox RNG: deterministic per seed, distributionally right (component frequencies, moments)
    Nbig = 20000
    p1, m1, s1 = pi[:1], mu[:1], sigma[:1]
    a, ka = gmm_sample(model._context(), p1, m1, s1, Nbig, seed=123, return_k=True)
    b2 = gmm_sample(model._context(), p1, m1, s1, Nbig, seed=123)
    c2 = gmm_sample(model._context(), p1, m1, s1, Nbig, seed=124)
    assert torch.equal(a, b2) and not torch.equal(a, c2)
    freq = torch.bincount(ka.long().cpu(), minlength=3).double() / Nbig
    assert (freq - pi[0].cpu().double()).abs().max() < 4 * 0.5 / np.sqrt(Nbig)
    mix_mean = (pi[0, :, None] * mu[0]).sum(0).cpu()
    mix_var = (pi[0, :, None] * (sigma[0] ** 2 + mu[0] ** 2)).sum(0).cpu() - mix_mean ** 2
    assert ((a.mean(0).cpu() - mix_mean).abs() < 5 * mix_var.sqrt() / np.sqrt(Nbig)).all()
    assert ((a.var(0).cpu() / mix_var) - 1).abs().max() < 0.15
    s = model.speaker_gen(metas)
    assert s.shape == (4, 256) and s.dtype == torch.float32
    # InterpolateGMM against the reference-generated fixture
    da, db = model.speaker_distribution(metas[0:1]), model.speaker_distribution(metas[3:4])
    ig = InterpolateGMM(da, db)
    assert np.abs(np.array(ig.ot_Cost) - gold["cost"]).max() / np.abs(gold["cost"]).max() < 1e-5
    assert np.abs(ig.mixture_distribution.probs[0].cpu().numpy() - gold["weights"]).max() < 1e-6
    assert np.abs(ig.component_distribution.base_dist.loc[0].cpu().numpy() - gold["mu_050"]).max() < 1e-6
    assert np.abs(ig.component_distribution.base_dist.scale[0].cpu().numpy() / gold["scale_050"] - 1).max() < 1e-5
    ig.interpolate_rate(0.25)
    assert np.abs(ig.component_distribution.base_dist.loc[0].cpu().numpy() - gold["mu_025"]).max() < 1e-6
    assert np.abs(ig.component_distribution.base_dist.scale[0].cpu().numpy() / gold["scale_025"] - 1).max() < 1e-5
    smp = ig.sample()
    assert smp.shape == (1, 256)
    import io
    buf = io.BytesIO(); torch.save(ig, buf); buf.seek(0)
    ig2 = torch.load(buf, weights_only=False)
    assert torch.equal(ig2.component_distribution.base_dist.loc, ig.component_distribution.base_dist.loc)
    # BarycenterGMM against the fixture written by the reference's OWN methods (oracle/ref_import.run_barycenter runs
    # model/distributions.py:79-192 unmodified apart from the one keyword that makes the constructor raise) ...
    gb = np.load(os.path.join(golden_dir, "barycenter.npz"))
    bg = BarycenterGMM(model)
    for tag in ("uniform", "rate"):
        if tag == "rate":
            bg.barycenter_rate([float(r) for r in gb["rate"]], _print=False)
        assert bg.mixture_distribution.probs.shape[-1] == gb["probs_" + tag].shape[0]
        assert np.abs(bg.mixture_distribution.probs[0].cpu().numpy() - gb["probs_" + tag]).max() < 1e-5
        assert np.abs(bg.component_distribution.base_dist.loc[0].cpu().numpy() - gb["mean_" + tag]).max() < 1e-5
        assert np.abs(bg.component_distribution.base_dist.scale[0].cpu().numpy() - gb["std_" + tag]).max() < 1e-5
    # ... and against the oracle's restatement of the same maths
    w, bm, bs, keys = restate.barycenter_gmm([p for p in pi.cpu()], [m for m in mu.cpu()], [s_ for s_ in sigma.cpu()],
                                             [float(r) for r in gb["rate"]])
    assert bg.mixture_distribution.probs.shape[-1] == len(keys)
    assert (bg.mixture_distribution.probs[0].cpu() - w / w.sum()).abs().max() < 1e-5
    assert (bg.component_distribution.base_dist.loc[0].cpu() - bm).abs().max() < 1e-5
    assert (bg.component_distribution.base_dist.scale[0].cpu() - bs).abs().max() < 1e-5
    assert bg.sample().shape == (1, 256)


# ------------------------------------------------------------------------------------------------
# end to end through the host-buffer C-ABI call; full-size properties
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_end_to_end_host_call_matches_staged_calls(precision):
    from masg_b200.synth import Synthesizer
    fs2, sd = helpers.fs2_on_gpu(1234, "fixed3", precision)
    voc, hsd = helpers.hifigan_on_gpu(4321, precision)
    syn_ = Synthesizer(fs2, voc, precision)
    b = syn.make_batch(3, 20, seed=12, lens=[20, 14, 6])
    g = torch.Generator().manual_seed(9)
    emb = torch.randn(3, 256, generator=g) * 0.5
    h_w, h_m, T, h2d, d2h = syn_.synthesize_host(b["texts"], b["accents"], b["src_lens"], emb, t_cap=64)
    assert T == 60 and h_m.tolist() == [60, 42, 18] and h2d > 0 and d2h == 3 * 60 * 256 * 2 + 24
    gb = helpers.to_gpu(b)
    w16, mel_lens, post = syn_.synthesize(gb["texts"], gb["accents"], gb["src_lens"], emb.to(DEV))
    assert torch.equal(h_w[:, : 256 * T], w16.cpu())                       # same kernels, same result, bit for bit
    ref = restate.fs2_forward(sd, b["texts"], b["accents"], b["src_lens"], 20, emb)
    ref_wav = restate.hifigan_forward(hsd, ref[1].transpose(1, 2))
    for i, ml in enumerate([60, 42, 18]):
        n = ml * 256
        snr = helpers.snr_db(ref_wav[i, 0, :n].numpy(), h_w[i, :n].float().numpy() / 32768.0)
        print("e2e %s utt %d SNR %.1f dB" % (precision, i, snr))
        assert snr >= 30.0                                                   # end-to-end waveform SNR over valid samples


@pytest.mark.parametrize("B,L", [(64, 200), (16, 1000)])
def test_full_size_properties_bf16(B, L):
    """BASELINE configs 3 and 5 at full size: properties that do not need the (slow) CPU oracle."""
    fs2, sd = helpers.fs2_on_gpu(1234, "fixed3", "bf16")
    voc, hsd = helpers.hifigan_on_gpu(4321, "bf16")
    b = syn.make_batch(B, L, seed=1234 + B)
    gb = helpers.to_gpu(b)
    g = torch.Generator().manual_seed(1)
    emb = (torch.randn(B, 256, generator=g) * 0.5).to(DEV)
    out = fs2.synthesize_from_speaker_emb(None, gb["texts"], gb["src_lens"], L, accents=gb["accents"], speaker_emb=emb)
    assert out[0].shape == (B, 3 * L, 80) and torch.equal(out[9].cpu(), torch.full((B,), 3 * L))
    assert torch.equal(out[5].cpu(), torch.full((B, L), 3.0)) and not out[7].any()
    assert torch.isfinite(out[1]).all()
    nb = min(B, 8)
    wav, w16 = voc(out[1][:nb], channels_last=True, return_int16=True)
    assert wav.shape == (nb, 1, 256 * 3 * L) and torch.isfinite(wav).all() and wav.abs().max() <= 1.0
    # batch-composition invariance for equal-length utterances: rows depend only on their own inputs.  A batch of 2 takes the
    # small-batch tiling (128-column fc / w_2 + separate LayerNorm), the full batch the fused 256-column one: same arithmetic,
    # another summation order in the LayerNorm statistics, so the fp16 hand-overs of the decoder may round differently — the rows
    # agree to within the mode's rounding, not bit for bit; two batches on the SAME path do agree exactly
    out2 = fs2.synthesize_from_speaker_emb(None, gb["texts"][:2], gb["src_lens"][:2], L, accents=gb["accents"][:2], speaker_emb=emb[:2])
    assert (out2[1] - out[1][:2]).abs().max() <= 5e-3
    nh = B // 2
    out3 = fs2.synthesize_from_speaker_emb(None, gb["texts"][:nh], gb["src_lens"][:nh], L, accents=gb["accents"][:nh], speaker_emb=emb[:nh])
    assert (out3[1] - out[1][:nh]).abs().max() <= 1e-5
    wav2 = voc(out[1][:2], channels_last=True)
    assert (wav2 - wav[:2]).abs().max() <= 1e-5


# ------------------------------------------------------------------------------------------------
# the dense-contraction kernel alone (tcgen05 implicit GEMM in bf16 mode) against torch conv1d on the CPU
# ------------------------------------------------------------------------------------------------
CONV_CASES = [
    # B, T, Cin, Cout, k, dil, transposed stride, in-lrelu slope, out act
    (2, 200, 256, 256, 1, 1, 0, 0.0, 0),      # fc / w_2-style linear
    (2, 150, 256, 768, 1, 1, 0, 0.0, 0),      # QKV
    (1, 300, 256, 1024, 9, 1, 0, 0.0, 1),     # FFN w_1 + ReLU
    (2, 131, 1024, 256, 1, 1, 0, 0.0, 0),     # FFN w_2, multi-chunk K
    (3, 77, 80, 512, 5, 1, 0, 0.0, 2),        # PostNet first conv + tanh, Cin = 80
    (2, 90, 512, 80, 5, 1, 0, 0.0, 0),        # PostNet last conv, N = 80
    (2, 1000, 128, 128, 11, 5, 0, 0.1, 0),    # HiFi-GAN dilated ResBlock conv
    (2, 1500, 32, 32, 3, 1, 0, 0.1, 0),       # HiFi-GAN stage 4
    (1, 2100, 64, 64, 7, 3, 0, 0.1, 0),
    (2, 37, 512, 256, 16, 1, 8, 0.1, 0),      # ConvTranspose1d ups[0]
    (2, 300, 64, 32, 4, 1, 2, 0.1, 0),        # ConvTranspose1d ups[3]
    (3, 700, 128, 128, 7, 3, 0, 0.1, 0),      # HiFi-GAN stage 2: 9 row blocks (odd: an empty peer tile in the CTA-pair form)
    (2, 520, 256, 256, 11, 5, 0, 0.1, 0),     # stage 1: two column tiles, multi-chunk K
    (1, 77, 256, 128, 16, 1, 8, 0.1, 0),      # ConvTranspose1d ups[1]: N = 1024
]


# "bf16-act-pair": the same planar layers through the CTA-pair form of the kernel (clusters of two CTAs, tcgen05 cta_group::2),
# which the library otherwise only picks for launches of at least two tiles per SM
@pytest.mark.parametrize("precision", ["fp32", "bf16", "bf16-split", "bf16-act", "bf16-act-pair"])
@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_kernel_matches_torch(precision, case):
    import torch.nn.functional as F
    from masg_b200 import _lib
    B, T, Cin, Cout, k, dil, ups, slope, act = case
    split = precision == "bf16-split"
    pair = precision == "bf16-act-pair"
    if pair:
        precision = "bf16-act"
        if ((ups or 1) * Cout) % 128:
            pytest.skip("the CTA-pair form exists for 128-column tiles")
    if precision == "bf16-act" and (Cout & (Cout - 1)):
        pytest.skip("planar bf16 activations (HiFi-GAN layout) need a power-of-two channel count")
    residual = precision == "bf16-act" and Cin == Cout and not ups       # bf16 activations in HBM + residual add (HiFi-GAN style)
    flags = (1 if split else 0) | (2 if precision == "bf16-act" else 0) | (4 if residual else 0) | (32 if pair else 0)
    cx = _lib.default_context(DEV, "bf16" if precision.startswith("bf16") else precision)
    g = torch.Generator().manual_seed(hash(case) % 1000)
    x = torch.randn(B, T, Cin, generator=g)
    w = torch.randn((Cin, Cout, k) if ups else (Cout, Cin, k), generator=g) / (Cin * (2 if ups else k)) ** 0.5
    bias = torch.randn(Cout, generator=g)
    xin = F.leaky_relu(x, slope) if slope else x
    if ups:
        ref = F.conv_transpose1d(xin.transpose(1, 2), w, bias, stride=ups, padding=(k - ups) // 2)
    else:
        ref = F.conv1d(xin.transpose(1, 2), w, bias, dilation=dil, padding=(k * dil - dil) // 2)
    ref = [ref, torch.relu(ref), torch.tanh(ref)][act].transpose(1, 2).contiguous()
    if residual:
        ref = ref + x
    xg, wg, bg = x.to(DEV), w.to(DEV), bias.to(DEV)
    y = torch.full(ref.shape, 7.0, device=DEV)
    ms = ctypes.c_float(0)
    cx.check(cx.lib.masg_conv1d(cx.h, cx.stream(), P(xg), B, T, Cin, P(wg), P(bg), Cout, k, dil, ups, slope, act, flags, 1, P(y),
                                ctypes.byref(ms)), "conv1d")
    err = (y.cpu() - ref).abs().max().item()
    print("conv %s %s: max-abs err %.3e, %.3f ms" % (precision, case, err, ms.value))
    assert err <= {"fp32": 2e-4, "bf16-split": 2e-4, "bf16": 4e-2, "bf16-act": 8e-2}[precision]


# ------------------------------------------------------------------------------------------------
# the fused ResBlock1 pair kernel (respair_tc.cu) against torch conv1d with the same bf16 rounding points
# (x, weights, the intermediate h and the output are bf16; accumulation fp32) — hifigan/models.py:96-103
# ------------------------------------------------------------------------------------------------
RESPAIR_CASES = [
    # B, T, C, k, dilation, with MRF accumulator, out_scale, post_slope
    (2, 1500, 32, 3, 1, False, 1.0, 1.0),
    (2, 1000, 32, 7, 3, False, 1.0, 1.0),
    (1, 777, 32, 11, 5, True, 1.0 / 3.0, 0.01),      # last pair of the last block of a stage
    (3, 100, 32, 11, 1, True, 1.0, 1.0),              # utterance shorter than one tile
    (1, 246 * 3, 32, 11, 3, False, 1.0, 1.0),         # T an exact multiple of the output rows per tile
    (2, 300, 64, 3, 5, False, 1.0, 0.1),
    (1, 2100, 64, 3, 1, True, 1.0, 1.0),
    (2, 1100, 64, 7, 5, True, 1.0 / 3.0, 0.1),       # weights fill shared memory: 2 activation slots, 1 h buffer
    (1, 250, 64, 7, 1, False, 1.0, 1.0),
    # the CTA-pair kernel (respair2_tc.cu: cluster of 2, tcgen05 cta_group::2, weight columns split over the pair)
    (2, 700, 128, 3, 1, False, 1.0, 1.0),
    (1, 1300, 128, 3, 5, True, 1.0 / 3.0, 0.1),
    (1, 126 * 3, 128, 3, 3, False, 1.0, 1.0),         # odd tile count: the last pair's second tile is empty
    (3, 126 * 2, 128, 3, 3, True, 1.0, 1.0),          # T an exact multiple of the output rows per tile
    (2, 1100, 64, 11, 5, True, 1.0 / 3.0, 0.1),
    (1, 246 * 3, 64, 11, 1, False, 1.0, 1.0),
    (1, 100, 64, 11, 3, False, 1.0, 1.0),             # one tile: the peer CTA has nothing to do
    (5, 3000, 64, 11, 3, False, 1.0, 1.0),            # more tile pairs than one wave of clusters walks in one go? no: 65 tiles, 33 pairs
    (40, 4000, 128, 3, 1, True, 1.0, 1.0),            # 1280 tiles = 640 pairs on 74 clusters: several pairs per cluster, slot reuse
]


@pytest.mark.parametrize("case", RESPAIR_CASES)
def test_fused_resblock_pair_matches_torch(case):
    import torch.nn.functional as F
    from masg_b200 import _lib
    B, T, C, k, dil, with_acc, scale, post = case
    cx = _lib.default_context(DEV, "bf16")
    g = torch.Generator().manual_seed(sum(int(1000 * v) for v in case) % 9973)
    bf = lambda t: t.to(torch.bfloat16).to(torch.float32)
    x = torch.randn(B, T, C, generator=g)
    acc = torch.randn(B, T, C, generator=g) if with_acc else None
    w1 = torch.randn(C, C, k, generator=g) / (C * k) ** 0.5
    w2 = torch.randn(C, C, k, generator=g) / (C * k) ** 0.5
    b1, b2 = torch.randn(C, generator=g) * 0.3, torch.randn(C, generator=g) * 0.3
    xb = bf(x)
    h = F.conv1d(bf(F.leaky_relu(xb, 0.1)).transpose(1, 2), bf(w1), b1, dilation=dil, padding=(k * dil - dil) // 2)
    h = bf(F.leaky_relu(h, 0.1))
    ref = F.conv1d(h, bf(w2), b2, padding=(k - 1) // 2).transpose(1, 2) + xb
    if with_acc:
        ref = ref + bf(acc)
    ref = ref * scale
    if post != 1.0:
        ref = F.leaky_relu(ref, post)
    to = lambda t: t.contiguous().to(DEV)
    xg, w1g, w2g, b1g, b2g = to(x), to(w1), to(w2), to(b1), to(b2)
    ag = to(acc) if with_acc else None
    y = torch.full((B, T, C), 7.0, device=DEV)
    ms = ctypes.c_float(0)
    cx.check(cx.lib.masg_respair(cx.h, cx.stream(), P(xg), P(ag) if with_acc else None, B, T, C, P(w1g), P(b1g), P(w2g), P(b2g), k, dil,
                                 scale, post, 1, P(y), ctypes.byref(ms)), "respair")
    err = (y.cpu() - ref).abs().max().item()
    print("respair %s: max-abs err %.3e (|ref| max %.2f), %.3f ms" % (case, err, ref.abs().max().item(), ms.value))
    assert err <= 4e-2                                 # one bf16 ulp of the output (|y| < 8) + re-rounding of h at ties


# ------------------------------------------------------------------------------------------------
# masked multi-head attention alone (tcgen05 kernel in bf16 mode) against the reference formula in torch fp64
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("B,S,lens", [(3, 200, [200, 77, 1]), (2, 600, [600, 389]), (2, 1300, [1300, 1025]), (2, 130, [129, 0])])
def test_attention_matches_reference_formula(precision, B, S, lens):
    """transformer/SubLayers.py:43-52 + Modules.py:14-25: softmax(q k^T / sqrt(128), -inf at pad keys) v per head; pad
    query rows compared as zeros (the reference zeroes them after the LayerNorm, Layers.py:25)."""
    from masg_b200 import _lib
    cx = _lib.default_context(DEV, precision)
    g = torch.Generator().manual_seed(S + B)
    qkv = torch.randn(B, S, 768, generator=g) * 1.5
    lens_t = torch.tensor(lens, dtype=torch.int64)
    q, k, v = [t.double().view(B, S, 2, 128).permute(0, 2, 1, 3) for t in qkv.split(256, dim=2)]
    att = q @ k.transpose(2, 3) / np.power(128, 0.5)
    key_pad = torch.arange(S)[None, :] >= lens_t[:, None]
    att = att.masked_fill(key_pad[:, None, None, :], float("-inf"))
    ref = (torch.softmax(att, dim=3) @ v).permute(0, 2, 1, 3).reshape(B, S, 256)
    ref = torch.nan_to_num(ref).masked_fill(key_pad[:, :, None], 0.0).float()
    qg, lg = qkv.to(DEV), lens_t.to(DEV)
    out = torch.full((B, S, 256), 7.0, device=DEV)
    cx.check(cx.lib.masg_attention(cx.h, cx.stream(), P(qg), P(lg), B, S, P(out)), "attention")
    err = (out.cpu() - ref).abs().max().item()
    print("attention %s B=%d S=%d: max-abs err %.3e" % (precision, B, S, err))
    assert err <= 2e-4


@pytest.mark.parametrize("B,S,lens", [(3, 200, [200, 77, 1]), (2, 600, [600, 389]), (2, 1300, [1300, 1025]), (2, 130, [129, 0]), (1, 128, [128])])
def test_attention_planar_fp16_matches_reference_formula(B, S, lens):
    """The decoder-side attention kernel (planar fp16 q|k|v in, planar fp16 out; V consumed as an MN-major operand straight
    from the planar layout) against the same reference formula evaluated in fp64 on the fp16-rounded inputs."""
    from masg_b200 import _lib
    cx = _lib.default_context(DEV, "bf16")
    g = torch.Generator().manual_seed(S + B)
    qkv = (torch.randn(B, S, 768, generator=g) * 1.5).half()
    lens_t = torch.tensor(lens, dtype=torch.int64)
    q, k, v = [t.double().view(B, S, 2, 128).permute(0, 2, 1, 3) for t in qkv.split(256, dim=2)]
    att = q @ k.transpose(2, 3) / np.power(128, 0.5)
    key_pad = torch.arange(S)[None, :] >= lens_t[:, None]
    att = att.masked_fill(key_pad[:, None, None, :], float("-inf"))
    ref = (torch.softmax(att, dim=3) @ v).permute(0, 2, 1, 3).reshape(B, S, 256)
    ref = torch.nan_to_num(ref).masked_fill(key_pad[:, :, None], 0.0).float()
    planar = qkv.view(B, S, 96, 8).permute(0, 2, 1, 3).contiguous().to(DEV)          # [B][96][S][8]
    lg = lens_t.to(DEV)
    out = torch.full((B, 32, S, 8), 7.0, device=DEV, dtype=torch.float16)
    cx.check(cx.lib.masg_attention_planar(cx.h, cx.stream(), P(planar), P(lg), B, S, P(out)), "attention_planar")
    got = out.permute(0, 2, 1, 3).reshape(B, S, 256).float().cpu()
    err = (got - ref).abs().max().item()
    print("attention planar fp16 B=%d S=%d: max-abs err %.3e (|ref| max %.2f)" % (B, S, err, ref.abs().max().item()))
    assert err <= 4e-3                                 # fp16 P (2^-11 relative) and the fp16 output ulp (|out| < 8: 2^-8 / 2)
