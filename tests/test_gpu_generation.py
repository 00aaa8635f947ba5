"""GPU tests (-m gpu) of the rows either side of the hot path: trimmed / packed int16 output and wav
emission (f1), the batched generation driver and reference-pickled distribution artefacts (f2), the synth_samples drop-in (f3)
and the tokeniser feeding real label sequences end to end (f4) — each against the oracle or the reference itself."""
import os
import types

import numpy as np
import pytest
import torch

import masg_b200  # noqa: F401
from masg_b200 import synthetic as syn
from oracle import ref_import, restate
from tests import helpers

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_packed_vocoder_output_equals_padded_then_sliced(precision):
    """utils/model.py:82-90: int16 cast + per-utterance trim fused into the generator's last kernel (packed rows) must equal the
    padded int16 output sliced on the host, sample for sample; vocoder_infer returns exactly those arrays."""
    from masg_b200.hifigan import vocoder_infer
    voc, hsd = helpers.hifigan_on_gpu(4321, precision)
    g = torch.Generator().manual_seed(9)
    mel = torch.randn(5, 80, 23, generator=g).to(DEV)
    lens = torch.tensor([23, 1, 0, 17, 8])
    wav, w16 = voc(mel, return_int16=True)
    only16 = voc(mel, int16_only=True)
    assert torch.equal(only16, w16)                                            # the int16-only path skips the fp32 store, nothing else
    packed, off = voc.forward_packed(mel, lens)
    off = off.cpu()
    assert off.tolist() == [0] + torch.cumsum(lens * 256, 0).tolist()
    for i in range(5):
        assert torch.equal(packed[off[i]:off[i + 1]].cpu(), w16[i, : int(lens[i]) * 256].cpu())
    wavs = vocoder_infer(mel, voc, syn.MODEL_CONFIG, syn.PREPROCESS_CONFIG, lengths=lens * 256)
    assert [len(w) for w in wavs] == (lens * 256).tolist() and wavs[0].dtype == np.int16
    assert np.array_equal(wavs[3], w16[3, : 17 * 256].cpu().numpy())
    odd = vocoder_infer(mel, voc, syn.MODEL_CONFIG, syn.PREPROCESS_CONFIG, lengths=[100, 5, 0, 4000, 256])   # not multiples of hop
    assert [len(w) for w in odd] == [100, 5, 0, 4000, 256] and np.array_equal(odd[3], w16[3, :4000].cpu().numpy())
    full = vocoder_infer(mel, voc, syn.MODEL_CONFIG, syn.PREPROCESS_CONFIG)
    assert len(full) == 5 and np.array_equal(full[0], w16[0].cpu().numpy())
    # against the oracle's own cast of the oracle's waveform (<= a few LSB in fp32 mode)
    if precision == "fp32":
        ref = restate.wav_to_int16(restate.hifigan_forward(hsd, mel.cpu()).squeeze(1))
        assert np.abs(wavs[0].astype(np.int32) - ref[0][: 23 * 256].astype(np.int32)).max() <= 4


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_host_packed_call_equals_padded_host_call(precision):
    from masg_b200.synth import Synthesizer
    fs2, sd = helpers.fs2_on_gpu(11, "spread", precision)
    voc, hsd = helpers.hifigan_on_gpu(4321, precision)
    s = Synthesizer(fs2, voc, precision)
    b = syn.make_batch(4, 30, seed=15, lens=[30, 22, 9, 3])
    emb = torch.randn(4, 256, generator=torch.Generator().manual_seed(1)) * 0.5
    h_w, h_m, T, h2d, d2h = s.synthesize_host(b["texts"], b["accents"], b["src_lens"], emb, t_cap=400)
    h_w, h_m = h_w.clone(), h_m.clone()
    p_w, p_o, p_m, T2, h2d2, d2h2 = s.synthesize_host_packed(b["texts"], b["accents"], b["src_lens"], emb, cap_samples=4 * T * 256)
    assert T2 == T and torch.equal(p_m, h_m) and p_o.tolist() == [0] + torch.cumsum(h_m * 256, 0).tolist()
    for i in range(4):
        n = int(h_m[i]) * 256
        assert torch.equal(p_w[p_o[i]:p_o[i + 1]], h_w[i, :n])
    assert d2h2 == int(h_m.sum()) * 512 + 5 * 8 and d2h2 < d2h                 # only the valid samples cross the bus
    with pytest.raises(RuntimeError, match="capacity"):
        s.synthesize_host_packed(b["texts"], b["accents"], b["src_lens"], emb, cap_samples=256)


def test_generation_driver_n_speakers_x_m_texts(tmp_path, golden_dir):
    """examples_gen.py:247-273 as ONE call: 6 generated speakers x 5 real sentences (jsut-lab labels through the native
    tokeniser) -> 30 wav files + 6 embedding files; every utterance equals the oracle's on its own padded batch composition
    is not required here (rows are independent): each is compared with a solo oracle run at the waveform level."""
    from masg_b200 import frontend
    from masg_b200.synth import Synthesizer, speaker_names, load_speaker_embeddings
    from scipy.io import wavfile
    g = np.load(os.path.join(golden_dir, "frontend_jsut.npz"))
    fe = frontend.Frontend([str(s_) for s_ in g["symbols"]])
    pick = [3, 20, 41, 60, 77]
    tok = fe.tokenize([str(g["labels"][i]).split("\n") for i in pick])
    assert np.array_equal(tok["src_lens"].numpy(), g["src_lens"][pick]) and tok["texts"].is_pinned()
    fs2, sd = helpers.fs2_on_gpu(1234, "fixed3", "fp32")
    voc, hsd = helpers.hifigan_on_gpu(4321, "fp32")
    s = Synthesizer(fs2, voc, "fp32")
    metas = torch.tensor([syn.META_COMBOS[i % 4] for i in range(6)], dtype=torch.float32)
    names = speaker_names("F", "ja", 6)
    out = tmp_path / "speakers_wsample"
    res = s.generate(tok, metas, out_dir=out, speaker_names=names, text_ids=[str(g["names"][i])[:-4] for i in pick], seed=3,
                     frame_budget=3000, max_batch=8)
    assert len(res) == 30 and res[7]["speaker"] == names[1] and res[7]["text_id"] == str(g["names"][pick[2]])[:-4]
    embs = load_speaker_embeddings([out / (n + ".npy") for n in names])
    assert embs.shape == (6, 256) and torch.equal(embs, fs2.speaker_gen(metas.to(DEV), seed=3).cpu())
    hsd_f = restate.fold_weight_norm(hsd)
    assert len({r["batch"] for r in res}) >= 4
    checked = 0
    for k, r in enumerate(res):
        rate, data = wavfile.read(r["path"])
        assert rate == 22050 and len(data) == r["samples"] == r["mel_len"] * 256
        si, ti = k // 5, k % 5
        L = int(tok["src_lens"][ti])
        assert r["mel_len"] == 3 * L and r["padded_to"] >= L
        # the reference's result depends on the padding of the batch an utterance sits in (pad leak into the last phonemes,
        # then global attention: SURVEY Appendix B-1), so a solo oracle run is comparable for the un-padded rows: the longest
        # utterance of each batch
        if r["padded_to"] != L or checked >= 3:
            continue
        ref = restate.fs2_forward(sd, tok["texts"][ti:ti + 1, :L], tok["accents"][ti:ti + 1, :L], tok["src_lens"][ti:ti + 1], L, embs[si:si + 1])
        ref_wav = restate.hifigan_forward(hsd_f, ref[1].transpose(1, 2))[0, 0]
        assert helpers.snr_db(ref_wav.numpy(), data.astype(np.float64) / 32768.0) >= 60.0
        checked += 1
    assert checked == 3
    # a distribution as the speaker source: one batched draw for all speakers, no files
    from masg_b200.distributions import InterpolateGMM
    da = fs2.speaker_distribution(metas[0:1].to(DEV)); db = fs2.speaker_distribution(metas[3:4].to(DEV))
    res2 = s.generate(tok, InterpolateGMM(da, db), n_speakers=3, seed=5)
    assert len(res2) == 15 and res2[0]["wav"].dtype == np.int16 and len(res2[4]["wav"]) == res2[4]["mel_len"] * 256


def test_speaker_sampler_is_batched_and_reproducible():
    from masg_b200.synth import SpeakerSampler, sample_speakers
    fs2, sd = helpers.fs2_on_gpu(1234, "fixed3", "fp32")
    ctx = fs2._context()
    sm = SpeakerSampler(fs2, 32, 32, seed=77)
    n0 = ctx.launch_count()
    e = sm.sample()
    assert ctx.launch_count() - n0 <= 3                                        # gmm_params + two gmm_sample launches for 64 speakers
    assert e.shape == (64, 256) and torch.isfinite(e).all()
    e2 = SpeakerSampler(fs2, 32, 32, seed=77).sample()
    assert torch.equal(e, e2) and not torch.equal(e, sm.sample())              # same seed, same draw; the next call draws anew
    assert sample_speakers(fs2, 4, 6, seed=1).shape == (10, 256)
    # the interpolated half really comes from the mid-attribute priors: speaker 32 is drawn from InterpolateGMM(0, 1) at t = .25
    pi, mu, sigma = restate.gmm_params(sd, torch.tensor(syn.META_COMBOS, dtype=torch.float32))
    w, cm, cs = restate.interpolate_gmm(pi[0], mu[0], sigma[0], pi[1], mu[1], sigma[1], 0.25)
    assert torch.allclose(sm.mu[0].cpu(), cm.float(), atol=1e-5) and torch.allclose(sm.pi[0].cpu(), w.float(), atol=1e-6)


def test_reference_pickled_distribution_loads_into_the_dropin(golden_dir):
    """examples_gen_distri.py:211: a distri_*.pth written by the REFERENCE classes becomes the drop-in on the GPU — same
    components, same rate, the reference's own transport plan — and keeps working (interpolate_rate / barycenter_rate / sample)."""
    from masg_b200.distributions import load_reference_distribution, read_reference_distribution, InterpolateGMM, BarycenterGMM
    ig = load_reference_distribution(os.path.join(golden_dir, "distri_ref_interp.pth"), DEV)
    parts = read_reference_distribution(os.path.join(golden_dir, "distri_ref_interp.pth"))
    assert isinstance(ig, InterpolateGMM) and ig.t == 0.75
    assert np.allclose(ig.mixture_distribution.probs[0].cpu().numpy(), parts["plan"].flatten(), atol=1e-7)
    w, cm, cs = restate.interpolate_gmm(*parts["a"], *parts["b"], 0.75)
    assert torch.allclose(ig.component_distribution.base_dist.loc[0].cpu(), cm.float(), atol=1e-6)
    assert torch.allclose(ig.component_distribution.base_dist.scale[0].cpu(), cs.float(), rtol=1e-5, atol=1e-7)
    ig.interpolate_rate(0.5)
    _, cm5, _ = restate.interpolate_gmm(*parts["a"], *parts["b"], 0.5)
    assert torch.allclose(ig.component_distribution.base_dist.loc[0].cpu(), cm5.float(), atol=1e-6)
    assert ig.sample().shape == (1, 256)
    bg = load_reference_distribution(os.path.join(golden_dir, "distri_ref_bary.pth"), DEV)
    gb = np.load(os.path.join(golden_dir, "barycenter.npz"))
    assert isinstance(bg, BarycenterGMM)
    assert np.abs(bg.mixture_distribution.probs[0].cpu().numpy() - gb["probs_rate"]).max() < 1e-5
    assert np.abs(bg.component_distribution.base_dist.loc[0].cpu().numpy() - gb["mean_rate"]).max() < 1e-5
    bg.barycenter_rate([0.25] * 4, _print=False)
    assert np.abs(bg.mixture_distribution.probs[0].cpu().numpy() - gb["probs_uniform"]).max() < 1e-5
    assert bg.sample().shape == (1, 256)
    import io
    buf = io.BytesIO(); torch.save(ig, buf); buf.seek(0)                       # and it pickles as the drop-in class from here on
    assert isinstance(torch.load(buf, weights_only=False), InterpolateGMM)


@pytest.mark.skipif(not ref_import.available(), reason="no reference tree on this box (oracle/_ref not staged)")
def test_synth_samples_dropin_against_the_reference_function(tmp_path):
    """utils/tools.py:228-274 end to end: the drop-in on the GPU (fp32 mode) and the unmodified reference function with the
    reference vocoder on the CPU, from the same weights and the same predictions: same file names, same lengths, samples within
    a few LSB."""
    import json
    from unittest import mock
    from scipy.io import wavfile
    from masg_b200 import tools
    fs2, sd = helpers.fs2_on_gpu(11, "spread", "fp32")
    voc, hsd = helpers.hifigan_on_gpu(4321, "fp32")
    b = syn.make_batch(3, 14, seed=33, lens=[14, 9, 4])
    emb = torch.randn(3, 256, generator=torch.Generator().manual_seed(2)) * 0.5
    gb = helpers.to_gpu(b)
    pred = fs2.synthesize_from_speaker_emb(None, gb["texts"], gb["src_lens"], 14, accents=gb["accents"], speaker_emb=emb.to(DEV))
    cfg = tmp_path / "cfg"; cfg.mkdir()
    json.dump(syn.STATS_JSON, open(cfg / "stats.json", "w"))
    args = types.SimpleNamespace(config=str(cfg))
    targets = (["a", "b", "c"],)
    ours = tmp_path / "ours"; ours.mkdir()
    paths = tools.synth_samples(args, targets, pred, voc, syn.MODEL_CONFIG, syn.PREPROCESS_CONFIG, str(ours))
    ref_import.load()
    import utils.tools as rt
    ref_voc = ref_import.build_hifigan(hsd)
    pred_cpu = tuple(p.cpu() if torch.is_tensor(p) else p for p in pred)
    refdir = tmp_path / "ref"; refdir.mkdir()
    pre, mod = ref_import.load_configs()
    with mock.patch.object(rt, "plot_mel", lambda *a, **k: None), mock.patch.object(rt, "plt", mock.MagicMock()):
        rt.synth_samples(args, targets, pred_cpu, ref_voc, mod, pre, str(refdir))
    for p in paths:
        r0, d0 = wavfile.read(p)
        r1, d1 = wavfile.read(str(refdir / os.path.basename(p)))
        assert r0 == r1 == 22050 and d0.shape == d1.shape and d0.dtype == d1.dtype == np.int16
        assert np.abs(d0.astype(np.int32) - d1.astype(np.int32)).max() <= 4


@pytest.mark.parametrize("packed", [False, True])
def test_small_batch_graph_replay_is_bit_identical(packed):
    """The latency path (host-buffer entry points, <= 512 phoneme rows): first call plain, second captured into CUDA graphs
    (phase A; phase B + vocoder), later calls replayed — same audio bit for bit, same kernel count claimed, and a change of
    inputs, controls or shape must not replay a stale graph."""
    from masg_b200.synth import Synthesizer
    fs2, sd = helpers.fs2_on_gpu(11, "spread", "bf16")
    voc, hsd = helpers.hifigan_on_gpu(4321, "bf16")
    s = Synthesizer(fs2, voc, "bf16")
    b = syn.make_batch(2, 24, seed=41, lens=[24, 13])
    emb = torch.randn(2, 256, generator=torch.Generator().manual_seed(4)) * 0.5

    def call(bb, e, **kw):
        n0 = s.ctx.launch_count()
        if packed:
            w, o, m, T, _, _ = s.synthesize_host_packed(bb["texts"], bb["accents"], bb["src_lens"], e, 2 * 600 * 256, **kw)
            return w[: int(o[-1])].clone(), m.clone(), s.ctx.launch_count() - n0
        w, m, T, _, _ = s.synthesize_host(bb["texts"], bb["accents"], bb["src_lens"], e, 600, **kw)
        return w[:, : T * 256].clone(), m.clone(), s.ctx.launch_count() - n0

    runs = [call(b, emb) for _ in range(4)]                                   # plain, capture, replay, replay
    for w, m, n in runs[1:]:
        assert torch.equal(w, runs[0][0]) and torch.equal(m, runs[0][1]) and n == runs[0][2]
    emb2 = emb * 0.5
    w2, m2, _ = call(b, emb2)                                                  # new inputs through the replayed graph
    gb = helpers.to_gpu(b)
    ref16, ref_m, _ = s.synthesize(gb["texts"], gb["accents"], gb["src_lens"], emb2.to(DEV))
    assert torch.equal(m2, ref_m.cpu())
    if packed:
        assert torch.equal(w2[: int(m2[0]) * 256], ref16[0, : int(m2[0]) * 256].cpu())
    else:
        assert torch.equal(w2, ref16.cpu())
    w3, m3, _ = call(b, emb, d_control=1.5)                                    # another control value: another graph key
    assert int(m3.sum()) > int(runs[0][1].sum())
    w4, m4, _ = call(b, emb)
    assert torch.equal(w4, runs[0][0])


def test_out_of_range_ids_raise_like_nn_embedding():
    """transformer/Models.py:101-103: nn.Embedding raises IndexError for an id outside its table; the kernel flags it and the
    host raises at phase A's synchronisation instead of reading out of bounds."""
    fs2, sd = helpers.fs2_on_gpu(11, "spread", "fp32")
    b = syn.make_batch(2, 9, seed=3)
    gb = helpers.to_gpu(b)
    emb = torch.zeros(2, 256, device=DEV)
    bad = gb["texts"].clone(); bad[1, 4] = 429
    with pytest.raises(RuntimeError, match="phoneme id"):
        fs2.synthesize_from_speaker_emb(None, bad, gb["src_lens"], 9, accents=gb["accents"], speaker_emb=emb)
    bad_a = gb["accents"].clone(); bad_a[0, 0] = -1
    with pytest.raises(RuntimeError, match="accent id"):
        fs2.synthesize_from_speaker_emb(None, gb["texts"], gb["src_lens"], 9, accents=bad_a, speaker_emb=emb)
    out = fs2.synthesize_from_speaker_emb(None, gb["texts"], gb["src_lens"], 9, accents=gb["accents"], speaker_emb=emb)   # still healthy
    assert torch.isfinite(out[1]).all()
    # torch.bucketize sends NaN to the last bucket (index 255)
    from masg_b200 import _lib
    cx = _lib.default_context(DEV, "fp32")
    v = torch.tensor([float("nan"), 0.0], device=DEV); bins = torch.linspace(-1, 1, 255, device=DEV)
    table = torch.zeros(256, 256, device=DEV); x = torch.zeros(2, 256, device=DEV); idx = torch.empty(2, dtype=torch.int64, device=DEV)
    cx.check(cx.lib.masg_bucketize_embed_add(cx.h, cx.stream(), _lib.ptr(v), _lib.ptr(bins), 255, _lib.ptr(table), 2, 256, _lib.ptr(x), _lib.ptr(idx)), "bucket")
    assert idx.tolist() == torch.bucketize(v.cpu(), bins.cpu()).tolist() == [255, 127]


def test_deferred_copy_calls_match_blocking_calls():
    """masg_synthesize_host_packed_async: three batches in flight-order (two output buffers alternate, the third call waits on
    the device for the first copy) give bit for bit the audio of the blocking call, whatever order the handles are waited in."""
    from masg_b200.synth import Synthesizer
    fs2, sd = helpers.fs2_on_gpu(11, "spread", "bf16")
    voc, hsd = helpers.hifigan_on_gpu(4321, "bf16")
    s = Synthesizer(fs2, voc, "bf16")
    batches = [syn.make_batch(3, 20, seed=50 + i, lens=[20, 12 + i, 5]) for i in range(3)]
    embs = [torch.randn(3, 256, generator=torch.Generator().manual_seed(60 + i)) * 0.5 for i in range(3)]
    want = []
    for b, e in zip(batches, embs):
        w, o, m, T, _, _ = s.synthesize_host_packed(b["texts"], b["accents"], b["src_lens"], e, 3 * 400 * 256)
        want.append((w[: int(o[-1])].clone(), o.clone(), m.clone()))
    for rep in range(2):                                                       # second round: replayed graphs on both buffers
        h0 = s.synthesize_host_packed_async(batches[0]["texts"], batches[0]["accents"], batches[0]["src_lens"], embs[0], 3 * 400 * 256)
        h1 = s.synthesize_host_packed_async(batches[1]["texts"], batches[1]["accents"], batches[1]["src_lens"], embs[1], 3 * 400 * 256)
        r0 = h0.wait()
        got0 = (r0[0][: int(r0[1][-1])].clone(), r0[1].clone(), r0[2].clone())  # its pinned set is reused by the third call
        h2 = s.synthesize_host_packed_async(batches[2]["texts"], batches[2]["accents"], batches[2]["src_lens"], embs[2], 3 * 400 * 256)
        r2, r1 = h2.wait(), h1.wait()
        got = [got0, (r1[0][: int(r1[1][-1])], r1[1], r1[2]), (r2[0][: int(r2[1][-1])], r2[1], r2[2])]
        for (gw, go, gm), (ww, wo, wm) in zip(got, want):
            assert torch.equal(go, wo) and torch.equal(gm, wm) and torch.equal(gw, ww)
