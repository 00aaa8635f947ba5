"""The oracle (oracle/restate.py) against golden vectors produced by the UNMODIFIED reference
(oracle/gen_golden.py), and — when /root/reference is present — against the live reference modules.
CPU only."""
import os

import numpy as np
import pytest
import torch

import masg_b200  # noqa: F401
from masg_b200 import synthetic as syn
from oracle import ref_import, restate


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"))


@pytest.mark.parametrize("name", ["fs2_ragged", "fs2_c1"])
def test_fs2_restatement_matches_reference_golden(golden_dir, name):
    g = _load(golden_dir, name)
    sd = syn.make_fs2_state_dict(seed=int(g["weights_seed"]), duration_mode=str(g["duration_mode"]))
    texts, acc, lens = (torch.from_numpy(g[k]) for k in ("texts", "accents", "src_lens"))
    out = restate.fs2_forward(sd, texts, acc, lens, texts.shape[1], torch.from_numpy(g["speaker_emb"]),
                              p_control=float(g["p_control"]), d_control=float(g["d_control"]))
    assert torch.equal(out[9], torch.from_numpy(g["mel_lens"]))
    assert torch.equal(out[5], torch.from_numpy(g["d_rounded"]))            # durations: bit-exact
    assert torch.equal(out[6], torch.from_numpy(g["src_mask"])) and torch.equal(out[7], torch.from_numpy(g["mel_mask"]))
    for i, key in ((0, "mel"), (1, "postnet_mel"), (2, "pitch"), (3, "energy"), (4, "log_d")):
        np.testing.assert_allclose(out[i].numpy(), g[key], atol=2e-5, rtol=0)


def test_hifigan_restatement_matches_reference_golden(golden_dir):
    g = _load(golden_dir, "hifigan_small")
    sd = syn.make_hifigan_state_dict(seed=int(g["weights_seed"]))
    wav = restate.hifigan_forward(sd, torch.from_numpy(g["mel"]))
    assert wav.shape == (2, 1, 20 * 256)                                       # length law len_out = 256*T
    np.testing.assert_allclose(wav.numpy(), g["wav"], atol=2e-5, rtol=0)
    assert float(np.std(g["wav"])) > 0.1                                       # the fixture is not a DC signal


def test_gmm_prior_restatement_matches_reference_golden(golden_dir):
    g = _load(golden_dir, "gmm_prior")
    sd = syn.make_fs2_state_dict(seed=int(g["weights_seed"]))
    metas = torch.tensor(syn.META_COMBOS, dtype=torch.float32)
    pi, mu, sigma = restate.gmm_params(sd, metas)
    np.testing.assert_allclose(pi.numpy(), g["pi"], atol=1e-6)
    np.testing.assert_allclose(mu.numpy(), g["mu"], atol=1e-6)
    np.testing.assert_allclose(sigma.numpy(), g["sigma"], atol=1e-6)
    for t, tag in ((0.5, "050"), (0.25, "025")):
        w, cm, cs = restate.interpolate_gmm(pi[0], mu[0], sigma[0], pi[3], mu[3], sigma[3], t)
        np.testing.assert_allclose(w.numpy(), g["weights"], atol=1e-6)
        np.testing.assert_allclose(cm.numpy(), g["mu_" + tag], atol=1e-6)
        np.testing.assert_allclose(cs.numpy(), g["scale_" + tag], rtol=1e-5, atol=1e-7)
    cost = [[restate.interp_w2sq(mu[0][i].numpy(), (sigma[0][i] ** 2).numpy(), mu[3][j].numpy(), (sigma[3][j] ** 2).numpy())
             for j in range(3)] for i in range(3)]
    np.testing.assert_allclose(np.array(cost), g["cost"], rtol=1e-5)


def test_emd_exact_known_answers():
    # marginals respected, cost optimal on hand-checkable cases
    G = restate.emd_exact([0.5, 0.5], [0.5, 0.5], [[0.0, 1.0], [1.0, 0.0]])
    np.testing.assert_allclose(G, [[0.5, 0.0], [0.0, 0.5]], atol=1e-12)
    G = restate.emd_exact([0.5, 0.5], [0.5, 0.5], [[1.0, 0.0], [0.0, 1.0]])
    np.testing.assert_allclose(G, [[0.0, 0.5], [0.5, 0.0]], atol=1e-12)
    rng = np.random.default_rng(3)
    from scipy.optimize import linprog
    for _ in range(20):
        a = rng.random(3); a /= a.sum(); b = rng.random(3); b /= b.sum(); M = rng.random((3, 3))
        G = restate.emd_exact(a, b, M)
        np.testing.assert_allclose(G.sum(1), a, atol=1e-9); np.testing.assert_allclose(G.sum(0), b, atol=1e-9)
        A_eq = np.zeros((6, 9))
        for i in range(3):
            for j in range(3):
                A_eq[i, i * 3 + j] = 1; A_eq[3 + j, i * 3 + j] = 1
        lp = linprog(M.ravel(), A_eq=A_eq, b_eq=np.concatenate([a, b]), bounds=(0, None), method="highs")
        assert abs((G * M).sum() - lp.fun) < 1e-9


def test_length_regulator_properties():
    torch.manual_seed(0)
    x = torch.randn(3, 7, 8)
    d = torch.tensor([[2.6, 0.0, 1.0, 3.9, 0.0, 0.0, 1.3], [0.0] * 7, [1.0] * 7])
    out, mel_len = restate.length_regulate(x, d)
    assert mel_len.tolist() == [7, 0, 7]                                       # int() truncation, sum conservation
    assert torch.equal(out[0, :2], x[0, 0].expand(2, -1)) and torch.equal(out[0, 2], x[0, 2])
    assert torch.equal(out[1], torch.zeros_like(out[1]))                       # all-zero durations → all pad
    assert torch.equal(restate.duration_round(torch.tensor([np.log(3.5), np.log(4.5), -5.0]), 1.0),
                       torch.tensor([2.0, 4.0, 0.0]))                          # half-to-even, clamp at 0


def test_algorithmic_flops_match_survey_totals():
    # C3 = 25.72 TFLOP (B=64, L=200, T=600), per-frame HiFi-GAN 614.105 MFLOP
    assert restate.M_HG * 2 == 614105088
    f = restate.algorithmic_flops([200] * 64, [600] * 64)
    assert abs(f / 1e12 - 25.72) < 0.01


@pytest.mark.skipif(not ref_import.available(), reason="reference tree not mounted (GPU box)")
def test_restatement_matches_live_reference():
    sd = syn.make_fs2_state_dict(seed=3, duration_mode="spread")
    ref = ref_import.build_fs2(sd, syn.speakers_json(), syn.STATS_JSON)
    b = syn.make_batch(2, 16, seed=9, lens=[16, 11])
    torch.manual_seed(1)
    emb = ref.speaker_gen(b["speaker_meta"])
    with torch.no_grad():
        want = ref.synthesize_from_speaker_emb(None, b["texts"], b["src_lens"], b["max_src_len"], accents=b["accents"], speaker_emb=emb)
    got = restate.fs2_forward(sd, b["texts"], b["accents"], b["src_lens"], b["max_src_len"], emb)
    for i in range(8):
        assert torch.allclose(want[i].float(), got[i].float(), atol=1e-5), i
    assert torch.equal(want[9], got[9])
    # long-sequence position table (Models.py:86-98 rebuild) is bit-identical
    from transformer.Models import get_sinusoid_encoding_table
    assert torch.equal(get_sinusoid_encoding_table(1200, 256), restate.sinusoid_table(1200))
    assert torch.equal(get_sinusoid_encoding_table(1200, 256), syn.sinusoid_table(1200))


def test_fs2_restatement_matches_reference_golden_on_real_sequences(golden_dir):
    """fs2_jsut.npz: the reference on four real (phoneme, accent) sequences tokenised from the vendored jsut-lab labels."""
    g = _load(golden_dir, "fs2_jsut")
    sd = syn.make_fs2_state_dict(seed=int(g["weights_seed"]), duration_mode=str(g["duration_mode"]))
    texts, acc, lens = (torch.from_numpy(g[k]) for k in ("texts", "accents", "src_lens"))
    out = restate.fs2_forward(sd, texts, acc, lens, texts.shape[1], torch.from_numpy(g["speaker_emb"]))
    assert torch.equal(out[9], torch.from_numpy(g["mel_lens"])) and torch.equal(out[5], torch.from_numpy(g["d_rounded"]))
    for i, key in ((0, "mel"), (1, "postnet_mel"), (2, "pitch"), (3, "energy"), (4, "log_d")):
        np.testing.assert_allclose(out[i].numpy(), g[key], atol=3e-5, rtol=0)
    # teacher-forced targets restated as the reference does (modules.py:80-100): forcing the reference's own outputs changes nothing
    out_f = restate.fs2_forward(sd, texts, acc, lens, texts.shape[1], torch.from_numpy(g["speaker_emb"]), p_target=torch.from_numpy(g["pitch"]),
                                e_target=torch.from_numpy(g["energy"]), d_target=torch.from_numpy(g["d_rounded"]))
    np.testing.assert_allclose(out_f[1].numpy(), g["postnet_mel"], atol=3e-5, rtol=0)
