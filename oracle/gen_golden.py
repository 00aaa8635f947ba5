"""Generate tests/golden/*.npz by running the UNMODIFIED reference (/root/reference, imported
read-only through oracle/ref_import.py) on seeded inputs and weights.  Run in the build container
only:  python -m oracle.gen_golden.   TEST INFRASTRUCTURE ONLY."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import masg_b200  # noqa: E402
from masg_b200 import synthetic as syn  # noqa: E402
from oracle import ref_import  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")

FS2_CASES = {
    # name: (weights seed, duration mode, batch seed, lens, p_control, d_control)
    "fs2_ragged": (11, "spread", 5, [24, 17, 9], 0.9, 1.3),
    "fs2_c1": (1234, "fixed3", 1235, [50], 1.0, 1.0),
}


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(0)
    torch.set_num_threads(1)          # deterministic reduction order for the fixtures
    ns = ref_import.load()
    for name, (wseed, dmode, bseed, lens, pc, dc) in FS2_CASES.items():
        sd = syn.make_fs2_state_dict(seed=wseed, duration_mode=dmode)
        ref = ref_import.build_fs2(sd, syn.speakers_json(), syn.STATS_JSON)
        b = syn.make_batch(len(lens), max(lens), seed=bseed, lens=lens)
        torch.manual_seed(bseed)
        emb = ref.speaker_gen(b["speaker_meta"])
        with torch.no_grad():
            out = ref.synthesize_from_speaker_emb(None, b["texts"], b["src_lens"], b["max_src_len"], accents=b["accents"],
                                                  speaker_emb=emb, p_control=pc, d_control=dc)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), texts=b["texts"].numpy(), accents=b["accents"].numpy(),
                            src_lens=b["src_lens"].numpy(), speaker_emb=emb.numpy(), p_control=pc, d_control=dc,
                            weights_seed=wseed, duration_mode=dmode,
                            mel=out[0].numpy(), postnet_mel=out[1].numpy(), pitch=out[2].numpy(), energy=out[3].numpy(),
                            log_d=out[4].numpy(), d_rounded=out[5].numpy(), src_mask=out[6].numpy(), mel_mask=out[7].numpy(),
                            mel_lens=out[9].numpy())
        print(name, "T =", out[0].shape[1], "mel_lens =", out[9].tolist())
    # HiFi-GAN: weight-normed checkpoint layout → reference generator
    hsd = syn.make_hifigan_state_dict(seed=4321)
    g = ref_import.build_hifigan(hsd)
    gen = torch.Generator().manual_seed(77)
    mel = torch.randn(2, 80, 20, generator=gen)
    with torch.no_grad():
        wav = g(mel)
    np.savez_compressed(os.path.join(OUT, "hifigan_small.npz"), mel=mel.numpy(), wav=wav.numpy(), weights_seed=4321)
    print("hifigan", wav.shape, float(wav.std()))
    # speaker prior: SpeakerMetaEncoder params and InterpolateGMM (with the oracle's exact EMD bound to ot.emd)
    sd = syn.make_fs2_state_dict(seed=1234)
    ref = ref_import.build_fs2(sd, syn.speakers_json(), syn.STATS_JSON)
    metas = torch.tensor(syn.META_COMBOS, dtype=torch.float32)
    dists = [ref.speaker_distribution(metas[i:i + 1]) for i in range(4)]
    pi = torch.cat([d.mixture_distribution.probs for d in dists]).numpy()
    mu = torch.cat([d.component_distribution.base_dist.loc for d in dists]).numpy()
    sigma = torch.cat([d.component_distribution.base_dist.scale for d in dists]).numpy()
    ig = ns.InterpolateGMM(dists[0], dists[3])
    rec = dict(pi=pi, mu=mu, sigma=sigma, cost=np.array(ig.ot_Cost, dtype=np.float64),
               weights=ig.mixture_distribution.probs[0].numpy(),
               mu_050=ig.component_distribution.base_dist.loc[0].numpy(),
               scale_050=ig.component_distribution.base_dist.scale[0].numpy())
    ig.interpolate_rate(0.25)
    rec["mu_025"] = ig.component_distribution.base_dist.loc[0].numpy()
    rec["scale_025"] = ig.component_distribution.base_dist.scale[0].numpy()
    np.savez_compressed(os.path.join(OUT, "gmm_prior.npz"), weights_seed=1234, **rec)
    print("gmm weights", rec["weights"])
    barycenter(ref)


def frontend_fixture():
    """tests/golden/frontend_jsut.npz: the reference's own tokenisation (pp_symbols -> openjtalk2julius -> symbol_to_id /
    accent_to_id, synthesize.py:96-101,215-234) of every 80th utterance of the vendored jsut-lab full-context labels
    (MIT-licensed label data, jsut-lab/LICENSE.md), with the symbol inventory (text/symbols.py) it was run with; and
    tests/golden/fs2_jsut.npz: the reference FastSpeech2 on a ragged batch of four of those real (phoneme, accent) sequences."""
    import glob
    ref_import.load()
    from prepare_tg_accent_jsut import pp_symbols
    from convert_label import openjtalk2julius
    from text import symbols
    files = sorted(glob.glob(os.path.join(ref_import.REF_ROOT, "jsut-lab", "*", "lab", "*.lab")))[::80]
    sym2id = {s_: i for i, s_ in enumerate(symbols)}
    acc2id = {"0": 0, "[": 1, "]": 2, "#": 3}
    labels, ids, accs, names = [], [], [], []
    for f in files:
        lab = [ln.strip().rsplit(None, 1)[-1] for ln in open(f) if ln.strip()]
        ph, ac = pp_symbols(lab)
        ph = [openjtalk2julius(p_) for p_ in ph if p_ != ""]
        labels.append("\n".join(lab)); names.append(os.path.basename(f))
        ids.append(np.array([sym2id[p_] for p_ in ph], dtype=np.int64)); accs.append(np.array([acc2id[a] for a in ac], dtype=np.int64))
    lens = np.array([len(x) for x in ids], dtype=np.int64)
    L = int(lens.max())
    pad = lambda rows: np.stack([np.pad(r, (0, L - len(r))) for r in rows])
    np.savez_compressed(os.path.join(OUT, "frontend_jsut.npz"), symbols=np.array(symbols), names=np.array(names), labels=np.array(labels),
                        texts=pad(ids), accents=pad(accs), src_lens=lens)
    print("frontend fixture:", len(files), "utterances, lengths", lens.min(), "-", lens.max())
    # FastSpeech2 on real sequences: the longest, the shortest and two mid-length utterances of the sample
    order = np.argsort(-lens)
    pick = [int(order[0]), int(order[len(order) // 3]), int(order[2 * len(order) // 3]), int(order[-1])]
    texts = torch.from_numpy(pad(ids)[pick]); accents = torch.from_numpy(pad(accs)[pick]); src_lens = torch.from_numpy(lens[pick])
    sd = syn.make_fs2_state_dict(seed=11, duration_mode="spread")
    ref = ref_import.build_fs2(sd, syn.speakers_json(), syn.STATS_JSON)
    torch.manual_seed(5)
    meta = torch.tensor([syn.META_COMBOS[i % 4] for i in range(4)], dtype=torch.float32)
    emb = ref.speaker_gen(meta)
    with torch.no_grad():
        out = ref.synthesize_from_speaker_emb(None, texts, src_lens, L, accents=accents, speaker_emb=emb)
    np.savez_compressed(os.path.join(OUT, "fs2_jsut.npz"), texts=texts.numpy(), accents=accents.numpy(), src_lens=src_lens.numpy(),
                        speaker_emb=emb.numpy(), p_control=1.0, d_control=1.0, weights_seed=11, duration_mode="spread", picked=np.array(pick),
                        mel=out[0].numpy(), postnet_mel=out[1].numpy(), pitch=out[2].numpy(), energy=out[3].numpy(), log_d=out[4].numpy(),
                        d_rounded=out[5].numpy(), src_mask=out[6].numpy(), mel_mask=out[7].numpy(), mel_lens=out[9].numpy())
    print("fs2_jsut: lens", src_lens.tolist(), "T =", out[0].shape[1], "mel_lens", out[9].tolist())


BARY_RATE = [0.5, 0.25, 0.125, 0.125]      # sums to 1 exactly in floating point (distributions.py:112 asserts equality)


def barycenter(ref=None):
    """tests/golden/barycenter.npz: BarycenterGMM run by the reference's own methods (ref_import.run_barycenter) at the
    default uniform rate and at BARY_RATE."""
    if ref is None:
        ref = ref_import.build_fs2(syn.make_fs2_state_dict(seed=1234), syn.speakers_json(), syn.STATS_JSON)
    rec = {}
    with torch.no_grad():
        for tag, rate in (("uniform", None), ("rate", BARY_RATE)):
            probs, mean, std, _ = ref_import.run_barycenter(ref, rate)
            rec["probs_" + tag], rec["mean_" + tag], rec["std_" + tag] = probs.numpy(), mean.numpy(), std.numpy()
            print("barycenter", tag, "components", probs.shape[0], "probs", probs.numpy().round(4))
    np.savez_compressed(os.path.join(OUT, "barycenter.npz"), weights_seed=1234, rate=np.array(BARY_RATE), **rec)
    distri_artefacts(ref)


def distri_artefacts(ref=None):
    """tests/golden/distri_ref_{interp,bary}.pth: ``torch.save`` of the REFERENCE's distribution objects, i.e. the on-disk
    artefacts examples_gen_distri.py:211 loads (class path model.distributions.*), for the loader tests."""
    ns = ref_import.load()
    if ref is None:
        ref = ref_import.build_fs2(syn.make_fs2_state_dict(seed=1234), syn.speakers_json(), syn.STATS_JSON)
    metas = torch.tensor(syn.META_COMBOS, dtype=torch.float32)
    with torch.no_grad():
        ig = ns.InterpolateGMM(ref.speaker_distribution(metas[1:2]), ref.speaker_distribution(metas[2:3]))
        ig.interpolate_rate(0.75)
        torch.save(ig, os.path.join(OUT, "distri_ref_interp.pth"))
        _, _, _, bg = ref_import.run_barycenter(ref, BARY_RATE)
        bg.__class__ = ns.BarycenterGMM                  # the harness subclass is local; the artefact carries the reference class
        torch.save(bg, os.path.join(OUT, "distri_ref_bary.pth"))
    print("distribution artefacts:", [os.path.getsize(os.path.join(OUT, f)) for f in ("distri_ref_interp.pth", "distri_ref_bary.pth")])


if __name__ == "__main__":
    if "--barycenter" in sys.argv:
        barycenter()
    elif "--distri" in sys.argv:
        distri_artefacts()
    elif "--frontend" in sys.argv:
        frontend_fixture()
    else:
        main()
