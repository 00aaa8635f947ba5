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


if __name__ == "__main__":
    if "--barycenter" in sys.argv:
        barycenter()
    else:
        main()
