"""Shared test helpers: build the drop-in modules on the GPU from seeded weights, run the oracle."""
import json
import os
import tempfile

import numpy as np
import torch

import masg_b200  # noqa: F401
from masg_b200 import synthetic as syn

_cfg_dir = None


def config_dir():
    """Fabricated config dir (the reference's speakers.json/stats.json are 0-byte files)."""
    global _cfg_dir
    if _cfg_dir is None:
        _cfg_dir = tempfile.mkdtemp(prefix="masg_cfg_")
        json.dump(syn.speakers_json(), open(os.path.join(_cfg_dir, "speakers.json"), "w"))
        json.dump(syn.STATS_JSON, open(os.path.join(_cfg_dir, "stats.json"), "w"))
    return _cfg_dir


_models = {}


def fs2_on_gpu(seed, duration_mode, precision):
    from masg_b200.fastspeech2 import FastSpeech2
    key = ("fs2", seed, duration_mode, precision)
    if key not in _models:
        sd = syn.make_fs2_state_dict(seed=seed, duration_mode=duration_mode)
        m = FastSpeech2(syn.PREPROCESS_CONFIG, syn.MODEL_CONFIG, config_dir(), precision=precision)
        m.load_state_dict(sd)
        m.eval().to("cuda:0")
        _models[key] = (m, sd)
    return _models[key]


def hifigan_on_gpu(seed, precision):
    from masg_b200.hifigan import Generator, AttrDict
    key = ("hg", seed, precision)
    if key not in _models:
        sd = syn.make_hifigan_state_dict(seed=seed)
        g = Generator(AttrDict(syn.HIFIGAN_CONFIG), precision=precision)
        g.load_state_dict(sd)
        g.eval()
        import contextlib, io
        with contextlib.redirect_stdout(io.StringIO()):
            g.remove_weight_norm()
        g.to("cuda:0")
        _models[key] = (g, sd)
    return _models[key]


def snr_db(ref, out):
    ref = np.asarray(ref, dtype=np.float64)
    out = np.asarray(out, dtype=np.float64)
    return 10.0 * np.log10((ref ** 2).sum() / max(((ref - out) ** 2).sum(), 1e-30))


def to_gpu(batch):
    return {k: (v.to("cuda:0") if torch.is_tensor(v) else v) for k, v in batch.items()}


def check_unforced_durations(out_log_d, ref_log_d, out_d, ref_d, tag=""):
    """Un-forced duration parity (north_star: durations bit-exact given identical predictor outputs).  The predictor outputs
    of a reduced-precision mode are not identical, so: the integer durations must still be EQUAL, except at positions where
    the reference's own exp(log_d) - 1 sits on a half-integer rounding boundary to within twice the measured log_d error
    (there the reference's decision is itself an accident of rounding).  Returns the number of such boundary flips."""
    out_log_d, ref_log_d = out_log_d.detach().cpu().double(), torch.as_tensor(ref_log_d).double()
    out_d, ref_d = out_d.detach().cpu(), torch.as_tensor(ref_d)
    err = (out_log_d - ref_log_d).abs().max().item()
    flips = out_d != ref_d
    n = int(flips.sum())
    print("un-forced durations %s: log_d max-abs err %.2e, flips %d of %d" % (tag, err, n, flips.numel()))
    if n:
        d = torch.exp(ref_log_d) - 1
        dist = ((d - torch.floor(d)) - 0.5).abs()[flips]
        assert (dist <= 2 * err * torch.exp(ref_log_d)[flips] + 1e-7).all(), \
            "%s: duration flip away from a rounding boundary (boundary distances %s, log_d err %.2e)" % (tag, dist.tolist()[:8], err)
    return n


def check_unforced_buckets(out_v, ref_v, bins, tag="", assert_boundary=True):
    """Un-forced pitch / energy bucket parity (torch.bucketize of the predictor output, model/modules.py:85-88): the bucket
    index must be EQUAL, except where the reference's own value sits within twice the measured prediction error of a bin
    boundary (an adjacent-bucket flip there is the reference's rounding accident, exactly like a duration on a .5 boundary).
    assert_boundary=False only reports (used for the un-forced ENERGY, whose predictor input already differs wherever a pitch
    bucket flipped upstream; it is checked again with the pitch teacher-forced).  Returns a bool mask [B, L] of the flips."""
    out_v, ref_v, bins = out_v.detach().cpu().float(), torch.as_tensor(ref_v).float(), torch.as_tensor(bins).float().cpu()
    err = (out_v - ref_v).abs().max().item()
    io, ir = torch.bucketize(out_v, bins), torch.bucketize(ref_v, bins)
    flips = io != ir
    n = int(flips.sum())
    print("un-forced buckets %s: max-abs err %.2e, flips %d of %d" % (tag, err, n, flips.numel()))
    if n and assert_boundary:
        assert int((io - ir).abs().max()) == 1, "%s: bucket flip by more than one bin" % tag
        dist = (ref_v[flips][:, None] - bins[None, :]).abs().min(1).values
        assert (dist <= 2 * err + 1e-7).all(), "%s: bucket flip away from a bin boundary (distances %s, err %.2e)" % (tag, dist.tolist()[:8], err)
    return flips
