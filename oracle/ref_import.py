"""Import the *unmodified* reference read-only with the five absent third-party modules stubbed:
from /root/reference in the build container, or from the byte-for-byte staged copy ``oracle/_ref`` (written by
oracle/make_ref.py, git-ignored, shipped with the snapshot) on the GPU box, where /root/reference does not exist.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
"""
import json
import os
import shutil
import sys
import tempfile
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
_STAGED = os.path.join(_HERE, "_ref")            # oracle/make_ref.py: byte-for-byte staged copy that travels to the GPU box
_CFG_REL = "config/JVS-VCTK_langemb_configs/JVS-VCTK_1"


def _resolve_root():
    env = os.environ.get("MASG_REFERENCE_ROOT")
    for cand in ([env] if env else []) + ["/root/reference", _STAGED]:
        if cand and os.path.isfile(os.path.join(cand, "model", "fastspeech2.py")):
            return cand
    return env or "/root/reference"


REF_ROOT = _resolve_root()


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "model", "fastspeech2.py"))


def kind() -> str:
    """'live' = /root/reference itself (build container), 'staged' = oracle/_ref (GPU box), 'absent'."""
    if not available():
        return "absent"
    return "staged" if os.path.abspath(REF_ROOT) == os.path.abspath(_STAGED) else "live"


def _stub(name, **attrs):
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules.setdefault(name, m)
    return sys.modules[name]


def install_stubs(emd=None):
    """Stub modules the reference imports but this image lacks.  None is touched by inference
    maths except ``ot.emd`` (model/distributions.py:22), which is bound to the oracle's exact
    transportation solver (POT is un-vendored and unpinned: absent from requirements.txt)."""
    if emd is None:
        from .restate import emd_exact as emd
    ot = _stub("ot")
    ot.emd = lambda a, b, M: emd(a, b, M)
    mpl = _stub("matplotlib", use=lambda *a, **k: None)
    plt = _stub("matplotlib.pyplot")
    mpl.pyplot = plt
    _stub("unidecode", unidecode=lambda s: s)
    _stub("inflect", engine=lambda: None)
    _stub("Multilingual-Speaker-Encoder-with-Domain-Adaptation", SpeechEmbedder=object, GE2ELoss=object)


_loaded = {}


def load():
    """Returns a namespace with the reference classes (FastSpeech2, Generator, AttrDict,
    InterpolateGMM, BarycenterGMM, get_mask_from_lengths)."""
    if _loaded:
        return _loaded["ns"]
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    install_stubs()
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    import model as ref_model                     # noqa: model/__init__.py
    import hifigan as ref_hifigan                 # noqa
    from model import distributions as ref_dist  # noqa
    from utils import tools as ref_tools          # noqa
    ns = types.SimpleNamespace(
        FastSpeech2=ref_model.FastSpeech2,
        Generator=ref_hifigan.Generator,
        AttrDict=ref_hifigan.AttrDict,
        InterpolateGMM=ref_dist.InterpolateGMM,
        BarycenterGMM=ref_dist.BarycenterGMM,
        get_mask_from_lengths=ref_tools.get_mask_from_lengths,
        vocoder_infer=__import__("utils.model", fromlist=["vocoder_infer"]).vocoder_infer,
        root=REF_ROOT,
    )
    _loaded["ns"] = ns
    return ns


def load_configs():
    import yaml
    d = os.path.join(REF_ROOT, _CFG_REL)
    pre = yaml.load(open(os.path.join(d, "preprocess.yaml")), Loader=yaml.FullLoader)
    mod = yaml.load(open(os.path.join(d, "model.yaml")), Loader=yaml.FullLoader)
    return pre, mod


def hifigan_config():
    return json.load(open(os.path.join(REF_ROOT, "hifigan", "config.json")))


def build_fs2(state_dict, speakers_json, stats_json):
    """Reference FastSpeech2 with a fabricated config dir (the shipped speakers.json/stats.json are
    0-byte files) and the given weights; eval mode, CPU."""
    ns = load()
    pre, mod = load_configs()
    tmp = tempfile.mkdtemp(prefix="masg_cfg_")
    try:
        json.dump(speakers_json, open(os.path.join(tmp, "speakers.json"), "w"))
        json.dump(stats_json, open(os.path.join(tmp, "stats.json"), "w"))
        m = ns.FastSpeech2(pre, mod, tmp)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    missing, unexpected = m.load_state_dict(state_dict, strict=False)
    bad = [k for k in missing if "num_batches_tracked" not in k]
    assert not bad and not unexpected, (bad, unexpected)
    return m.eval()


def build_hifigan(state_dict_weight_normed):
    """Reference generator loaded the way utils/model.py:61-69 does: load weight-normed keys,
    eval(), remove_weight_norm()."""
    import contextlib
    import io
    ns = load()
    g = ns.Generator(ns.AttrDict(hifigan_config()))
    g.load_state_dict(state_dict_weight_normed)
    g.eval()
    with contextlib.redirect_stdout(io.StringIO()):
        g.remove_weight_norm()
    return g


def run_barycenter(model, rate=None):
    """The reference's BarycenterGMM (model/distributions.py:79-192) on `model`, executed by the reference's own code.

    As written the constructor raises TypeError: :92 calls ``self._barycenter_gaussians(_print=False)`` while :136 defines
    ``_barycenter_gaussians(self)``.  The subclass below overrides that ONE method to swallow the
    keyword and forward to the reference's implementation; ``__init__``, ``_product``, ``_one_barycenter_gaussian``,
    ``_determine_pi``, ``_w2sq`` and ``barycenter_rate`` all run unmodified.  This is what pins a20.
    Returns (probs [n], mean [n, D], std [n, D], used positions) of the resulting mixture."""
    import contextlib
    import io
    import tqdm as _tqdm
    ns = load()

    class _Runnable(ns.BarycenterGMM):
        def _barycenter_gaussians(self, _print=True):
            return ns.BarycenterGMM._barycenter_gaussians(self)

    with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
        bg = _Runnable(model)
        if rate is not None:
            bg.barycenter_rate(list(rate), _print=False)
    comp = bg.component_distribution.base_dist
    return bg.mixture_distribution.probs[0], comp.loc[0], comp.scale[0], bg
