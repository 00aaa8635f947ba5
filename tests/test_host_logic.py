"""CPU tests: the C-ABI library loads and exports every symbol include/masg.h declares, the product
path fails loudly without a GPU (no CPU fallback), drop-in state_dict layouts match the reference's,
and the batch scheduler / multi-rank sharding logic (gloo, world_size 2)."""
import ctypes
import json
import os
import re
import socket
import sys

import numpy as np
import pytest
import torch

import masg_b200  # noqa: F401
from masg_b200 import synthetic as syn
from masg_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib_path():
    import __graft_entry__ as ge
    ge.build()
    from masg_b200 import _lib
    return _lib.LIB_PATH


def test_library_exports_every_header_symbol(lib_path):
    header = open(os.path.join(ROOT, "include", "masg.h")).read()
    declared = set(re.findall(r"\b(masg_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 20
    lib = ctypes.CDLL(lib_path)
    for name in declared:
        assert hasattr(lib, name), "libmasg.so does not export %s" % name
    from masg_b200 import _lib
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    lib.masg_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.masg_version()


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(lib_path):
    from masg_b200 import _lib
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.Context("cuda:0")
    from tests.helpers import config_dir
    from masg_b200.fastspeech2 import FastSpeech2
    m = FastSpeech2(syn.PREPROCESS_CONFIG, syn.MODEL_CONFIG, config_dir()).eval()
    b = syn.make_batch(1, 5, seed=0)
    with pytest.raises(RuntimeError, match="CUDA"):
        m.synthesize_from_speaker_emb(None, b["texts"], b["src_lens"], 5, accents=b["accents"], speaker_emb=torch.zeros(1, 256))
    from masg_b200.hifigan import Generator, AttrDict
    with pytest.raises(RuntimeError, match="CUDA"):
        Generator(AttrDict(syn.HIFIGAN_CONFIG))(torch.zeros(1, 80, 4))


def test_dropin_state_dict_layouts():
    from tests.helpers import config_dir
    from masg_b200.fastspeech2 import FastSpeech2
    from masg_b200.hifigan import Generator, AttrDict
    m = FastSpeech2(syn.PREPROCESS_CONFIG, syn.MODEL_CONFIG, config_dir())
    sd = syn.make_fs2_state_dict(seed=0)
    msd = m.state_dict()
    assert set(msd) == set(sd) and len(sd) == 242
    for k in sd:
        assert tuple(msd[k].shape) == tuple(sd[k].shape), k
    assert sum(p.numel() for p in m.parameters()) == 35239248
    m.load_state_dict(sd)
    g = Generator(AttrDict(syn.HIFIGAN_CONFIG))
    hsd = syn.make_hifigan_state_dict(seed=0)
    assert set(g.state_dict()) == set(hsd) and len(hsd) == 234
    g.load_state_dict(hsd)
    import contextlib, io
    with contextlib.redirect_stdout(io.StringIO()):
        g.remove_weight_norm()
    folded = syn.fold_weight_norm(hsd)
    gsd = g.state_dict()
    assert set(gsd) == set(folded) and len(folded) == 156
    for k in folded:
        assert torch.allclose(gsd[k], folded[k], atol=1e-6), k
    assert sum(v.numel() for v in folded.values()) == 13926017


def test_unsupported_config_is_an_error():
    import copy
    from tests.helpers import config_dir
    from masg_b200.fastspeech2 import FastSpeech2
    cfg = copy.deepcopy(syn.MODEL_CONFIG)
    cfg["jdit"]["use_jdit"] = True
    with pytest.raises(NotImplementedError):
        FastSpeech2(syn.PREPROCESS_CONFIG, cfg, config_dir())


def test_scheduler_covers_all_utterances_and_balances():
    lens = syn.c4_lengths()
    assert lens.sum() * 3 == 1978896
    batches = synth.plan_batches(lens, frame_budget=40000, max_batch=64)
    flat = np.concatenate(batches)
    assert sorted(flat.tolist()) == list(range(4096))
    for b in batches:
        assert len(b) <= 64 and len(b) * lens[b].max() * 3 <= 40000
    for ws in (1, 2, 4, 8):
        plan = synth.assign_batches(batches, lens, ws)
        assert sorted(sum(plan, [])) == list(range(len(batches)))
        loads = [sum(synth.utterance_cost(lens[i]) for bi in p for i in batches[bi]) for p in plan]
        assert max(loads) / (sum(loads) / ws) < 1.05                         # LPT keeps ranks within 5 %
    assert [synth.shard_batch(64, r, 8) for r in range(8)] == [(8 * r, 8 * r + 8) for r in range(8)]
    assert synth.shard_batch(16, 7, 8) == (14, 16) and synth.shard_batch(3, 3, 4) == (3, 3)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    lens = syn.c4_lengths(256)
    batches = synth.plan_batches(lens, frame_budget=20000, max_batch=32)
    mine = synth.assign_batches(batches, lens, world)[rank]
    frames = torch.tensor([float(sum(int(lens[i]) * 3 for bi in mine for i in batches[bi]))])
    count = torch.tensor([float(sum(len(batches[bi]) for bi in mine))])
    dist.all_reduce(frames); dist.all_reduce(count)
    # max-over-ranks timing reduction used by bench.py
    t = torch.tensor([1.0 + rank]); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    q.put((rank, frames.item(), count.item(), t.item()))
    dist.destroy_process_group()


def test_two_rank_sharding_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue(); port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = [q.get(timeout=120) for _ in ps]
    [p.join(timeout=60) for p in ps]
    lens = syn.c4_lengths(256)
    for rank, frames, count, tmax in res:
        assert frames == float(lens.sum() * 3) and count == 256.0 and tmax == 2.0


def test_speaker_embedding_artefacts_round_trip(tmp_path):
    """Reference on-disk format of generated speakers: one float32 [1, 256] .npy per speaker (examples_gen.py:272)."""
    import numpy as np
    import torch
    from masg_b200 import synth
    names = synth.speaker_names("F", "ja", 3)
    assert names == ["F_jagen001", "F_jagen002", "F_jagen003"]
    g = torch.Generator().manual_seed(3)
    embs = torch.randn(3, 256, generator=g)
    paths = synth.save_speaker_embeddings(tmp_path / "speakers_wsample", names, embs)
    one = np.load(paths[1])
    assert one.shape == (1, 256) and one.dtype == np.float32                 # what the reference's loaders expect
    back = synth.load_speaker_embeddings(paths)
    assert back.shape == (3, 256) and torch.equal(back, embs)
    np.save(tmp_path / "flat.npy", embs[0].numpy())                            # a [256] array is accepted too
    assert torch.equal(synth.load_speaker_embeddings([tmp_path / "flat.npy"])[0], embs[0])


def test_bench_reference_arm_prints_the_contract_line():
    """bench.py --impl reference (the unmodified reference when a tree is present, else the CPU oracle port, timed on the host cores) prints one JSON line with the driver's keys;
    the sharded workloads are selectable without a GPU being touched at argument-parsing time."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-sample", "1",
                          "--latency-reps", "1"],
                         capture_output=True, text=True, timeout=600, cwd=root)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "audio-s/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert line["gpu_launches"] == 0 and "C3" in line["config"]["workload"]
    import bench
    assert line["config"] == bench.c3_config(1)                              # the reference arm runs on OUR arm's config object
    assert line["cpu_baseline"]["c1_latency_p50_ms"]["value"] > 0 and line["cpu_baseline"]["c2_latency_p50_ms"]["value"] > 0
    helptext = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--help"], capture_output=True, text=True, timeout=600, cwd=root).stdout
    assert "--workload" in helptext and "c4" in helptext and "c5" in helptext


# ------------------------------------------------------------------------------------------------
# round 2: wav emission, synth_samples drop-in, distribution artefacts, waveform gather, signatures
# ------------------------------------------------------------------------------------------------
def test_wav_files_are_byte_identical_to_scipy(tmp_path):
    """utils/tools.py:272-274 writes with scipy.io.wavfile.write; the background writer's files must be the same bytes."""
    from scipy.io import wavfile
    from masg_b200 import wavio
    rng = np.random.default_rng(0)
    arrays = [rng.integers(-32768, 32768, n).astype(np.int16) for n in (0, 1, 255, 22050, 153600)]
    with wavio.WavWriter(n_threads=3) as w:
        for i, a in enumerate(arrays):
            w.submit(tmp_path / "sub" / ("o%d.wav" % i), 22050, a)
    assert w.files == len(arrays)
    for i, a in enumerate(arrays):
        wavfile.write(tmp_path / ("r%d.wav" % i), 22050, a)
        assert (tmp_path / "sub" / ("o%d.wav" % i)).read_bytes() == (tmp_path / ("r%d.wav" % i)).read_bytes()
    with pytest.raises(ValueError):
        wavio.write_wav(tmp_path / "x.wav", 22050, np.zeros(4, dtype=np.float32))


class _FakeVocoder(torch.nn.Module):
    """A deterministic stand-in with BOTH call surfaces (test double, not a fallback): the reference's ``vocoder(mels[B,80,T])
    -> [B,1,256T]`` and the drop-in's ``forward_packed(mel[B,T,80], mel_lens, channels_last=True)``."""

    def forward(self, mels):
        B, C, T = mels.shape
        ramp = torch.linspace(-1.0, 1.0, 256)
        return torch.tanh(mels.mean(1)[:, :, None] * 0.5 + ramp[None, None, :] * mels[:, 3, :, None]).reshape(B, 1, T * 256)

    def forward_packed(self, mel_btc, mel_lens, channels_last=True, hop=256):
        from oracle import restate
        wav = self.forward(mel_btc.transpose(1, 2))
        w16 = restate.wav_to_int16(wav.squeeze(1))
        lens = torch.as_tensor(mel_lens).to(torch.int64)
        off = torch.zeros(len(lens) + 1, dtype=torch.int64)
        off[1:] = torch.cumsum(lens * hop, 0)
        packed = np.concatenate([w16[i][: int(lens[i]) * hop] for i in range(len(lens))])
        return torch.from_numpy(packed), off


def _fake_predictions(B=3, L=7, seed=0):
    g = torch.Generator().manual_seed(seed)
    src_lens = torch.tensor([7, 5, 2])
    d = torch.randint(0, 4, (B, L), generator=g).float()
    d[torch.arange(L)[None, :] >= src_lens[:, None]] = 0
    d[0, 0] = 2.0
    mel_lens = d.long().sum(1)
    T = int(mel_lens.max())
    post = torch.randn(B, T, 80, generator=g)
    return (post * 0.9, post, torch.randn(B, L, generator=g), torch.randn(B, L, generator=g), torch.randn(B, L, generator=g), d,
            None, None, src_lens, mel_lens)


def test_synth_samples_dropin_writes_the_reference_files(tmp_path):
    """utils/tools.py:228-274: same wav files (bytes) and the same arrays handed to plot_mel as the reference function produces
    from the same predictions — with one stats.json read and no per-utterance .item()."""
    import types
    from unittest import mock
    from masg_b200 import tools
    from oracle import ref_import
    cfg = tmp_path / "cfg"; cfg.mkdir()
    json.dump(syn.STATS_JSON, open(cfg / "stats.json", "w"))
    args = types.SimpleNamespace(config=str(cfg))
    pred = _fake_predictions()
    targets = (["utt_a", "utt_b", "utt_c"],)
    voc = _FakeVocoder()
    ours = tmp_path / "ours"; ours.mkdir()
    calls = []

    def rec_plot(data, stats, titles):
        calls.append((data, stats, titles))
        return None

    opened = []
    real_open = open

    def counting_open(p, *a, **k):
        if str(p).endswith("stats.json"):
            opened.append(p)
        return real_open(p, *a, **k)

    with mock.patch("builtins.open", counting_open):
        paths = tools.synth_samples(args, targets, pred, voc, syn.MODEL_CONFIG, syn.PREPROCESS_CONFIG, str(ours), plot=True, plot_fn=rec_plot)
    assert len(opened) == 1 and [os.path.basename(p) for p in paths] == ["utt_a.wav", "utt_b.wav", "utt_c.wav"]
    assert len(calls) == 3 and calls[0][2] == ["Synthetized Spectrogram"] and calls[0][1] == syn.STATS_JSON["pitch"] + syn.STATS_JSON["energy"][:2]
    lens = pred[9].tolist()
    from scipy.io import wavfile
    for i, p in enumerate(paths):
        rate, data = wavfile.read(p)
        assert rate == 22050 and data.dtype == np.int16 and len(data) == lens[i] * 256
        mel_i, pitch_i, energy_i = calls[i][0][0]
        assert mel_i.shape == (80, lens[i]) and len(pitch_i) == lens[i] == len(energy_i)
    no_plot = tools.synth_samples(args, targets, pred, voc, syn.MODEL_CONFIG, syn.PREPROCESS_CONFIG, str(ours), basename="same")
    assert [os.path.basename(p) for p in no_plot] == ["same.wav"] * 3                 # examples_gen.py:68-109 variant
    with pytest.raises(ValueError):
        tools.synth_samples(args, targets, pred, voc, syn.MODEL_CONFIG, syn.PREPROCESS_CONFIG, str(ours), plot=True)
    if not ref_import.available():
        return
    # ---- the reference function itself on the same predictions (matplotlib is a stub here: capture plot_mel's arguments) ----
    ref_import.load()
    import utils.tools as rt
    refdir = tmp_path / "ref"; refdir.mkdir()
    ref_calls = []
    with mock.patch.object(rt, "plot_mel", lambda data, stats, titles: ref_calls.append((data, stats, titles))), \
            mock.patch.object(rt, "plt", mock.MagicMock()), mock.patch.object(rt, "device", torch.device("cpu")):
        rt.synth_samples(args, targets, pred, voc, syn.MODEL_CONFIG, syn.PREPROCESS_CONFIG, str(refdir))
    for name in ("utt_a", "utt_b", "utt_c"):
        assert (refdir / (name + ".wav")).read_bytes() == (ours / (name + ".wav")).read_bytes()
    assert len(ref_calls) == 3
    for (d0, s0, t0), (d1, s1, t1) in zip(ref_calls, calls):
        assert s0 == s1 and t0 == t1
        for a, b in zip(d0[0], d1[0]):
            assert np.array_equal(np.asarray(a), np.asarray(b))
    # expand(): the vectorised form against the reference's loop, negative and fractional durations included
    rng = np.random.default_rng(1)
    for _ in range(20):
        v = rng.standard_normal(9).astype(np.float32); dd = rng.uniform(-2, 4, 9).astype(np.float32)
        assert np.array_equal(tools.expand(v, dd), rt.expand(v, dd))


def test_reference_pickled_distributions_are_readable_without_the_reference(golden_dir):
    """examples_gen_distri.py:211 unpickles model.distributions.InterpolateGMM; read_reference_distribution must not need the
    reference package (run in a child interpreter whose sys.path has no reference tree)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "import masg_b200, numpy as np\n"
        "from masg_b200 import distributions as d\n"
        "assert 'model' not in sys.modules\n"
        "p = d.read_reference_distribution(%r)\n"
        "assert p['kind'] == 'interpolate' and p['t'] == 0.75 and p['plan'].shape == (3, 3) and abs(p['plan'].sum() - 1) < 1e-6\n"
        "assert p['a'][1].shape == (3, 256) and p['b'][2].shape == (3, 256)\n"
        "q = d.read_reference_distribution(%r)\n"
        "assert q['kind'] == 'barycenter' and q['pi'].shape == (4, 3) and q['mu'].shape == (4, 3, 256) and q['rate'] == [0.5, 0.25, 0.125, 0.125]\n"
        "assert 'model' not in sys.modules\n"
        "print('ok')\n") % (root, os.path.join(golden_dir, "distri_ref_interp.pth"), os.path.join(golden_dir, "distri_ref_bary.pth"))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stderr[-2000:]
    # the parts equal the SpeakerMetaEncoder parameters the artefact was built from (seed-1234 weights, combos 1 and 2)
    from masg_b200 import distributions as d
    from oracle import restate
    sd = syn.make_fs2_state_dict(seed=1234)
    pi, mu, sigma = restate.gmm_params(sd, torch.tensor(syn.META_COMBOS, dtype=torch.float32))
    p = d.read_reference_distribution(os.path.join(golden_dir, "distri_ref_interp.pth"))
    assert torch.allclose(p["a"][1], mu[1], atol=1e-6) and torch.allclose(p["b"][2], sigma[2], atol=1e-6) and torch.allclose(p["a"][0], pi[1], atol=1e-6)


def _gather_worker(rank, world, port, q):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    g = torch.Generator().manual_seed(100 + rank)
    lens = [3, 1, 2] if rank == 0 else [4]                                     # utterances per rank and their frame counts differ
    off = torch.zeros(len(lens) + 1, dtype=torch.int64)
    off[1:] = torch.cumsum(torch.tensor(lens) * 256, 0)
    packed = torch.randint(-32768, 32767, (int(off[-1]),), generator=g, dtype=torch.int16)
    got = synth.gather_waveforms(packed, off)
    got2 = synth.WaveformGather().start(packed, off, n_max=6 * 256, b_max=3).wait()   # sizes known up front: no counts exchange
    assert all(torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) for a, b in zip(got, got2))
    q.put((rank, packed.numpy(), off.numpy(), [(w.numpy().copy(), o.numpy().copy()) for w, o in got]))
    dist.destroy_process_group()


def test_waveform_gather_two_ranks_gloo():
    """The optional waveform gather: ragged per-rank audio, padded all_gather + counts, trimmed on arrival."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue(); port = _free_port()
    ps = [ctx.Process(target=_gather_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted([q.get(timeout=120) for _ in ps], key=lambda r: r[0])
    [p.join(timeout=60) for p in ps]
    sent = [(r[1], r[2]) for r in res]
    for rank, _, _, got in res:
        assert len(got) == 2
        for r in range(2):
            assert np.array_equal(got[r][0], sent[r][0]) and np.array_equal(got[r][1], sent[r][1])


@pytest.mark.skipif(not __import__("oracle.ref_import", fromlist=["x"]).available(), reason="no reference tree")
def test_dropin_signatures_equal_the_reference():
    """model/fastspeech2.py:52-69,187-204 and hifigan/models.py:149: parameter names, order and defaults of the drop-ins."""
    import inspect
    from oracle import ref_import
    from masg_b200.fastspeech2 import FastSpeech2
    from masg_b200.hifigan import Generator
    from masg_b200 import distributions, tools
    ns = ref_import.load()
    sig = lambda f: [(p.name, p.default) for p in inspect.signature(f).parameters.values()]
    assert sig(FastSpeech2.forward) == sig(ns.FastSpeech2.forward)
    assert sig(FastSpeech2.synthesize_from_speaker_emb) == sig(ns.FastSpeech2.synthesize_from_speaker_emb)
    assert sig(FastSpeech2.speaker_distribution) == sig(ns.FastSpeech2.speaker_distribution)
    assert sig(FastSpeech2.speaker_gen)[:2] == sig(ns.FastSpeech2.speaker_gen)
    assert sig(FastSpeech2.__init__)[:4] == sig(ns.FastSpeech2.__init__)
    assert sig(Generator.forward)[:2] == sig(ns.Generator.forward) and sig(Generator.__init__)[:2] == sig(ns.Generator.__init__)
    assert [n for n, _ in sig(distributions.InterpolateGMM.__init__)][:3] == [n for n, _ in sig(ns.InterpolateGMM.__init__)]
    assert sig(distributions.InterpolateGMM.interpolate_rate) == sig(ns.InterpolateGMM.interpolate_rate)
    assert sig(distributions.BarycenterGMM.__init__) == sig(ns.BarycenterGMM.__init__)
    assert sig(distributions.BarycenterGMM.barycenter_rate) == sig(ns.BarycenterGMM.barycenter_rate)
    import utils.tools as rt
    import utils.model as rm
    from masg_b200.hifigan import vocoder_infer
    assert sig(tools.synth_samples)[:7] == sig(rt.synth_samples) and sig(tools.expand) == sig(rt.expand)
    assert sig(tools.get_mask_from_lengths) == sig(rt.get_mask_from_lengths) and sig(vocoder_infer) == sig(rm.vocoder_infer)
