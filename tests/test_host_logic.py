"""CPU tests: the C-ABI library loads and exports every symbol include/masg.h declares, the product
path fails loudly without a GPU (no CPU fallback), drop-in state_dict layouts match the reference's,
and the batch scheduler / multi-rank sharding logic (gloo, world_size 2)."""
import ctypes
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
