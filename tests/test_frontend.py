"""Text front-end, CPU only: the native batched tokenizer (csrc/frontend.cu through the C ABI) against
the reference's own pp_symbols -> openjtalk2julius -> symbol_to_id / accent_to_id on real full-context labels — the committed
jsut-lab sample (tests/golden/frontend_jsut.npz, written by oracle/gen_golden.py --frontend) and, when /root/reference is
mounted, all 7,696 vendored .lab files."""
import glob
import os

import numpy as np
import pytest
import torch

import masg_b200  # noqa: F401
from masg_b200 import frontend
from oracle import ref_import


@pytest.fixture(scope="module")
def fixture(golden_dir):
    return np.load(os.path.join(golden_dir, "frontend_jsut.npz"))


def test_tokenizer_matches_reference_fixture(fixture):
    g = fixture
    fe = frontend.Frontend([str(s) for s in g["symbols"]], n_threads=4)
    utts = [str(lab).split("\n") for lab in g["labels"]]
    b = fe.tokenize(utts, pin=False)
    assert b["texts"].dtype == torch.int64 and b["max_src_len"] == int(g["src_lens"].max())
    assert np.array_equal(b["src_lens"].numpy(), g["src_lens"])
    assert np.array_equal(b["texts"].numpy(), g["texts"]) and np.array_equal(b["accents"].numpy(), g["accents"])
    # single-threaded and per-utterance calls agree with the batched one; L_cap larger than needed is trimmed
    fe1 = frontend.Frontend([str(s) for s in g["symbols"]], n_threads=1)
    b1 = fe1.tokenize(utts[:7], L_cap=400, pin=False)
    n = b1["max_src_len"]
    assert np.array_equal(b1["texts"].numpy(), g["texts"][:7, :n]) and b1["texts"].shape[1] == int(g["src_lens"][:7].max())
    ph, ac = fe.labels_to_phonemes(utts[3])
    assert len(ph) == len(ac) == int(g["src_lens"][3]) and set(ac) <= set("0[]#")
    assert [{s: i for i, s in enumerate(fe.symbols)}[p] for p in ph] == g["texts"][3, :len(ph)].tolist()
    na = frontend.without_accents(b)
    assert int(na["accents"][0, 0]) == 4 and int(na["accents"][-1, -1]) in (0, 4)
    assert (na["accents"][torch.arange(b["texts"].shape[1])[None, :] >= b["src_lens"][:, None]] == 0).all()


def test_tokenizer_error_behaviour(fixture):
    g = fixture
    syms = [str(s) for s in g["symbols"]]
    fe = frontend.Frontend(syms)
    utt = str(g["labels"][0]).split("\n")
    with pytest.raises(AssertionError):                       # 'sil' inside an utterance: pp_symbols asserts
        fe.tokenize([utt[:3] + [utt[0]] + utt[3:]], pin=False)
    with pytest.raises(IndexError):                           # no trailing 'sil': labels[n + 1] raises in the reference
        fe.tokenize([utt[:-1]], pin=False)
    with pytest.raises(KeyError):                             # symbol outside the inventory
        frontend.Frontend([s for s in syms if s != "a"]).tokenize([utt], pin=False)
    with pytest.raises(ValueError):
        fe.tokenize([], pin=False)
    with pytest.raises(RuntimeError):                         # pyopenjtalk is not installed here
        fe.preprocess_japanese("これはボイスサンプルです。")


@pytest.mark.skipif(ref_import.kind() != "live", reason="needs /root/reference (jsut-lab is not staged to the GPU box)")
def test_tokenizer_matches_live_reference_on_all_jsut_labels():
    ref_import.load()
    from prepare_tg_accent_jsut import pp_symbols
    from convert_label import openjtalk2julius
    from text import symbols
    files = sorted(glob.glob(os.path.join(ref_import.REF_ROOT, "jsut-lab", "*", "lab", "*.lab")))
    assert len(files) == 7696 and len(symbols) == 428
    utts = [frontend.read_lab(f) for f in files]
    b = frontend.Frontend(symbols).tokenize(utts, pin=False)
    sym2id = {s: i for i, s in enumerate(symbols)}
    acc2id = {"0": 0, "[": 1, "]": 2, "#": 3}
    for i, u in enumerate(utts):
        ph, ac = pp_symbols(u)
        ph = [openjtalk2julius(p) for p in ph if p != ""]
        n = int(b["src_lens"][i])
        assert n == len(ph) == len(ac), files[i]
        assert b["texts"][i, :n].tolist() == [sym2id[p] for p in ph] and b["accents"][i, :n].tolist() == [acc2id[a] for a in ac], files[i]
        assert not b["texts"][i, n:].any()
