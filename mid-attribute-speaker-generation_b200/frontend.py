"""Batched text front-end: HTS full-context labels -> padded (phoneme id, accent id) batches in
pinned host memory, ready for ``Synthesizer.synthesize_host``.

Stands behind the Japanese tokenisation the reference runs per sentence in Python (synthesize.py:96-101,215-234):
``pyopenjtalk.extract_fullcontext(text)`` -> ``pp_symbols`` (prepare_tg_accent_jsut.py:18-70) -> ``openjtalk2julius``
(convert_label.py:52-59) -> ``symbol_to_id`` / ``accent_to_id`` lookups -> ``np.array([...])`` per utterance.
Here the label scan and the lookups run in libmasg.so (csrc/frontend.cu: one forward pass per label instead of seven regex
searches, utterances split over host threads) and write whole padded batches straight into pinned int64 buffers.

pyopenjtalk (the grapheme-to-label step) and espeak-ng (the English branch, synthesize.py:35-66) are not installed in this
image; this module starts from full-context labels — what ``extract_fullcontext`` returns, and what the vendored
``jsut-lab/*.lab`` files hold.  ``preprocess_japanese`` is provided with the reference's signature and imports pyopenjtalk
lazily.
"""
import ctypes
import os

import numpy as np
import torch

from . import _lib

ACCENT_MARKS = ("0", "[", "]", "#")                # accent_to_id of synthesize.py:218
NO_ACCENT_ID = 4                                   # "no accent information" (synthesize.py:234)


def read_lab(path):
    """Full-context labels of one HTS .lab file: the last whitespace-separated field of every non-empty line (the jsut-lab
    files carry `start end label`; a file of bare labels works too)."""
    out = []
    with open(path, "r", encoding="utf-8") as f:
        for line in f:
            line = line.strip()
            if line:
                out.append(line.rsplit(None, 1)[-1])
    return out


class Frontend:
    """One native tokenizer bound to a symbol inventory (``text.symbols.symbols``: index = id, later duplicates win)."""

    def __init__(self, symbols, n_threads=None):
        self.lib = _lib.load_library()
        self.symbols = [str(s) for s in symbols]
        arr = (ctypes.c_char_p * len(self.symbols))(*[s.encode("utf-8") for s in self.symbols])
        h = ctypes.c_void_p()
        rc = self.lib.masg_frontend_create(arr, len(self.symbols), ctypes.byref(h))
        if rc != 0:
            raise RuntimeError("masg_frontend_create: " + self.lib.masg_frontend_last_error(None).decode())
        self.h = h
        self.n_threads = int(n_threads or min(16, os.cpu_count() or 1))

    def close(self):
        if getattr(self, "h", None):
            self.lib.masg_frontend_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def tokenize(self, utterances, L_cap=None, pin=None, drop_unvoiced_vowels=True):
        """utterances: list of label lists (one list of full-context label strings per utterance).
        Returns {"texts", "accents": int64 [N, L_max] zero padded, "src_lens": int64 [N], "max_src_len"} as host tensors
        (pinned when CUDA is available, so the H2D copy of ``synthesize_host`` is asynchronous).
        Errors mirror the reference: AssertionError for a 'sil' inside an utterance, KeyError for a symbol outside the
        inventory, IndexError for an utterance that does not end in 'sil'."""
        n = len(utterances)
        if n == 0:
            raise ValueError("empty batch")
        counts = np.fromiter((len(u) for u in utterances), dtype=np.int64, count=n)
        if L_cap is None:
            L_cap = int(counts.max())                  # every label yields at most one phoneme
        L_cap = max(int(L_cap), 1)
        blob = "\n".join("\n".join(u) for u in utterances if len(u)).encode("utf-8")
        pin = torch.cuda.is_available() if pin is None else pin
        mk = (lambda *s: torch.empty(*s, dtype=torch.int64).pin_memory()) if pin else (lambda *s: torch.empty(*s, dtype=torch.int64))
        texts, accents, lens = mk(n, L_cap), mk(n, L_cap), mk(n)
        rc = self.lib.masg_frontend_tokenize(self.h, blob, len(blob), counts.ctypes.data_as(ctypes.c_void_p), n, L_cap,
                                             1 if drop_unvoiced_vowels else 0, self.n_threads, _lib.ptr(texts), _lib.ptr(accents),
                                             _lib.ptr(lens))
        if rc != 0:
            msg = self.lib.masg_frontend_last_error(self.h).decode()
            raise {11: AssertionError, 12: KeyError, 13: IndexError}.get(rc, RuntimeError)(msg)
        L = int(lens.max())
        if L < L_cap:
            texts, accents = texts[:, :L].contiguous(), accents[:, :L].contiguous()
            if pin:
                texts, accents = texts.pin_memory(), accents.pin_memory()
        return {"texts": texts, "accents": accents, "src_lens": lens, "max_src_len": L}

    def labels_to_phonemes(self, labels):
        """One utterance, the reference's return type: (phoneme strings, accent marks) =
        ``pp_symbols`` + ``[openjtalk2julius(p) for p in phonemes if p != '']`` (synthesize.py:96-101)."""
        b = self.tokenize([list(labels)], pin=False)
        n = int(b["src_lens"][0])
        return [self.symbols[i] for i in b["texts"][0, :n].tolist()], [ACCENT_MARKS[a] for a in b["accents"][0, :n].tolist()]

    def preprocess_japanese(self, text):
        """synthesize.py:96-101 with the reference's signature; needs pyopenjtalk for the grapheme -> label step."""
        try:
            import pyopenjtalk
        except ImportError as e:
            raise RuntimeError("preprocess_japanese needs pyopenjtalk (not installed in this image); feed full-context labels to "
                               "Frontend.tokenize instead") from e
        return self.labels_to_phonemes(pyopenjtalk.extract_fullcontext(text))


def without_accents(batch):
    """The 'no accent information' variant of a tokenised batch: accents = 4 on valid positions (synthesize.py:234)."""
    acc = torch.full_like(batch["accents"], NO_ACCENT_ID)
    pad = torch.arange(acc.shape[1])[None, :] >= batch["src_lens"][:, None]
    acc[pad] = 0
    return dict(batch, accents=acc)
