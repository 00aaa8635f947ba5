"""Recipe that stages the UNMODIFIED reference path into the git-ignored ``oracle/_ref/`` so that it travels to the
GPU box with the snapshot (like the built ``libmasg.so`` does) and can be imported there as the checker of the
restatement and as the timed CPU / eager-GPU baseline (``cpu_baseline.kind = "reference"``).

The reference is pure Python (no build step): "compiling" it is a byte-for-byte copy of the files the hot path imports,
from where they lie under /root/reference, into ``oracle/_ref/`` — nothing is edited, nothing enters the git history
(``oracle/_ref/`` is listed in .gitignore, not in .gpurunignore).  ``MANIFEST.json`` records every file with its sha256
so a reader can check that the staged tree equals the reference.

Run in the build container only (``__graft_entry__.build()`` calls it when /root/reference exists):
    python -m oracle.make_ref

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("MASG_REFERENCE_SRC", "/root/reference")
DST = os.path.join(HERE, "_ref")

# what `import model`, `import hifigan`, `utils.tools`, `utils.model` and the text front-end pull in
DIRS = ["model", "transformer", "hifigan", "utils", "text"]
FILES = ["prepare_tg_accent_jsut.py", "convert_label.py"]
CFG = "config/JVS-VCTK_langemb_configs/JVS-VCTK_1"
CFG_FILES = ["model.yaml", "preprocess.yaml", "train.yaml"]
KEEP_EXT = (".py", ".json", ".yaml", ".txt")


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def make(force=False):
    """Stage the reference files; returns the destination, or None when /root/reference is absent (GPU box: the
    prebuilt copy that travelled with the snapshot is used as is)."""
    if not os.path.isfile(os.path.join(SRC, "model", "fastspeech2.py")):
        return DST if os.path.isfile(os.path.join(DST, "MANIFEST.json")) else None
    pairs = []
    for d in DIRS:
        for root, _, names in os.walk(os.path.join(SRC, d)):
            for n in sorted(names):
                if n.endswith(KEEP_EXT):
                    p = os.path.join(root, n)
                    pairs.append((p, os.path.relpath(p, SRC)))
    pairs += [(os.path.join(SRC, f), f) for f in FILES]
    pairs += [(os.path.join(SRC, CFG, f), os.path.join(CFG, f)) for f in CFG_FILES]
    manifest = {rel: _sha(src) for src, rel in pairs}
    mpath = os.path.join(DST, "MANIFEST.json")
    if not force and os.path.isfile(mpath):
        try:
            old = json.load(open(mpath))
            if old.get("files") == manifest and all(os.path.isfile(os.path.join(DST, r)) for r in manifest):
                return DST
        except Exception:
            pass
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    for src, rel in pairs:
        out = os.path.join(DST, rel)
        os.makedirs(os.path.dirname(out), exist_ok=True)
        shutil.copyfile(src, out)
    json.dump({"source": SRC, "note": "byte-for-byte copies of the reference files on the hot path; not part of the repository",
               "files": manifest}, open(mpath, "w"), indent=1, sort_keys=True)
    return DST


if __name__ == "__main__":
    out = make(force="--force" in sys.argv)
    print(out if out else "reference tree not present at %s and no staged copy under %s" % (SRC, DST))
