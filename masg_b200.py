"""Import alias: ``import masg_b200`` → the package directory ``mid-attribute-speaker-generation_b200``
(whose name is not a valid Python identifier).  Every sub-module is registered under both names so
``from masg_b200.fastspeech2 import FastSpeech2`` returns the one real module object."""
import importlib
import os
import sys

_REAL = "mid-attribute-speaker-generation_b200"
_here = os.path.dirname(os.path.abspath(__file__))
if _here not in sys.path:
    sys.path.insert(0, _here)
_pkg = importlib.import_module(_REAL)
for _sub in ("synthetic", "_lib", "fastspeech2", "hifigan", "distributions", "wavio", "synth", "build", "frontend", "tools"):
    _m = importlib.import_module(_REAL + "." + _sub)
    sys.modules["masg_b200." + _sub] = _m
sys.modules["masg_b200"] = _pkg
