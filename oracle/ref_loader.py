"""TEST INFRASTRUCTURE ONLY -- loader for the *unmodified* reference scripts.

Imports /root/reference/scripts/NN_*.py by path so that the oracle restatements in
``oracle/vae_ref.py`` / ``oracle/cluster_ref.py`` can be pinned against the reference's
own classes, and so that ``oracle/make_golden.py`` can generate the committed fixtures
under ``tests/golden/``.  /root/reference exists only in the build container, never on
the GPU box: nothing in ``-m gpu`` tests, ``smoke()`` or ``bench.py`` may import this
module at run time (tests that use it skip when the directory is absent).

Loader details follow SURVEY.md section 8(c):
  * file names start with a digit -> importlib.util.spec_from_file_location
  * 06/07/12 combine ``from __future__ import annotations`` with @dataclass, so the
    module must be in sys.modules before exec_module
  * 06/19/22 import librosa / matplotlib at module top -> empty stub modules
  * 19/20/22 mkdir relative to cwd at import -> import from a scratch cwd
"""
from __future__ import annotations

import contextlib
import importlib.util
import os
import sys
import tempfile
import types
from pathlib import Path

REFERENCE_ROOT = Path(os.environ.get("VAECLUST_REFERENCE_ROOT", "/root/reference"))
_SCRIPTS = {
    "06": "06_train_basic_vae_easy.py",
    "07": "07_clustering_easy.py",
    "12": "12_train_conv_multimodal_vae_medium.py",
    "13": "13_cluster_and_evaluate_medium.py",
    "16": "16_full_clustering_sweep_medium.py",
    "19": "19_train_beta_cvae_multimodal_hard.py",
    "20": "20_cluster_and_evaluate_hard.py",
    "22": "22_compare_with_baselines_hard.py",
}
_cache: dict = {}


def available() -> bool:
    return (REFERENCE_ROOT / "scripts" / _SCRIPTS["12"]).is_file()


def _install_stubs() -> None:
    for name in ("matplotlib", "matplotlib.pyplot", "librosa", "librosa.display", "umap", "seaborn"):
        if name in sys.modules:
            continue
        try:
            importlib.import_module(name)
        except Exception:
            mod = types.ModuleType(name)
            mod.__path__ = []  # behave like a package for dotted imports
            sys.modules[name] = mod
    if "matplotlib" in sys.modules and "matplotlib.pyplot" in sys.modules:
        setattr(sys.modules["matplotlib"], "pyplot", sys.modules["matplotlib.pyplot"])


@contextlib.contextmanager
def _scratch_cwd():
    old = os.getcwd()
    with tempfile.TemporaryDirectory(prefix="vaeclust_ref_") as d:
        os.chdir(d)
        try:
            yield d
        finally:
            os.chdir(old)


def load(script: str):
    """Return the module object of reference script ``script`` ("12", "19", ...)."""
    if script in _cache:
        return _cache[script]
    if not available():
        raise FileNotFoundError(f"reference not mounted at {REFERENCE_ROOT}")
    _install_stubs()
    path = REFERENCE_ROOT / "scripts" / _SCRIPTS[script]
    name = f"_vaeclust_ref_{script}"
    spec = importlib.util.spec_from_file_location(name, str(path))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    with _scratch_cwd():
        spec.loader.exec_module(mod)
    _cache[script] = mod
    return mod
