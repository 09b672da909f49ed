"""Import the UNMODIFIED reference modules from /root/reference/src.
TEST INFRASTRUCTURE ONLY.  /root/reference exists only in the build container, so nothing under
``-m gpu`` or ``smoke()`` calls this; the one use on the GPU box is the timed reference arm of
``bench.py``, which loads the copy that ``oracle/build_ref.py`` staged in ``oracle/_ref`` (STAGED_SRC).

``src/config.py`` is git-ignored upstream (SURVEY.md Fact 3), so a stub module with
the attributes the phase scripts read at import time is injected first.
``phase3_final.py`` / ``phase3_aggressive.py`` run API calls at import (Fact 4): only the
``def jaccard`` / ``def measure_fairness`` blocks are exec'd, straight from the file.
"""
from __future__ import annotations

import ast
import importlib
import os
import sys
import types

REF_SRC = "/root/reference/src"
STAGED_SRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")   # oracle/build_ref.py


def available(src: str = REF_SRC) -> bool:
    return os.path.isfile(os.path.join(src, "utils.py"))


def _stub_config() -> types.ModuleType:
    cfg = types.ModuleType("config")
    cfg.OPENAI_API_KEY = "offline"
    cfg.RESULTS_DIR = "/tmp/flr_ref_results"
    cfg.RANDOM_SEED = 42
    cfg.API_RATE_LIMIT = 60
    cfg.CONFORMAL_ALPHA = 0.1
    cfg.DATASETS = {"movielens": {"path": "/root/reference/ml-1m"}}
    cfg.SENSITIVE_ATTRIBUTES = {"gender": ["male", "female", "non-binary"],
                                "age": ["18-24", "25-34", "35-44", "45-54", "55+"]}
    cfg.MODELS = {"gpt-3.5-turbo": {"temperature": 0.7, "max_tokens": 500}}
    cfg.DEFAULT_MODELS = {"phase1": "gpt-3.5-turbo", "phase2": ["gpt-3.5-turbo"],
                          "phase3": "gpt-3.5-turbo"}
    cfg.SUCCESS_THRESHOLDS = {"bias_reduction_target": 50, "accuracy_preservation_min": 80}
    cfg.create_directories = lambda: None
    return cfg


def load(name: str, src: str = REF_SRC):
    """Return reference module ``utils`` / ``phase1_bias_detection`` /
    ``phase3_facter_mitigation`` under the private name ``_flr_ref_<name>``, from the reference
    tree (build container) or from ``STAGED_SRC``."""
    if not available(src):
        raise RuntimeError("reference modules not found under %s" % src)
    sys.dont_write_bytecode = True
    saved_path = list(sys.path)
    saved = {k: sys.modules.get(k) for k in ("config", "utils")}
    try:
        sys.path.insert(0, src)
        sys.modules["config"] = _stub_config()
        sys.modules.pop("utils", None)
        mod = importlib.import_module(name)
        sys.modules["_flr_ref_" + name] = sys.modules.pop(name)
        return mod
    finally:
        sys.path[:] = saved_path
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def load_script_functions(filename: str, names=("jaccard", "measure_fairness")):
    """Exec only the named top-level ``def`` blocks of a script-style reference file."""
    path = os.path.join(REF_SRC, filename)
    tree = ast.parse(open(path, encoding="utf-8").read(), path)
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    ns = {"np": __import__("numpy")}
    exec(compile(ast.Module(body=body, type_ignores=[]), path, "exec"), ns)
    return types.SimpleNamespace(**{n: ns[n] for n in names})
