"""Recipe for the CPU-side artefacts of the oracle.  TEST INFRASTRUCTURE ONLY.

    python -m oracle.build_ref            (also run by __graft_entry__.build())

* ``oracle/_build/libsynth_ref.so`` -- gcc build of ``oracle/synth_ref.c`` (input generator for the
  CPU arms; no CUDA, no product code).
* ``oracle/_ref/`` -- the reference's own modules for the timed reference arm of ``bench.py``.  The
  reference is Python, so "building" it is copying the two files the hot path lives in
  (``src/utils.py`` and ``src/phase3_facter_mitigation.py``) from ``/root/reference`` and writing the stub
  ``config.py`` they import (upstream's is git-ignored, SURVEY.md Fact 3).  Done only where
  ``/root/reference`` exists (the build container); the directory is git-ignored, so the sources stay out
  of this repository's history, and it travels to the GPU box with the snapshot like a built ``.so``.
"""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
BUILD_DIR = os.path.join(HERE, "_build")
REF_DIR = os.path.join(HERE, "_ref")
SYNTH_LIB = os.path.join(BUILD_DIR, "libsynth_ref.so")
REFERENCE_SRC = "/root/reference/src"
REFERENCE_FILES = ("utils.py", "phase3_facter_mitigation.py")

STUB_CONFIG = '''"""Stub of the reference's git-ignored src/config.py: the attributes its modules read at import."""
OPENAI_API_KEY = "offline"
RESULTS_DIR = "/tmp/flr_ref_results"
RANDOM_SEED = 42
API_RATE_LIMIT = 60
CONFORMAL_ALPHA = 0.1
DATASETS = {"movielens": {"path": "ml-1m"}}
SENSITIVE_ATTRIBUTES = {"gender": ["male", "female", "non-binary"], "age": ["18-24", "25-34", "35-44", "45-54", "55+"]}
MODELS = {"gpt-3.5-turbo": {"temperature": 0.7, "max_tokens": 500}}
DEFAULT_MODELS = {"phase1": "gpt-3.5-turbo", "phase2": ["gpt-3.5-turbo"], "phase3": "gpt-3.5-turbo"}
SUCCESS_THRESHOLDS = {"bias_reduction_target": 50, "accuracy_preservation_min": 80}


def create_directories():
    pass
'''


def build_synth_lib(force: bool = False) -> str:
    src = os.path.join(HERE, "synth_ref.c")
    if not force and os.path.exists(SYNTH_LIB) and os.path.getmtime(SYNTH_LIB) >= os.path.getmtime(src):
        return SYNTH_LIB
    os.makedirs(BUILD_DIR, exist_ok=True)
    cmd = ["gcc", "-O2", "-shared", "-fPIC", "-ffp-contract=off", "-o", SYNTH_LIB, src]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("gcc failed on %s:\n%s" % (src, r.stderr))
    return SYNTH_LIB


def stage_reference() -> bool:
    """Copy the reference's hot-path modules into oracle/_ref (build container only)."""
    if not os.path.isfile(os.path.join(REFERENCE_SRC, "utils.py")):
        return os.path.isfile(os.path.join(REF_DIR, "utils.py"))
    os.makedirs(REF_DIR, exist_ok=True)
    for name in REFERENCE_FILES:
        shutil.copyfile(os.path.join(REFERENCE_SRC, name), os.path.join(REF_DIR, name))
    with open(os.path.join(REF_DIR, "config.py"), "w") as f:
        f.write(STUB_CONFIG)
    return True


def build_all() -> dict:
    return {"synth_lib": build_synth_lib(), "reference_staged": stage_reference()}


if __name__ == "__main__":
    print(build_all())
