"""ctypes front of ``oracle/synth_ref.c`` (CPU input generator).  TEST INFRASTRUCTURE ONLY:
the CPU arms of ``bench.py`` and the tests use it; it never loads the CUDA library."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import build_ref

_lib = None
_cdf = {}


def _load():
    global _lib
    if _lib is None:
        lib = C.CDLL(build_ref.build_synth_lib())
        vp, i32, i64, u64 = C.c_void_p, C.c_int, C.c_int64, C.c_uint64
        lib.synth_ref_zipf_cdf.argtypes = [vp, i32]
        lib.synth_ref_lists.argtypes = [u64, i64, i64, i32, i32, i32, vp, vp, vp, vp, vp]
        lib.synth_ref_embeddings.argtypes = [u64, i64, i64, i32, i32, vp, vp]
        for f in (lib.synth_ref_zipf_cdf, lib.synth_ref_lists, lib.synth_ref_embeddings):
            f.restype = None
        _lib = lib
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def zipf_cdf(vocab: int) -> np.ndarray:
    if vocab not in _cdf:
        cdf = np.empty(vocab, np.float32)
        _load().synth_ref_zipf_cdf(_p(cdf), vocab)
        _cdf[vocab] = cdf
    return _cdf[vocab]


def lists(seed: int, user0: int, U: int, V: int, K: int, vocab: int):
    """(neutral [U,K], neutral_len [U], lists [U,V,K], lens [U,V]), int32."""
    assert 1 <= K <= 64 and vocab >= 2 * K + 2
    cdf = zipf_cdf(vocab)
    neutral = np.empty((U, K), np.int32)
    nlen = np.empty((U,), np.int32)
    out = np.empty((U, V, K), np.int32)
    lens = np.empty((U, V), np.int32)
    _load().synth_ref_lists(seed, user0, U, V, K, vocab, _p(cdf), _p(neutral), _p(nlen), _p(out), _p(lens))
    return neutral, nlen, out, lens


def embeddings(seed: int, user0: int, U: int, V: int, D: int):
    """(emb_neutral [U,D], emb [U,V,D]), float32."""
    assert D % 2 == 0
    en = np.empty((U, D), np.float32)
    e = np.empty((U, V, D), np.float32)
    _load().synth_ref_embeddings(seed, user0, U, V, D, _p(en), _p(e))
    return en, e
