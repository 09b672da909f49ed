"""The C-ABI library loads without a GPU and exports every symbol include/flr_b200.h declares;
host-only entry points (the synthetic generator) are exercised.  No compute kernels run here."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "flr_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(flr_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    from fairness_llm_replication_b200 import _lib
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/flr_b200.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature"
    assert set(_lib.SIGNATURES) == set(names)
    assert lib.flr_version() >= 100


def test_header_constants_match_python():
    from fairness_llm_replication_b200 import _lib
    text = open(os.path.join(ROOT, "include", "flr_b200.h")).read()
    d = {k: int(v) for k, v in re.findall(r"#define\s+(FLR_[A-Z0-9_]+)\s+(-?\d+)", text)}
    assert d["FLR_SELECT_BINS"] == _lib.SELECT_BINS and d["FLR_MAX_ROUNDS"] == _lib.MAX_ROUNDS
    assert d["FLR_SELECT_STATE_WORDS"] == _lib.SELECT_STATE_WORDS and d["FLR_MAX_RANK_K"] == _lib.MAX_RANK_K
    assert (d["FLR_ST_N"], d["FLR_ST_K"], d["FLR_ST_LESS"], d["FLR_ST_EQUAL"]) == (
        _lib.ST_N, _lib.ST_K, _lib.ST_LESS, _lib.ST_EQUAL)
    assert d["FLR_EMPTY_BOTH_IS_ONE"] == _lib.EMPTY_BOTH_IS_ONE == 1


def test_argument_errors_are_reported(lib):
    from fairness_llm_replication_b200 import _lib
    rc = lib.flr_synth_zipf_cdf_host(None, 10)
    assert rc == _lib.E_INVALID and b"null" in lib.flr_last_error()
    with pytest.raises(_lib.FlrError):
        _lib.call("flr_synth_lists_host", 1, 0, 4, 3, 100, 50, None, None, None, None, None)   # K > 64


def test_synth_lists_host_properties(lib):
    from fairness_llm_replication_b200 import ops
    U, V, K, vocab = 3000, 31, 20, 3706
    neutral, nlen, lists, lens = ops.synth_lists_host(20251203, 0, U, V, K, vocab)
    again = ops.synth_lists_host(20251203, 0, U, V, K, vocab)
    assert all(np.array_equal(a, b) for a, b in zip((neutral, nlen, lists, lens), again))
    shifted = ops.synth_lists_host(20251203, 1000, 10, V, K, vocab)          # user0 offsets the stream
    assert np.array_equal(shifted[2], lists[1000:1010]) and np.array_equal(shifted[0], neutral[1000:1010])
    pos = np.arange(K)
    live = pos[None, None, :] < lens[:, :, None]
    assert lists[live].min() >= 0 and lists[live].max() < vocab
    assert np.all(lists[~live] == -1)
    short = (lens < K).mean()
    assert 0.002 < short < 0.03                                             # ~1% short rows
    srt = np.sort(np.where(live, lists, -np.arange(1, K + 1)[None, None, :]), axis=-1)
    has_dup = (srt[..., 1:] == srt[..., :-1]).any(-1)
    assert 0.0 < has_dup.mean() < 0.005                                     # ~0.1% planted duplicates
    # similarity falls as the replacement probability rises with the variant index
    from oracle import np_oracle as O
    jac = O.list_metrics(neutral[:400], nlen[:400], lists[:400], lens[:400])["jaccard"].mean(0)
    assert jac[0] > jac[15] > jac[30] and 0.05 < jac[30] < jac[0] < 0.95
    # Zipf head: item 0 is the most frequent neutral item
    counts = np.bincount(neutral[neutral >= 0].ravel(), minlength=vocab)
    assert counts[0] == counts.max()


def test_synth_embeddings_host_properties(lib):
    from fairness_llm_replication_b200 import ops
    en, e = ops.synth_embeddings_host(20251203, 5, 64, 31, 384)
    en2, e2 = ops.synth_embeddings_host(20251203, 0, 69, 31, 384)
    assert np.array_equal(en, en2[5:]) and np.array_equal(e, e2[5:])
    assert abs(en.mean()) < 0.02 and 0.95 < en.std() < 1.05
    from oracle import np_oracle as O
    s = O.cosine_nonconformity(en, e).mean(0)
    assert np.all(np.diff(s) > -1e-3) and s[0] < 0.003 and 0.05 < s[-1] < 0.2
