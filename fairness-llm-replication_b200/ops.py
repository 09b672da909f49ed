"""Tensor-level entry points over the C ABI (include/flr_b200.h).

PyTorch is used for device memory, streams and (in ``pipeline``) ``torch.distributed``
only; every computation below is one of the library's own sm_100a kernels, launched on the
current CUDA stream.  Nothing here synchronises unless stated.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from . import _lib
from ._lib import (DTYPE_F32, DTYPE_F64, EMPTY_ANY_IS_ZERO, EMPTY_BOTH_IS_ONE,  # noqa: F401 (re-exported)
                   MAX_RANK_K, SELECT_BINS, SELECT_STATE_WORDS)

__all__ = [
    "list_metrics", "ListMetrics", "list_metrics_kernel_name", "cosine_kernel_name", "pair_jaccard", "cosine_nonconformity", "radix_select",
    "threshold_schedule", "violation_detect", "conformal_keep_len", "normalize_rows_bf16",
    "pair_similarity", "group_item_histogram", "js_distance_matrix", "equal_opportunity_counts", "ndcg", "exposure_by_group", "synth_lists", "synth_embeddings", "synth_lists_host",
    "synth_embeddings_host", "zipf_cdf_host", "device_info",
    "EMPTY_ANY_IS_ZERO", "EMPTY_BOTH_IS_ONE",
]


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dev(t: torch.Tensor, dtype, name: str) -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TypeError(f"{name} must be a CUDA tensor (this path has no CPU implementation)")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    return t


_scratch = {}


def _workspace(device, nbytes: int) -> torch.Tensor:
    """Per-device scratch, grown on demand and reused by later calls on the same stream order."""
    key = (device.type, device.index)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


def device_info():
    sm, maj, mnr = C.c_int(), C.c_int(), C.c_int()
    _lib.call("flr_device_info", C.byref(sm), C.byref(maj), C.byref(mnr))
    return {"sm_count": sm.value, "cc": (maj.value, mnr.value)}


# ------------------------------------------------------------------------------ list metrics
@dataclass
class ListMetrics:
    inter: torch.Tensor                  # int32 [U, V]  |A & B|, exact
    union: Optional[torch.Tensor]        # int32 [U, V]  |A | B|, exact
    jaccard: torch.Tensor                # float32 [U, V]
    serp: Optional[torch.Tensor]         # float32 [U, V]
    prag: Optional[torch.Tensor]         # float32 [U, V]
    variant_sums: Optional[torch.Tensor]  # float64 [V, 4]: sum jaccard, sum serp_num, sum prag_num, users
    status: torch.Tensor                 # int32 [1]: ids found outside [0, vocab)
    k_cap: int

    def variant_means(self) -> torch.Tensor:
        """float64 [V, 3]: mean Jaccard, SERP*, PRAG* per variant (divides the exact sums)."""
        s = self.variant_sums
        n = s[:, 3:4].clamp_min(1.0)
        kn = float(self.k_cap)
        scale = torch.tensor([1.0, 1.0 / (kn * (kn + 1) / 2.0), 1.0 / (kn * (kn + 1.0))],
                             dtype=torch.float64, device=s.device)
        return s[:, :3] / n * scale


def list_metrics(neutral, neutral_len, lists, lens, *, vocab: int, k_cap: Optional[int] = None,
                 empty_policy: int = EMPTY_BOTH_IS_ONE, rank_metrics: bool = True,
                 want_union: bool = False, want_sums: bool = True) -> ListMetrics:
    """Jaccard / SERP* / PRAG* of lists[u, v, :] against neutral[u, :]  (flr_list_metrics).

    neutral int32 [U, K], neutral_len int32 [U], lists int32 [U, V, K], lens int32 [U, V].
    """
    lists = _dev(lists, torch.int32, "lists")
    if lists.dim() != 3:
        raise ValueError("lists must be [U, V, K]")
    U, V, K = lists.shape
    neutral = _dev(neutral, torch.int32, "neutral")
    neutral_len = _dev(neutral_len, torch.int32, "neutral_len")
    lens = _dev(lens, torch.int32, "lens")
    if tuple(neutral.shape) != (U, K) or neutral_len.numel() != U or lens.numel() != U * V:
        raise ValueError("shape mismatch between neutral/neutral_len/lists/lens")
    k_cap = K if k_cap is None else int(k_cap)
    dev = lists.device
    inter = torch.empty((U, V), dtype=torch.int32, device=dev)
    union = torch.empty((U, V), dtype=torch.int32, device=dev) if want_union else None
    jac = torch.empty((U, V), dtype=torch.float32, device=dev)
    serp = torch.empty((U, V), dtype=torch.float32, device=dev) if rank_metrics else None
    prag = torch.empty((U, V), dtype=torch.float32, device=dev) if rank_metrics else None
    sums = torch.empty((V, 4), dtype=torch.float64, device=dev) if want_sums else None
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    nbytes = _lib.load().flr_list_metrics_scratch_bytes(U, V, K, vocab)
    ws = _workspace(dev, nbytes)
    _lib.call("flr_list_metrics", _ptr(neutral), _ptr(neutral_len), _ptr(lists), _ptr(lens),
              U, V, K, k_cap, int(vocab), int(empty_policy), int(bool(rank_metrics)),
              _ptr(inter), _ptr(union), _ptr(jac), _ptr(serp), _ptr(prag), _ptr(sums), _ptr(status),
              _ptr(ws), ws.numel(), _stream())
    return ListMetrics(inter, union, jac, serp, prag, sums, status, min(k_cap, K))


def list_metrics_counts(neutral, neutral_len, lists, lens, *, vocab: int, k_cap: int, out_inter, out_union) -> None:
    """The lean form behind the string API: exact |A & B| and |A | B| only, written into caller-provided
    int32 views; no ratios, no sums, no status word, nothing allocated beyond the cached workspace."""
    U, V, K = lists.shape
    ws = _workspace(lists.device, _lib.load().flr_list_metrics_scratch_bytes(U, V, K, vocab))
    _lib.call("flr_list_metrics", _ptr(neutral), _ptr(neutral_len), _ptr(lists), _ptr(lens), U, V, K, int(k_cap),
              int(vocab), EMPTY_BOTH_IS_ONE, 0, _ptr(out_inter), _ptr(out_union), None, None, None, None, None,
              _ptr(ws), ws.numel(), _stream())


def list_metrics_kernel_name(V: int, K: int, vocab: int, k_cap: Optional[int] = None, rank_metrics: bool = True) -> str:
    """Name of the kernel ``list_metrics`` launches for a shape (mirrors csrc/list_metrics.cu:make_plan),
    for labelling measurements: what ncu and the profiler list is this name."""
    k_cap = K if k_cap is None else min(k_cap, K)
    r = 1 if rank_metrics else 0
    a16 = lambda x: (x + 15) & ~15
    if 0 < vocab <= 16384 and k_cap <= MAX_RANK_K and 2 <= V <= 128:
        cap = next((c for c in (10, 20, 32, 50, 62) if K <= c), 0)
        wcap = 0 if K > 64 else 64 if K > 50 else cap
        rs = K + 4 if (K % 8 == 0 and K >= 32) else K
        Vw = min(V, 32)                    # variants per launch; more than 32 go in chunks of 32
        warp_bytes = 16 + a16(Vw * rs * 4 + 16) + a16(vocab + 65) + a16(((vocab >> 5) + 3) * 128) + a16(wcap * 4)
        warp_fits = bool(wcap) and a16((2 * wcap + 1) * 8) + 4 * ((warp_bytes + 127) & ~127) <= 226 * 1024
        warp_name = "list_metrics_warp_kernel<%d,%d,%d>" % (wcap, r, 1 if V > 32 else 0)
        if warp_fits and (V >= 17 or (V >= 4 and V * K >= 80)):
            return warp_name
        vocab_pad, W = (vocab + 16) & ~15, (vocab + 31) // 32

        def tile_bytes(T):
            n = 16
            for part in (T * V * K * 4, T * K * 4, T * V * 4, T * 4, T * 4, T * vocab_pad, (W + 1) * 128 * 4):
                n = a16(n + part)
            return n + (a16(T * (cap * 4 + 4)) + a16((2 * cap + 1) * 8) if cap else 0)
        T = 128 // V
        if T >= 4:
            T &= ~3
        while T >= 1 and tile_bytes(T) > 110 * 1024:
            T = ((T - 1) & ~3) if T > 4 else T - 1
        if T >= 1 and T * V * 2 >= 128:
            if cap:
                return "list_metrics_reg_kernel<%d,%d>" % (cap, r)
            return "list_metrics_table_kernel<%d,%d>" % (4 if K % 4 == 0 else 2 if K % 2 == 0 else 1, r)
        if warp_fits:
            return warp_name
    return "list_metrics_cmp_kernel<%d>" % r


def cosine_kernel_name(V: int, D: int) -> str:
    """Name of the kernel ``cosine_nonconformity`` launches (csrc/cosine.cu:launch_cos)."""
    nc = (D + 127) // 128
    if (V + 1) * D * 4 * 2 <= 200 * 1024:
        return "cosine_tma_kernel<%d,%d>" % (nc, 1 if D == nc * 128 else 0)
    return "cosine_kernel<%d>" % nc


def pair_jaccard(lists, lens, *, vocab: int):
    """All-pairs intersections among each user's M lists: int32 [U, M, M], sizes int32 [U, M]."""
    lists = _dev(lists, torch.int32, "lists")
    U, M, K = lists.shape
    lens = _dev(lens, torch.int32, "lens")
    inter = torch.empty((U, M, M), dtype=torch.int32, device=lists.device)
    sizes = torch.empty((U, M), dtype=torch.int32, device=lists.device)
    _lib.call("flr_pair_jaccard", _ptr(lists), _ptr(lens), U, M, K, int(vocab), _ptr(inter), _ptr(sizes),
              _stream())
    return inter, sizes


# ------------------------------------------------------------------------------ conformal
def cosine_nonconformity(a, b, *, hist0: Optional[torch.Tensor] = None, out=None) -> torch.Tensor:
    """s[u, v] = max(0, 1 - cos(a[u], b[u, v]));  a float32 [U, D], b float32 [U, V, D] (or [U, D]).

    ``hist0`` (int64 [2048], zeroed by the caller) receives the pass-0 select histogram."""
    a = _dev(a, torch.float32, "a")
    b = _dev(b, torch.float32, "b")
    squeeze = b.dim() == 2
    if squeeze:
        b = b.unsqueeze(1)
    U, V, D = b.shape
    if tuple(a.shape) != (U, D):
        raise ValueError("a must be [U, D] matching b [U, V, D]")
    if out is None:
        out = torch.empty((U, V), dtype=torch.float32, device=a.device)
    else:
        _dev(out, torch.float32, "out")
    if hist0 is not None:
        _dev(hist0, torch.int64, "hist0")
        if hist0.numel() != SELECT_BINS:
            raise ValueError("hist0 must hold 2048 int64 bins")
    _lib.call("flr_cosine_nonconformity", _ptr(a), _ptr(b), U, V, D, _ptr(out), _ptr(hist0), _stream())
    return out.squeeze(1) if squeeze else out


def radix_select(scores, *, alpha: float = 0.1, groups: Optional[torch.Tensor] = None, n_groups: int = 1,
                 k: Optional[torch.Tensor] = None, hist0: Optional[torch.Tensor] = None,
                 all_reduce=None):
    """Exact per-group order statistic at the reference's conformal index
    (src/phase3_facter_mitigation.py:165-178).  Returns (thresholds [G], state int64 [G, 8]).

    ``scores`` float32 or float64 (any shape), ``groups`` uint8 ids (same numel) or None.
    ``hist0``: a pass-0 histogram already built (cosine kernel), skips the first read.
    ``all_reduce``: callable applied to each pass's int64 histogram before it is scanned
    (multi-GPU: sum over ranks); the result is then the order statistic of the union.
    """
    if scores.dtype == torch.float32:
        dtype = DTYPE_F32
    elif scores.dtype == torch.float64:
        dtype = DTYPE_F64
    else:
        raise TypeError("scores must be float32 or float64")
    scores = _dev(scores, scores.dtype, "scores")
    N = scores.numel()
    G = int(n_groups)
    dev = scores.device
    if groups is not None:
        groups = _dev(groups, torch.uint8, "groups")
        if groups.numel() != N:
            raise ValueError("groups must have one id per score")
    elif G != 1:
        raise ValueError("n_groups > 1 needs group ids")
    n_pass = _lib.load().flr_select_num_passes(dtype)
    hist = torch.zeros((n_pass, G, SELECT_BINS), dtype=torch.int64, device=dev)
    state = torch.zeros((G, SELECT_STATE_WORDS), dtype=torch.int64, device=dev)
    thr = torch.empty(G, dtype=scores.dtype, device=dev)
    if k is not None:
        k = _dev(k, torch.int64, "k")
    for p in range(n_pass):
        h = hist[p]
        if p == 0 and hist0 is not None:
            if G != 1:
                raise ValueError("a fused pass-0 histogram exists only for one group")
            h = hist0.view(1, SELECT_BINS)
        else:
            _lib.call("flr_select_hist", _ptr(scores), dtype, _ptr(groups), N, G, p, _ptr(state), _ptr(h),
                      _stream())
        if all_reduce is not None:
            all_reduce(h)
        _lib.call("flr_select_scan", dtype, G, p, float(1 - alpha), _ptr(k), _ptr(h), _ptr(state), _ptr(thr),
                  _stream())
    return thr, state


def threshold_schedule(q0, state, *, gamma: float = 0.95, q_min: float = 0.0, rounds: int = 4):
    """thr[0..rounds] of the round-synchronous update (oracle/np_oracle.py:threshold_schedule)."""
    q0 = _dev(q0, torch.float32, "q0")
    state = _dev(state, torch.int64, "state")
    thr = torch.empty(rounds + 1, dtype=torch.float32, device=q0.device)
    _lib.call("flr_threshold_schedule", _ptr(q0), _ptr(state), float(gamma), float(q_min), int(rounds), _ptr(thr),
              _stream())
    return thr


def violation_detect(scores, thr, *, rounds: int, n_variants: int = 1, per_variant: bool = True,
                     counts: Optional[torch.Tensor] = None, variant_counts: Optional[torch.Tensor] = None):
    """flags[i] bit r = scores[i] > thr[r]; returns (flags uint8, counts int64 [R], variant_counts
    int64 [V, R] or None).  Counters are added to when passed in."""
    scores = _dev(scores, torch.float32, "scores")
    thr = _dev(thr, torch.float32, "thr")
    if thr.numel() < rounds:
        raise ValueError("thr must hold at least `rounds` thresholds")
    dev = scores.device
    flags = torch.empty(scores.shape, dtype=torch.uint8, device=dev)
    if counts is None:
        counts = torch.zeros(rounds, dtype=torch.int64, device=dev)
    if per_variant and variant_counts is None:
        variant_counts = torch.zeros((n_variants, rounds), dtype=torch.int64, device=dev)
    _lib.call("flr_violation_detect", _ptr(scores), scores.numel(), int(n_variants), _ptr(thr), int(rounds),
              _ptr(flags), _ptr(counts), _ptr(variant_counts if per_variant else None), _stream())
    return flags, counts, (variant_counts if per_variant else None)


def conformal_keep_len(lens, thr, groups: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Prefix length kept by apply_conformal_filtering (src/phase3_facter_mitigation.py:205-216)."""
    lens = _dev(lens, torch.int32, "lens")
    thr = _dev(thr, torch.float64, "thr")
    if groups is not None:
        groups = _dev(groups, torch.uint8, "groups")
    out = torch.empty(lens.shape, dtype=torch.int32, device=lens.device)
    _lib.call("flr_conformal_keep_len", _ptr(lens), _ptr(groups), _ptr(thr), lens.numel(), _ptr(out), _stream())
    return out


# ------------------------------------------------------------------------------ demographic parity
def group_item_histogram(lists, lens, *, vocab: int, n_groups: int, groups: Optional[torch.Tensor] = None,
                         k_cap: Optional[int] = None, hist: Optional[torch.Tensor] = None) -> torch.Tensor:
    """hist[g, x] += occurrences of item x in the rows of group g (src/utils.py:180-187).
    lists int32 [..., K] (rows = all leading dims); groups uint8 per row, or None when the rows are laid
    out [U, V, K] and row r belongs to group r % n_groups.  Returns int64 [n_groups, vocab]."""
    lists = _dev(lists, torch.int32, "lists")
    lens = _dev(lens, torch.int32, "lens")
    K = lists.shape[-1]
    R = lists.numel() // K
    if lens.numel() != R:
        raise ValueError("lens must hold one length per row")
    if groups is not None:
        groups = _dev(groups, torch.uint8, "groups")
        if groups.numel() != R:
            raise ValueError("groups must hold one id per row")
    if hist is None:
        hist = torch.zeros((n_groups, vocab), dtype=torch.int64, device=lists.device)
    else:
        _dev(hist, torch.int64, "hist")
    _lib.call("flr_group_item_histogram", _ptr(lists), _ptr(lens), _ptr(groups), R, K, K if k_cap is None else int(k_cap),
              int(n_groups), int(vocab), _ptr(hist), _stream())
    return hist


def js_distance_matrix(hist):
    """Jensen-Shannon distances (src/utils.py:89-102) between every pair of group histograms, i < j in
    row-major order: (float64 [G(G-1)/2], uint8 validity -- 0 where a group is empty)."""
    hist = _dev(hist, torch.int64, "hist")
    G, vocab = hist.shape
    n_pairs = G * (G - 1) // 2
    dist = torch.zeros(max(n_pairs, 1), dtype=torch.float64, device=hist.device)
    valid = torch.zeros(max(n_pairs, 1), dtype=torch.uint8, device=hist.device)
    _lib.call("flr_js_distance_matrix", _ptr(hist), G, vocab, _ptr(dist), _ptr(valid), _stream())
    return dist[:n_pairs], valid[:n_pairs]


def equal_opportunity_counts(hist, relevant):
    """Per group: (distinct relevant items recommended, recommendations made) -- src/utils.py:254-266.
    hist int64 [G, vocab] from ``group_item_histogram``; relevant uint8 [vocab] (non-zero = relevant)."""
    hist = _dev(hist, torch.int64, "hist")
    relevant = _dev(relevant, torch.uint8, "relevant")
    G, vocab = hist.shape
    if relevant.numel() != vocab:
        raise ValueError("relevant must hold one flag per vocabulary id")
    rel = torch.zeros(G, dtype=torch.int64, device=hist.device)
    tot = torch.zeros(G, dtype=torch.int64, device=hist.device)
    _lib.call("flr_equal_opportunity_counts", _ptr(hist), _ptr(relevant), G, vocab, _ptr(rel), _ptr(tot), _stream())
    return rel, tot


# ------------------------------------------------------------------------------ ranking metrics (phase 2)
def ndcg(rel, rel_len, ideal, ideal_len, k: int) -> torch.Tensor:
    """NDCG@k of B rankings (src/utils.py:113-132): rel float64 [B, n] relevances in rank order, ideal
    float64 [B, m] ground-truth relevances sorted descending; returns float64 [B]."""
    rel = _dev(rel, torch.float64, "rel")
    ideal = _dev(ideal, torch.float64, "ideal")
    rel_len = _dev(rel_len, torch.int32, "rel_len")
    ideal_len = _dev(ideal_len, torch.int32, "ideal_len")
    B, n = rel.shape
    m = ideal.shape[1]
    log2tab = torch.from_numpy(np.log2(np.arange(max(n, m, 1)) + 2)).to(rel.device)      # the reference's np.log2
    out = torch.empty(B, dtype=torch.float64, device=rel.device)
    _lib.call("flr_ndcg", _ptr(rel), _ptr(rel_len), n, _ptr(ideal), _ptr(ideal_len), m, int(k), _ptr(log2tab), B,
              _ptr(out), _stream())
    return out            # log2tab goes back to torch's stream-ordered allocator: reuse is ordered after this launch


def exposure_by_group(group, lens, n_groups: int):
    """Mean exposure 1/log2(pos + 2) per protected group for B rankings
    (src/phase2_cross_model_eval.py:228-243): group uint8 [B, n] (255 = skip); returns
    (float64 [B, G] means, int32 [B, G] counts)."""
    group = _dev(group, torch.uint8, "group")
    lens = _dev(lens, torch.int32, "lens")
    B, n = group.shape
    weight = torch.from_numpy(1.0 / np.log2(np.arange(n) + 2)).to(group.device)          # the reference's expression
    mean = torch.empty((B, n_groups), dtype=torch.float64, device=group.device)
    cnt = torch.empty((B, n_groups), dtype=torch.int32, device=group.device)
    _lib.call("flr_exposure_by_group", _ptr(group), _ptr(lens), n, int(n_groups), _ptr(weight), B, _ptr(mean), _ptr(cnt),
              _stream())
    return mean, cnt      # (weight: same stream-ordered release as in ndcg)


# ------------------------------------------------------------------------------ config 4
def normalize_rows_bf16(x) -> torch.Tensor:
    x = _dev(x, torch.float32, "x")
    D = x.shape[-1]
    out = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    _lib.call("flr_normalize_rows_bf16", _ptr(x), x.numel() // D, D, _ptr(out), _stream())
    return out


def pair_similarity(x_bf16) -> torch.Tensor:
    """S[u] = X[u] X[u]^T, x bf16 [U, M, D] -> float32 [U, M, M] (tcgen05 / TMEM / TMA)."""
    x = _dev(x_bf16, torch.bfloat16, "x_bf16")
    U, M, D = x.shape
    out = torch.empty((U, M, M), dtype=torch.float32, device=x.device)
    nbytes = _lib.load().flr_pair_similarity_scratch_bytes(U, M, D)
    ws = _workspace(x.device, nbytes)
    _lib.call("flr_pair_similarity_gemm", _ptr(x), U, M, D, _ptr(out), _ptr(ws), ws.numel(), _stream())
    return out


# ------------------------------------------------------------------------------ synthetic inputs
_cdf_cache = {}


def zipf_cdf_host(vocab: int) -> np.ndarray:
    cdf = _cdf_cache.get(vocab)
    if cdf is None:
        cdf = np.empty(vocab, np.float32)
        _lib.call("flr_synth_zipf_cdf_host", cdf.ctypes.data_as(C.c_void_p), int(vocab))
        _cdf_cache[vocab] = cdf
    return cdf


def synth_lists_host(seed: int, user0: int, U: int, V: int, K: int, vocab: int):
    """numpy (neutral [U,K], neutral_len [U], lists [U,V,K], lens [U,V]); runs on host cores."""
    cdf = zipf_cdf_host(vocab)
    neutral = np.empty((U, K), np.int32)
    nlen = np.empty((U,), np.int32)
    lists = np.empty((U, V, K), np.int32)
    lens = np.empty((U, V), np.int32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    _lib.call("flr_synth_lists_host", int(seed), int(user0), U, V, K, int(vocab), p(cdf), p(neutral), p(nlen),
              p(lists), p(lens))
    return neutral, nlen, lists, lens


def synth_embeddings_host(seed: int, user0: int, U: int, V: int, D: int):
    en = np.empty((U, D), np.float32)
    e = np.empty((U, V, D), np.float32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    _lib.call("flr_synth_embeddings_host", int(seed), int(user0), U, V, D, p(en), p(e))
    return en, e


def synth_lists(seed: int, user0: int, U: int, V: int, K: int, vocab: int, device):
    device = torch.device(device)
    cdf = torch.from_numpy(zipf_cdf_host(vocab)).to(device)
    neutral = torch.empty((U, K), dtype=torch.int32, device=device)
    nlen = torch.empty((U,), dtype=torch.int32, device=device)
    lists = torch.empty((U, V, K), dtype=torch.int32, device=device)
    lens = torch.empty((U, V), dtype=torch.int32, device=device)
    with torch.cuda.device(device):
        _lib.call("flr_synth_lists", int(seed), int(user0), U, V, K, int(vocab), _ptr(cdf), _ptr(neutral),
                  _ptr(nlen), _ptr(lists), _ptr(lens), _stream())
        torch.cuda.current_stream().synchronize()      # cdf is freed on return
    return neutral, nlen, lists, lens


def synth_embeddings(seed: int, user0: int, U: int, V: int, D: int, device, chunk_users: int = 1 << 16):
    device = torch.device(device)
    en = torch.empty((U, D), dtype=torch.float32, device=device)
    e = torch.empty((U, V, D), dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        for s in range(0, U, chunk_users):
            n = min(chunk_users, U - s)
            _lib.call("flr_synth_embeddings", int(seed), int(user0 + s), n, V, D, _ptr(en[s:]), _ptr(e[s:]),
                      _stream())
    return en, e
