"""numpy float64 restatement of the scoring hot path.  TEST INFRASTRUCTURE ONLY
(see ``oracle/__init__.py`` for who may import this).

Every function cites the reference lines it follows (paths relative to
``/root/reference``).  Where the reference has no code (SERP*, PRAG*,
SNSR/SNSV, cosine nonconformity, violation rounds, similarity matrices) the
docstring here IS the normative definition -- parity for those is "unpinned".

Conventions shared by all list functions
----------------------------------------
A list row is ``ids[:min(len, k_cap)]``.  A row is read as a *sequence of
first occurrences*: a later duplicate of an id is ignored (the reference
collapses duplicates with ``set()``, src/utils.py:228-229) and the rank of an
item is the 0-based position of its first occurrence in the raw row, so ranks
are positional (a dropped duplicate leaves a hole, it does not shift later
items).  ``Kn`` is the nominal cut ``k_cap`` and is the ``K`` of the SERP*/PRAG*
normalisers.
"""
from __future__ import annotations

import numpy as np

# 31 counterfactual variants = 3 gender + 7 age + 21 occupation values
# (ml-1m/README:103-135 for the age/occupation codes).
ATTRIBUTE_SLICES = {"gender": (0, 3), "age": (3, 10), "occupation": (10, 31)}

EMPTY_BOTH_IS_ONE = 1   # src/utils.py:232-237 and src/phase3_facter_mitigation.py:358-363
EMPTY_ANY_IS_ZERO = 0   # src/phase3_final.py:112-117 (and src/phase3_aggressive.py:174-179)


# --------------------------------------------------------------------------
# list metrics
# --------------------------------------------------------------------------
def _first_occurrence(ids: np.ndarray, valid: np.ndarray) -> np.ndarray:
    """``first[..., j]`` is True when position j is valid and no valid i<j holds the same id."""
    K = ids.shape[-1]
    eq = ids[..., :, None] == ids[..., None, :]            # [..., j, i]
    eq &= valid[..., None, :]
    lower = np.tril(np.ones((K, K), dtype=bool), k=-1)      # i < j
    dup = (eq & lower).any(axis=-1)
    return valid & ~dup


def list_metrics(neutral, neutral_len, lists, lens, k_cap=None, vocab=None,
                 empty_policy=EMPTY_BOTH_IS_ONE, chunk_users=2048):
    """Per (user, variant) list similarity against the user's neutral list.

    Jaccard follows src/utils.py:228-237 (sets, |A∩B| / |A∪B|, both empty ->
    1.0) or, with ``empty_policy=EMPTY_ANY_IS_ZERO``, src/phase3_final.py:112-117
    (either empty -> 0.0).  With ``k_cap=10`` it is the quality-preservation
    overlap of src/phase3_facter_mitigation.py:354-363.

    Not in the reference (normative here; FaiRLLM, arXiv 2305.07609, as
    restated in SURVEY.md section 8(c)), with A the neutral sequence, B the
    variant sequence, r(x) the rank of x in A (+inf when absent), positions
    0-based, Kn = k_cap:

      serp_num = sum over first occurrences j of B with B[j] in A of (Kn - j)
      SERP*    = serp_num / (Kn (Kn + 1) / 2)
      prag_num = #{ (i, j) first occurrences of B, i < j, B[i] in A,
                    r(B[i]) < r(B[j]) }          (B[j] not in A counts: r = +inf)
      PRAG*    = prag_num / (Kn (Kn + 1))

    Returns a dict of arrays shaped [U, V]: integer ``inter``, ``union``,
    ``n_a`` (shape [U]), ``n_b``, ``serp_num``, ``prag_num`` (exact), and
    float64 ``jaccard``, ``serp``, ``prag``.
    """
    neutral = np.asarray(neutral)
    lists = np.asarray(lists)
    U, V, K = lists.shape
    assert neutral.shape == (U, K)
    neutral_len = np.asarray(neutral_len).reshape(U)
    lens = np.asarray(lens).reshape(U, V)
    Kn = K if k_cap is None else int(k_cap)
    if vocab is not None:
        pos = np.arange(K)
        live_a = pos[None, :] < np.minimum(neutral_len, Kn)[:, None]
        live_b = pos[None, None, :] < np.minimum(lens, Kn)[:, :, None]
        bad = ((neutral < 0) | (neutral >= vocab)) & live_a
        badb = ((lists < 0) | (lists >= vocab)) & live_b
        if bad.any() or badb.any():
            raise ValueError("item id outside [0, vocab)")

    out = {
        "inter": np.zeros((U, V), np.int64), "union": np.zeros((U, V), np.int64),
        "n_a": np.zeros((U,), np.int64), "n_b": np.zeros((U, V), np.int64),
        "serp_num": np.zeros((U, V), np.int64), "prag_num": np.zeros((U, V), np.int64),
    }
    pos = np.arange(K)
    upper = np.triu(np.ones((K, K), dtype=bool), k=1)        # i < j
    BIG = K + 1
    for s in range(0, U, chunk_users):
        e = min(U, s + chunk_users)
        A = neutral[s:e]
        B = lists[s:e]
        va = pos[None, :] < np.minimum(neutral_len[s:e], Kn)[:, None]
        vb = pos[None, None, :] < np.minimum(lens[s:e], Kn)[:, :, None]
        fa = _first_occurrence(A, va)                         # [u, K]
        fb = _first_occurrence(B, vb)                         # [u, V, K]
        match = (B[:, :, :, None] == A[:, None, None, :]) & fa[:, None, None, :]   # [u,V,j,i]
        hit = match.any(axis=-1)
        rank = np.where(hit, match.argmax(axis=-1), BIG)      # rank in A, BIG = +inf
        present = hit & fb
        out["n_a"][s:e] = fa.sum(-1)
        out["n_b"][s:e] = fb.sum(-1)
        out["inter"][s:e] = present.sum(-1)
        out["serp_num"][s:e] = (present * (Kn - pos)[None, None, :]).sum(-1)
        # pairs (i<j): first-occurrence j, present i, r_i < r_j
        lt = rank[:, :, :, None] < rank[:, :, None, :]        # [u,V,i,j]
        ok = present[:, :, :, None] & fb[:, :, None, :] & lt & upper[None, None]
        out["prag_num"][s:e] = ok.sum(axis=(-1, -2))
    out["union"] = out["n_a"][:, None] + out["n_b"] - out["inter"]
    out.update(ratios_from_counts(out["inter"], out["union"], out["n_a"][:, None], out["n_b"],
                                  out["serp_num"], out["prag_num"], Kn, empty_policy))
    return out


def ratios_from_counts(inter, union, n_a, n_b, serp_num, prag_num, Kn, empty_policy):
    """float64 ratios from the exact integer counts (same expression the reference
    evaluates: ``intersection / union if union > 0 else 0.0``, src/utils.py:237)."""
    inter = np.asarray(inter, np.float64)
    union = np.asarray(union, np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        jac = np.where(union > 0, inter / np.where(union > 0, union, 1.0), 0.0)
    n_a = np.broadcast_to(np.asarray(n_a), jac.shape)
    n_b = np.broadcast_to(np.asarray(n_b), jac.shape)
    if empty_policy == EMPTY_BOTH_IS_ONE:
        jac = np.where((n_a == 0) & (n_b == 0), 1.0, jac)
    else:
        jac = np.where((n_a == 0) | (n_b == 0), 0.0, jac)
    serp = np.asarray(serp_num, np.float64) / (Kn * (Kn + 1) / 2.0)
    prag = np.asarray(prag_num, np.float64) / (Kn * (Kn + 1.0))
    return {"jaccard": jac, "serp": serp, "prag": prag}


def list_metrics_pair_python(a, la, b, lb, Kn):
    """Scalar, set/dict based restatement of one pair; cross-checks the vectorised code."""
    A = list(a[:min(la, Kn)])
    B = list(b[:min(lb, Kn)])
    rank = {}
    for i, x in enumerate(A):
        rank.setdefault(int(x), i)
    seen, seq = set(), []
    for j, x in enumerate(B):
        x = int(x)
        if x in seen:
            continue
        seen.add(x)
        seq.append((j, x))
    inter = sum(1 for _, x in seq if x in rank)
    serp_num = sum(Kn - j for j, x in seq if x in rank)
    prag_num = 0
    for p, (i, x) in enumerate(seq):
        if x not in rank:
            continue
        for (j, y) in seq[p + 1:]:
            if rank[x] < rank.get(y, 1 << 30):
                prag_num += 1
    n_a, n_b = len(rank), len(seen)
    return dict(inter=inter, union=n_a + n_b - inter, n_a=n_a, n_b=n_b,
                serp_num=serp_num, prag_num=prag_num)


def variant_means(metric_uv):
    """Sim(a) = mean over users of metric[u, a]  (float64)."""
    return np.asarray(metric_uv, np.float64).mean(axis=0)


def snsr_snsv(sim_by_variant, slices=ATTRIBUTE_SLICES):
    """SNSR(attr) = max_a Sim(a) - min_a Sim(a); SNSV(attr) = population std over the
    attribute's values (ddof=0, consistent with ``np.std`` at
    src/phase1_bias_detection.py:238).  Not in the reference; normative here."""
    sim = np.asarray(sim_by_variant, np.float64)
    res = {}
    for name, (lo, hi) in slices.items():
        seg = sim[lo:hi]
        res[name] = {"snsr": float(seg.max() - seg.min()), "snsv": float(np.std(seg))}
    return res


# --------------------------------------------------------------------------
# cosine nonconformity
# --------------------------------------------------------------------------
def cosine_nonconformity(a, b):
    """s = 1 - <x,y> / (|x| |y|), clamped below at 0; a zero-norm row gives s = 1.

    ``a`` is [U, D] (one anchor row per user), ``b`` is [U, V, D] (its V
    counterfactual rows); the calibration form is V = 1.  float64 throughout,
    result float64 [U, V].  Not in the reference (which simulates a scalar
    score, src/phase3_facter_mitigation.py:125-139); normative here.
    """
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    dot = np.einsum("ud,uvd->uv", a, b)
    na = np.einsum("ud,ud->u", a, a)[:, None]
    nb = np.einsum("uvd,uvd->uv", b, b)
    den = np.sqrt(na * nb)
    with np.errstate(divide="ignore", invalid="ignore"):
        s = 1.0 - dot / den
    s = np.where(den > 0, np.maximum(s, 0.0), 1.0)
    return s


def cosine_nonconformity_mirror(a, b):
    """Bit-exact mirror of the CUDA kernel's arithmetic, for flag/threshold parity.

    The kernel accumulates exact fp32 x fp32 products in fp64: lane l of a warp
    takes float4 chunks c = l, l+32, l+64, ... (elements 4c..4c+3, in order),
    then the 32 lane sums are combined by an xor butterfly (offsets 16,8,4,2,1).
    cos = dot / sqrt(na*nb) in fp64, s = max(0, 1-cos), rounded once to fp32.
    D must be a multiple of 4.
    """
    a = np.asarray(a, np.float32)
    b = np.asarray(b, np.float32)
    U, V, D = b.shape
    assert D % 4 == 0
    nchunk = D // 4

    def lane_sums(x, y):                      # x,y float64 [..., D]
        acc = np.zeros(x.shape[:-1] + (32,), np.float64)
        prod = x * y                           # exact: 24+24 bits < 53
        for c0 in range(0, nchunk, 32):
            for e in range(4):
                idx = (np.arange(c0, min(c0 + 32, nchunk))) * 4 + e
                lanes = np.arange(idx.size)
                acc[..., lanes] = acc[..., lanes] + prod[..., idx]
        for off in (16, 8, 4, 2, 1):
            acc = acc + acc[..., np.arange(32) ^ off]
        return acc[..., 0]

    a64 = a.astype(np.float64)
    b64 = b.astype(np.float64)
    na = lane_sums(a64, a64)[:, None]
    nb = lane_sums(b64, b64)
    dot = lane_sums(np.broadcast_to(a64[:, None, :], b64.shape), b64)
    den = np.sqrt(na * nb)
    with np.errstate(divide="ignore", invalid="ignore"):
        s = 1.0 - dot / den
    s = np.where(den > 0, np.maximum(s, 0.0), 1.0)
    return s.astype(np.float32)


# --------------------------------------------------------------------------
# conformal threshold (order statistic)
# --------------------------------------------------------------------------
def conformal_index(n: int, alpha: float) -> int:
    """src/phase3_facter_mitigation.py:174-175, the expression verbatim in meaning:
    ``int(np.ceil((n + 1) * (1 - alpha))) - 1`` clamped to [0, n-1]."""
    idx = int(np.ceil((n + 1) * (1 - alpha))) - 1
    return max(0, min(idx, n - 1))


def conformal_threshold(scores, alpha=0.1):
    """Order statistic of src/phase3_facter_mitigation.py:165-178 for one group;
    empty -> 0.5 (:169-171).  Returns the selected element unchanged (same dtype)."""
    scores = np.asarray(scores)
    n = scores.size
    if n == 0:
        return 0.5
    return np.sort(scores.ravel())[conformal_index(n, alpha)]


def conformal_thresholds_by_group(scores, groups, n_groups, alpha=0.1):
    """Per-group thresholds (src/phase3_facter_mitigation.py:159-178) over integer group ids."""
    scores = np.asarray(scores).ravel()
    groups = np.asarray(groups).ravel()
    return [conformal_threshold(scores[groups == g], alpha) for g in range(n_groups)]


def select_counts(scores, thr):
    """(#less, #equal, #greater) around a threshold -- what the radix select reports."""
    scores = np.asarray(scores).ravel()
    less = int((scores < thr).sum())
    eq = int((scores == thr).sum())
    return less, eq, scores.size - less - eq


# --------------------------------------------------------------------------
# violation detection with round-synchronous threshold updates
# --------------------------------------------------------------------------
def violation_rounds(scores, q0, gamma=0.95, rounds=4, q_min=0.0):
    """FACTER-style violation detection.  Not in the reference; normative here
    (SURVEY.md section 8(c)).

    Round r = 0..R-1, threshold Q_r (fp32), over ALL scores of the batch (all
    shards): flag_r[u,v] = scores[u,v] > Q_r ; c_r = number of flags;
    Q_{r+1} = max(fl32(fl32(gamma) * Q_r), fl32(q_min)) if c_r > 0 else Q_r.

    Returns ``flags`` uint8 [U,V] with bit r = flag_r, ``counts`` int64 [R],
    ``variant_counts`` int64 [V,R], ``thresholds`` float32 [R+1].
    """
    s = np.asarray(scores, np.float32)
    U, V = s.shape
    q = np.float32(q0)
    g32 = np.float32(gamma)
    qmin = np.float32(q_min)
    flags = np.zeros((U, V), np.uint8)
    counts = np.zeros(rounds, np.int64)
    vcounts = np.zeros((V, rounds), np.int64)
    thr = np.zeros(rounds + 1, np.float32)
    for r in range(rounds):
        thr[r] = q
        f = s > q
        flags |= (f.astype(np.uint8) << r).astype(np.uint8)
        counts[r] = int(f.sum())
        vcounts[:, r] = f.sum(axis=0)
        if counts[r] > 0:
            q = np.maximum(np.float32(g32 * q), qmin)
    thr[rounds] = q
    return {"flags": flags, "counts": counts, "variant_counts": vcounts, "thresholds": thr}


def threshold_schedule(q0, n_greater, gamma=0.95, rounds=4, q_min=0.0):
    """Closed form of the trajectory in ``violation_rounds`` when 0 < gamma <= 1 and
    q0 >= 0: the thresholds never rise, so c_r > 0 for every r once c_0 > 0, and
    c_0 = n_greater (the count above the selected order statistic)."""
    q = np.float32(q0)
    thr = np.zeros(rounds + 1, np.float32)
    for r in range(rounds):
        thr[r] = q
        if n_greater > 0:
            q = np.maximum(np.float32(np.float32(gamma) * q), np.float32(q_min))
    thr[rounds] = q
    return thr


# --------------------------------------------------------------------------
# conformal filtering (prefix length)
# --------------------------------------------------------------------------
def conformal_keep_len(list_len, thr):
    """src/phase3_facter_mitigation.py:205-216: keep ranks with ``1.0 - rank*0.05 >= thr``
    (evaluated in float64 exactly as the reference does), falling back to the first 3
    when fewer than 3 survive and at least 3 exist.  Because the confidence falls with
    rank the kept set is a prefix; returns its length."""
    list_len = np.asarray(list_len, np.int64)
    thr = np.broadcast_to(np.asarray(thr, np.float64), list_len.shape)
    kmax = int(list_len.max()) if list_len.size else 0
    ranks = np.arange(kmax)
    conf = 1.0 - (ranks * 0.05)
    ok = (conf[None, :] >= thr.reshape(-1, 1)) & (ranks[None, :] < list_len.reshape(-1, 1))
    kept = ok.sum(axis=1).reshape(list_len.shape)
    return np.where((kept < 3) & (list_len >= 3), 3, kept)


# --------------------------------------------------------------------------
# pairwise similarity matrices (config 4)
# --------------------------------------------------------------------------
def bf16_round(x):
    """Round-to-nearest-even fp32 -> bf16, returned as fp32 holding the bf16 value."""
    x = np.ascontiguousarray(x, np.float32)
    u = x.view(np.uint32).astype(np.uint64)
    r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16
    return r.astype(np.uint32).view(np.float32).reshape(x.shape)


def normalize_rows_bf16(x):
    """Rows L2-normalised in fp32 then rounded to bf16."""
    x = np.asarray(x, np.float32)
    n = np.sqrt((x.astype(np.float64) ** 2).sum(-1, keepdims=True))
    y = np.where(n > 0, x / np.where(n > 0, n, 1.0), 0.0).astype(np.float32)
    return bf16_round(y)


def pair_similarity(xhat_bf16):
    """S[u] = X[u] X[u]^T in float64 on the bf16-rounded inputs; tolerance 2e-3 abs."""
    x = np.asarray(xhat_bf16, np.float64)
    return np.einsum("umd,und->umn", x, x)


def pair_jaccard(lists, lens, empty_policy=EMPTY_BOTH_IS_ONE):
    """All-pairs Jaccard between the M lists of each user: [U, M, K] -> [U, M, M]
    (integer intersections exact).  Same set semantics as ``list_metrics``."""
    lists = np.asarray(lists)
    U, M, K = lists.shape
    lens = np.asarray(lens).reshape(U, M)
    inter = np.zeros((U, M, M), np.int64)
    sizes = np.zeros((U, M), np.int64)
    for u in range(U):
        sets = [set(int(x) for x in lists[u, m, :lens[u, m]]) for m in range(M)]
        for m in range(M):
            sizes[u, m] = len(sets[m])
            for n in range(M):
                inter[u, m, n] = len(sets[m] & sets[n])
    union = sizes[:, :, None] + sizes[:, None, :] - inter
    r = ratios_from_counts(inter, union, sizes[:, :, None], sizes[:, None, :], 0, 0, 1, empty_policy)
    return {"inter": inter, "sizes": sizes, "jaccard": r["jaccard"]}


# --------------------------------------------------------------------------
# radix select restated (digit structure of the CUDA select; used to test sharded runs)
# --------------------------------------------------------------------------
def radix_key_f32(x):
    """Order-preserving uint32 keys: unsigned key order == np.sort order (NaN last, -0 == +0)."""
    x = np.ascontiguousarray(x, np.float32)
    u = x.view(np.uint32)
    key = np.where(u & np.uint32(0x80000000), ~u, u | np.uint32(0x80000000)).astype(np.uint32)
    key = np.where(x == 0, np.uint32(0x80000000), key)
    return np.where(np.isnan(x), np.uint32(0xFFFFFFFF), key).astype(np.uint32)


def radix_unkey_f32(k):
    k = np.asarray(k, np.uint32)
    u = np.where(k & np.uint32(0x80000000), k & np.uint32(0x7FFFFFFF), ~k).astype(np.uint32)
    return u.view(np.float32)


def radix_select_sharded(shards, alpha=0.1, all_reduce=lambda h: h):
    """Three 11/11/10-bit passes over float32 score shards; ``all_reduce`` sums a histogram
    over ranks.  Returns (threshold, n, k, less, equal) exactly as the CUDA select does."""
    keys = [radix_key_f32(np.asarray(s).ravel()) for s in shards]
    prefix, less, krem, n, k = 0, 0, 0, 0, 0
    for p, (shift, nbits) in enumerate(((21, 11), (10, 11), (0, 10))):
        hist = np.zeros(2048, np.int64)
        for kk in keys:
            sel = kk if p == 0 else kk[(kk >> np.uint32(shift + nbits)) == (prefix >> (shift + nbits))]
            hist += np.bincount(((sel >> np.uint32(shift)) & np.uint32((1 << nbits) - 1)).astype(np.int64),
                                minlength=2048)
        hist = all_reduce(hist)
        if p == 0:
            n = int(hist.sum())
            if n == 0:
                return np.float32(0.5), 0, 0, 0, 0
            k = conformal_index(n, alpha)
            krem = k
        cum = np.cumsum(hist)
        b = int(np.searchsorted(cum, krem, side="right"))
        before = int(cum[b] - hist[b])
        prefix |= b << shift
        krem -= before
        less += before
        equal = int(hist[b])
    return radix_unkey_f32(np.uint32(prefix)), n, k, less, equal


# --------------------------------------------------------------------------
# demographic parity
# --------------------------------------------------------------------------
def group_item_counts(lists, lens, groups, n_groups, vocab, k_cap=None):
    """counts[g, x] = occurrences of item x over all rows of group g, WITH multiplicity
    (src/utils.py:180-187 counts every occurrence; nothing is de-duplicated here)."""
    lists = np.asarray(lists)
    R, K = lists.shape
    lens = np.minimum(np.asarray(lens).reshape(R), K if k_cap is None else k_cap)
    groups = np.asarray(groups).reshape(R)
    live = np.arange(K)[None, :] < lens[:, None]
    live &= (lists >= 0) & (lists < vocab)
    flat = (groups[:, None].astype(np.int64) * vocab + np.where(live, lists, 0))[live]
    return np.bincount(flat, minlength=n_groups * vocab).reshape(n_groups, vocab).astype(np.int64)


def js_distance_from_counts(ci, cj):
    """calculate_js_divergence (src/utils.py:89-102) on two count vectors over a shared item
    vocabulary: keys = items either group holds; p = count/total or 1e-10 when missing; renormalise;
    scipy's jensenshannon (normalise again, natural log, square root)."""
    ci = np.asarray(ci, np.float64)
    cj = np.asarray(cj, np.float64)
    keys = (ci > 0) | (cj > 0)
    p = np.where(ci > 0, ci / ci.sum(), 1e-10)[keys]
    q = np.where(cj > 0, cj / cj.sum(), 1e-10)[keys]
    p = p / p.sum()
    q = q / q.sum()
    p = p / np.sum(p)
    q = q / np.sum(q)
    m = (p + q) / 2.0
    return float(np.sqrt((np.sum(p * np.log(p / m)) + np.sum(q * np.log(q / m))) / 2.0))


def demographic_parity_from_counts(counts):
    """Score and pairwise distances (i < j, both groups non-empty, row-major order) as
    calculate_demographic_parity (src/utils.py:196-208) produces them."""
    counts = np.asarray(counts)
    G = counts.shape[0]
    nonempty = counts.sum(1) > 0
    div = [js_distance_from_counts(counts[i], counts[j])
           for i in range(G) for j in range(i + 1, G) if nonempty[i] and nonempty[j]]
    avg = float(np.mean(div)) if div else 0.0
    return 1.0 - avg, div
