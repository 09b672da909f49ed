"""Host-side boundary between the reference's Python objects and dense id tensors.

The reference compares raw title strings with ``set`` semantics: exact, case- and
whitespace-sensitive equality (src/utils.py:228-229).  Here every distinct string of a call
gets a dense int32 id, so equality of ids == equality of strings; anything hashable works
the same way, as in the reference.
"""
from __future__ import annotations

from typing import Dict, Hashable, Iterable, List, Sequence, Tuple

import numpy as np


class Vocabulary:
    """Dense ids for the distinct items seen so far (insertion order)."""

    def __init__(self):
        self.index: Dict[Hashable, int] = {}

    def __len__(self):
        return len(self.index)

    def encode(self, items: Iterable[Hashable]) -> List[int]:
        idx = self.index
        out = []
        for it in items:
            i = idx.get(it)
            if i is None:
                i = len(idx)
                idx[it] = i
            out.append(i)
        return out


def pack_rows(rows: Sequence[Sequence[int]], k: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """Ragged id rows -> (int32 [R, K] padded with -1, int32 lengths [R]); K >= 1."""
    lens = np.fromiter((len(r) for r in rows), dtype=np.int32, count=len(rows))
    K = max(int(lens.max()) if len(rows) else 0, int(k), 1)
    out = np.full((len(rows), K), -1, dtype=np.int32)
    for i, r in enumerate(rows):
        if r:
            out[i, :len(r)] = r
    return out, lens


def encode_pair_lists(lists_a: Sequence[Sequence[Hashable]], lists_b: Sequence[Sequence[Hashable]]):
    """Two aligned sequences of item lists -> (a [P,K], len_a [P], b [P,K], len_b [P], vocab size).
    Lists that are the same Python object are encoded once."""
    vocab = Vocabulary()
    cache: Dict[int, List[int]] = {}

    def enc(lst):
        key = id(lst)
        r = cache.get(key)
        if r is None:
            r = vocab.encode(lst)
            cache[key] = r
        return r

    rows_a = [enc(l) for l in lists_a]
    rows_b = [enc(l) for l in lists_b]
    kmax = max([len(r) for r in rows_a] + [len(r) for r in rows_b] + [1])
    a, la = pack_rows(rows_a, kmax)
    b, lb = pack_rows(rows_b, kmax)
    return a, la, b, lb, max(len(vocab), 1)
