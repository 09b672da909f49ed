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


def _dense_codes(flat: list) -> Tuple[np.ndarray, int]:
    """Dense int32 ids for a flat list of hashables, equal items <=> equal ids (Python ``==``/``hash``
    semantics, as in the reference's sets).  All-string input -- the reference's case -- goes through
    pandas' C hash table; anything else through a dict."""
    if not flat:
        return np.zeros(0, np.int32), 0
    arr = np.empty(len(flat), dtype=object)
    arr[:] = flat
    try:
        import pandas as pd
        if pd.api.types.infer_dtype(arr, skipna=False) == "string":
            codes, uniques = pd.factorize(arr, use_na_sentinel=False)
            return codes.astype(np.int32), len(uniques)
    except ImportError:
        pass
    index: Dict[Hashable, int] = {}
    codes = np.fromiter((index.setdefault(x, len(index)) for x in flat), dtype=np.int32, count=len(flat))
    return codes, len(index)


def encode_pair_lists(lists_a: Sequence[Sequence[Hashable]], lists_b: Sequence[Sequence[Hashable]]):
    """Two aligned sequences of item lists -> (a [P,K], len_a [P], b [P,K], len_b [P], vocab size).
    Lists that are the same Python object are encoded once; everything after the flattening is
    vectorised (one hash pass over all items, one scatter into the padded table, two row gathers)."""
    slot: Dict[int, int] = {}
    order: List[Sequence[Hashable]] = []

    def row_of(lst) -> int:
        key = id(lst)
        i = slot.get(key)
        if i is None:
            i = len(order)
            slot[key] = i
            order.append(lst)
        return i

    ia = np.fromiter((row_of(l) for l in lists_a), dtype=np.int64, count=len(lists_a))
    ib = np.fromiter((row_of(l) for l in lists_b), dtype=np.int64, count=len(lists_b))
    lens = np.fromiter((len(l) for l in order), dtype=np.int32, count=len(order))
    flat = [x for l in order for x in l]
    codes, n_vocab = _dense_codes(flat)
    K = max(int(lens.max()) if len(order) else 0, 1)
    table = np.full((len(order), K), -1, dtype=np.int32)
    if len(flat):
        starts = np.cumsum(lens, dtype=np.int64) - lens
        rows = np.repeat(np.arange(len(order), dtype=np.int64), lens)
        table[rows, np.arange(len(flat), dtype=np.int64) - np.repeat(starts, lens)] = codes
    return table[ia], lens[ia], table[ib], lens[ib], max(n_vocab, 1)


class EncodedRecommendations:
    """Recommendation lists encoded ONCE: one vocabulary, one padded id table, resident on the device.

    The metric step of a phase (src/phase1_bias_detection.py:343-405) looks at the same lists four times
    -- demographic parity by gender and by age, individual fairness, equal opportunity.  Encoded per
    call, the string hashing costs more than everything the GPU does; encoded here, each metric only
    sends row indices.  A row is found again by the identity of its Python list (``row_of``) or by the
    key it was registered under (``row_of_key``)."""

    def __init__(self, lists: Sequence[Sequence[Hashable]], keys: Sequence[Hashable] = None):
        order: List[Sequence[Hashable]] = []
        row_of: Dict[int, int] = {}
        seq: List[int] = []                      # row of every input list, in input order
        for lst in lists:
            k = id(lst)
            r = row_of.get(k)
            if r is None:
                r = row_of[k] = len(order)
                order.append(lst)
            seq.append(r)
        self._row_of_obj, self._keep, self.input_rows = row_of, order, seq   # _keep holds the lists alive: ids stay unique
        self.row_of_key: Dict[Hashable, int] = dict(zip(keys, seq)) if keys is not None else {}
        lens_list = [len(l) for l in order]
        n_items = sum(lens_list)
        if n_items <= 65536:                     # small inputs: a dict beats the set-up cost of a hash-table library
            index: Dict[Hashable, int] = {}
            sd = index.setdefault
            codes = [sd(x, len(index)) for l in order for x in l]
            n_vocab, uniques = len(index), list(index)
        else:
            codes, n_vocab, uniques = _dense_codes_with_items([x for l in order for x in l])
        K = max(max(lens_list, default=0), 1)
        lens = np.array(lens_list, dtype=np.int32)
        if order and min(lens_list) == K:        # every list full: the codes ARE the table
            table = np.array(codes, dtype=np.int32).reshape(len(order), K)
        else:
            table = np.full((len(order), K), -1, dtype=np.int32)
            if n_items:
                starts = np.cumsum(lens, dtype=np.int64) - lens
                rows = np.repeat(np.arange(len(order), dtype=np.int64), lens)
                table[rows, np.arange(n_items, dtype=np.int64) - np.repeat(starts, lens)] = codes
        self.table, self.lens, self.vocab_size, self.items = table, lens, max(n_vocab, 1), uniques
        self._device_copy = None

    def __len__(self):
        return len(self._keep)

    def row_of(self, lst) -> int:
        """Row of a list registered at construction (by object identity); -1 if unknown."""
        return self._row_of_obj.get(id(lst), -1)

    def rows_of(self, lists) -> np.ndarray:
        """int64 rows of the given list objects, or None if one of them was not registered."""
        get = self._row_of_obj.get
        rows = np.fromiter((get(id(l), -1) for l in lists), dtype=np.int64, count=len(lists))
        return None if (len(rows) and rows.min() < 0) else rows

    def on_device(self, device):
        """(table int32 [R, K], lens int32 [R]) on ``device``; uploaded on first use, in one copy."""
        import torch
        if self._device_copy is None or self._device_copy[0].device != device:
            R, K = self.table.shape
            packed = np.concatenate([self.table.reshape(-1), self.lens])
            d = torch.from_numpy(packed).to(device)
            self._device_copy = (d[:R * K].view(R, K), d[R * K:])
        return self._device_copy


def _dense_codes_with_items(flat: list):
    """``_dense_codes`` that also returns the distinct items in code order."""
    if not flat:
        return np.zeros(0, np.int32), 0, []
    arr = np.empty(len(flat), dtype=object)
    arr[:] = flat
    try:
        import pandas as pd
        if pd.api.types.infer_dtype(arr, skipna=False) == "string":
            codes, uniques = pd.factorize(arr, use_na_sentinel=False)
            return codes.astype(np.int32), len(uniques), list(uniques)
    except ImportError:
        pass
    index: Dict[Hashable, int] = {}
    codes = np.fromiter((index.setdefault(x, len(index)) for x in flat), dtype=np.int32, count=len(flat))
    return codes, len(index), list(index.keys())
