"""The benchmarked pipeline: list metrics + conformal scoring over a shard of users.

Users are independent, so a job is sharded by contiguous user ranges, one process per GPU.
Nothing bulky ever crosses NVLink: per step the ranks all-reduce (sum) only
  * the 2048-bin radix histograms of the select (3 passes; pass 0 is built inside the cosine
    kernel), so every rank derives the same GLOBAL quantile threshold, and
  * one small float64 buffer holding the per-variant metric sums and the violation counts
    (integers below 2^53, exact in float64).
Definitions of every quantity: oracle/np_oracle.py.  SURVEY.md section 8(e).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import ops

ATTRIBUTE_SLICES = {"gender": (0, 3), "age": (3, 10), "occupation": (10, 31)}
SEED = 20251203


@dataclass
class PipelineConfig:
    n_variants: int = 31
    list_len: int = 50            # K
    emb_dim: int = 384            # D
    vocab: int = 3706
    alpha: float = 0.1
    gamma: float = 0.95
    rounds: int = 4
    q_min: float = 0.0
    k_cap: Optional[int] = None
    empty_policy: int = ops.EMPTY_BOTH_IS_ONE
    attribute_slices: Dict[str, Tuple[int, int]] = field(default_factory=lambda: dict(ATTRIBUTE_SLICES))

    def bytes_per_user(self) -> int:
        V, K, D = self.n_variants, self.list_len, self.emb_dim
        return (V + 1) * K * 4 + (V + 1) * 4 + (V + 1) * D * 4


@dataclass
class Batch:
    """One shard's inputs (device tensors, or pinned host tensors for ``run_host``)."""
    neutral: torch.Tensor        # int32 [U, K]
    neutral_len: torch.Tensor    # int32 [U]
    lists: torch.Tensor          # int32 [U, V, K]
    lens: torch.Tensor           # int32 [U, V]
    emb_neutral: torch.Tensor    # float32 [U, D]
    emb: torch.Tensor            # float32 [U, V, D]

    @property
    def n_users(self) -> int:
        return self.lists.shape[0]

    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in
                   (self.neutral, self.neutral_len, self.lists, self.lens, self.emb_neutral, self.emb))

    def tensors(self):
        return (self.neutral, self.neutral_len, self.lists, self.lens, self.emb_neutral, self.emb)


@dataclass
class StepResult:
    """Device-side outputs of one step (per-pair arrays cover this rank's shard only)."""
    inter: torch.Tensor           # int32 [U, V]
    jaccard: torch.Tensor         # float32 [U, V]
    serp: torch.Tensor            # float32 [U, V]
    prag: torch.Tensor            # float32 [U, V]
    scores: torch.Tensor          # float32 [U, V]
    flags: torch.Tensor           # uint8 [U, V], bit r = violation in round r
    thresholds: torch.Tensor      # float32 [R+1], global
    select_state: torch.Tensor    # int64 [1, 8], global (n, k, less, equal)
    reduced: torch.Tensor         # float64 [V*4 + R + V*R], global after all-reduce
    status: torch.Tensor          # int32 [1]


def shard_range(n_users: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous user range [lo, hi) of ``rank``."""
    base, rem = divmod(n_users, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def synth_batch(cfg: PipelineConfig, user0: int, n_users: int, device, seed: int = SEED) -> Batch:
    """Generate a shard on the device (Philox; identical bytes to ``synth_batch_host``)."""
    neutral, nlen, lists, lens = ops.synth_lists(seed, user0, n_users, cfg.n_variants, cfg.list_len, cfg.vocab,
                                                 device)
    en, e = ops.synth_embeddings(seed, user0, n_users, cfg.n_variants, cfg.emb_dim, device)
    return Batch(neutral, nlen, lists, lens, en, e)


def synth_batch_host(cfg: PipelineConfig, user0: int, n_users: int, seed: int = SEED, pin: bool = False) -> Batch:
    arrs = ops.synth_lists_host(seed, user0, n_users, cfg.n_variants, cfg.list_len, cfg.vocab)
    embs = ops.synth_embeddings_host(seed, user0, n_users, cfg.n_variants, cfg.emb_dim)
    ts = [torch.from_numpy(a) for a in (*arrs, *embs)]
    if pin:
        ts = [t.pin_memory() for t in ts]
    return Batch(*ts)


class ScoringPipeline:
    """list_metrics -> cosine (+pass-0 histogram) -> exact quantile -> threshold schedule ->
    violation flags, with the two small all-reduces when ``process_group``/world > 1."""

    def __init__(self, cfg: PipelineConfig, distributed: bool = False, process_group=None):
        self.cfg = cfg
        self.distributed = distributed
        self.pg = process_group
        self.n_launches = 0          # kernels of this library enqueued so far
        self.kernel_events = None    # bench.py sets a list to collect (name, start, end) CUDA events

    def _timed(self, name, fn):
        if self.kernel_events is None:
            return fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = fn()
        b.record()
        self.kernel_events.append((name, a, b))
        return out

    def kernel_names(self) -> Tuple[str, str]:
        """The two kernels that carry a step's bytes, by the names the profiler shows."""
        cfg = self.cfg
        return (ops.list_metrics_kernel_name(cfg.n_variants, cfg.list_len, cfg.vocab, cfg.k_cap),
                ops.cosine_kernel_name(cfg.n_variants, cfg.emb_dim))

    # -- collectives -------------------------------------------------------------------
    def _all_reduce(self, t: torch.Tensor) -> None:
        if self.distributed:
            import torch.distributed as dist
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.pg)

    # -- one step over device-resident inputs ---------------------------------------------
    def run(self, b: Batch) -> StepResult:
        cfg = self.cfg
        V, R = cfg.n_variants, cfg.rounds
        dev = b.lists.device
        hist0 = torch.zeros(ops.SELECT_BINS, dtype=torch.int64, device=dev)
        lm = self._timed(self.kernel_names()[0], lambda: ops.list_metrics(
            b.neutral, b.neutral_len, b.lists, b.lens, vocab=cfg.vocab, k_cap=cfg.k_cap,
            empty_policy=cfg.empty_policy, rank_metrics=True, want_union=False, want_sums=True))
        scores = self._timed(self.kernel_names()[1], lambda: ops.cosine_nonconformity(b.emb_neutral, b.emb, hist0=hist0))
        self.n_launches += 3
        return self._finish(lm.inter, lm.jaccard, lm.serp, lm.prag, lm.variant_sums, lm.status, scores, hist0)

    def _finish(self, inter, jac, serp, prag, variant_sums, status, scores, hist0) -> StepResult:
        cfg = self.cfg
        V, R = cfg.n_variants, cfg.rounds
        dev = scores.device
        thr0, state = ops.radix_select(scores, alpha=cfg.alpha, hist0=hist0, all_reduce=self._all_reduce)
        thr = ops.threshold_schedule(thr0, state, gamma=cfg.gamma, q_min=cfg.q_min, rounds=R)
        reduced = torch.zeros(V * 4 + R + V * R, dtype=torch.float64, device=dev)
        counts = torch.zeros(R + V * R, dtype=torch.int64, device=dev)
        flags, _, _ = ops.violation_detect(scores, thr, rounds=R, n_variants=V, per_variant=True,
                                           counts=counts[:R], variant_counts=counts[R:].view(V, R))
        self.n_launches += 2 + 3 + 1 + 1      # 2 hist + 3 scan + schedule + detect
        reduced[:V * 4] = variant_sums.reshape(-1)
        reduced[V * 4:] = counts.to(torch.float64)
        self._all_reduce(reduced)
        return StepResult(inter, jac, serp, prag, scores, flags, thr, state, reduced, status)

    # -- the call a user makes with HOST buffers -----------------------------------------
    def run_host(self, hb: Batch, device, chunk_users: int = 16384) -> Tuple[StepResult, dict]:
        """Stream a pinned host batch through the GPU in user chunks (H2D on a copy stream,
        double-buffered against the kernels), finish the global steps, and read the summary
        back.  Returns (device results, host summary)."""
        cfg = self.cfg
        V, K, D, R = cfg.n_variants, cfg.list_len, cfg.emb_dim, cfg.rounds
        U = hb.n_users
        dev = torch.device(device)
        chunk = max(1, min(chunk_users, U))
        n_chunks = (U + chunk - 1) // chunk
        main = torch.cuda.current_stream(dev)
        copy = self._copy_stream(dev)
        bufs = self._staging(dev, chunk)
        i32, f32 = torch.int32, torch.float32
        inter = torch.empty((U, V), dtype=i32, device=dev)
        jac = torch.empty((U, V), dtype=f32, device=dev)
        serp = torch.empty((U, V), dtype=f32, device=dev)
        prag = torch.empty((U, V), dtype=f32, device=dev)
        scores = torch.empty((U, V), dtype=f32, device=dev)
        sums = torch.zeros((n_chunks, V, 4), dtype=torch.float64, device=dev)
        status = torch.zeros(1, dtype=i32, device=dev)
        hist0 = torch.zeros(ops.SELECT_BINS, dtype=torch.int64, device=dev)
        ws_bytes = ops._lib.load().flr_list_metrics_scratch_bytes(chunk, V, K, cfg.vocab)
        ws = ops._workspace(dev, ws_bytes)
        k_cap = K if cfg.k_cap is None else cfg.k_cap
        freed = [None, None]
        copy.wait_stream(main)
        for c in range(n_chunks):
            s, e = c * chunk, min(U, (c + 1) * chunk)
            n = e - s
            slot = bufs[c & 1]
            with torch.cuda.stream(copy):
                if freed[c & 1] is not None:
                    copy.wait_event(freed[c & 1])
                for dst, src in zip(slot, hb.tensors()):
                    dst[:n].copy_(src[s:e], non_blocking=True)
                ready = copy.record_event()
            main.wait_event(ready)
            d_neutral, d_nlen, d_lists, d_lens, d_en, d_e = slot
            P = ops._ptr
            ops._lib.call("flr_list_metrics", P(d_neutral), P(d_nlen), P(d_lists), P(d_lens), n, V, K, k_cap,
                          cfg.vocab, cfg.empty_policy, 1, P(inter[s:]), None, P(jac[s:]), P(serp[s:]), P(prag[s:]),
                          P(sums[c]), P(status), P(ws), ws.numel(), ops._stream())
            ops._lib.call("flr_cosine_nonconformity", P(d_en), P(d_e), n, V, D, P(scores[s:]), P(hist0),
                          ops._stream())
            self.n_launches += 3
            freed[c & 1] = main.record_event()
        res = self._finish(inter, jac, serp, prag, sums.sum(dim=0), status, scores, hist0)
        return res, self.summarize(res)

    _streams: dict = {}
    _stage: dict = {}

    def _copy_stream(self, dev):
        key = (dev.type, dev.index)
        if key not in self._streams:
            self._streams[key] = torch.cuda.Stream(dev)
        return self._streams[key]

    def _staging(self, dev, chunk):
        cfg = self.cfg
        key = (dev.type, dev.index, chunk, cfg.n_variants, cfg.list_len, cfg.emb_dim)
        if key not in self._stage:
            V, K, D = cfg.n_variants, cfg.list_len, cfg.emb_dim
            mk = lambda shape, dt: torch.empty(shape, dtype=dt, device=dev)
            self._stage[key] = [(mk((chunk, K), torch.int32), mk((chunk,), torch.int32),
                                 mk((chunk, V, K), torch.int32), mk((chunk, V), torch.int32),
                                 mk((chunk, D), torch.float32), mk((chunk, V, D), torch.float32))
                                for _ in range(2)]
        return self._stage[key]

    # -- host summary (one small D2H; synchronises) ------------------------------------------
    def summarize(self, res: StepResult) -> dict:
        cfg = self.cfg
        V, R = cfg.n_variants, cfg.rounds
        red = res.reduced.cpu().numpy()
        thr = res.thresholds.cpu().numpy()
        st = res.select_state.cpu().numpy()[0]
        sums = red[:V * 4].reshape(V, 4)
        kn = float(cfg.list_len if cfg.k_cap is None else min(cfg.k_cap, cfg.list_len))
        users = np.maximum(sums[:, 3], 1.0)
        means = np.stack([sums[:, 0] / users, sums[:, 1] / users / (kn * (kn + 1) / 2.0),
                          sums[:, 2] / users / (kn * (kn + 1.0))], axis=1)          # [V, 3]
        spreads = {}
        for name, (lo, hi) in cfg.attribute_slices.items():
            if hi <= V:
                seg = means[lo:hi]
                spreads[name] = {m: {"snsr": float(seg[:, i].max() - seg[:, i].min()),
                                     "snsv": float(np.std(seg[:, i]))}
                                 for i, m in enumerate(("jaccard", "serp", "prag"))}
        counts = red[V * 4:V * 4 + R].astype(np.int64)
        vcounts = red[V * 4 + R:].reshape(V, R).astype(np.int64)
        return {"variant_means": means, "spreads": spreads, "thresholds": thr, "violation_counts": counts,
                "variant_violation_counts": vcounts, "n_scores": int(st[ops._lib.ST_N]),
                "select_rank": int(st[ops._lib.ST_K]), "n_users": int(sums[0, 3]),
                "oov_ids": int(res.status.item())}


def multi_gpu_parity(cfg: PipelineConfig, rank: int, world: int, device, users_per_rank: int = 8192,
                     process_group=None) -> Optional[dict]:
    """Correctness witness of the sharded path on real hardware.

    Every rank scores its own shard of ``users_per_rank`` users with the collectives on; rank 0 then
    scores the union of all shards alone on its GPU.  The global quantities must agree exactly --
    thresholds, select state (n, k, less, equal), violation counts, the integer SERP*/PRAG* sums --
    the Jaccard sums to 1e-12, and rank 0's per-pair outputs must equal the first slice of the union run.
    Returns the verdict on rank 0, None elsewhere.  Collective: every rank must call it."""
    V = cfg.n_variants
    shard = synth_batch(cfg, rank * users_per_rank, users_per_rank, device)
    res = ScoringPipeline(cfg, distributed=world > 1, process_group=process_group).run(shard)
    torch.cuda.synchronize(device)
    if rank != 0:
        return None
    union = synth_batch(cfg, 0, world * users_per_rank, device)
    one = ScoringPipeline(cfg).run(union)
    torch.cuda.synchronize(device)
    red_n, red_1 = res.reduced.cpu().numpy(), one.reduced.cpu().numpy()
    sums_n, sums_1 = red_n[:V * 4].reshape(V, 4), red_1[:V * 4].reshape(V, 4)
    n = users_per_rank
    checks = {
        "thresholds": bool(torch.equal(res.thresholds, one.thresholds)),
        "select_state": bool(torch.equal(res.select_state, one.select_state)),
        "violation_counts": bool(np.array_equal(red_n[V * 4:], red_1[V * 4:])),
        "integer_sums": bool(np.array_equal(sums_n[:, 1:], sums_1[:, 1:])),
        "jaccard_sums_1e-12": bool(np.allclose(sums_n[:, 0], sums_1[:, 0], rtol=1e-12, atol=0)),
        "shard0_pairs": bool(torch.equal(res.inter, one.inter[:n]) and torch.equal(res.scores, one.scores[:n])
                             and torch.equal(res.flags, one.flags[:n]) and torch.equal(res.prag, one.prag[:n])),
    }
    return {"multi_gpu_parity": all(checks.values()), "ranks": world, "users_per_rank": users_per_rank,
            "checks": checks}
