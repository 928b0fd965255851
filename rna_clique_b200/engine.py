"""In-memory all-vs-all engine: the device-resident form of the reference's phase 1 + phase 2.

  select_top_genes()  host, bit-exact restatement target of select_top_genes.py:108-127
  Engine.upload()     kernels (1)+(2)            (replaces makeblastdb, find_all_pairs.py:120-159)
  Engine.search()     kernel (3) per unordered pair (replaces 2x blastn + find_homologs.py:215-302)
  Engine.distances()  kernel (4) on the device-resident tables + exact Fractions on the host
                      (build_graph.py, filtered_distance.py, similarity_computer.py)

Multi-GPU (one process per GPU): pairs are sharded round-robin by estimated cost; the only
collective is an all-gather of the flat hit tables (torch.distributed, NCCL) before phase 2.
"""
from __future__ import annotations

import itertools
import os
import time
from concurrent.futures import ThreadPoolExecutor
from fractions import Fraction
from typing import Optional, Sequence

import numpy as np

from . import _native


class NoIdealComponentsError(Exception):
    """Reference: filtered_distance.py:126-127."""


from .select_top_genes import top_gene_mask as select_top_genes  # noqa: E402  (select_top_genes.py:108-127)


def schedule_pairs(n: int, totals: Sequence[int], world: int):
    """All unordered pairs in itertools.combinations order, and the rank that owns each:
    greedy longest-processing-time assignment by estimated cost L_i + L_j."""
    pairs = list(itertools.combinations(range(n), 2))
    cost = [totals[i] + totals[j] for i, j in pairs]
    owner = [0] * len(pairs)
    load = [0] * world
    for k in sorted(range(len(pairs)), key=lambda k: (-cost[k], k)):
        r = min(range(world), key=lambda r: (load[r], r))
        owner[k] = r
        load[r] += cost[k]
    return pairs, owner


class Engine:
    """`streams` CUDA contexts (one stream + scratch each) on one GPU, each driven by its own host
    thread: while one pair's host code waits for a device counter, the other streams keep the SMs
    busy.  ctypes releases the GIL during library calls, so the threads overlap for real."""

    def __init__(self, device: int = 0, streams: int = 3):
        self.ctxs = [_native.Context(device) for _ in range(max(1, streams))]
        self.ctx = self.ctxs[0]
        self.device = device
        self._pool = ThreadPoolExecutor(len(self.ctxs)) if len(self.ctxs) > 1 else None
        self._meta_group = None       # second process group for the metadata of the N > 1 sample exchange

    def _fan_out(self, n_items: int, fn):
        """out[k] = fn(ctx, k) for k < n_items; item k runs on context k % streams, and each context
        is only ever used by one thread at a time."""
        out = [None] * n_items
        S = len(self.ctxs)

        def work(w):
            for k in range(w, n_items, S):
                out[k] = fn(self.ctxs[w], k)

        if self._pool is None or n_items <= 1:
            for w in range(S):
                work(w)
        else:
            for f in [self._pool.submit(work, w) for w in range(S)]:
                f.result()
        return out

    def _prefetch(self, inputs, which):
        """Queues the host->device copy of every pinned `ascii` buffer among inputs[which] on the engine's
        copy stream; returns {k: (device tensor, event)}.  Pageable buffers are left to the upload itself
        (their copies would block this thread)."""
        staged = {}
        try:
            import torch
        except ImportError:
            return staged
        todo = [k for k in which if torch.is_tensor(inputs[k]["ascii"]) and not inputs[k]["ascii"].is_cuda
                and inputs[k]["ascii"].is_pinned()]
        if not todo:
            return staged
        dev = torch.device("cuda", self.device)
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(dev)
        with torch.cuda.stream(self._copy_stream):
            for k in todo:
                src = inputs[k]["ascii"].reshape(-1)
                dst = torch.empty(src.numel(), dtype=torch.uint8, device=dev)
                dst.copy_(src, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self._copy_stream)
                staged[k] = (dst, ev)
        return staged

    def _run_tasks(self, tasks, ready, run, poll: Optional[float] = None):
        """Workers (one per context) drain `tasks` in list order: worker w takes the first open task with
        ready(task, w) (evaluated under the lock) and calls run(ctx, task), which stores its own result.
        An exception in any worker is re-raised here."""
        import threading
        cv = threading.Condition()
        state = {"next": [True] * len(tasks), "left": len(tasks), "err": None}

        S = len(self.ctxs)
        # RCQ_TRACE_TASKS=1: (worker, task, start ms, end ms) of every task of the last call in self.task_trace
        trace = [] if os.environ.get("RCQ_TRACE_TASKS") else None
        self.task_trace = trace
        t_base = time.perf_counter()

        def work(w):
            ctx = self.ctxs[w]
            while True:
                with cv:
                    pick = None
                    while pick is None:
                        if state["err"] is not None or state["left"] == 0:
                            return
                        pending = False
                        for t, open_ in enumerate(state["next"]):
                            if not open_:
                                continue
                            pending = True
                            if ready(tasks[t], w):
                                pick = t
                                break
                        if pick is None:
                            if not pending:
                                return
                            cv.wait(poll)     # poll: readiness may change without a notify (a collective finishing)
                    state["next"][pick] = False
                try:
                    if trace is not None:
                        t0 = time.perf_counter()
                    run(ctx, tasks[pick])
                    if trace is not None:
                        trace.append((w, tasks[pick], t0 - t_base, time.perf_counter() - t_base))
                except BaseException as e:  # noqa: BLE001
                    with cv:
                        state["err"] = e
                        cv.notify_all()
                    return
                with cv:
                    state["left"] -= 1
                    cv.notify_all()

        if self._pool is None:
            work(0)
        else:
            for f in [self._pool.submit(work, w) for w in range(S)]:
                f.result()
        if state["err"] is not None:
            raise state["err"]

    @property
    def launches(self) -> int:
        return sum(c.launches for c in self.ctxs)

    def stage_times(self, reset: bool = False) -> dict:
        tot = None
        for c in self.ctxs:
            st = c.stage_times(reset)
            if tot is None:
                tot = st
            else:
                for k, v in st.items():
                    tot[k]["ms"] += v["ms"]
                    tot[k]["calls"] += v["calls"]
        return tot

    # ---- kernels (1)+(2)
    def upload(self, ascii_, starts, lengths, gene, iso, name: str = "", ctx=None) -> _native.Sample:
        return _native.Sample(ctx or self.ctx, ascii_, starts, gene, iso, name=name, lengths=lengths)

    # ---- kernel (3)
    def search(self, samples, pairs=None, evalue: float = 1e-99, top_n: int = 1, keep_all: bool = True):
        """Returns (pairs, tables): tables[k] joins samples pairs[k] = (i, j), i = transcripts1."""
        if pairs is None:
            pairs = list(itertools.combinations(range(len(samples)), 2))
        tables = self._fan_out(len(pairs), lambda ctx, k: _native.search_pair(
            ctx, samples[pairs[k][0]], samples[pairs[k][1]], evalue, top_n, keep_all))
        return pairs, tables

    # ---- kernel (4)
    def phase2(self, n_samples: int, pairs, tables, stride: int, sample_count: Optional[int] = None,
               want_valid: bool = False):
        return _native.phase2_tables(self.ctx, n_samples, stride, tables, [p[0] for p in pairs],
                                     [p[1] for p in pairs], sample_count, want_valid)

    @staticmethod
    def matrix_from_sums(labels: Sequence[str], pairs, nident_sum, denom_sum):
        """Exact rational similarity per pair, d = 1 - sim, float() last; rows/cols sorted by label
        (similarity_computer.py:216-251,285-312).  Raises NoIdealComponentsError on an empty table."""
        present = sorted({labels[i] for p in pairs for i in p})
        pos = {s: k for k, s in enumerate(present)}
        m = np.zeros((len(present), len(present)), dtype=np.float64)
        for (i, j), ni, de in zip(pairs, nident_sum, denom_sum):
            if int(de) == 0:
                raise NoIdealComponentsError()
            d = float(1 - Fraction(int(ni), int(de)))
            m[pos[labels[i]], pos[labels[j]]] = d
            m[pos[labels[j]], pos[labels[i]]] = d
        return present, m

    def all_vs_all(self, inputs, labels, top_genes: Optional[int] = None, evalue: float = 1e-99,
                   top_n: int = 1, keep_all: bool = True, dist=None, keep_tables: bool = False):
        """inputs[k] = dict(ascii=uint8 array/tensor, starts, lengths, coverage, gene, iso) for sample k.
        Returns dict(labels, matrix, sums, pairs, stats[, tables, samples])."""
        world = dist.get_world_size() if dist is not None else 1
        rank = dist.get_rank() if dist is not None else 0
        t_start = time.perf_counter()

        # pinned host buffers are prefetched: every host->device copy is queued on one copy stream right
        # away (PCIe is the bottleneck of this leg: 1.24 GB per step at configs[1]), so all but the first
        # copy hide behind the searches of earlier samples; an upload only waits for its own copy's event
        staged = self._prefetch(inputs, [k for k in range(len(inputs)) if world == 1 or k % world == rank])

        # top-gene selection (host code, 5 ms per 100,000 transcripts) runs ahead of the uploads on its own threads:
        # with the samples shared out over N GPUs a rank builds its few samples back to back at the start of the
        # step, and every millisecond in front of a build delays the round the other ranks wait for
        own = [k for k in range(len(inputs)) if world == 1 or k % world == rank]
        selected = {}
        if top_genes is not None:
            if getattr(self, "_select_pool", None) is None:
                from concurrent.futures import ThreadPoolExecutor
                self._select_pool = ThreadPoolExecutor(4)
            for k in own:
                selected[k] = self._select_pool.submit(select_top_genes, inputs[k]["coverage"], np.asarray(inputs[k]["gene"]), top_genes)

        def up(ctx, k):
            s = inputs[k]
            starts, lengths, gene, iso = (np.asarray(s[c]) for c in ("starts", "lengths", "gene", "iso"))
            if top_genes is not None:
                keep = selected.pop(k).result() if k in selected else select_top_genes(s["coverage"], gene, top_genes)
                starts, lengths, gene, iso = starts[keep], lengths[keep], gene[keep], iso[keep]
            ascii_ = s["ascii"]
            if k in staged:
                ascii_, ev = staged.pop(k)
                ev.synchronize()
            return self.upload(ascii_, starts, lengths, gene, iso, name=labels[k], ctx=ctx)

        # One task list, dependency-driven: [upload 0, upload 1, search (0,1), upload 2, search (0,2), ...].
        # A worker (one per stream) takes the first task whose samples are resident, so host->device copies
        # and index builds of later samples overlap the searches of earlier ones instead of forming a
        # barrier before the first search.
        n = len(inputs)
        # cost estimate for the pair schedule: all bases of a sample (identical on every rank)
        totals = [int(np.sum(np.asarray(inputs[k]["lengths"], dtype=np.int64))) for k in range(n)]
        # (drawing the pairs at run time from a counter in the process group's store - the cost of a pair follows the
        # divergence of its samples, which the length-based assignment cannot know - was measured and lost: 654-711 ms
        # per step against 601 ms on two GPUs at configs[2]; the trips to the store cost more than the imbalance)
        pairs, owner = schedule_pairs(n, totals, world)
        mine = [k for k in range(len(pairs)) if owner[k] == rank]
        samples = [None] * n
        my_tables = [None] * len(mine)
        tasks = []
        S = len(self.ctxs)
        share = None
        if world > 1:
            # N ranks: every rank builds the samples k % N == rank (pack + index are not repeated on every
            # GPU) and broadcasts them as flat device buffers over NCCL, one round of N samples at a time:
            # task ("b", j) builds this rank's sample of round j and issues the round's broadcasts (always on
            # worker 0, so the collectives of all ranks line up), ("u", k) waits for sample k to arrive and
            # unpacks it, and searches start as soon as their two samples are resident - the transfers of
            # later rounds run behind the searches of earlier ones
            import torch
            from .exchange import PipelinedShare
            device = torch.device("cuda", self.device)
            if getattr(self, "_meta_group", None) is None:
                self._meta_group = dist.new_group()      # collective: every rank makes its first N > 1 call here
            share = PipelinedShare(dist, n, device, meta_group=self._meta_group)
            # list order is priority: builds, then arrivals (cheap, and they unlock searches), then searches by
            # the round of their later sample
            tasks = [("b", j) for j in range(share.rounds)]
            tasks.extend(("u", k) for k in range(n) if k % world != rank)
            tasks.extend(("s", slot) for slot in sorted(range(len(mine)), key=lambda x: (max(pairs[mine[x]]) // world, x)))
        else:
            by_last = {}
            for slot, k in enumerate(mine):
                by_last.setdefault(max(pairs[k]), []).append(slot)
            for k in range(n):
                tasks.append(("u", k))
                tasks.extend(("s", slot) for slot in by_last.get(k, []))

        def ready(task, w):
            kind, x = task
            if kind == "b":
                return w == 0
            if kind == "u":
                # uploads are pinned to a context (sample k -> context k % S): every context's memory
                # pool then sees the same allocation sequence each step and stops growing after warm-up
                if share is None:
                    return x % S == w
                # arrivals go to the workers that do not build (worker 0 is busy with its "b" tasks first)
                # (taken only once the data is there: a worker never sits in a wait while searches are ready)
                return share.arrived(x) and (S == 1 or 1 + x % (S - 1) == w)
            i, j = pairs[mine[x]]
            return samples[i] is not None and samples[j] is not None

        def run(ctx, task):
            kind, x = task
            if kind == "b":
                torch.cuda.set_device(self.device)
                k = share.own_item(x)
                obj = up(ctx, k) if k is not None else None
                share.issue_round(x, obj, lambda s: s.meta(), lambda s, buf: s.export_flat(buf.data_ptr()))
                if k is not None:
                    samples[k] = obj
            elif kind == "u" and share is not None:
                torch.cuda.set_device(self.device)
                samples[x] = share.fetch(x, lambda k, meta, buf: _native.Sample.import_flat(ctx, meta, buf.data_ptr(), labels[k]))
            elif kind == "u":
                samples[x] = up(ctx, x)
            else:
                i, j = pairs[mine[x]]
                my_tables[x] = _native.search_pair(ctx, samples[i], samples[j], evalue, top_n, keep_all)

        self._run_tasks(tasks, ready, run, poll=0.0003 if share is not None else None)
        if share is not None:
            share.finish()
        t_upload = time.perf_counter()
        stats = {}
        for t in my_tables:
            for key, v in t.stats.items():
                stats[key] = stats.get(key, 0) + v
        t_search = time.perf_counter()
        stride = 1 + max([0] + [int(np.max(s["gene"], initial=0)) for s in inputs])
        if world > 1 and not keep_tables:
            # the one exchange of results: the five columns phase 2 reads (20 bytes per row), all-gathered as one
            # block per rank and reduced in place on every rank (exchange.phase2_exchanged)
            from .exchange import phase2_exchanged
            tables = my_tables
            t_exchange = time.perf_counter()
            res = phase2_exchanged(self.ctx, dist, n, stride, pairs, owner, mine, my_tables)
            res.pop("_keep", None)
        else:
            if world > 1:       # whole tables on every rank (subset re-runs keep them)
                from .exchange import allgather_tables
                tables = allgather_tables(self.ctx, dist, pairs, owner, mine, my_tables)
            else:
                tables = my_tables
            t_exchange = time.perf_counter()
            res = self.phase2(n, pairs, tables, stride)
        t_phase2 = time.perf_counter()
        out = dict(host_ms=dict(upload=(t_upload - t_start) * 1e3, search=(t_search - t_upload) * 1e3,
                                exchange=(t_exchange - t_search) * 1e3, phase2=(t_phase2 - t_exchange) * 1e3),
                   pairs=pairs, nident_sum=res["nident_sum"], denom_sum=res["denom_sum"], counts=res["counts"],
                   stats=stats, no_ideal=res["no_ideal"], n_local_pairs=len(mine),
                   lengths=[s.total for s in samples], n_index=[s.n_index for s in samples], stride=stride)
        if not res["no_ideal"]:
            out["labels"], out["matrix"] = self.matrix_from_sums(labels, pairs, res["nident_sum"], res["denom_sum"])
        if keep_tables:
            out["tables"], out["samples"] = tables, samples
        else:
            for t in tables:
                t.close()
            for s in samples:
                s.close()
        return out

    def subset_distances(self, result: dict, labels: Sequence[str], keep: Sequence[int], stride: Optional[int] = None):
        """Phase 2 again on the device-resident tables of `result` (all_vs_all(..., keep_tables=True)),
        restricted to the samples `keep`: the reference's subset analysis (make_subset.py:166-201 links the
        tables of the kept samples and rebuilds the graph; subset_comparisons.py:45-73) without touching
        the searches.  Returns dict(labels, matrix, pairs, nident_sum, denom_sum, counts, no_ideal)."""
        keep_set = set(int(k) for k in keep)
        sel = [k for k, (i, j) in enumerate(result["pairs"]) if i in keep_set and j in keep_set]
        pairs = [result["pairs"][k] for k in sel]
        tables = [result["tables"][k] for k in sel]
        if stride is None:
            stride = result["stride"]
        res = self.phase2(len(labels), pairs, tables, stride)
        out = dict(pairs=pairs, nident_sum=res["nident_sum"], denom_sum=res["denom_sum"], counts=res["counts"],
                   no_ideal=res["no_ideal"])
        if not res["no_ideal"]:
            out["labels"], out["matrix"] = self.matrix_from_sums(labels, pairs, res["nident_sum"], res["denom_sum"])
        return out

    def close(self):
        if self._pool is not None:
            self._pool.shutdown()
            self._pool = None
        if getattr(self, "_select_pool", None) is not None:
            self._select_pool.shutdown()
            self._select_pool = None
        for c in self.ctxs:
            c.close()
