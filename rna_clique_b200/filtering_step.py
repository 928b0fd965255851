"""Phase 1 driver.  Mirrors the reference's filtering_step.py:61-164: top genes per sample ->
all-pairs tables -> gene-matches graph -> graph pickle; returns lazy table readers.

The reference runs three joblib stages one after the other (select_top_and_save per sample :128-143,
makeblastdb per sample and blastn per pair inside find_all_pairs :144-156, build_graph :157).  Here they are
one dependency-driven schedule on the multi-stream engine (engine.Engine._run_tasks):

  host pool (up to `jobs` threads)    map the file, decode ids, select, extract the selected records into page-locked
                                       memory; write od1/<sample>_top.fasta  (native, GIL released; parses are served first)
  engine workers (one per CUDA stream) upload sample k as soon as its records exist (pack, DUST, seed index),
                                       search pair (i, j) as soon as both samples are resident, download the table
  helper processes (_od2_worker.py)    table -> DataFrame -> od2/<S1>--<S2>.<ext>, while later pairs search (pandas and
                                       pickle hold the interpreter lock: processes, not threads)

and the device tables stay resident: graph.pkl is prepared on the device (node numbering, distinct edges, adjacency
lists) and only serialised on the host, and SampleSimilarity reduces the device tables directly (rcq_phase2_tables)
instead of reading od2/ back."""
from __future__ import annotations

import itertools
import os
import threading
from pathlib import Path
from typing import Callable, Iterable, Optional

import numpy as np

from . import _native
from .build_graph import EdgeArrays, GeneMatchGraph, dump_graph
from .filtered_distance import SampleSimilarity
from .find_all_pairs import _label_and_save, make_output_path
from .find_homologs import table_to_frame
from .gene_matches_tables import default_extension
from .select_top_genes import _parse_and_select_lean
from .transcripts import default_parser

_engine = None
_engine_lock = threading.Lock()


def shared_engine(jobs: Optional[int] = None, device: int = 0):
    """The process-wide engine of the drop-in API: 2-6 CUDA streams (contexts) on one GPU, kept between calls so
    that scratch slabs and memory pools are warm."""
    global _engine
    from .engine import Engine
    with _engine_lock:
        if _engine is None:
            want = os.environ.get("RCQ_API_STREAMS")
            _engine = Engine(device, streams=int(want) if want else max(2, min(6, int(jobs or 4))))
        return _engine


def close_shared_engine() -> None:
    global _engine
    with _engine_lock:
        if _engine is not None:
            _engine.close()
            _engine = None
    _pinned.clear()
    _close_table_writers()


class _PinnedPool:
    """Page-locked host buffers for the letters on their way to the device, kept between samples and between calls
    (cudaHostAlloc of 100 MB costs tens of milliseconds, the copy from it 4 ms; from pageable memory it costs 40).
    Beyond `limit` bytes in use the letters go through ordinary numpy memory."""

    def __init__(self, limit: int = 6 << 30):
        self.lock = threading.Lock()
        self.free = []            # torch uint8 tensors, pinned
        self.out = {}             # id(numpy base) -> tensor
        self.held = 0
        self.limit = limit

    def get(self, nbytes: int) -> np.ndarray:
        import torch
        with self.lock:
            best = None
            for k, t in enumerate(self.free):
                if t.numel() >= nbytes and (best is None or t.numel() < self.free[best].numel()):
                    best = k
            t = self.free.pop(best) if best is not None else None
            if t is None and self.held + nbytes > self.limit:
                return np.empty(nbytes, dtype=np.uint8)
            if t is None:
                cap = (nbytes + (16 << 20) - 1) >> 24 << 24
                self.held += cap
        if t is None:
            try:
                t = torch.empty(cap, dtype=torch.uint8, pin_memory=True)
            except RuntimeError:
                with self.lock:
                    self.held -= cap
                return np.empty(nbytes, dtype=np.uint8)
        a = t.numpy()
        with self.lock:
            self.out[a.ctypes.data] = t
        return a

    def put(self, a: np.ndarray) -> None:
        """a: the array get() returned, or a view that starts at its first byte."""
        with self.lock:
            t = self.out.pop(a.ctypes.data, None)
            if t is not None:
                self.free.append(t)

    def clear(self) -> None:
        with self.lock:
            self.free.clear()
            self.out.clear()
            self.held = 0


_pinned = _PinnedPool()


class _TableWriters:
    """Helper PROCESSES for the od2/ files (_od2_worker.py): frame building and pickling hold the interpreter lock, so
    threads do not scale (seven threads write 1.5 times the tables of one) and at BASELINE configs[2] the 435 tables
    were what the call waited for.  Each process is fed by one thread of this process: the columns go through a file
    under /dev/shm (or the temp directory), the request through the process's stdin.  Whatever goes wrong with a
    process, its task is finished by the feeding thread itself, so the files are always written."""

    def __init__(self, n_procs: int):
        import queue
        import subprocess
        import sys
        import tempfile
        from concurrent.futures import Future
        self._Future = Future
        self.q = queue.Queue()
        self.seq = itertools.count()
        base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
        self.dir = tempfile.mkdtemp(prefix="rcq_od2_", dir=base)
        self.procs, self.threads = [], []
        pkg_parent = str(Path(__file__).resolve().parent.parent)
        for _ in range(n_procs):
            try:
                p = subprocess.Popen([sys.executable, "-m", "rna_clique_b200._od2_worker"], stdin=subprocess.PIPE,
                                     stdout=subprocess.PIPE, cwd=pkg_parent, text=True, bufsize=1)
            except OSError:
                break
            self.procs.append(p)
            t = threading.Thread(target=self._feed, args=(p,), daemon=True)
            t.start()
            self.threads.append(t)

    def __bool__(self):
        return bool(self.procs)

    def _feed(self, proc):
        import json
        alive = True
        try:
            alive = proc.stdout.readline().strip() == "ready"       # the process has imported pandas
        except Exception:  # noqa: BLE001
            alive = False
        while True:
            item = self.q.get()
            if item is None:
                break
            fut, cols, db, t1, t2, out = item
            if not fut.set_running_or_notify_cancel():
                continue
            try:
                done = False
                if alive:
                    try:
                        n = len(cols["length"])
                        block = os.path.join(self.dir, f"t{next(self.seq)}.bin")
                        with open(block, "wb") as f:
                            for c in _native.TABLE_COLS:
                                f.write(np.ascontiguousarray(cols[c], dtype=np.int32).data)
                            f.write(np.ascontiguousarray(cols["minus"], dtype=np.uint8).data)
                            f.write(np.ascontiguousarray(cols["reverse"], dtype=np.uint8).data)
                        proc.stdin.write(json.dumps(dict(block=block, n=n, db=[int(x) for x in db], t1=str(t1), t2=str(t2),
                                                         out=os.path.abspath(str(out)))) + "\n")
                        proc.stdin.flush()
                        reply = proc.stdout.readline().strip()
                        done = reply == "ok"
                        if not done and not reply:
                            alive = False                          # the process is gone
                    except Exception:  # noqa: BLE001 - broken pipe, full disk, ...: this thread does the task itself
                        alive = False
                if not done:
                    _frame_and_save(cols, db, t1, t2, out)
                fut.set_result(None)
            except BaseException as e:  # noqa: BLE001 - handed to whoever asks for the result
                fut.set_exception(e)

    def submit(self, cols, db, t1, t2, out):
        fut = self._Future()
        self.q.put((fut, cols, db, t1, t2, out))
        return fut

    def close(self):
        import shutil
        for _ in self.threads:
            self.q.put(None)
        for t in self.threads:
            t.join(timeout=30)
        for p in self.procs:
            try:
                p.stdin.close()
                p.wait(timeout=5)
            except Exception:  # noqa: BLE001
                p.kill()
        self.procs, self.threads = [], []
        shutil.rmtree(self.dir, ignore_errors=True)


_writers = None


def _table_writers(workers: int):
    """The process-wide helper processes, started at the first call that has enough host threads to feed them
    (RCQ_TABLE_PROCS overrides the count; 0 keeps the work on threads)."""
    global _writers
    with _engine_lock:
        if _writers is None:
            want = os.environ.get("RCQ_TABLE_PROCS")
            n = int(want) if want is not None else (min(8, workers // 2) if workers >= 4 else 0)
            if n <= 0:
                return None                   # this call keeps the work on its threads; a later call may start them
            w = _TableWriters(n)
            if not w:
                w.close()
                return None
            _writers = w
            import atexit
            atexit.register(_close_table_writers)
        return _writers


def _close_table_writers() -> None:
    global _writers
    with _engine_lock:
        w, _writers = _writers, None
    if w:
        w.close()


class _PriorityPool:
    """Host threads that always take the most urgent task first: parsing a transcriptome (a search is waiting for
    it) before writing od1/ files before turning finished tables into od2/ files.  With one FIFO queue a parse
    submitted late waited behind dozens of table tasks while the GPU ran out of resident samples."""

    def __init__(self, workers: int):
        import queue
        from concurrent.futures import Future
        self._Future = Future
        self.q = queue.PriorityQueue()
        self.seq = itertools.count()
        self.threads = [threading.Thread(target=self._run, daemon=True) for _ in range(workers)]
        for t in self.threads:
            t.start()

    def _run(self):
        while True:
            _, _, item = self.q.get()
            if item is None:
                return
            fut, fn, args = item
            if not fut.set_running_or_notify_cancel():
                continue
            try:
                fut.set_result(fn(*args))
            except BaseException as e:      # noqa: BLE001 - handed to whoever asks for the result
                fut.set_exception(e)

    def submit(self, priority: int, fn, *args):
        fut = self._Future()
        self.q.put((priority, next(self.seq), (fut, fn, args)))
        return fut

    def shutdown(self, wait: bool = True):
        for _ in self.threads:
            self.q.put((1 << 30, next(self.seq), None))
        if wait:
            for t in self.threads:
                t.join()


def _frame_and_save(cols, db, t1, t2, out_path):
    _label_and_save(table_to_frame(cols, *db), t1, t2, out_path)


def filtering_step(dirs: Iterable[Path], out_dir_1: Path, out_dir_2: Path, cache_dir: Path, output_graph: Path,
                   top_genes: int, transcripts: str = "transcripts.fasta", top_matches: int = 1,
                   id_parser: Callable = default_parser, evalue: float = 1e-99, keep_all: bool = True,
                   jobs: int = max((os.cpu_count() or 2) - 1, 1)):
    """Returns (tables, table_paths, graph, num_tables, path_to_sample) like the reference; `tables` reads the
    od2/ files lazily, `graph` is a GeneMatchGraph (an nx.Graph filled on first use) that also carries the
    device-resident tables for SampleSimilarity.  cache_dir is accepted and unused: the databases live in HBM."""
    import time
    trace = os.environ.get("RCQ_TRACE_API")
    t_0 = time.perf_counter()
    dirs = [Path(d) for d in dirs]
    n = len(dirs)
    eng = shared_engine(jobs)
    S = len(eng.ctxs)
    workers = max(1, min(int(jobs or 1), 16))
    ext = default_extension()
    pairs = list(itertools.combinations(range(n), 2))
    out_paths = [Path(out_dir_1) / (d.stem + "_top.fasta") for d in dirs]
    names = [d.stem for d in dirs]
    table_paths = [make_output_path(out_dir_2, out_paths[i], out_paths[j], dict(zip(out_paths, names)).__getitem__, ext)
                   for i, j in pairs]
    samples = [None] * n
    tables = [None] * len(pairs)
    edges = [None] * len(pairs)
    writes = [None] * len(pairs)
    host = _PriorityPool(workers)
    writers = _table_writers(workers)        # helper processes for the od2/ files (None: the host pool does them)
    # the engine workers need the interpreter lock only for microseconds between native calls, the host pool holds it
    # for whole pandas operations: a short switch interval keeps the GPU from waiting for the lock
    import sys
    switch = sys.getswitchinterval()
    sys.setswitchinterval(min(switch, 0.0002))
    try:
        # at most `window` parsed files are alive (a parsed 100,000-transcript file is ~0.4 GB of host memory)
        window = workers + 2
        lock = threading.Lock()
        parsed = {}
        state = {"next": 0}

        od1_writes = [None] * n

        split = []                # RCQ_TRACE_API: (search, download) seconds of every search task
        uses = {}                 # sample -> users of its letter buffer still running (upload, od1 write)

        def release(k, letters):
            with lock:
                uses[k] -= 1
                last = uses[k] == 0
            if last:
                _pinned.put(letters)

        def parse_task(k):
            # the device needs the selected records only, in page-locked memory: od1/<sample>_top.fasta is formatted
            # and written from the same buffer as a task of its own, behind the parses already queued
            out, stem, rec, write = _parse_and_select_lean(out_dir_1, transcripts, dirs[k], top_genes, id_parser, _pinned.get)
            uses[k] = 2

            def write_and_release():
                try:
                    write()
                finally:
                    release(k, rec[1])

            od1_writes[k] = host.submit(1, write_and_release)
            return out, stem, rec

        def submit_parse():
            with lock:
                k = state["next"]
                if k >= n:
                    return
                state["next"] = k + 1
                parsed[k] = host.submit(0, parse_task, k)

        for _ in range(min(window, n)):
            submit_parse()
        by_last = {}
        for slot, (i, j) in enumerate(pairs):
            by_last.setdefault(j, []).append(slot)
        tasks = []
        for k in range(n):
            tasks.append(("u", k))
            tasks.extend(("s", slot) for slot in by_last.get(k, []))

        def ready(task, w):
            kind, x = task
            if kind == "u":
                f = parsed.get(x)
                return x % S == w and f is not None and f.done()
            i, j = pairs[x]
            return samples[i] is not None and samples[j] is not None

        def run(ctx, task):
            kind, x = task
            if kind == "u":
                out, _, rec = parsed.pop(x).result()
                try:
                    samples[x] = _native.Sample(ctx, rec[1], rec[2], rec[3], rec[4], name=str(out), lengths=rec[5])
                finally:
                    release(x, rec[1])
                del rec
                submit_parse()
                return
            i, j = pairs[x]
            s1, s2 = samples[i], samples[j]
            t_a = time.perf_counter()
            t = _native.search_pair(ctx, s1, s2, evalue, top_matches, keep_all)
            t_b = time.perf_counter()
            cols = t.download()
            if trace:
                split.append((t_b - t_a, time.perf_counter() - t_b))
            tables[x] = t
            edges[x] = (cols["sgene"], cols["qgene"])
            db = (s1.total, s1.n_seq, s2.total, s2.n_seq)
            writes[x] = (writers.submit(cols, db, out_paths[i], out_paths[j], table_paths[x]) if writers else
                         host.submit(2, _frame_and_save, cols, db, out_paths[i], out_paths[j], table_paths[x]))

        eng._run_tasks(tasks, ready, run, poll=0.0005)
        t_1 = time.perf_counter()
        e = EdgeArrays()
        for slot, (i, j) in enumerate(pairs):
            if len(edges[slot][0]):
                e.add(str(out_paths[i]), str(out_paths[j]), *edges[slot])
        graph = GeneMatchGraph(e)
        # the device tables travel with the graph: SampleSimilarity reduces them in place (rna_clique.py hands them on)
        graph.__dict__["_rcq_device"] = dict(engine=eng, tables=tables, pairs=pairs, labels=[str(p) for p in out_paths],
                                             empty=[len(t) == 0 for t in tables])
        if output_graph is not None:
            # graph.pkl: prepared on the device from the resident tables (csrc/graph_device.cu), by the host writer
            # when the tables do not qualify (sparse custom gene ids)
            try:
                _native.graph_pickle_tables(eng.ctxs[0], output_graph, [str(p) for p in out_paths], tables,
                                            [i for i, _ in pairs], [j for _, j in pairs])
            except _native.RcqError as err:
                if err.code not in (-2, -3):
                    raise
                dump_graph(graph, output_graph)
        t_2 = time.perf_counter()
        for f in writes + od1_writes:
            if f is not None:
                f.result()
        if trace and split:
            print(f"[filtering_step] search tasks: search_pair {1e3 * sum(a for a, _ in split) / len(split):.1f} ms, "
                  f"download {1e3 * sum(b for _, b in split) / len(split):.1f} ms on average", file=__import__("sys").stderr)
        if trace:
            print(f"[filtering_step] parse+upload+search {t_1 - t_0:.3f} s, graph {t_2 - t_1:.3f} s, "
                  f"waiting for table files {time.perf_counter() - t_2:.3f} s", file=__import__("sys").stderr)
    finally:
        sys.setswitchinterval(switch)
        host.shutdown(wait=True)
        for s in samples:
            if s is not None:
                s.close()
    path_to_sample = dict(zip(out_paths, names))
    p1, p2 = iter(table_paths), iter(table_paths)
    return map(SampleSimilarity._read_table, p1), p2, graph, len(pairs), path_to_sample
