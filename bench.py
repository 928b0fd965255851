#!/usr/bin/env python
"""bench.py — RNA-clique hot path on B200: all-vs-all top-gene matching -> gene-match graph -> ideal components
-> distance matrix, on transcriptomes synthesised like the reference's distance_sequence_simulator.
One JSON line on stdout (rank 0).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config 0|1|2]

Workload (default): BASELINE.json configs[2], the configuration the metric is quoted on — 30 simulated samples x
100,000 transcripts, top_genes = 50,000, 435 sample pairs.  --config 1 / 0 select the smaller BASELINE configs.

A step = one full pass: pack + DUST + seed index of every sample, every unordered pair searched (both blastn
directions), phase 2, distance matrix.
  value  device-timed, ASCII transcripts already resident in HBM (Engine.all_vs_all).
  e2e    N = 1: the reference-facing call itself, rna_clique() from FASTA directories on disk to od1/, od2/,
         graph.pkl and distance_matrix.h5 (wall clock of >= 3 calls after one warm-up call; host->device copies of
         the letters and device->host copies of every table inside).  N > 1: Engine.all_vs_all from pinned host
         buffers with the pairs sharded over the ranks (the drop-in call drives one GPU).
Metric: GCUPS = base comparisons actually evaluated by the gapped extension ("extension steps", counted alike by
the kernels and by the CPU oracle: tests/test_search_gpu.py) per second.  The Smith-Waterman-equivalent figure
(sum over ordered searches of L_q x L_d) is reported beside it, labelled.
--impl reference: BLAST+ does not exist on this box, so the reference arm is the C oracle (oracle/, "port") on all
host cores, every step a bounded sample of the same workload (as many sample pairs as cores, --cpu-genes top genes
per sample); it imports nothing of the product library.
"""
import argparse
import itertools
import json
import os
import shutil
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {0: (3, 5000, 2000), 1: (10, 60000, 50000), 2: (30, 100000, 50000)}
OPS_PER_STEP = 12      # integer lane-operations one extension step needs at the least


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=[0, 1, 2], help="BASELINE.json configs[k]")
    ap.add_argument("--taxa", type=int, default=0)
    ap.add_argument("--count", type=int, default=0)
    ap.add_argument("--top", type=int, default=0)
    ap.add_argument("--seed", type=int, default=487)
    ap.add_argument("--cpu-genes", type=int, default=1200, help="top genes per sample in one CPU-oracle step")
    ap.add_argument("--cpu-pairs", type=int, default=0, help="sample pairs per CPU-oracle step (0 = one per core)")
    ap.add_argument("--e2e-calls", type=int, default=3, help="timed rna_clique() calls (N = 1)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-api-leg", action="store_true", help="e2e from pinned host buffers instead of rna_clique() on FASTA directories")
    ap.add_argument("--streams", type=int, default=0, help="CUDA streams (host threads) per GPU; 0 = 6 up to four GPUs, 3 beyond (measured)")
    ap.add_argument("--clock-interval", type=float, default=0.2, help="seconds between NVML clock samples (0 = off)")
    a = ap.parse_args()
    t, c, n = CONFIGS[a.config]
    a.taxa, a.count, a.top = a.taxa or t, a.count or c, a.top or n
    return a


def workload_name(a):
    cfg = {v: f"BASELINE.json configs[{k}]" for k, v in CONFIGS.items()}.get((a.taxa, a.count, a.top), "custom size")
    return (f"{a.taxa} simulated samples x {a.count} transcripts, top_genes={a.top} "
            f"({cfg}; birth-death tree, JC/HKY85 mu=0.01, seed {a.seed})")


def equiv_cells(lengths):
    """Smith-Waterman-equivalent cells of the all-vs-all: sum over ORDERED searches of L_q * L_d."""
    tot = float(sum(lengths))
    return tot * tot - float(sum(float(x) * x for x in lengths))


class ClockSampler(threading.Thread):
    def __init__(self, device, interval=0.05):
        super().__init__(daemon=True)
        self.interval = interval
        self.device, self.samples, self.reasons, self.stop_flag, self.max_mhz = device, [], set(), False, None

    def run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.device)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            names = {pynvml.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                     pynvml.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                     pynvml.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                     pynvml.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown"}
            while not self.stop_flag:
                self.samples.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, n in names.items():
                    if r & bit:
                        self.reasons.add(n)
                time.sleep(self.interval)
        except Exception as e:  # noqa: BLE001
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def make_inputs(a, device):
    from rna_clique_b200.simulate import simulate
    tree, ts = simulate(a.taxa, a.count, seed=a.seed, device=device)
    inputs, labels = [], []
    for t in ts:
        offs = t.offsets.numpy()
        inputs.append(dict(ascii=t.ascii, starts=offs[:-1].copy(), lengths=np.diff(offs).astype(np.int32),
                           coverage=t.coverage, gene=t.gene, iso=t.iso))
        labels.append(f"od1/{t.name}_top.fasta")
    return tree, inputs, labels, ts


# ------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port on the host cores.  Imports oracle/ and the (pure torch / numpy) simulator only.
def cpu_pair_job(args):
    """One unordered pair through the CPU oracle (runs in a worker process): samples built (codes, reverse
    complement, DUST) and searched.  Returns (seconds, table rows, extension steps)."""
    sys.path.insert(0, ROOT)
    from oracle import oracle as O
    (a1, o1, g1, i1), (a2, o2, g2, i2) = args
    t0 = time.perf_counter()
    s1, s2 = O.Sample(a1, o1, g1, i1), O.Sample(a2, o2, g2, i2)
    r = O.search_pair(s1, s2)
    return time.perf_counter() - t0, r.counts["rows"], r.counts["dp_cells"]


def cpu_sample(a, cores):
    """A bounded sample of the workload for the CPU oracle: the same simulator, tree and seed with --cpu-genes
    genes per sample (coverage is independent of the sequences, so any G of the top genes are distributed like G
    simulated genes), and as many of the sample pairs, in the order the workload lists them, as there are cores."""
    from rna_clique_b200.simulate import simulate
    n_genes = min(a.cpu_genes, a.top)
    _, ts = simulate(a.taxa, n_genes, seed=a.seed, device="cpu")
    subs = [(t.ascii.numpy(), t.offsets.numpy(), t.gene.astype(np.int32), t.iso.astype(np.int32)) for t in ts]
    pairs = list(itertools.combinations(range(a.taxa), 2))
    n_pairs = min(a.cpu_pairs or cores, len(pairs))
    jobs = [(subs[i], subs[j]) for i, j in pairs[:n_pairs]]
    desc = (f"CPU oracle (C restatement of the search, -O2; BLAST+ is unavailable on this box), {n_pairs} of the "
            f"{len(pairs)} sample pairs with {n_genes} of the {a.top} top genes per sample, one pair per core; "
            f"rate = extension steps of the sample / its wall time")
    return jobs, desc


def run_cpu_steps(jobs, cores, steps, warmup):
    import multiprocessing as mp
    times, cells = [], 0
    with mp.get_context("fork").Pool(cores) as pool:
        for s in range(warmup + steps):
            t0 = time.perf_counter()
            out = pool.map(cpu_pair_job, jobs, chunksize=1)
            if s >= warmup:
                times.append(time.perf_counter() - t0)
                cells = sum(o[2] for o in out)
    return times, cells


def host_cores():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def reference_arm(a):
    cores = host_cores()
    jobs, desc = cpu_sample(a, cores)
    times, cells = run_cpu_steps(jobs, cores, a.steps, a.warmup)
    t = float(np.mean(times))
    gcups = cells / t / 1e9
    print(json.dumps({
        "impl": "reference", "metric": "all_vs_all_alignment_gcups", "value": gcups, "unit": "GCUPS",
        "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": t * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": workload_name(a)},
        "cpu_baseline": {"value": gcups, "unit": "GCUPS", "cores": cores, "kind": "port", "sample": desc,
                         "extension_steps_per_step": cells},
        "e2e": {"value": gcups, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "BLAST+ (the reference's native component) is unavailable on this box; the CPU arm is the C oracle, "
                "a rate measured on the bounded sample (no extrapolated times)",
    }))


# ------------------------------------------------------------------------------------------------------------
def solo_kernel_times(eng, resident, labels, pairs):
    """The gapped stage (k_gapped_reg passes + k_gapped) and k_seed_probe timed alone (one stream, nothing else on the GPU) on several sample pairs:
    per-launch CUDA-event times of the stage timers and the extension steps / index sizes of the same launches."""
    from rna_clique_b200 import _native
    ctx = eng.ctxs[0]
    need = sorted({k for p in pairs for k in p})
    ss = {k: eng.upload(resident[k]["ascii"], resident[k]["starts"], resident[k]["lengths"], resident[k]["gene"],
                        resident[k]["iso"], name=labels[k], ctx=ctx) for k in need}
    _native.search_pair(ctx, ss[pairs[0][0]], ss[pairs[0][1]]).close()       # warm-up
    ctx.stage_times(reset=True)
    cells, verified, n_index = 0, 0, 0
    for i, j in pairs:
        t = _native.search_pair(ctx, ss[i], ss[j])
        cells += t.stats["dp_cells"]
        verified += t.stats["n_verified"]
        n_index += ss[i].n_index + ss[j].n_index
        t.close()
    st = ctx.stage_times(reset=True)
    for x in ss.values():
        x.close()
    n = len(pairs)
    return {"gapped_ms": st["gapped_dp"]["ms"] / max(st["gapped_dp"]["calls"], 1), "cells_per_launch": cells / n,
            "seed_probe_ms": st["seed_probe"]["ms"] / max(st["seed_probe"]["calls"], 1),
            "probe_bytes_per_launch": 16.0 * n_index / n + 4.0 * ((1 << 22) + 1) + 8.0 * verified / n, "launches": n}


def scratch_dir(need_bytes):
    for base in ("/dev/shm", tempfile.gettempdir()):
        try:
            if shutil.disk_usage(base).free > need_bytes:
                return tempfile.mkdtemp(prefix="rcq_bench_", dir=base)
        except OSError:
            pass
    return tempfile.mkdtemp(prefix="rcq_bench_")


def run_api_leg(transcriptomes, a, engine_matrix, calls, selected_letters):
    """The reference-facing call itself: rna_clique() from transcriptome directories (FASTA on disk) to od1/, od2/,
    graph.pkl and distance_matrix.h5.  One warm-up call, then `calls` timed calls (output directories emptied
    between calls, untimed).  Returns dict(seconds=[...], same=matrix equals the engine's, h2d, d2h)."""
    from pathlib import Path
    from rna_clique_b200 import h5fixed
    from rna_clique_b200.filtering_step import close_shared_engine
    from rna_clique_b200.rna_clique import rna_clique
    from rna_clique_b200.simulate import write_dataset
    letters = sum(int(t.ascii.numel()) for t in transcriptomes)
    tmp = Path(scratch_dir(4 * letters))
    try:
        dirs = write_dataset(tmp / "in", transcriptomes)
        secs, same, d2h = [], True, 0
        for rep in range(1 + calls):
            for d in ("od1", "od2"):
                shutil.rmtree(tmp / d, ignore_errors=True)
                (tmp / d).mkdir()
            t0 = time.perf_counter()
            sim, _ = rna_clique(dirs, tmp / "od1", tmp / "od2", tmp / "db", tmp / "graph.pkl", tmp / "distance_matrix.h5", a.top,
                                jobs=max(host_cores() - 1, 1))
            dt = time.perf_counter() - t0
            if rep > 0:
                secs.append(dt)
            m = h5fixed.read_frame(tmp / "distance_matrix.h5", "matrix")
            same = same and bool(np.array_equal(m.to_numpy(), engine_matrix))
            rows = sum(len(t) for t in sim._device["tables"])
            # tables (17 columns), the sums, and the graph arrays of graph.pkl (two adjacency entries per distinct edge =
            # one per row when every reciprocal hit has its forward and its backward row, 8 bytes each; 16 bytes per node)
            d2h = 68 * rows + 16 * len(sim._device["tables"]) + 8 * rows + 16 * a.top * len(transcriptomes)
            del sim
        close_shared_engine()
        h2d = selected_letters + 28 * a.top * len(transcriptomes)      # the letters of the selected records + their tables
        return dict(seconds=secs, same=same, h2d=int(h2d), d2h=int(d2h))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def main():
    a = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if a.impl == "reference":
        if rank == 0:
            reference_arm(a)
        return

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's version banner and any other NCCL logging go to stderr
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ.pop("NCCL_DEBUG")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import torch.distributed as dist_mod
        opts = None
        if os.environ.get("RCQ_NCCL_PRIORITY", "1") != "0":
            try:
                opts = dist_mod.ProcessGroupNCCL.Options(is_high_priority_stream=True)
            except Exception:  # noqa: BLE001
                opts = None
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local), pg_options=opts)
        dist = dist_mod
    from rna_clique_b200.engine import Engine, select_top_genes
    if a.streams <= 0:
        # six streams per GPU keep one GPU busy; with the pairs shared out over eight GPUs a rank has 54 searches and
        # four sample builds that every other rank waits for: three streams leave the builds more of the GPU
        # (configs[2], 8 GPUs: 156 ms per step with three streams, 182 ms with six; 4 GPUs: 284 / 280; 2 GPUs: 527 / 515)
        a.streams = 6 if world <= 4 else 3
    eng = Engine(local, streams=a.streams)
    stream = torch.cuda.ExternalStream(eng.ctx.stream, device=torch.device("cuda", local))

    _, inputs_dev, labels, transcriptomes = make_inputs(a, f"cuda:{local}")
    torch.cuda.synchronize()
    # HBM-resident variant: top-gene selection already applied (starts/lengths of the kept transcripts)
    resident = []
    for s in inputs_dev:
        keep = select_top_genes(s["coverage"], s["gene"], a.top)
        resident.append(dict(ascii=s["ascii"], starts=s["starts"][keep], lengths=s["lengths"][keep],
                             coverage=s["coverage"][keep], gene=s["gene"][keep], iso=s["iso"][keep]))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    step_walls = []

    def timed(fn, steps, warmup):
        out = None
        for _ in range(warmup):
            out = fn()
        barrier()
        l0 = eng.launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        walls = []
        for _ in range(steps):
            t0 = time.perf_counter()
            out = fn()
            walls.append((time.perf_counter() - t0) * 1e3)
        e1.record(stream)
        step_walls.append(walls)
        barrier()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device=f"cuda:{local}", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms / steps, out, (eng.launches - l0) / steps

    sampler = ClockSampler(local, a.clock_interval or 0.05)
    if a.clock_interval > 0:
        sampler.start()
    ms_dev, res, launches = timed(lambda: eng.all_vs_all(resident, labels, None, dist=dist), a.steps, a.warmup)
    engine_host = None
    if world > 1 or a.no_api_leg:
        # engine from pinned host buffers: top-gene selection, H2D of every letter, D2H of the sums inside
        host = [dict(s, ascii=s["ascii"].cpu().pin_memory()) for s in inputs_dev]
        h2d = int(sum(s["ascii"].numel() for s in host) + sum(16 * len(s["starts"]) for s in resident))
        e_steps = max(3, min(a.steps, 5))
        ms_host, res2, _ = timed(lambda: eng.all_vs_all(host, labels, a.top, dist=dist), e_steps, min(a.warmup, 3))
        assert np.array_equal(res["matrix"], res2["matrix"])
        engine_host = dict(ms=ms_host, steps=e_steps, h2d=h2d, d2h=int(16 * len(res["pairs"]) + 64))
        del host
    sampler.stop_flag = True
    if a.clock_interval > 0:
        sampler.join(timeout=2)

    lengths = res["lengths"]
    stats = res["stats"]
    agg = {k: float(v) for k, v in stats.items()}
    if dist is not None:
        keys = sorted(agg)
        t = torch.tensor([agg[k] for k in keys], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(t)
        agg = dict(zip(keys, t.tolist()))
    cells = agg.get("dp_cells", 0.0)          # extension steps of one step, all ranks
    out = None
    if rank == 0:
        clocks = sampler.summary()
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:  # noqa: BLE001
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        n_pairs = len(res["pairs"])
        solo_pairs = [res["pairs"][k] for k in np.linspace(0, n_pairs - 1, num=min(6, n_pairs)).astype(int)]
        solo = solo_kernel_times(eng, resident, labels, solo_pairs)
        int_peak = eng.ctx.int32_issue_peak(5)            # lane-operations per second, measured now
        gapped_ops = solo["cells_per_launch"] * OPS_PER_STEP / max(solo["gapped_ms"] * 1e-3, 1e-9)
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        except Exception:  # noqa: BLE001
            traffic = {}
        out = {
            "metric": "all_vs_all_alignment_gcups", "value": cells / (ms_dev * 1e-3) / 1e9, "unit": "GCUPS",
            "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_dev,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": workload_name(a), "pairs": n_pairs,
                       "searches": f"{n_pairs} symmetric searches, one per unordered pair; both blastn direction frames are "
                                   f"derived from each (the reference runs {2 * n_pairs} blastn processes)",
                       "nt_per_sample": int(np.mean(lengths)),
                       "cells": "GCUPS counts base comparisons actually evaluated by the gapped extension; "
                                "gcups_equiv_ordered_pairs is the Smith-Waterman-equivalent area sum(L_q x L_d) over ordered pairs",
                       "l2": f"inputs larger than L2 ({sum(int(s['ascii'].numel()) for s in inputs_dev) / 1e9:.1f} GB of ASCII per step)",
                       "parallelism": f"pairs sharded over {world} rank(s)", "streams_per_gpu": a.streams},
            "gcups_equiv_ordered_pairs": equiv_cells(lengths) / (ms_dev * 1e-3) / 1e9,
            "extension_steps_per_step": cells,
            "seconds_to_matrix": {"hbm_resident": ms_dev * 1e-3},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"kernel": "k_gapped_reg (gapped stage: k_gapped_reg<8,6> + <8,8> + k_gapped)", "bound": "int32_issue", "unit": "G lane-ops/s",
                         "achieved": gapped_ops / 1e9, "peak": int_peak / 1e9, "frac": gapped_ops / int_peak,
                         "traffic": traffic.get("k_gapped"),
                         "work_unit": f"one lane-operation = one thread executing one 32-bit integer instruction; an extension "
                                      f"step (one base comparison of the greedy recurrence) needs {OPS_PER_STEP} of them at the least "
                                      "(DESIGN.md section 4); achieved = extension steps per launch x that / launch time",
                         "extension_steps_per_launch": solo["cells_per_launch"], "launch_ms": solo["gapped_ms"],
                         "gcups_alone": solo["cells_per_launch"] / max(solo["gapped_ms"] * 1e-3, 1e-9) / 1e9,
                         "timing": f"CUDA events around the kernel on its own stream, {solo['launches']} sample pairs spread over the "
                                   "workload, searched one at a time after the timed steps (inside a step six streams overlap, so "
                                   "event times there measure queueing)",
                         "peak_source": "measured in-run: rcq_int32_issue_peak (independent add/xor/shift chains at full occupancy)"},
            "roofline_ext": {"kernel": "k_seed_probe", "bound": "hbm", "unit": "GB/s",
                             "achieved": solo["probe_bytes_per_launch"] / max(solo["seed_probe_ms"] * 1e-3, 1e-9) / 1e9,
                             "peak": hbm_peak,
                             "frac": solo["probe_bytes_per_launch"] / max(solo["seed_probe_ms"] * 1e-3, 1e-9) / 1e9 / hbm_peak,
                             "traffic": traffic.get("k_seed_probe"), "algorithmic_bytes_per_launch": solo["probe_bytes_per_launch"],
                             "launch_ms": solo["seed_probe_ms"], "timing": "alone, as above",
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s"},
            "search_stats_per_step": agg,
            "phase2_counts": res["counts"], "host_wall_ms_per_step": step_walls,
            "host_ms_last_step": res["host_ms"],
        }
        if engine_host is not None:
            out["seconds_to_matrix"]["engine_from_host_buffers"] = engine_host["ms"] * 1e-3
    if dist is not None:
        dist.barrier()
    if rank == 0:
        if world == 1 and not a.no_api_leg:
            api = run_api_leg(transcriptomes, a, res["matrix"], max(a.e2e_calls, 3), int(sum(int(x) for x in lengths)))
            t_api = float(np.median(api["seconds"]))          # every call is listed; one slow call (host page cache) does not move the figure
            out["e2e"] = {"value": cells / t_api / 1e9, "unit": "GCUPS", "ms_per_step": t_api * 1e3, "calls": len(api["seconds"]),
                          "seconds_each_call": api["seconds"], "h2d_bytes_per_step": api["h2d"], "d2h_bytes_per_step": api["d2h"],
                          "what": "rna_clique() (the reference's entry point, rna_clique.py:28-178) from FASTA directories on disk to "
                                  "od1/*_top.fasta, od2/*.pkl, graph.pkl and distance_matrix.h5 (files under /dev/shm when it has room, else the temp dir); "
                                  "wall clock per call, median of the timed calls",
                          "matrix_equals_engine": api["same"]}
            out["seconds_to_matrix"]["rna_clique_api_from_fasta_dirs"] = t_api
        else:
            out["e2e"] = {"value": cells / (engine_host["ms"] * 1e-3) / 1e9, "unit": "GCUPS", "ms_per_step": engine_host["ms"],
                          "calls": engine_host["steps"], "h2d_bytes_per_step": engine_host["h2d"], "d2h_bytes_per_step": engine_host["d2h"],
                          "what": "Engine.all_vs_all from pinned host buffers (top-gene selection, host->device copy of every letter, "
                                  "device->host copy of the sums), pairs sharded over the ranks; the drop-in rna_clique() drives one GPU"}
        if not a.no_cpu_baseline and world == 1:
            cores = host_cores()
            jobs, desc = cpu_sample(a, cores)
            times, ccells = run_cpu_steps(jobs, cores, 1, 0)
            out["cpu_baseline"] = {"value": ccells / times[0] / 1e9, "unit": "GCUPS", "cores": cores, "kind": "port", "sample": desc,
                                   "seconds_sample": times[0], "extension_steps_sample": ccells}
        else:
            out["cpu_baseline"] = None
        print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
