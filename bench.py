#!/usr/bin/env python
"""bench.py — RNA-clique hot path on B200: all-vs-all top-gene matching -> gene-match graph ->
ideal components -> distance matrix, on transcriptomes synthesised like the reference's
distance_sequence_simulator.  One JSON line on stdout (rank 0).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                  [--taxa 10 --count 60000 --top 50000]        # BASELINE.json configs[1]

A step = one full pass: pack + seed index of every sample, every unordered pair searched
(both blastn directions), phase 2, distance matrix.  `value` starts with the ASCII transcripts
resident in HBM; `e2e` starts from pinned host buffers (top-gene selection, H2D, D2H inside).
N > 1: the same workload with the sample pairs sharded over the ranks (strong scaling); the only
collective is the all-gather of hit tables.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

INT32_LANES_PER_SM = 128   # Blackwell SM: 128 FP32/INT32 lanes
SMS = 148


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--taxa", type=int, default=10)
    ap.add_argument("--count", type=int, default=60000)
    ap.add_argument("--top", type=int, default=50000)
    ap.add_argument("--seed", type=int, default=487)
    ap.add_argument("--cpu-sample", type=int, default=1500, help="top-gene transcripts per sample in the CPU legs")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-api-leg", action="store_true", help="skip timing rna_clique() from FASTA directories (N = 1 only)")
    ap.add_argument("--streams", type=int, default=0,
                    help="CUDA streams (host threads) per GPU; 0 = 6 on one GPU (measured 86.7 -> 83.8 ms per step against 3), "
                         "3 under torchrun, where the multi-GPU schedule was tuned")
    ap.add_argument("--clock-interval", type=float, default=0.2, help="seconds between nvidia-smi/NVML clock samples (0 = off)")
    return ap.parse_args()


def workload_name(a):
    cfg = {(3, 5000, 2000): "BASELINE.json configs[0]", (10, 60000, 50000): "BASELINE.json configs[1]",
           (30, 100000, 50000): "BASELINE.json configs[2]"}.get((a.taxa, a.count, a.top), "custom size")
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


def cpu_pair_job(args):
    """One unordered pair through the CPU oracle (runs in a worker process)."""
    sys.path.insert(0, ROOT)
    from oracle import oracle as O
    (a1, o1, g1, i1), (a2, o2, g2, i2) = args
    t0 = time.perf_counter()
    s1, s2 = O.Sample(a1, o1, g1, i1), O.Sample(a2, o2, g2, i2)
    r = O.search_pair(s1, s2)
    return time.perf_counter() - t0, r.counts["rows"], r.counts["dp_cells"]


def cpu_sample_inputs(inputs, a, n_sub):
    """The same workload cut down to the n_sub highest-coverage genes of every sample."""
    from rna_clique_b200.engine import select_top_genes
    out = []
    for s in inputs:
        keep = select_top_genes(s["coverage"], s["gene"], min(n_sub, a.top))
        asc = s["ascii"].cpu().numpy() if hasattr(s["ascii"], "cpu") else np.asarray(s["ascii"])
        st, ln = s["starts"][keep], s["lengths"][keep]
        offs = np.concatenate([[0], np.cumsum(ln)]).astype(np.int64)
        buf = np.empty(int(offs[-1]), dtype=np.uint8)
        for k in range(len(keep)):
            buf[offs[k]:offs[k + 1]] = asc[st[k]:st[k] + ln[k]]
        out.append((buf, offs, s["gene"][keep].astype(np.int32), s["iso"][keep].astype(np.int32)))
    return out


def run_cpu_leg(inputs, a, lengths_full, steps, warmup):
    """CPU oracle over every pair of the cut-down workload, one pair per host core at a time.
    Returns (seconds per step, cores, description, GCUPS-equivalent scaled to the full workload)."""
    import itertools
    import multiprocessing as mp
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    n_sub = min(a.cpu_sample, a.top)
    sub = cpu_sample_inputs(inputs, a, n_sub)
    jobs = [(sub[i], sub[j]) for i, j in itertools.combinations(range(len(sub)), 2)]
    times = []
    with mp.get_context("fork").Pool(cores) as pool:
        for s in range(warmup + steps):
            t0 = time.perf_counter()
            pool.map(cpu_pair_job, jobs, chunksize=1)
            if s >= warmup:
                times.append(time.perf_counter() - t0)
    t = float(np.mean(times))
    scale = a.top / n_sub
    gcups = equiv_cells(lengths_full) / (t * scale) / 1e9
    desc = (f"CPU oracle (C restatement, -O2), all {len(jobs)} pairs with the {n_sub} highest-coverage top genes of "
            f"each sample ({n_sub}/{a.top} of the workload), one pair per core; time scaled x{scale:.1f} (work is "
            f"linear in transcripts) to the full workload")
    return t, cores, desc, gcups


def solo_kernel_times(eng, resident, labels, reps=3):
    """k_seed_probe and k_gapped timed alone (one stream, nothing else on the GPU): the first two samples
    are searched `reps` times on context 0 and the per-stage CUDA-event times are read back."""
    ctx = eng.ctxs[0]
    ss = [eng.upload(s["ascii"], s["starts"], s["lengths"], s["gene"], s["iso"], name=labels[k], ctx=ctx)
          for k, s in enumerate(resident[:2])]
    from rna_clique_b200 import _native
    _native.search_pair(ctx, ss[0], ss[1]).close()
    ctx.stage_times(reset=True)
    for _ in range(reps):
        _native.search_pair(ctx, ss[0], ss[1]).close()
    st = ctx.stage_times(reset=True)
    for x in ss:
        x.close()
    return {"seed_probe_ms": st["seed_probe"]["ms"] / max(st["seed_probe"]["calls"], 1),
            "gapped_ms": st["gapped_dp"]["ms"] / max(st["gapped_dp"]["calls"], 1)}


def run_api_leg(transcriptomes, a, engine_matrix):
    """The reference-facing call itself: rna_clique() from transcriptome directories (FASTA on disk) to
    od1/, od2/, graph.pkl and the matrix file, one warm-up call and one timed call.  Returns
    (seconds, matrix equals the engine's)."""
    import shutil
    import tempfile
    from pathlib import Path
    from rna_clique_b200.find_homologs import default_cache
    from rna_clique_b200.rna_clique import rna_clique
    from rna_clique_b200.simulate import write_dataset
    tmp = Path(tempfile.mkdtemp(prefix="rcq_api_"))
    try:
        dirs = write_dataset(tmp / "in", transcriptomes)
        secs, same = None, None
        for rep in range(2):
            for d in ("od1", "od2"):
                shutil.rmtree(tmp / d, ignore_errors=True)
                (tmp / d).mkdir()
            t0 = time.perf_counter()
            sim, _ = rna_clique(dirs, tmp / "od1", tmp / "od2", tmp / "db", tmp / "graph.pkl", tmp / "matrix.pkl", a.top)
            secs = time.perf_counter() - t0
            same = bool(np.array_equal(sim.get_dissimilarity_df().to_numpy(), engine_matrix))
        default_cache().close()
        return secs, same
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def main():
    a = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if a.streams <= 0:
        a.streams = 6 if world == 1 else 3

    if a.impl == "reference":
        if rank != 0:
            return
        import torch
        _, inputs, labels, _ = make_inputs(a, "cpu" if not torch.cuda.is_available() else "cuda")
        from rna_clique_b200.engine import select_top_genes
        lengths_full = [int(s["lengths"][select_top_genes(s["coverage"], s["gene"], a.top)].sum()) for s in inputs]
        t, cores, desc, gcups = run_cpu_leg(inputs, a, lengths_full, a.steps, min(a.warmup, 1))
        print(json.dumps({
            "impl": "reference", "metric": "all_vs_all_alignment_gcups_equiv", "value": gcups, "unit": "GCUPS",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": min(a.warmup, 1), "ms_per_step": t * 1e3 * (a.top / min(a.cpu_sample, a.top)),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": workload_name(a)},
            "cpu_baseline": {"value": gcups, "unit": "GCUPS", "cores": cores, "kind": "port", "sample": desc},
            "e2e": {"value": gcups, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "BLAST+ (the reference's native component) is unavailable on this box; the CPU arm is the C oracle",
        }))
        return

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's version banner (NCCL_DEBUG=VERSION) and any other
        # NCCL logging go to stderr
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ.pop("NCCL_DEBUG")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import torch.distributed as dist_mod
        # NCCL's kernels on a high-priority stream: the sample exchange runs behind the searches, and its thread
        # blocks should be the first to get an SM when one frees up
        opts = None
        if os.environ.get("RCQ_NCCL_PRIORITY", "1") != "0":
            try:
                opts = dist_mod.ProcessGroupNCCL.Options(is_high_priority_stream=True)
            except Exception:  # noqa: BLE001
                opts = None
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local), pg_options=opts)
        dist = dist_mod
    from rna_clique_b200.engine import Engine, select_top_genes
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
    # host variant: pinned ASCII + unselected id arrays
    host = [dict(s, ascii=s["ascii"].cpu().pin_memory()) for s in inputs_dev]
    h2d = int(sum(s["ascii"].numel() for s in host) + sum(16 * len(s["starts"]) for s in resident))

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
        eng.stage_times(reset=True)
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
        return ms / steps, out, eng.stage_times(reset=True), (eng.launches - l0) / steps

    sampler = ClockSampler(local, a.clock_interval or 0.05)
    if a.clock_interval > 0:
        sampler.start()
    ms_dev, res, stages, launches = timed(lambda: eng.all_vs_all(resident, labels, None, dist=dist), a.steps, a.warmup)
    ms_e2e, res2, _, _ = timed(lambda: eng.all_vs_all(host, labels, a.top, dist=dist), a.steps, a.warmup)
    sampler.stop_flag = True
    if a.clock_interval > 0:
        sampler.join(timeout=2)
    assert np.array_equal(res["matrix"], res2["matrix"])

    lengths = res["lengths"]
    cells_eq = equiv_cells(lengths)
    stats = res["stats"]
    agg = {k: float(v) for k, v in stats.items()}
    if dist is not None:
        keys = sorted(agg)
        t = torch.tensor([agg[k] for k in keys], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(t)
        agg = dict(zip(keys, t.tolist()))
    out = None
    if rank == 0:
        n_pairs_local = max(res["n_local_pairs"], 1)
        dp_ms = stages["gapped_dp"]["ms"] / a.steps           # this rank, per step
        dp_calls = max(stages["gapped_dp"]["calls"] / a.steps, 1)
        dp_cells_rank = stats.get("dp_cells", 0)
        clocks = sampler.summary()
        mhz = clocks["sm_max_mhz"] or 1965
        int_ops_peak = SMS * INT32_LANES_PER_SM * mhz * 1e6
        OPS_PER_CELL = 12   # minimum int32 instructions of one cell of this recurrence
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:  # noqa: BLE001
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        # ---- dominant HBM-streaming kernel of the per-pair path: k_seed_probe.
        # Algorithmic bytes per launch: entry + flank context of both sorted indexes read once
        # (16 B x I_A + 16 B x I_B), the 2^22+1-word directory of B, 8 B per surviving candidate written.
        n_pairs_all = max(len(res["pairs"]), 1)
        i_mean = float(np.mean(res["n_index"]))
        cand_per_launch = agg.get("n_verified", 0.0) / n_pairs_all
        probe_bytes = 32.0 * i_mean + 4.0 * ((1 << 22) + 1) + 8.0 * cand_per_launch
        probe_calls = max(stages["seed_probe"]["calls"], 1)
        probe_ms = stages["seed_probe"]["ms"] / probe_calls
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("k_seed_probe")
        except Exception:  # noqa: BLE001
            pass
        solo = solo_kernel_times(eng, resident, labels)
        out = {
            "metric": "all_vs_all_alignment_gcups_equiv", "value": cells_eq / (ms_dev * 1e-3) / 1e9, "unit": "GCUPS",
            "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_dev,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": workload_name(a), "pairs": len(res["pairs"]), "ordered_searches": 2 * len(res["pairs"]),
                       "nt_per_sample": int(np.mean(lengths)), "l2": "inputs larger than L2 (1 GB of ASCII per step)",
                       "parallelism": f"pairs sharded over {world} rank(s)", "streams_per_gpu": a.streams},
            "seconds_to_matrix": {"hbm_resident": ms_dev * 1e-3, "from_host_buffers": ms_e2e * 1e-3},
            "gcups_actual": agg.get("dp_cells", 0) / (ms_dev * 1e-3) / 1e9,
            "e2e": {"value": cells_eq / (ms_e2e * 1e-3) / 1e9, "unit": "GCUPS", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(16 * len(res["pairs"]) + 64)},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"kernel": "k_seed_probe", "bound": "hbm",
                         "achieved": probe_bytes / max(probe_ms * 1e-3, 1e-9) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": probe_bytes / max(probe_ms * 1e-3, 1e-9) / 1e9 / hbm_peak, "traffic": traffic,
                         "algorithmic_bytes_per_launch": probe_bytes, "launch_ms": probe_ms,
                         "share_of_step": stages["seed_probe"]["ms"] / a.steps / ms_dev,
                         "achieved_alone": probe_bytes / max(solo["seed_probe_ms"] * 1e-3, 1e-9) / 1e9,
                         "launch_ms_alone": solo["seed_probe_ms"],
                         "frac_alone": probe_bytes / max(solo["seed_probe_ms"] * 1e-3, 1e-9) / 1e9 / hbm_peak,
                         "note": f"launch_ms is measured inside the timed steps, where {a.streams} streams share the GPU; "
                                 "*_alone is the same kernel timed on one stream after the steps "
                                 "(more streams shorten the step and lengthen every overlapped launch: 3 streams 86.7 ms per step "
                                 "and frac 0.29, 6 streams 84.0 ms and frac 0.15, frac_alone 0.46 either way)",
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s"},
            "roofline_ext": {"kernel": "k_gapped", "bound": "int_issue",
                             "achieved": dp_cells_rank / max(dp_ms * 1e-3, 1e-9) / 1e9,
                             "peak": int_ops_peak / OPS_PER_CELL / 1e9, "unit": "G extension steps/s",
                             "frac": (dp_cells_rank / max(dp_ms * 1e-3, 1e-9)) / (int_ops_peak / OPS_PER_CELL),
                             "launch_ms": dp_ms / dp_calls, "share_of_step": dp_ms / ms_dev,
                             "launch_ms_alone": solo["gapped_ms"],
                             "peak_source": f"148 SMs x 128 int32 lanes x {mhz} MHz / {OPS_PER_CELL} ops per step (nominal)"},
            "stages_ms_per_step": {k: v["ms"] / a.steps for k, v in stages.items()},
            "search_stats_per_step": agg,
            "phase2_counts": res["counts"], "host_wall_ms_per_step": step_walls, "host_ms_last_step": {"hbm_resident": res["host_ms"], "from_host_buffers": res2["host_ms"]},
        }
    if dist is not None:
        dist.barrier()
    if rank == 0:
        if not a.no_cpu_baseline and world == 1:
            t, cores, desc, gcups = run_cpu_leg(inputs_dev, a, lengths, 1, 0)
            out["cpu_baseline"] = {"value": gcups, "unit": "GCUPS", "cores": cores, "kind": "port", "sample": desc,
                                   "seconds_sample": t}
        else:
            out["cpu_baseline"] = None
        if not a.no_api_leg and world == 1:
            secs, same = run_api_leg(transcriptomes, a, res["matrix"])
            out["seconds_to_matrix"]["rna_clique_api_from_fasta_dirs"] = secs
            out["seconds_to_matrix"]["api_matrix_equals_engine"] = same
        print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
