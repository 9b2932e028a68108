#!/usr/bin/env python
"""bench.py -- GCUPS of the database-search hot path on N B200s (BASELINE.json metric).

Default workload (config 2 of BASELINE.json): Swiss-Prot-shaped synthetic database
(570k proteins, ~2.04e8 residues, 32 planted giants up to 35213 aa), BLOSUM62,
gap open 11 / extend 1, Smith-Waterman score mode, queries of 100, 300, 1000,
2000 and 5000 aa.  One STEP = the five queries searched against the whole
database through ONE batched call of the C-ABI (searches stream-ordered back to back;
the reference's pattern for several queries is a ThreadPool over Aligner.align,
README.md:116-144).  N > 1: the database is dealt by cumulative cell count over
the ranks (one process per GPU), each rank scores its shard and the per-shard
score arrays are all-gathered with NCCL by libpyopal_cuda.so itself -- the total
work is fixed, so scaling is "strong".  torch appears here only to barrier and
max-reduce the timings of the ranks and to hand the 128-byte NCCL id around.

Other workloads: --workload config1 (100k proteins, one 300-aa query; also a batch
of 64), --algorithm nw|hw|ov (config 3), --mode end (score + end cell over the whole
database), --workload config4 (top-1000 hits re-aligned in end and full mode),
--workload config5 (UniRef50-shaped blocks, batches of 64 queries, device top-k).

GCUPS = 1e-9 * sum_q Lq * total residues * steps / seconds  (algorithmic cells).
"""

import argparse
import array
import importlib.util
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

QUERY_LENGTHS = (100, 300, 1000, 2000, 5000)
GAP_OPEN, GAP_EXTEND, MATRIX = 11, 1, "BLOSUM62"
METRIC = "GCUPS (whole box, device-timed) at 1/2/4/8 B200 vs PyOpal CPU"
ALU_INSTR_PER_CELL_PAIR = 5.5  # PRMT, E, M, G, F + half a VIMNMX3 for the running best (kernels.cuh)
ALGO = {"nw": 0, "hw": 1, "ov": 2, "sw": 3}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="config2", choices=["config1", "config2", "config4", "config5"])
    ap.add_argument("--targets", type=int, default=None, help="database size (config 2: 570000, config 1: 100000)")
    ap.add_argument("--algorithm", default="sw", choices=["sw", "nw", "hw", "ov"], help="nw / hw / ov = config 3")
    ap.add_argument("--mode", default="score", choices=["score", "end"], help="end: score + end cell over the whole database")
    ap.add_argument("--serial", action="store_true", help="one blocking search per query instead of the batched call")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the CPU baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-verify", action="store_true", help="skip the oracle check of a subsample (outside the timed region)")
    ap.add_argument("--blocks-per-gpu", type=int, default=8, help="config5: blocks of --block proteins per GPU (8 = 1/8 of C5)")
    ap.add_argument("--block", type=int, default=781_250, help="config5: proteins per block")
    ap.add_argument("--queries", type=int, default=64, help="config5 / config1 batch: queries per batch")
    ap.add_argument("--topk", type=int, default=1000, help="config4 / config5: hits kept per query")
    ap.add_argument("--verify", action="store_true", help="config5: compare 1e5 random targets with the oracle")
    return ap.parse_args()


def int_matrix():
    """BLOSUM62 as the row-major int matrix, read from the data module by path: the CPU arm must not import the
    GPU package (its shared libraries would be mapped into the reference process)."""
    spec = importlib.util.spec_from_file_location("_matrix_data", os.path.join(ROOT, "pyopal_b200", "_matrix_data.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return array.array("i", [int(v) for row in mod.MATRICES[MATRIX] for v in row])


def workload(args):
    import numpy as np
    import synthetic

    if args.workload == "config1":
        q, db = synthetic.config1(args.targets or 100_000)
        return db, [np.ascontiguousarray(q)]
    n = args.targets or 570_000
    db = synthetic.config2(n, giants=32 if n >= 10_000 else 0)
    queries = synthetic.config2_queries(db, QUERY_LENGTHS)
    return db, [np.ascontiguousarray(q) for q in queries]


def config_dict(args, db, queries):
    """The same dict in both arms (GPU and CPU reference): what is searched, not how."""
    residues, offsets, lengths = db
    qlens = [len(q) for q in queries]
    name = {"config1": "config1: synthetic db", "config2": "config2: Swiss-Prot-shaped synthetic db",
            "config4": "config4: Swiss-Prot-shaped synthetic db, top hits re-aligned in end and full mode"}[args.workload]
    what = "score + end cell" if args.mode == "end" else "score"
    return {
        "workload": f"{name}, {len(lengths)} proteins, {int(lengths.sum())} residues, {args.algorithm.upper()} {what} mode, "
                    f"{MATRIX} open {GAP_OPEN} ext {GAP_EXTEND}, queries {qlens} aa per step",
        "targets": int(len(lengths)),
        "residues": int(lengths.sum()),
        "query_lengths": qlens,
        "cells_per_step": int(sum(qlens)) * int(lengths.sum()),
        "l2": "inputs (2 B per pair column of the interleaved residue stream, re-read for every query) exceed the 126 MB L2"
              if int(lengths.sum()) * 2 > 126e6 else "L2 flushed by the 126+ MB of score rows / residue stream between queries: no",
    }


# --------------------------------------------------------------------------
# CPU arm: the oracle port (the reference's own Opal sources are absent from
# /root/reference, so it cannot be compiled: kind = "port").


def cpu_sample(db, queries, algorithm, budget_s, threads):
    """Time the CPU restatement on a bounded sample of the workload.

    SW (the benchmark's algorithm): oracle/opal_simd.c, an inter-sequence SIMD kernel in the style of
    the reference's engine (16-bit lanes, overflow recomputed at full width), all host threads; queries in
    workload order, each against the first n targets that fit the remaining budget.  Other algorithms: the
    scalar oracle."""
    import oracle

    oracle.build()
    residues, offsets, lengths = db
    M = int_matrix()

    def run(q, n):
        n = max(1, min(n, len(lengths)))
        off = array.array("q", offsets[:n].tolist())
        ln = array.array("i", lengths[:n].tolist())
        buf = residues[: int(offsets[n - 1] + lengths[n - 1])].tobytes() + b"\0"
        t0 = time.perf_counter()
        if algorithm == "sw":
            oracle.sw_scores_simd(q, buf, off, ln, M, 24, GAP_OPEN, GAP_EXTEND, threads)
        else:
            oracle.search_flat(q, buf, off, ln, M, 24, GAP_OPEN, GAP_EXTEND, algorithm, "score", threads)
        return len(q) * int(lengths[:n].sum()), time.perf_counter() - t0, n

    qs = [q.tobytes() for q in queries]
    probe = qs[min(1, len(qs) - 1)]
    cells, dt, n = run(probe, 20000)
    rate = cells / max(dt, 1e-6)
    per_target = float(lengths.mean())
    total_cells, total_dt, parts = 0, 0.0, []
    order = sorted(range(len(qs)), key=lambda i: abs(len(qs[i]) - 650))  # mid-length queries first
    for qi in order * 8:  # keep going until the budget is used
        left = budget_s - total_dt
        if left <= 0.5:
            break
        q = qs[qi]
        n_budget = int(min(len(lengths), max(20000, rate * left / (len(q) * per_target))))
        cells, dt, n = run(q, n_budget)
        total_cells += cells
        total_dt += dt
        parts.append(f"{len(q)} aa x first {n} targets")
    seen = {}
    for p_ in parts:
        seen[p_] = seen.get(p_, 0) + 1
    parts = [p_ if c == 1 else f"{c} x ({p_})" for p_, c in seen.items()]
    impl = "inter-sequence SIMD port (oracle/opal_simd.c, 16-bit lanes only: Opal's first rung is 8-bit, so this may understate it)" \
        if algorithm == "sw" else "scalar oracle port"
    return total_cells / total_dt / 1e9, f"{impl}; queries {' + '.join(parts)} ({total_cells:.3g} cells in {total_dt:.1f} s)"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.workload == "config5":
        return run_config5_reference(args)
    db, queries = workload(args)
    threads = os.cpu_count() or 1
    per_step = max(2.0, min(args.cpu_seconds, 150.0 / max(1, args.steps + args.warmup)))
    vals, sample, t_total = [], "", 0.0
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        v, sample = cpu_sample(db, queries, args.algorithm, per_step, threads)
        if i >= args.warmup:
            vals.append(v)
            t_total += time.perf_counter() - t0
    value = sum(vals) / len(vals)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / max(1, args.steps),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "s16 SIMD lanes (s64 rescore)",
        "data": "synthetic", "config": config_dict(args, db, queries),
        "cpu_baseline": {"value": value, "unit": "GCUPS", "cores": threads, "kind": "port",
                         "sample": sample + f"; one step here = {per_step:.1f} s of that sample, ms_per_step is its wall time"},
        "e2e": {"value": value, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "CPU restatement of the reference's inter-sequence SIMD scheme (oracle/opal_simd.c, gcc auto-vectorised "
                "16-bit lanes, pthreads), not Opal itself: the Opal sources are absent from /root/reference "
                "(vendor/opal submodule empty)",
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons of one GPU during the timed region, through NVML in-process
    (a looping `nvidia-smi` process measurably perturbs the kernels it is watching)."""

    def __init__(self, pci_bus_id, period=0.2):
        super().__init__(daemon=True)
        self.pci = pci_bus_id
        self.period = period
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self.stop_flag = threading.Event()
        self.ready = threading.Event()

    def begin(self):
        """Start the thread and wait until NVML is initialised: nvmlInit takes tens of milliseconds during which
        CUDA launches stall, which must not fall inside the timed region."""
        self.start()
        self.ready.wait(20)

    def run(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByPciBusId(self.pci.encode() if isinstance(self.pci, str) else self.pci)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            masks = {
                "hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap,
            }
            nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            self.ready.set()
            last = False
            while not last:
                # the last sample is taken when the timed region closes, so that a region shorter than one period
                # still has one taken under load
                last = self.stop_flag.wait(self.period)
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for name, m in masks.items():
                    if r & m:
                        self.reasons.add(name)
        except Exception as err:  # NVML missing: report no clocks rather than die
            self.error = repr(err)
        finally:
            self.ready.set()

    def finish(self):
        self.stop_flag.set()
        self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "error": getattr(self, "error", None)}
        sm = sorted(self.samples)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(sm)}


class Ranks:
    """Plumbing between the ranks of a one-process-per-GPU launch: barrier, max / sum of a number, the NCCL id.
    torch.distributed when WORLD_SIZE > 1 (as the driver launches it), nothing at all otherwise."""

    def __init__(self, args):
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world > 1:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={self.world}")
        self.dist = None
        if self.world > 1:
            import torch
            import torch.distributed as dist

            torch.cuda.set_device(self.local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist, self.torch = dist, torch

    def barrier(self):
        if self.dist:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def reduce(self, x, op="max"):
        if not self.dist:
            return float(x)
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return float(t.item())

    def nccl_id(self, cuda):
        if not self.dist:
            return None
        box = [cuda.nccl_unique_id() if self.rank == 0 else None]
        self.dist.broadcast_object_list(box, src=0)
        return box[0]

    def close(self):
        if self.dist:
            self.dist.barrier()
            self.dist.destroy_process_group()


def checksums(scores_row):
    """(int64 sum, position-weighted hash) of one database-order score row: equal at every N or the run is wrong."""
    import numpy as np

    s = np.asarray(scores_row, dtype=np.int64)
    w = (np.arange(len(s), dtype=np.int64) % 1_000_003) + 1
    return int(s.sum()), int((s * w).sum() % 2_147_483_647)


def verify_rows(db, qbytes, rows, algorithm, M, mode="score", extra=None, sample=2000):
    """Compare a random subsample + every giant with the oracle (outside any timed region)."""
    import numpy as np
    import oracle
    import synthetic

    oracle.build()
    residues, offsets, lengths = db
    rng = np.random.Generator(np.random.PCG64(2024))
    giants = np.nonzero(lengths >= 10_000)[0]
    idx = np.unique(np.concatenate([rng.choice(len(lengths), min(sample, len(lengths)), replace=False), giants]))
    sr, so, sl = synthetic.pack_subset(db, idx)
    flat = (sr.tobytes() + b"\0", array.array("q", so.tolist()), array.array("i", sl.tolist()))
    ok = True
    for qi, q in enumerate(qbytes):
        want = oracle.search_flat(q, *flat, M, 24, GAP_OPEN, GAP_EXTEND, algorithm, mode, os.cpu_count() or 8)
        got = np.asarray(rows[qi])[idx]
        ok = ok and got.tolist() == list(want["score"])
        if extra is not None:
            ok = ok and np.asarray(extra[qi][0])[idx].tolist() == list(want["query_end"])
            ok = ok and np.asarray(extra[qi][1])[idx].tolist() == list(want["target_end"])
    return bool(ok), int(len(idx))


def roofline_parts(cuda, local, per_q, store_bytes, residues_local, args, world):
    import ctypes

    rate = None
    r, c = ctypes.c_double(0), ctypes.c_double(0)
    if cuda.load_library().pyopal_cuda_int_issue_rate(local, ctypes.byref(r), ctypes.byref(c)) == 0:
        rate = r.value
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    tot_us = sum(p["k16_us"] for p in per_q) or 1
    cells_k = sum(p["Lq"] for p in per_q) * residues_local
    gcups_k = cells_k / tot_us / 1e3
    bytes_k = sum(p["passes"] * store_bytes for p in per_q)
    roof_gcups = rate * 2 / ALU_INSTR_PER_CELL_PAIR / 1e9 if rate else None
    traffic = traffic_note = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))["launches"]
        if args.workload == "config2" and world == 1 and (args.targets or 570_000) == 570_000 and args.mode == "score":
            t = max(tr, key=lambda x: x["query_length"])
            pq = [p for p in per_q if p["Lq"] == t["query_length"]][0]
            algo_bytes = pq["passes"] * store_bytes
            traffic = t["dram_bytes_per_launch"]
            traffic_note = (f"bulk launch of the {t['query_length']}-aa query: {traffic} B measured by ncu vs {algo_bytes} B "
                            f"algorithmic (residue stream x {pq['passes']} passes)")
    except Exception:
        pass
    roof = {"bound": "int_alu", "achieved": gcups_k, "peak": roof_gcups, "unit": "GCUPS",
            "frac": (gcups_k / roof_gcups) if roof_gcups else None, "traffic": traffic,
            "alu_thread_instr_per_s": rate, "alu_instr_per_cell_pair": ALU_INSTR_PER_CELL_PAIR,
            "peak_source": "pyopal_cuda_int_issue_rate (measured on this GPU in this run)",
            "note": "the binding roof of max-plus DP is the integer ALU pipe (64 lanes/clk/SM), neither HBM nor tensor cores "
                    "(DESIGN.md 4.1): achieved = algorithmic cell updates of the 16-bit kernels / their CUDA-event time"}
    hbm = {"bound": "hbm", "achieved": bytes_k / tot_us / 1e3, "peak": hbm_peak, "unit": "GB/s",
           "frac": bytes_k / tot_us / 1e3 / hbm_peak, "traffic": traffic, "traffic_note": traffic_note,
           "note": "residue stream (2 B per pair column per pass) of the 16-bit kernel; 1/Lq byte per cell, so HBM is not the binding roof",
           "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback"}
    return roof, hbm


def run_b200(args):
    import numpy as np

    import pyopal_b200 as pyopal
    from pyopal_b200 import sharding
    from pyopal_b200.platform import cuda

    ranks = Ranks(args)
    world, rank, local = ranks.world, ranks.rank, ranks.local
    db, queries = workload(args)
    residues, offsets, lengths = db
    total_res = int(lengths.sum())
    n_all = len(lengths)
    M = int_matrix()
    algo = ALGO[args.algorithm]
    qbytes = [q.tobytes() for q in queries]
    cells_step = sum(len(q) for q in qbytes) * total_res

    # the repo's multi-GPU handle: this rank's cell-balanced shard on its GPU + the NCCL gather (csrc/multi.cu)
    # (a one-target store first, so that creating the CUDA context is not charged to the build)
    cuda.DeviceStore(np.zeros(16, np.uint8), np.zeros(1, np.int64), np.full(1, 16, np.int32), 24, device=local).close()
    t_build = time.perf_counter()
    sdb = sharding.ShardedDatabase.deal(residues, offsets, np.ascontiguousarray(lengths), 24, rank=rank, world=world,
                                        nccl_id=ranks.nccl_id(cuda), device=local)
    t_build = time.perf_counter() - t_build  # one-off: sort + pair on the host, one H2D of the residues, interleave on the GPU
    res_local = sdb.local_residues(0)
    end_mode = args.mode == "end"

    host_rows = array.array("i", bytes(4 * len(qbytes) * n_all))  # the caller's result buffer of the e2e leg, reused

    def step(to_host=False):
        if end_mode:  # score + end cell: the blocking call, host arrays out
            return [sdb.search(q, GAP_OPEN, GAP_EXTEND, M, 1, algo, lengths=lengths) for q in qbytes]
        if args.serial:
            return [sdb.search(q, GAP_OPEN, GAP_EXTEND, M, 0, algo)[0] for q in qbytes]
        return sdb.search_batch(qbytes, GAP_OPEN, GAP_EXTEND, M, algo, to_host=to_host, out=host_rows if to_host else None)

    for _ in range(args.warmup):
        step()
    ranks.barrier()
    sampler = ClockSampler(cuda.device_pci_bus_id(local))
    sampler.begin()
    launches = 0
    ranks.barrier()
    sdb.time_begin()          # cudaDeviceSynchronize + event on the handle's stream
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
        launches += sdb.stats()["launches"]
    ms = sdb.time_end()       # device clock: events on the streams the kernels ran on, everything complete
    wall = time.perf_counter() - t0
    ranks.barrier()
    clocks = sampler.finish()
    ms = ranks.reduce(ms, "max")
    value = cells_step * args.steps / (ms * 1e-3) / 1e9
    launches = int(ranks.reduce(launches, "sum"))

    # end to end: the same call with HOST buffers (queries H2D, every score D2H to rank 0's host inside the timed region)
    ranks.barrier()
    t0 = time.perf_counter()
    rows = None
    for _ in range(args.steps):
        rows = step(to_host=True)
    dt = ranks.reduce(time.perf_counter() - t0, "max")
    e2e_value = cells_step * args.steps / dt / 1e9
    h2d = sum(len(q) for q in qbytes) + len(qbytes) * (24 * 24 * 4 + 64 + 4096)
    d2h = len(qbytes) * n_all * 4 * (3 if end_mode else 1)

    # database-order results of every query: checksums (equal at every N) and the oracle on a subsample
    if end_mode:
        score_rows = [np.frombuffer(r[0], dtype=np.int32) for r in rows]
        extra = [(np.frombuffer(r[1], dtype=np.int32), np.frombuffer(r[2], dtype=np.int32)) for r in rows]
    elif args.serial:
        score_rows, extra = [np.frombuffer(r, dtype=np.int32) for r in rows], None
    else:
        flat = np.frombuffer(rows, dtype=np.int32)
        score_rows, extra = [flat[i * n_all: (i + 1) * n_all] for i in range(len(qbytes))], None
    sums = [checksums(r) for r in score_rows]
    if end_mode:
        sums = [s + checksums(e[0]) + checksums(e[1]) for s, e in zip(sums, extra)]

    # per-kernel pass (outside the timed region): CUDA events around the 16-bit kernel of each search, one at a time
    per_q = []
    for q in qbytes:
        sdb.search(q, GAP_OPEN, GAP_EXTEND, M, 1 if end_mode else 0, algo, lengths=lengths)
        st = sdb.stats()
        per_q.append({"Lq": len(q), "G": st["G"], "K": st["K"], "passes": st["passes"], "k16_us": st["k16_us"],
                      "k32_us": st["k32_us"], "rescore_us": st["rescore_us"], "device_us": st["device_us"],
                      "gcups_k16": round(len(q) * res_local / max(st["k16_us"], 1) / 1e3, 1)})
    if end_mode:  # stats[5..7] describe pairdp.cu's spill launch there; the packed kernel's time is what matters
        for p in per_q:
            p["passes"] = max(1, p["passes"])

    line = None
    if rank == 0:
        roof, hbm = roofline_parts(cuda, local, per_q, int(sdb.device_bytes(0)), res_local, args, world)
        cfg = config_dict(args, db, queries)
        line = {
            "metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u16x2 (s32 rescore)", "data": "synthetic", "config": cfg,
            "run": {"parallelism": f"db dealt by cells over {world} GPU(s), ncclAllGather of the per-shard scores inside libpyopal_cuda.so",
                    "call": "one blocking search per query" if (args.serial or end_mode) else "one batched call per step, searches stream-ordered back to back on every GPU",
                    "store_build_s": round(t_build, 3), "wall_ms_per_step": round(1e3 * wall / args.steps, 3),
                    "nccl_version": cuda.load_library().pyopal_cuda_nccl_version() if world > 1 else None},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "GCUPS", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "pyopal_cuda_mdb_search_batch / pyopal_cuda_mdb_search (host queries in, database-order host scores "
                           "out on rank 0; shard searches + NCCL gather inside), wall clock, max over ranks"},
            "roofline": roof, "roofline_hbm": hbm, "kernels": per_q,
            "gpu_launches": launches,
            "checksums": {str(len(q)): list(s) for q, s in zip(qbytes, sums)},
        }
        if not args.no_verify:
            ok, nver = verify_rows(db, qbytes, score_rows, args.algorithm, M, "end" if end_mode else "score", extra)
            line["verified"] = ok
            line["verified_targets"] = nver
    if world == 1 and rank == 0 and not end_mode and args.workload == "config2" and not args.serial:
        # Aligner.align end to end, result objects included (the reference-facing call)
        pdb = pyopal.Database.from_packed(residues, lengths, validate=False)
        aligner = pyopal.Aligner(MATRIX, GAP_OPEN, GAP_EXTEND)
        qstr = [aligner.alphabet.decode(q) for q in qbytes]
        aligner.align(qstr[0], pdb, algorithm=args.algorithm)  # builds the store
        t0 = time.perf_counter()
        n_obj = 0
        for q in qstr:
            n_obj += len(aligner.align(q, pdb, algorithm=args.algorithm))
        dta = time.perf_counter() - t0
        line["e2e_align"] = {"value": cells_step / dta / 1e9, "unit": "GCUPS", "ms_per_step": 1e3 * dta,
                             "result_objects": n_obj, "api": "Aligner.align on a pyopal.Database.from_packed database, wall clock, "
                             "one ScoreResult per target built natively (csrc/_results.c)"}
    if args.workload == "config1" and rank == 0 and not end_mode:
        # the same database against a batch of --queries random 300-aa queries: the launch-bound case
        import synthetic

        rng = np.random.Generator(np.random.PCG64(22))
        bq = [synthetic.random_residues(rng, 300).tobytes() for _ in range(args.queries)]
        sdb.search_batch(bq, GAP_OPEN, GAP_EXTEND, M, algo, to_host=False)
        sdb.time_begin()
        sdb.search_batch(bq, GAP_OPEN, GAP_EXTEND, M, algo, to_host=False)
        bms = sdb.time_end()
        g = len(bq) * 300 * total_res / (bms * 1e-3) / 1e9
        line["batch"] = {"queries": len(bq), "ms": bms, "gcups": g, "frac_of_roof": g / line["roofline"]["peak"] if line["roofline"]["peak"] else None}
    if rank == 0:
        if not args.no_cpu_baseline and world == 1:
            v, sample = cpu_sample(db, queries, args.algorithm, args.cpu_seconds, os.cpu_count() or 1)
            line["cpu_baseline"] = {"value": v, "unit": "GCUPS", "cores": os.cpu_count() or 1, "kind": "port", "sample": sample}
        print(json.dumps(line), flush=True)
    sdb.close()
    ranks.close()


# --------------------------------------------------------------------------
# Config 4 of BASELINE.json: SW score pass, top-1000 hits per query, `end` and `full` on database.extract(top).


def run_config4(args):
    import numpy as np

    import pyopal_b200 as pyopal
    import synthetic
    from pyopal_b200.platform import cuda

    args.workload = "config4"
    db, queries = workload(args)
    residues, offsets, lengths = db
    M = int_matrix()
    qbytes = [q.tobytes() for q in queries]
    store = cuda.DeviceStore(residues, offsets, np.ascontiguousarray(lengths), 24, device=0)
    k = min(args.topk, len(lengths))
    subs = []
    for q in qbytes:
        _, idx = store.search_topk(q, GAP_OPEN, GAP_EXTEND, M, 3, k)
        top = np.sort(np.frombuffer(idx, dtype=np.int32))
        sr, so, sl = synthetic.pack_subset(db, top)
        subs.append((cuda.DeviceStore(sr, so, sl, 24, device=0), sl, (sr, so, sl)))

    def step():
        out = []
        for q, (sub, sl, _) in zip(qbytes, subs):
            out.append((sub.search(q, GAP_OPEN, GAP_EXTEND, M, 1, 3, lengths=sl), sub.search(q, GAP_OPEN, GAP_EXTEND, M, 2, 3, lengths=sl)))
        return out

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(cuda.device_pci_bus_id(0))
    sampler.begin()
    dev_us = 0
    launches = 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        for q, (sub, sl, _) in zip(qbytes, subs):
            for st_ in (1, 2):
                sub.search(q, GAP_OPEN, GAP_EXTEND, M, st_, 3, lengths=sl)
                s = sub.stats()
                dev_us += s["device_us"]
                launches += s["launches"]
    wall = time.perf_counter() - t0
    clocks = sampler.finish()
    cells_step = 2 * sum(len(q) * int(sl.sum()) for q, (_, sl, _) in zip(qbytes, subs))
    value = cells_step * args.steps / (dev_us * 1e-6) / 1e9
    res = step()
    ok = None
    if not args.no_verify:
        import oracle

        oracle.build()
        ok = True
        for q, (sub, sl, packed), (e, f) in zip(qbytes, subs, res):
            flat = (packed[0].tobytes() + b"\0", array.array("q", packed[1].tolist()), array.array("i", packed[2].tolist()))
            want = oracle.search_flat(q, *flat, M, 24, GAP_OPEN, GAP_EXTEND, "sw", "full", os.cpu_count() or 8)
            ok = ok and list(e[0]) == want["score"] and list(e[1]) == want["query_end"] and list(e[2]) == want["target_end"]
            ok = ok and list(f[3]) == want["query_start"] and list(f[4]) == want["target_start"]
            ops, oo, ol = f[5], f[6], f[7]
            ok = ok and all(bytes(ops[oo[j]: oo[j] + ol[j]]) == want["alignment"][j] for j in range(len(sl)))
    cfg = config_dict(args, db, queries)
    cfg["workload"] += f"; top {k} hits of every query (SW score pass + device top-k outside the timed region), then end + full mode on them"
    cfg["cells_per_step"] = cells_step
    line = {
        "metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_us / 1e3 / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u16x2 forward (s32 reverse / traceback)", "data": "synthetic", "config": cfg, "clocks": clocks,
        "alignments_per_s": 2 * k * len(qbytes) * args.steps / wall,
        "e2e": {"value": cells_step * args.steps / wall / 1e9, "unit": "GCUPS", "h2d_bytes_per_step": sum(len(q) for q in qbytes) * 2,
                "d2h_bytes_per_step": int(sum(5 * 4 * len(sl) * 2 + len(q) * len(sl) + int(sl.sum()) for q, (_, sl, _) in zip(qbytes, subs))),
                "api": "pyopal_cuda_search end + full on the extracted hits (host arrays in / out), wall clock"},
        "roofline": {"bound": "latency", "achieved": value, "peak": None, "unit": "GCUPS", "frac": None, "traffic": None,
                     "note": "bound by the serial walk of the longest hit (31-35 k columns), not by a throughput roof: DESIGN.md 4.3"},
        "gpu_launches": launches, "verified": ok,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------
# Config 5 of BASELINE.json: UniRef50-shaped database (5e7 proteins at 8 GPUs), batched multi-query
# search, block-dealt shards, NCCL gather of every query's scores.  One STEP = one batch of --queries
# queries.  Per-GPU work is fixed (--blocks-per-gpu blocks), so this workload scales WEAKLY: N GPUs
# search N x 6.25e6 proteins; 8 GPUs = the full 5e7-protein C5.


def config5_cfg(args, world, qbytes, total_res, extra=None):
    n_blocks = args.blocks_per_gpu * max(1, world)
    sum_lq = sum(len(q) for q in qbytes)
    c = {"workload": f"config5: UniRef50-shaped synthetic db, {n_blocks} blocks x {args.block} proteins "
                     f"({n_blocks * args.block} proteins, {total_res} residues), SW score mode, {MATRIX} open {GAP_OPEN} "
                     f"ext {GAP_EXTEND}, batch of {len(qbytes)} queries of {min(map(len, qbytes))}..{max(map(len, qbytes))} aa "
                     f"per step, top-{args.topk} hits per query",
         "targets": n_blocks * args.block, "residues": total_res, "queries": len(qbytes), "query_residues": sum_lq,
         "cells_per_step": sum_lq * total_res,
         "l2": "inputs (3.5 GB pair-interleaved residue stream per GPU) exceed the 126 MB L2"}
    if extra:
        c.update(extra)
    return c


def run_config5_reference(args):
    import numpy as np
    import synthetic

    queries = synthetic.config5_queries(args.queries)
    qbytes = [np.ascontiguousarray(q).tobytes() for q in queries]
    res, lens = synthetic.config5_block(0, args.block)
    db = synthetic.pack(lens, res)
    threads = os.cpu_count() or 1
    per_step = max(2.0, min(args.cpu_seconds, 150.0 / max(1, args.steps + args.warmup)))
    vals, t_total, sample = [], 0.0, ""
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        v, sample = cpu_sample(db, queries, "sw", per_step, threads)
        if i >= args.warmup:
            vals.append(v)
            t_total += time.perf_counter() - t0
    value = sum(vals) / len(vals)
    world = max(1, args.gpus)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "s16 SIMD lanes (s64 rescore)",
        "data": "synthetic", "config": config5_cfg(args, world, qbytes, int(lens.sum()) * args.blocks_per_gpu * world,
                                                   {"note": "residues extrapolated from block 0"}),
        "cpu_baseline": {"value": value, "unit": "GCUPS", "cores": threads, "kind": "port", "sample": sample + " of block 0"},
        "e2e": {"value": value, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)


def run_config5(args):
    import numpy as np

    import synthetic
    from pyopal_b200 import sharding
    from pyopal_b200.platform import cuda

    ranks = Ranks(args)
    world, rank, local = ranks.world, ranks.rank, ranks.local
    queries = synthetic.config5_queries(args.queries)
    qbytes = [np.ascontiguousarray(q).tobytes() for q in queries]
    sum_lq = sum(len(q) for q in qbytes)
    n_blocks = args.blocks_per_gpu * world
    t0 = time.perf_counter()
    sr, so, sl, mine = synthetic.config5_shard(rank, world, n_blocks, args.block)
    t_gen = time.perf_counter() - t0
    res_mine = int(sl.sum())
    M = int_matrix()
    t0 = time.perf_counter()
    # the repo's multi-GPU API: this rank's blocks on its GPU + the NCCL gather (pyopal_b200/sharding.py, csrc/multi.cu)
    sdb = sharding.ShardedDatabase.from_blocks((sr, so, np.ascontiguousarray(sl)), args.block, args.blocks_per_gpu, 24,
                                               rank=rank, world=world, nccl_id=ranks.nccl_id(cuda), device=local)
    t_build = time.perf_counter() - t0
    total_res = int(ranks.reduce(res_mine, "sum"))
    imbalance = ranks.reduce(res_mine, "max") * world / total_res - 1.0
    k = min(args.topk, sdb.size)

    def batch():
        return sdb.search_batch_topk(qbytes, GAP_OPEN, GAP_EXTEND, M, 3, k)

    for _ in range(args.warmup):
        batch()
    ranks.barrier()
    sampler = ClockSampler(cuda.device_pci_bus_id(local))
    sampler.begin()
    launches = 0
    ranks.barrier()
    sdb.time_begin()
    t0 = time.perf_counter()
    hits = None
    for _ in range(args.steps):
        hits = batch()
        launches += sdb.stats()["launches"]
    ms = ranks.reduce(sdb.time_end(), "max")
    wall = ranks.reduce(time.perf_counter() - t0, "max")
    ranks.barrier()
    clocks = sampler.finish()
    launches = int(ranks.reduce(launches, "sum"))
    cells_step = sum_lq * total_res
    value = cells_step * args.steps / (ms * 1e-3) / 1e9
    e2e_value = cells_step * args.steps / wall / 1e9
    checksum = int(np.frombuffer(hits[0], dtype=np.int32).astype(np.int64).sum())

    # roofline pass on rank 0's shard: the eight longest queries as four two-query batches (the unit the batched call
    # searches: two queries per pass of dual_kernel when the shard allows it), CUDA events around the 16-bit kernel
    k16_us, passes_cols, roof_res, dual_units = 0.0, 0.0, 0, 0
    by_len = sorted(qbytes, key=len, reverse=True)[:8]
    if rank == 0 or world > 1:
        for i in range(0, len(by_len), 2):
            unit = by_len[i: i + 2]
            sdb.search_batch(unit, GAP_OPEN, GAP_EXTEND, M, 3, to_host=False)
            st = sdb.stats()
            if st["profile_in_smem"] == 2:      # one dual launch scored both queries
                k16_us += st["k16_us"]
                passes_cols += st["passes"]
                dual_units += 1
            else:                                # searched one after the other: time them one by one
                for q in unit:
                    sdb.search(q, GAP_OPEN, GAP_EXTEND, M, 0, 3)
                    st = sdb.stats()
                    k16_us += st["k16_us"]
                    passes_cols += st["passes"]
            roof_res += sum(len(q) for q in unit)
    alu_per_pair = 4.5 if dual_units * 2 == len(by_len) else ALU_INSTR_PER_CELL_PAIR
    verified = None
    if args.verify:
        # 1e5 random targets of the real blocks against the oracle (full score rows of two queries)
        rows = sdb.search_batch(qbytes[:2], GAP_OPEN, GAP_EXTEND, M, 3, to_host=True)
        if rank == 0:
            import oracle

            oracle.build()
            flat = np.frombuffer(rows, dtype=np.int32).reshape(2, -1)
            rng = np.random.Generator(np.random.PCG64(77))
            verified = True
            for b in sorted(set([0, 1 % n_blocks, n_blocks - 1, n_blocks // 2])):
                res_b, len_b = synthetic.config5_block(b, args.block)
                pick = np.sort(rng.choice(args.block, min(25_000, args.block), replace=False))
                sub = synthetic.pack_subset(synthetic.pack(len_b, res_b), pick)
                fl = (sub[0].tobytes() + b"\0", array.array("q", sub[1].tolist()), array.array("i", sub[2].tolist()))
                for qi in range(2):
                    want = oracle.search_flat(qbytes[qi], *fl, M, 24, GAP_OPEN, GAP_EXTEND, "sw", "score", os.cpu_count() or 8)["score"]
                    verified = verified and flat[qi][b * args.block + pick].tolist() == list(want)
    if rank == 0:
        import ctypes

        rate = None
        r, c = ctypes.c_double(0), ctypes.c_double(0)
        if cuda.load_library().pyopal_cuda_int_issue_rate(local, ctypes.byref(r), ctypes.byref(c)) == 0:
            rate = r.value
        roof = rate * 2 / alu_per_pair / 1e9 if rate else None
        gcups_k = roof_res * res_mine / max(k16_us, 1) / 1e3
        try:
            hbm_peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0)
        except Exception:
            hbm_peak = 6650.0
        bytes_k = passes_cols * int(sdb.device_bytes(0))
        line = {
            "metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u16x2 (s32 rescore)", "data": "synthetic",
            "config": config5_cfg(args, world, qbytes, total_res),
            "run": {"parallelism": f"blocks dealt round-robin over {world} GPU(s), ncclAllGather of every query's scores inside "
                                   f"libpyopal_cuda.so, top-{k} on the device; queries paired by length, two per pass of the score kernel",
                    "generate_s": round(t_gen, 2), "store_build_s": round(t_build, 2),
                    "store_bytes_per_gpu": int(sdb.device_bytes(0)), "shard_residue_imbalance": round(imbalance, 5)},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "GCUPS", "h2d_bytes_per_step": sum_lq + len(qbytes) * (24 * 24 * 4 + 64 + 4096),
                    "d2h_bytes_per_step": len(qbytes) * k * 8,
                    "api": "pyopal_cuda_mdb_search_batch_topk (host queries in, top-k hits out), wall clock of the same steps"},
            "roofline": {"bound": "int_alu", "achieved": gcups_k, "peak": roof, "unit": "GCUPS",
                         "frac": (gcups_k / roof) if roof else None, "traffic": None,
                         "alu_thread_instr_per_s": rate, "alu_instr_per_cell_pair": alu_per_pair,
                         "peak_source": "pyopal_cuda_int_issue_rate (measured on this GPU in this run)",
                         "note": "batched queries are searched two per pass (dual_kernel: 4.5 ALU-pipe instructions per cell pair, "
                                 "no widening of the profile entries); achieved = the eight longest queries' cells / the CUDA-event "
                                 "time of their four dual launches" if alu_per_pair == 4.5 else
                                 "queries searched one per pass (5.5 ALU-pipe instructions per cell pair)"},
            "roofline_hbm": {"bound": "hbm", "achieved": bytes_k / max(k16_us, 1) / 1e3, "peak": hbm_peak, "unit": "GB/s",
                             "frac": bytes_k / max(k16_us, 1) / 1e3 / hbm_peak, "traffic": None,
                             "note": "residue stream once per pass of the 16-bit kernel on rank 0 (the eight longest queries)"},
            "top_hits_checksum": checksum, "verified": verified,
            "gpu_launches": launches,
        }
        if not args.no_cpu_baseline and world == 1:
            bres, blens = synthetic.config5_block(0, args.block)
            v, sample = cpu_sample(synthetic.pack(blens, bres), queries, "sw", args.cpu_seconds, os.cpu_count() or 1)
            line["cpu_baseline"] = {"value": v, "unit": "GCUPS", "cores": os.cpu_count() or 1, "kind": "port",
                                    "sample": sample + " of block 0"}
        print(json.dumps(line), flush=True)
    sdb.close()
    ranks.close()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "config5":
        run_config5(args)
    elif args.workload == "config4":
        run_config4(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
