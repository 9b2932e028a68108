#!/usr/bin/env python
"""bench.py -- GCUPS of the database-search hot path on N B200s (BASELINE.json metric).

Workload (config 2 of BASELINE.json): Swiss-Prot-shaped synthetic database
(570k proteins, ~2.04e8 residues, 32 planted giants up to 35213 aa), BLOSUM62,
gap open 11 / extend 1, Smith-Waterman score mode, queries of 100, 300, 1000,
2000 and 5000 aa.  One STEP = the five queries searched one after the other
against the whole database.  N > 1: the database is sharded by cumulative
cell count over the ranks (one process per GPU), each rank scores its shard
and the per-shard score arrays are gathered with NCCL -- the total work is
fixed, so scaling is "strong".

GCUPS = 1e-9 * sum_q Lq * total residues * steps / seconds  (algorithmic cells).
"""

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

QUERY_LENGTHS = (100, 300, 1000, 2000, 5000)
GAP_OPEN, GAP_EXTEND, MATRIX = 11, 1, "BLOSUM62"
METRIC = "GCUPS (whole box, device-timed) at 1/2/4/8 B200 vs PyOpal CPU"
ALU_INSTR_PER_CELL_PAIR = 5.5  # PRMT, E, M, G, F + half a VIMNMX3 for the running best (kernels.cuh)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--targets", type=int, default=570_000, help="database size (config 2: 570000)")
    ap.add_argument("--algorithm", default="sw", choices=["sw", "nw", "hw", "ov"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the CPU baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--debug-timing", action="store_true", help="print a search / gather breakdown to stderr")
    ap.add_argument("--workload", default="config2", choices=["config2", "config5"],
                    help="config2 (default, the metric's configuration) or config5: UniRef50-shaped blocks, batched queries")
    ap.add_argument("--blocks-per-gpu", type=int, default=8, help="config5: blocks of --block proteins per GPU (8 = 1/8 of C5)")
    ap.add_argument("--block", type=int, default=781_250, help="config5: proteins per block")
    ap.add_argument("--queries", type=int, default=64, help="config5: queries per batch (one step = one batch)")
    ap.add_argument("--topk", type=int, default=1000, help="config5: hits kept per query")
    return ap.parse_args()


def workload(args):
    import numpy as np
    import synthetic

    db = synthetic.config2(args.targets, giants=32 if args.targets >= 10_000 else 0)
    queries = synthetic.config2_queries(db, QUERY_LENGTHS)
    return db, [np.ascontiguousarray(q) for q in queries]


def config_dict(args, db, extra=None):
    residues, offsets, lengths = db
    cfg = {
        "workload": f"config2: Swiss-Prot-shaped synthetic db, {len(lengths)} proteins, {int(lengths.sum())} residues, "
                    f"{args.algorithm.upper()} score mode, {MATRIX} open {GAP_OPEN} ext {GAP_EXTEND}, "
                    f"queries {list(QUERY_LENGTHS)} aa per step",
        "targets": int(len(lengths)),
        "residues": int(lengths.sum()),
        "query_lengths": list(QUERY_LENGTHS),
        "cells_per_step": int(sum(QUERY_LENGTHS)) * int(lengths.sum()),
        "l2": "inputs (>= 400 MB pair-interleaved residue stream per step at N=1) exceed the 126 MB L2",
    }
    if extra:
        cfg.update(extra)
    return cfg


# --------------------------------------------------------------------------
# CPU arm: the oracle port (the reference's own Opal sources are absent from
# /root/reference, so it cannot be compiled: kind = "port").


def cpu_sample(db, queries, algorithm, budget_s, threads):
    """Time the CPU restatement on a bounded sample of the workload.

    SW (the benchmark's algorithm): oracle/opal_simd.c, an inter-sequence SIMD kernel in the style of
    the reference's engine (16-bit lanes, overflow recomputed at full width), all host threads, the
    300-aa and then the 1000-aa query against the first targets of the database.  Other algorithms:
    the scalar oracle."""
    import array
    import oracle
    import pyopal_b200 as pyopal

    oracle.build()
    residues, offsets, lengths = db
    M = pyopal.Aligner(MATRIX)._int_matrix

    def run(q, n):
        n = min(n, len(lengths))
        off = array.array("q", offsets[:n].tolist())
        ln = array.array("i", lengths[:n].tolist())
        buf = residues[: int(offsets[n - 1] + lengths[n - 1])].tobytes() + b"\0"
        t0 = time.perf_counter()
        if algorithm == "sw":
            oracle.sw_scores_simd(q, buf, off, ln, M, 24, GAP_OPEN, GAP_EXTEND, threads)
        else:
            oracle.search_flat(q, buf, off, ln, M, 24, GAP_OPEN, GAP_EXTEND, algorithm, "score", threads)
        dt = time.perf_counter() - t0
        return len(q) * int(lengths[:n].sum()), dt, n

    # probe the rate on a slice, then size the sample for the budget
    q300, q1000 = queries[1].tobytes(), queries[2].tobytes()
    cells, dt, n = run(q300, 20000)
    rate = cells / max(dt, 1e-6)
    per_target = float(lengths.mean())
    total_cells, total_dt, parts = 0, 0.0, []
    extra = [q.tobytes() for q in queries[3:5]] if len(queries) >= 5 else []
    for q in [q300, q1000] + extra + [q1000, q300] * 8:  # keep going until the budget is used
        left = budget_s - total_dt
        if left <= 0.5:
            break
        n_budget = int(min(len(lengths), max(20000, rate * left / (len(q) * per_target))))
        cells, dt, n = run(q, n_budget)
        total_cells += cells
        total_dt += dt
        parts.append(f"{len(q)} aa x first {n} targets")
    seen = {}
    for p_ in parts:
        seen[p_] = seen.get(p_, 0) + 1
    parts = [p_ if c == 1 else f"{c} x ({p_})" for p_, c in seen.items()]
    impl = "inter-sequence SIMD port (oracle/opal_simd.c)" if algorithm == "sw" else "scalar oracle port"
    return total_cells / total_dt / 1e9, f"{impl}, queries {' + '.join(parts)} ({total_cells:.3g} cells in {total_dt:.1f} s)"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    db, queries = workload(args)
    threads = os.cpu_count() or 1
    per_step = max(2.0, min(args.cpu_seconds, 150.0 / max(1, args.steps + args.warmup)))
    vals = []
    sample = ""
    t_total = 0.0
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
        "data": "synthetic", "config": config_dict(args, db),
        "cpu_baseline": {"value": value, "unit": "GCUPS", "cores": threads, "kind": "port", "sample": sample},
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
            while not self.stop_flag.is_set():
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for name, m in masks.items():
                    if r & m:
                        self.reasons.add(name)
                self.stop_flag.wait(self.period)
        except Exception as err:  # NVML missing: report no clocks rather than die
            self.error = repr(err)

    def finish(self):
        self.stop_flag.set()
        self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "error": getattr(self, "error", None)}
        sm = sorted(self.samples)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(sm)}


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import pyopal_b200 as pyopal
    from pyopal_b200.platform import cuda

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    db, queries = workload(args)
    residues, offsets, lengths = db
    total_res = int(lengths.sum())
    n_all = len(lengths)
    from pyopal_b200 import sharding

    mine = sharding.shard_indices(lengths, world, rank) if world > 1 else np.arange(n_all)
    # pack this rank's shard
    if world == 1:
        sr, so, sl = residues, offsets, lengths
    else:
        sr, so, sl = sharding.pack_shard(residues, offsets, lengths, mine)
    t_build = time.perf_counter()
    store = cuda.DeviceStore(sr, so, np.ascontiguousarray(sl), 24, device=local)
    t_build = time.perf_counter() - t_build  # one-off: sort, pair, interleave on the host, upload
    aligner = pyopal.Aligner(MATRIX, GAP_OPEN, GAP_EXTEND)
    M = aligner._int_matrix
    algo = {"nw": 0, "hw": 1, "ov": 2, "sw": 3}[args.algorithm]
    qbytes = [q.tobytes() for q in queries]

    n_mine = len(mine)
    # a non-blocking stream of our own: on torch's legacy default stream the library's
    # high-priority side stream (longest targets) does not overlap the bulk kernel
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    gather = sharding.ScoreGather(mine, n_all, world, "cuda")
    d_scores = [gather.new_local() for _ in qbytes]
    gathered = [torch.empty(world * gather.n_pad, dtype=torch.int32, device="cuda") for _ in qbytes] if world > 1 else None
    final = [torch.zeros(n_all, dtype=torch.int32, device="cuda") for _ in qbytes] if world > 1 else None

    launches = [0]
    k16_us = [0.0]
    k16_n = [0]

    def step(collect=False, do_gather=True):
        for qi, q in enumerate(qbytes):
            store.search_scores_device(q, GAP_OPEN, GAP_EXTEND, M, algo, d_scores[qi].data_ptr(), stream.cuda_stream)
            # one query = one API call: wait for it like Aligner.align does before the next one is issued
            stream.synchronize()
            if collect:
                st = store.stats()
                launches[0] += st["launches"]
                k16_us[0] += st["k16_us"]
                k16_n[0] += 1
                step.last = st
            if world > 1 and do_gather:
                if do_gather == "ag":
                    dist.all_gather_into_tensor(gathered[qi], d_scores[qi])
                elif do_gather == "idx":
                    final[qi][gather.dest] = gathered[qi][gather.valid]
                elif do_gather == "sync":
                    gather.gather(d_scores[qi], gathered[qi], final[qi])
                    stream.synchronize()
                else:
                    gather.gather(d_scores[qi], gathered[qi], final[qi])
                    # the query is complete when its scores are gathered: wait for it like Aligner.align would
                    # (an un-synchronised NCCL all-gather followed directly by the next search's persistent
                    # kernels measured +21 ms per query on 2 B200s; synchronised it costs ~0.1 ms)
                    stream.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    if args.debug_timing:
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        for qi, q in enumerate(qbytes):
            barrier()
            t0 = time.perf_counter()
            ev[0].record()
            store.search_scores_device(q, GAP_OPEN, GAP_EXTEND, M, algo, d_scores[qi].data_ptr(), stream.cuda_stream)
            ev[1].record()
            stream.synchronize()
            t1 = time.perf_counter()
            if world > 1:
                gather.gather(d_scores[qi], gathered[qi], final[qi])
            ev[2].record()
            torch.cuda.synchronize()
            t2 = time.perf_counter()
            print(f"[rank {rank}] Lq={len(q)} search dev {ev[0].elapsed_time(ev[1]):.3f} ms wall {1e3*(t1-t0):.3f} ms; "
                  f"gather dev {ev[1].elapsed_time(ev[2]):.3f} ms wall {1e3*(t2-t1):.3f} ms", file=sys.stderr, flush=True)
    if args.debug_timing:
        for g in (True, False, "ag", "idx", "sync", True):
            barrier()
            t0 = time.perf_counter()
            step(do_gather=g)
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            barrier()
            print(f"[rank {rank}] step gather={g}: {1e3*(t1-t0):.2f} ms (incl. barrier {1e3*(time.perf_counter()-t0):.2f})", file=sys.stderr, flush=True)
    props = torch.cuda.get_device_properties(local)
    sampler = ClockSampler("%08x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id))
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.finish()
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    cells_step = sum(QUERY_LENGTHS) * total_res
    value = cells_step * args.steps / (ms * 1e-3) / 1e9

    # per-kernel pass (outside the timed region): CUDA events around the 16-bit kernel of each search
    step(collect=True)
    barrier()
    per_q = []
    for qi, q in enumerate(qbytes):
        store.search_scores_device(q, GAP_OPEN, GAP_EXTEND, M, algo, d_scores[qi].data_ptr(), stream.cuda_stream)
        stream.synchronize()
        st = store.stats()
        per_q.append({"Lq": len(q), "G": st["G"], "K": st["K"], "passes": st["passes"], "k16_us": st["k16_us"],
                      "k32_us": st["k32_us"], "rescore_us": st["rescore_us"], "escalated": None,
                      "gcups_k16": round(len(q) * int(sl.sum()) / max(st["k16_us"], 1) / 1e3, 1)})

    # end-to-end: the public C-ABI call with HOST buffers (query H2D + scores D2H inside the timed region)
    e2e_value = None
    h2d = d2h = 0
    if rank == 0 or world > 1:
        barrier()
        e2e_q = [0.0] * len(qbytes)
        e2e_dev = [0.0] * len(qbytes)
        for q in qbytes:  # one untimed pass: the host-buffer path allocates its result row on first use
            store.search(q, GAP_OPEN, GAP_EXTEND, M, 0, algo)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            for qi, q in enumerate(qbytes):
                tq = time.perf_counter()
                store.search(q, GAP_OPEN, GAP_EXTEND, M, 0, algo)
                e2e_q[qi] += time.perf_counter() - tq
                e2e_dev[qi] += store.stats()["device_us"] * 1e-6
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        for qi in range(len(qbytes)):
            per_q[qi]["e2e_wall_ms"] = round(1e3 * e2e_q[qi] / args.steps, 3)
            per_q[qi]["e2e_device_ms"] = round(1e3 * e2e_dev[qi] / args.steps, 3)
        tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_value = cells_step * args.steps / float(tt.item()) / 1e9
        h2d = sum(len(q) for q in qbytes) + len(qbytes) * (24 * 24 * 4 + 64)
        d2h = len(qbytes) * n_mine * 4

    if rank == 0:
        # integer-ALU roofline: measured ALU-pipe rate of the kernel's own cell mix
        rate = None
        try:
            import ctypes
            r, c = ctypes.c_double(0), ctypes.c_double(0)
            if cuda.load_library().pyopal_cuda_int_issue_rate(local, ctypes.byref(r), ctypes.byref(c)) == 0:
                rate = r.value
        except Exception:
            rate = None
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        # dominant kernel = the 16-bit search kernel; weight by time over the five queries
        tot_us = sum(p["k16_us"] for p in per_q) or 1
        cells_k = sum(p["Lq"] for p in per_q) * int(sl.sum())
        gcups_k = cells_k / tot_us / 1e3
        bytes_k = sum(p["passes"] * int(store.device_bytes()) for p in per_q)
        roof_gcups = rate * 2 / ALU_INSTR_PER_CELL_PAIR / 1e9 if rate else None
        # DRAM traffic of the bulk kernel from the committed ncu --set full capture (config 2, one GPU):
        # the launch of the longest query captured, next to its algorithmic bytes
        traffic = traffic_note = None
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "r1_traffic.json")))["launches"]
            if args.workload == "config2" and world == 1 and args.targets == 570_000:
                t = max(tr, key=lambda x: x["query_length"])
                pq = [p for p in per_q if p["Lq"] == t["query_length"]][0]
                algo_bytes = pq["passes"] * int(store.device_bytes())
                traffic = t["dram_bytes_per_launch"]
                traffic_note = (f"bulk launch of the {t['query_length']}-aa query: {traffic} B measured by ncu vs {algo_bytes} B "
                                f"algorithmic (residue stream x {pq['passes']} passes); the excess is the pass-boundary rows "
                                f"(16 B per column per extra pass) that spill out of L2")
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u16x2 (s32 rescore)", "data": "synthetic",
            "config": config_dict(args, db, {"parallelism": f"db sharded by cells over {world} GPU(s), NCCL all-gather of scores",
                                             "store_build_s": round(t_build, 3)}),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "GCUPS", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "pyopal_cuda_search (host query in, host scores out), wall clock"},
            # the binding roof of max-plus DP is the integer ALU pipe (64 lanes/clk/SM), neither HBM nor tensor
            # cores: achieved = algorithmic cell updates of the 16-bit kernel / its CUDA-event
            # time; peak = ALU-pipe instruction rate measured in this run (VIADDMNMX.U16x2 stream) x 2 cells / 5.5
            "roofline": {"bound": "int_alu", "achieved": gcups_k, "peak": roof_gcups, "unit": "GCUPS",
                         "frac": (gcups_k / roof_gcups) if roof_gcups else None, "traffic": traffic,
                         "alu_thread_instr_per_s": rate, "alu_instr_per_cell_pair": ALU_INSTR_PER_CELL_PAIR,
                         "peak_source": "pyopal_cuda_int_issue_rate (measured on this GPU in this run)"},
            "roofline_hbm": {"bound": "hbm", "achieved": bytes_k / tot_us / 1e3, "peak": hbm_peak, "unit": "GB/s",
                             "frac": bytes_k / tot_us / 1e3 / hbm_peak, "traffic": traffic, "traffic_note": traffic_note,
                             "note": "residue stream (2 B per pair column per pass) of the 16-bit kernel; "
                                     "1/Lq byte per cell, so HBM is not the binding roof",
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback"},
            "kernels": per_q,
        }
        # launches inside the timed region: counted from the library's own counter
        line["gpu_launches"] = int(launches[0]) * args.steps
        if not args.no_cpu_baseline and world == 1:
            v, sample = cpu_sample(db, queries, args.algorithm, args.cpu_seconds, os.cpu_count() or 1)
            line["cpu_baseline"] = {"value": v, "unit": "GCUPS", "cores": os.cpu_count() or 1, "kind": "port",
                                    "sample": sample}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# --------------------------------------------------------------------------
# Config 5 of BASELINE.json: UniRef50-shaped database (5e7 proteins at 8 GPUs), batched multi-query
# search, block-dealt shards, NCCL gather of every query's scores.  One STEP = one batch of --queries
# queries.  Per-GPU work is fixed (--blocks-per-gpu blocks), so this workload scales WEAKLY: N GPUs
# search N x 6.25e6 proteins; 8 GPUs = the full 5e7-protein C5.


def run_config5(args):
    import numpy as np
    import synthetic

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    queries = synthetic.config5_queries(args.queries)
    qbytes = [np.ascontiguousarray(q).tobytes() for q in queries]
    sum_lq = sum(len(q) for q in qbytes)
    n_blocks = args.blocks_per_gpu * max(1, world)

    def cfg(total_res, extra=None):
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

    if args.impl == "reference":
        if rank != 0:
            return
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
        total_res = None
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / max(1, args.steps),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "s16 SIMD lanes (s64 rescore)",
            "data": "synthetic", "config": cfg(int(lens.sum()) * n_blocks, {"note": "residues extrapolated from block 0"}),
            "cpu_baseline": {"value": value, "unit": "GCUPS", "cores": threads, "kind": "port", "sample": sample + " of block 0"},
            "e2e": {"value": value, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)
        return

    import torch
    import torch.distributed as dist

    import pyopal_b200 as pyopal
    from pyopal_b200 import sharding
    from pyopal_b200.platform import cuda

    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    t0 = time.perf_counter()
    sr, so, sl, mine = synthetic.config5_shard(rank, world, n_blocks, args.block)
    t_gen = time.perf_counter() - t0
    res_mine = int(sl.sum())
    t0 = time.perf_counter()
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    # the repo's multi-GPU API: this rank's blocks on its GPU + the block gather (pyopal_b200/sharding.py)
    sdb = sharding.ShardedDatabase.from_blocks((sr, so, sl), args.block, args.blocks_per_gpu, 24, device=local, stream=stream)
    store = sdb.store
    t_build = time.perf_counter() - t0
    del sr
    M = pyopal.Aligner(MATRIX, GAP_OPEN, GAP_EXTEND)._int_matrix
    final = torch.empty(sdb.n_all, dtype=torch.int32, device="cuda") if world > 1 else None
    k = min(args.topk, sdb.n_all)
    top_v = torch.empty((len(qbytes), k), dtype=torch.int32, device="cuda")
    top_i = torch.empty((len(qbytes), k), dtype=torch.int32, device="cuda")
    host_v = torch.empty((len(qbytes), k), dtype=torch.int32).pin_memory()
    host_i = torch.empty((len(qbytes), k), dtype=torch.int32).pin_memory()
    tot = torch.tensor([float(res_mine)], dtype=torch.float64, device="cuda")
    mx = tot.clone()
    if world > 1:
        dist.all_reduce(tot)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    total_res = int(tot.item())
    imbalance = float(mx.item()) * world / total_res - 1.0

    acc = {"k16_us": 0.0, "launches": 0, "passes_cols": 0.0}

    def batch(to_host=False, collect=False):
        for qi, q in enumerate(qbytes):
            # search + all-gather + database order, complete on return (like Aligner.align)
            scores = sdb.search_scores(q, GAP_OPEN, GAP_EXTEND, M, 3, out=final)
            if collect:
                st = store.stats()
                acc["k16_us"] += st["k16_us"]
                acc["launches"] += st["launches"]
                acc["passes_cols"] += st["passes"]
            # exact device-side selection of the k best (csrc/topk.cu), unsorted; the host orders the k hits
            store.topk_device(scores.data_ptr(), sdb.n_all, k, top_v[qi].data_ptr(), top_i[qi].data_ptr(), stream.cuda_stream)
            if collect:
                acc["launches"] += 9
            if to_host:
                host_v[qi].copy_(top_v[qi], non_blocking=True)
                host_i[qi].copy_(top_i[qi], non_blocking=True)
        stream.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        batch()
    barrier()
    props = torch.cuda.get_device_properties(local)
    sampler = ClockSampler("%08x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id))
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        batch(collect=True)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.finish()
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    cells_step = sum_lq * total_res
    value = cells_step * args.steps / (ms * 1e-3) / 1e9
    # end to end: host queries in (H2D inside the C-ABI call), top-k hits out to pinned host memory
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        batch(to_host=True)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_value = cells_step * args.steps / float(dt.item()) / 1e9
    checksum = int(host_v.to(torch.int64).sum().item())

    if rank == 0:
        import ctypes
        rate = None
        r, c = ctypes.c_double(0), ctypes.c_double(0)
        if cuda.load_library().pyopal_cuda_int_issue_rate(local, ctypes.byref(r), ctypes.byref(c)) == 0:
            rate = r.value
        roof = rate * 2 / ALU_INSTR_PER_CELL_PAIR / 1e9 if rate else None
        gcups_k = sum_lq * res_mine * args.steps / max(acc["k16_us"], 1) / 1e3
        try:
            hbm_peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0)
        except Exception:
            hbm_peak = 6650.0
        bytes_k = acc["passes_cols"] * int(store.device_bytes())
        line = {
            "metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u16x2 (s32 rescore)", "data": "synthetic",
            "config": cfg(total_res, {"parallelism": f"blocks dealt round-robin over {world} GPU(s), NCCL all-gather of every "
                                                     f"query's scores, top-{k} on the device",
                                      "generate_s": round(t_gen, 2), "store_build_s": round(t_build, 2),
                                      "store_bytes_per_gpu": int(store.device_bytes()),
                                      "shard_residue_imbalance": round(imbalance, 5)}),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "GCUPS", "h2d_bytes_per_step": sum_lq + len(qbytes) * (24 * 24 * 4 + 64),
                    "d2h_bytes_per_step": len(qbytes) * k * 8,
                    "api": "sharding.ShardedDatabase.search_scores (host query in, NCCL gather) + device top-k, hits to pinned host; wall clock"},
            "roofline": {"bound": "int_alu", "achieved": gcups_k, "peak": roof, "unit": "GCUPS",
                         "frac": (gcups_k / roof) if roof else None, "traffic": None,
                         "alu_thread_instr_per_s": rate, "alu_instr_per_cell_pair": ALU_INSTR_PER_CELL_PAIR,
                         "peak_source": "pyopal_cuda_int_issue_rate (measured on this GPU in this run)"},
            "roofline_hbm": {"bound": "hbm", "achieved": bytes_k / max(acc["k16_us"], 1) / 1e3, "peak": hbm_peak, "unit": "GB/s",
                             "frac": bytes_k / max(acc["k16_us"], 1) / 1e3 / hbm_peak, "traffic": None,
                             "note": "residue stream once per pass of the 16-bit kernel on rank 0"},
            "top_hits_checksum": checksum,
            "gpu_launches": int(acc["launches"]),
        }
        if not args.no_cpu_baseline and world == 1:
            bres, blens = synthetic.config5_block(0, args.block)
            v, sample = cpu_sample(synthetic.pack(blens, bres), queries, "sw", args.cpu_seconds, os.cpu_count() or 1)
            line["cpu_baseline"] = {"value": v, "unit": "GCUPS", "cores": os.cpu_count() or 1, "kind": "port",
                                    "sample": sample + " of block 0"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse()
    if args.workload == "config5":
        run_config5(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
