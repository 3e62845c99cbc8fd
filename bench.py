#!/usr/bin/env python3
"""PairHMM GCUPS benchmark (BASELINE.json metric) for the B200-native engine and for the reference CPU arm.

  python bench.py --gpus N --steps K --warmup W             # this repo's CUDA engine
  python bench.py --impl reference --gpus N --steps K ...    # the reference's own AVX PairHMM on the host cores

A step = one pass of the hot path (FP32 forward -> <1e-28 -> FP64 rescue -> log10) over one batch of the
synthetic WES-like workload (BASELINE.json configs[1]: 10k regions, ~200 reads x 4-16 haplotypes, 150 bp reads,
250-500 bp haplotypes, ~2e7 pairs, ~1.1e12 DP cells). With N > 1 (torchrun, one rank per GPU) every rank owns
its own shard of that size: pairs are independent, there is no collective on the data path (weak scaling);
torch.distributed is used only for the barrier and the max-over-ranks of the device time.

  value    : GCUPS with the batch resident in HBM (CUDA events around the kernels of every step)
  e2e      : GCUPS through the C ABI from host arrays to host results (pack + H2D + kernels + D2H + scatter)
  roofline : FP32 forward kernel against the FP32 FMA-pipe peak (8 pipe lane-ops per cell)
  cpu_baseline : oracle/_ref (the reference's own sources compiled unmodified) on a bounded sample
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

METRIC = "PairHMM GCUPS (DP cell updates per second, float pass + double rescue)"
FP32_LANES_PER_SM = 128
PIPE_OPS_PER_CELL = 8      # 4 FMUL + 4 FFMA (computeMXY contracted), SURVEY.md 8(d)
FLOP_PER_CELL = 12


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--regions", type=int, default=0, help="c2: regions per rank (default 10000 = BASELINE configs[1]); "
                                                            "c5: regions of the whole job (default 25000 = configs[4])")
    ap.add_argument("--workload", default="c2", choices=["c2", "c5", "snv", "c1", "c3", "c4"],
                    help="c2 = BASELINE configs[1] per GPU (weak scaling, the metric's config); c5 = configs[4], the "
                         "whole-box chr20-scale batch LPT-sharded over the ranks (strong scaling); snv = the workload of "
                         "the reference's trie path (same-length haplotypes that differ by SNVs), engine with prefix "
                         "sharing on: value / e2e are EFFECTIVE GCUPS (R x H of every pair), the roofline counts only "
                         "the cells actually swept (SURVEY.md 8f N4); c1 / c3 / c4 = BASELINE configs[0] / [2] / [3] (one region; "
                         "one rescue-heavy region; long haplotypes), every rank the same batch")
    ap.add_argument("--cpu-sample-regions", type=int, default=0, help="regions in the CPU sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--strong-regions", type=int, default=25000,
                    help="regions of the strong-scaling job in the `strong` block (BASELINE configs[4] = 25000); 0 = no strong block")
    ap.add_argument("--strong-steps", type=int, default=2)
    return ap.parse_args()


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.rows = []
        self.stop_flag = False
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                self.rows.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def stop(self):
        self.stop_flag = True
        if self.proc:
            self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def cpu_reference_run(batch, threads, want_seconds=None):
    """times oracle/_ref (kind 'reference') or the C port (kind 'port') on `batch`; returns dict."""
    import oracle_py
    if oracle_py.reference_available():
        lib, kind = oracle_py.load_reference(), "reference"
        isa = "AVX-512" if lib.isa() == 512 else "AVX-256"
    else:
        lib, kind, isa = oracle_py.load_oracle(), "port", "scalar C"
    flat = batch.as_flat()
    out, _, secs = lib.compute(flat, threads=threads, tables_in_timed_region=True, want_rescued=False)
    cells = batch.cells()
    return {"value": cells / secs / 1e9, "unit": "GCUPS", "cores": threads, "kind": kind, "isa": isa,
            "seconds": secs, "cells": cells, "log10": out, "lib": lib, "flat": flat,
            "sample": "%d regions of the same workload (%d pairs, %.3g cells), per-read tables built inside the timed region"
                      % (batch.n_regions, batch.n_pairs, cells)}


def parity_gate(batch, n_reg, out, rs, cpu, threads):
    """the correctness gate of this very run. The engine's end-to-end results on the CPU sample's pairs
    (the first n_reg regions of the batch) against the CPU reference's: |dlog10| over every pair, the rescue set over
    EVERY pair of the sample, and the threshold-ambiguous pairs -- those on which the reference built with its default
    -ffp-contract (FMA in the AVX-512 objects: the canonical build) and built -ffp-contract=off disagree about
    `float sum < 1e-28f` (MIN_ACCEPTED, pairhmm_common.h:41); for each, which side the GPU took."""
    import oracle_py
    n_s = cpu["log10"].shape[0]
    sub_flat = cpu["flat"]
    _, want_resc, _ = cpu["lib"].compute(sub_flat, threads=threads)
    fin = np.isfinite(cpu["log10"])
    d = np.abs(out[:n_s][fin] - cpu["log10"][fin])
    parity = {"pairs": int(n_s), "max_abs_dlog10": float(d.max()) if d.size else 0.0,
              "mean_abs_dlog10": float(d.mean()) if d.size else 0.0, "tolerance": 1e-4,
              "nonfinite_pattern_identical": bool(np.array_equal(np.isfinite(out[:n_s]), fin)),
              "rescue_set_identical": bool(np.array_equal(rs[:n_s], want_resc)),
              "rescue_set_pairs_checked": int(n_s), "rescued_pairs": int(want_resc.sum()), "against": cpu["kind"]}
    if cpu["kind"] == "reference" and oracle_py.reference_nofma_available():
        _, resc_nofma, _ = oracle_py.load_reference_nofma().compute(sub_flat, threads=threads)
        amb = np.nonzero(want_resc != resc_nofma)[0]
        parity["threshold_ambiguous"] = {"pairs": int(amb.size),
                                         "gpu_sides_with_canonical_build": int((rs[:n_s][amb] == want_resc[amb]).sum()),
                                         "definition": "reference default build (FMA-contracted) vs -ffp-contract=off disagree on float sum < 1e-28f"}
    return parity


def strong_block(args, world, peaks_sm_mhz):
    """BASELINE configs[4] as north_star states it: the WHOLE chr20-scale job (regions x 200 reads x 10 haplotypes) handed to
    the library in ONE call on ONE multi-device handle (config.device_mask over the `world` GPUs of the box); the library
    LPT-splits the regions over the devices by DP cells, every device's results land in the caller's one host matrix. One
    process, T host threads -- Mutect2Cpp's own execution model (M2/main.cpp:199,570-574). Run by rank 0 alone; under
    torchrun the other ranks wait on a CPU (gloo) barrier meanwhile."""
    import hashlib
    import fgdp_b200
    from fgdp_b200 import synth
    t0 = time.perf_counter()
    job = synth.generate_parallel("c5", args.strong_regions)
    gen_s = time.perf_counter() - t0
    cells, pairs = job.cells(), job.n_pairs
    eng = fgdp_b200.Engine(device_mask=(1 << world) - 1)
    eng.initialize(int(job.read_len.max()), int(job.hap_len.max()))
    out = np.zeros(pairs, np.float64); rs = np.zeros(pairs, np.uint8)
    eng.compute_batch(job, out, rs)                               # warm-up: buffers, pinned arenas
    eng.stats_reset()
    t0 = time.perf_counter()
    for _ in range(args.strong_steps):
        eng.compute_batch(job, out, rs)
    e2e_s = (time.perf_counter() - t0) / args.strong_steps
    st = eng.stats()
    # device-resident replay of the same split: kernels only, max over the devices (they run side by side)
    eng.resident_load(job)
    eng.resident_run()
    ms = 0.0
    for _ in range(args.strong_steps):
        a, _ = eng.resident_run()
        ms += a
    ms /= args.strong_steps
    got, gresc = eng.resident_fetch()
    same = bool(np.array_equal(got, out) and np.array_equal(gresc, rs))
    eng.close()
    # identity with a single-device handle on the first regions of the job (the whole job when it is small)
    n_chk = min(job.n_regions, 1500)
    sub = job.subset(range(n_chk))
    one = fgdp_b200.Engine(device=0)
    o1, r1 = one.compute_batch(sub)
    one.close()
    n1 = sub.n_pairs
    return {"workload": "BASELINE configs[4]: chr20-scale synthetic batch, %d regions x 200 reads x 10 haps, 150 bp reads, "
                        "300-500 bp haps (%d pairs, %.4g cells): ONE pairhmm_cuda_compute_flat call on ONE multi-device handle, "
                        "regions LPT-split over %d GPU(s) inside the library, results gathered in one host matrix"
                        % (job.n_regions, pairs, cells, world),
            "scaling": "strong", "n_gpus": world, "steps": args.strong_steps,
            "value": cells / (ms * 1e-3) / 1e9, "ms_per_step": ms,
            "e2e": {"value": cells / e2e_s / 1e9, "unit": "GCUPS", "ms_per_step": e2e_s * 1e3,
                    "h2d_bytes_per_step": st["h2d_bytes"] // args.strong_steps, "d2h_bytes_per_step": st["d2h_bytes"] // args.strong_steps,
                    "host_pack_ms_per_step_max_over_devices": st["host_ms_pack"] / args.strong_steps,
                    "gpu_ms_per_step_max_over_devices": st["gpu_ms_total"] / args.strong_steps},
            "unit": "GCUPS", "rescued_frac": float(rs.mean()),
            "resident_run_identical_to_e2e": same,
            "identical_to_single_device_handle": bool(np.array_equal(out[:n1], o1) and np.array_equal(rs[:n1], r1)),
            "single_device_check_pairs": int(n1),
            "result_sha256": hashlib.sha256(out.tobytes() + rs.tobytes()).hexdigest()[:32],
            "host_threads": os.cpu_count(), "generate_s": gen_s}


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's CPU implementation on this box's host cores, bounded sample per step."""
    if rank != 0:
        return
    from fgdp_b200 import synth
    threads = os.cpu_count() or 1
    n_reg = args.cpu_sample_regions or 160       # ~1.8e10 cells: a few seconds per step on 16 AVX-512 threads
    sample = synth.config2(args.regions).subset(range(n_reg)) if args.regions <= 2000 else synth.config2(n_reg)
    for _ in range(max(0, args.warmup)):          # exactly the W warm-up passes the line reports
        cpu_reference_run(sample, threads)
    secs, res = 0.0, None
    for _ in range(args.steps):
        res = cpu_reference_run(sample, threads)
        secs += res["seconds"]
    value = sample.cells() * args.steps / secs / 1e9
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32+f64-rescue",
            "data": "synthetic",
            "config": {"workload": "BASELINE configs[1] synthetic WES-like batch (bounded CPU sample: %s)" % res["sample"],
                       "regions_per_rank": args.regions},
            "cpu_baseline": {"value": value, "unit": "GCUPS", "cores": threads, "kind": res["kind"], "isa": res["isa"],
                             "sample": res["sample"]},
            "e2e": {"value": value, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    if args.regions <= 0:
        args.regions = 25000 if args.workload == "c5" else 3000 if args.workload == "snv" else 10000
    rank, world, local = dist_env()
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import fgdp_b200
    from fgdp_b200 import synth
    # stdout carries exactly one JSON line: anything libraries print meanwhile (NCCL's version banner) goes to stderr
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        # control-plane barrier on the CPU: ranks that wait while rank 0 runs the single-process strong block must not sit
        # in a kernel on their GPU
        cpu_group = dist.new_group(backend="gloo")

    def barrier():
        if dist is not None:
            dist.barrier()

    def reduce_max(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def reduce_sum(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # weak scaling: the job is world x BASELINE configs[1]; its regions are split over the ranks by cost
    # (DP cells) with the LPT rule -- every rank derives the same split from the region shapes
    # and generates only its own regions. With one rank this is exactly configs[1].
    if args.workload == "c5":
        # strong scaling: the job is fixed (configs[4]), its regions are LPT-split over the ranks
        _, cost = synth.config5_shapes(args.regions)
        batch = synth.config5(args.regions, select=synth.lpt_shard(cost, world)[rank] if world > 1 else None)
        workload = ("BASELINE configs[4]: chr20-scale synthetic batch, %d regions x 200 reads x 10 haps, 150 bp reads, "
                    "300-500 bp haps, whole job LPT-sharded over %d GPU(s)" % (args.regions, world))
    elif args.workload in ("c1", "c3", "c4"):
        batch = {"c1": synth.config1, "c3": synth.config3, "c4": synth.config4}[args.workload]()
        args.regions = batch.n_regions
        workload = {"c1": "BASELINE configs[0]: one active region, 64 reads x 8 haplotypes, 150 bp reads, 300 bp haps (launch-bound: one step = one region)",
                    "c3": "BASELINE configs[2]: high-depth region, 2000 reads x 32 haplotypes, 150 bp reads, ~600 bp haps, low qualities (rescue-heavy)",
                    "c4": "BASELINE configs[3]: 64 regions x 128 reads x 8 haplotypes, 250 bp reads, 1-2 kb haplotypes of mixed lengths"}[args.workload]
    elif args.workload == "snv":
        batch = synth.snv_family(args.regions, 200, 12, 150, 300, dup_frac=0.0, seed_extra=rank)
        workload = ("SNV family (the reference's trie workload, VectorLoglessPairHMM.cpp:35): %d regions x 200 reads x 12 "
                    "haplotypes of 300 bases that differ by 1-3 SNVs, 150 bp reads, per GPU; prefix sharing on" % args.regions)
    elif world > 1:
        _, _, cost = synth.config2_shapes(args.regions * world)
        batch = synth.generate_parallel("c2", args.regions * world, procs=max(1, min(16, (os.cpu_count() or 8) // world)),
                                        select=synth.lpt_shard(cost, world)[rank])
    else:
        batch = synth.generate_parallel("c2", args.regions)
    if args.workload == "c2":
        workload = ("BASELINE configs[1]: synthetic WES-like batch, %d regions, ~200 reads x 4-16 haps, "
                    "150 bp reads, 250-500 bp haps, per GPU" % args.regions)
    cells, pairs = batch.cells(), batch.n_pairs
    # several ranks share the box's host threads: split them instead of 16 packing threads per rank
    per_rank = int(os.environ.get("BENCH_PACK_THREADS", "0")) or max(2, min(16, (os.cpu_count() or 16) // world))
    eng = fgdp_b200.Engine(device=local, pack_threads=0 if world == 1 else per_rank, prefix_sharing=args.workload == "snv")
    eng.initialize(int(batch.read_len.max()), int(batch.hap_len.max()))

    # ---- device-resident replay: `value` and the roofline of the FP32 forward kernel
    eng.resident_load(batch)
    for _ in range(args.warmup):
        eng.resident_run()
    sampler = ClockSampler(local); sampler.start()
    eng.stats_reset()
    barrier()
    ms_total = ms_float = 0.0
    for _ in range(args.steps):
        a, f = eng.resident_run()          # synchronises its stream; CUDA-event times of the step's kernels
        ms_total += a; ms_float += f
    barrier()
    launches = eng.stats()["kernel_launches"]
    swept_cells = eng.stats()["cells_swept"] // max(1, args.steps)      # unique reads x haplotype columns not shared
    got, resc = eng.resident_fetch()
    ms_step = reduce_max(ms_total / args.steps)
    total_cells = reduce_sum(float(cells))
    value = total_cells / (ms_step * 1e-3) / 1e9

    # ---- end to end through the C ABI: host arrays in, host results out, every step
    out = np.zeros(pairs, np.float64); rs = np.zeros(pairs, np.uint8)
    for _ in range(min(args.warmup, 2)):
        eng.compute_batch(batch, out, rs)
    eng.stats_reset()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        eng.compute_batch(batch, out, rs)
    e2e_s = reduce_max((time.perf_counter() - t0) / args.steps)
    barrier()
    st = eng.stats()
    # a step shorter than the 200 ms sampling tick (configs[0] is 0.08 ms) leaves no clock record: keep the same step
    # running, untimed, until the sampler has seen the device under this load
    clock_extra = 0
    t_clk = time.perf_counter()
    while len(sampler.rows) < 5 and time.perf_counter() - t_clk < 3.0:
        eng.compute_batch(batch, out, rs); clock_extra += 1
    sampler.stop()
    same = bool(np.array_equal(out, got) and np.array_equal(rs, resc))
    e2e_value = total_cells / e2e_s / 1e9
    eng.close()

    # ---- strong scaling as north_star states it: one job, one multi-device handle, one host matrix (rank 0 only)
    strong = None
    if args.workload == "c2" and args.strong_regions > 0:
        if rank == 0:
            try:
                strong = strong_block(args, world, None)
            except Exception as ex:      # noqa: BLE001 -- the headline line must still be printed
                strong = {"error": "%s: %s" % (type(ex).__name__, ex)}
        if dist is not None:
            dist.barrier(group=cpu_group)

    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    if rank != 0:
        return
    # ---- roofline: FP32 FMA pipe (compute bound); HBM traffic only to show it is not the limit
    import ctypes
    peaks = measured_peaks()
    sm_max_mhz = float(peaks.get("sm_max_mhz", 1965.0))
    n_sm = 148
    try:
        import torch
        n_sm = torch.cuda.get_device_properties(local).multi_processor_count
    except Exception:
        pass
    clocks = sampler.summary()
    peak_ops = n_sm * FP32_LANES_PER_SM * sm_max_mhz * 1e6                    # FMA-pipe lane-ops / s
    float_ms_step = ms_float / args.steps
    # the pipe works on the cells the kernel sweeps: with read de-duplication and prefix sharing fewer than R x H per pair
    pipe_cells = swept_cells if args.workload == "snv" else cells
    achieved_ops = pipe_cells * PIPE_OPS_PER_CELL / (float_ms_step * 1e-3)
    # 4 read planes (the uniform-gcp kernel never touches the gcp plane), one 4-byte column word per haplotype base,
    # one float per pair
    algo_bytes = int(batch.read_len.astype(np.int64).sum()) * 4 + int(batch.hap_len.astype(np.int64).sum()) * 4 + pairs * 4
    roofline = {"bound": "fp32_fma", "kernel": "forward2_kernel<W, K, KL> (packed FP32: two reads per lane, FFMA2/FMUL2)",
                "achieved": achieved_ops / 1e12, "peak": peak_ops / 1e12, "unit": "T FMA-pipe lane-ops/s",
                "frac": achieved_ops / peak_ops,
                "peak_source": "derived: %d SMs x 128 FP32 lanes x %.0f MHz (sm_max_mhz of MEASURED_PEAKS.json); "
                               "MEASURED_PEAKS.json has no FP32 entry" % (n_sm, sm_max_mhz),
                "frac_at_measured_clock": (achieved_ops / (n_sm * FP32_LANES_PER_SM * clocks["sm_mhz"] * 1e6)) if clocks.get("sm_mhz") else None,
                "tflops_flop_form": cells * FLOP_PER_CELL / (float_ms_step * 1e-3) / 1e12,
                "float_kernel_gcups": cells / (float_ms_step * 1e-3) / 1e9, "float_kernel_ms": float_ms_step,
                "algorithmic_bytes": algo_bytes, "algorithmic_hbm_GBps": algo_bytes / (float_ms_step * 1e-3) / 1e9,
                "hbm_peak_GBps_measured": peaks.get("hbm_gbs"), "traffic": None,
                # what the pipe gives to the kernel's own instruction mix in a microbenchmark (no dependencies, no loads):
                # FFMA2 (pair x uniform + pair) + FMUL2, 0.466 packed instr/clk/SMSP = 29.8 of the 32 lane-ops/clk/SMSP
                "peak_attainable_measured": peak_ops / 1e12 * (29.8 / 32.0),
                "peak_attainable_source": "profiles/r2_ubench_fma2_rate.txt (tools/ubench/fma2_rate.cu on B200 @1965 MHz): FFMA2c + FMUL2 mix "
                                          "29.8 of 32 FP32 lane-ops/clk/SMSP = 93 % of the derived peak",
                "frac_of_attainable": achieved_ops / (peak_ops * 29.8 / 32.0)}
    if args.workload == "snv":
        roofline["kernel"] = "forward2p_kernel<W, K, KL> (packed FP32 with prefix sharing between same-length haplotypes)"
        roofline["swept_frac"] = swept_cells / cells
        roofline["note"] = "achieved counts swept cells only; value / e2e / float_kernel_gcups are effective (R x H of every pair)"
    try:    # DRAM bytes of one launch of that kernel over this workload, from the committed ncu capture (profiles/)
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if tr.get("regions") == args.regions and world == 1 and args.workload == "c2":
            roofline["traffic"] = tr["dram_bytes_read"] + tr["dram_bytes_write"]
            roofline["traffic_source"] = tr["source"]
    except Exception:
        pass
    cpu = None
    if not args.no_cpu_baseline:
        n_reg = args.cpu_sample_regions or min(args.regions, 160)
        threads = os.cpu_count() or 1
        cpu = cpu_reference_run(batch.subset(range(n_reg)), threads)
        # the same sample with the per-read tables built before the timed region (the GPU's kernel-only
        # number has its counterpart), and the correctness gate of this very run: the engine's results on the sample's
        # pairs (the first regions of the batch) against the CPU reference's
        _, _, secs_pre = cpu["lib"].compute(cpu["flat"], threads=threads, tables_in_timed_region=False, want_rescued=False)
        parity = parity_gate(batch, n_reg, out, rs, cpu, threads)
        cpu["value_tables_prebuilt"] = cpu["cells"] / secs_pre / 1e9
        cpu = {k: cpu[k] for k in ("value", "value_tables_prebuilt", "unit", "cores", "kind", "isa", "sample")}
    in_bytes = batch.read_bases.nbytes * 5 + batch.hap_bases.nbytes * 4
    line = {"metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if args.workload == "c5" else "weak",
            "vs_baseline": None,
            "dtype": "f32+f64-rescue", "data": "synthetic",
            "config": {"workload": workload, "pairs_per_gpu": pairs, "cells_per_gpu": cells, "rescued_frac": float(resc.mean()),
                       "l2": ("inputs (%.3f GB of read planes + haplotype words per GPU) exceed the 126 MB L2" if in_bytes > 126e6 else
                              "inputs (%.3f GB) fit the 126 MB L2 and are not flushed between steps: the path is FMA-pipe-bound at "
                              "~0.002 B per cell (roofline.algorithmic_hbm_GBps), L2 residency does not change the step time") % (in_bytes / 1e9),
                       "parallelism": "%d rank(s), regions LPT-sharded by DP cells, no collective on the data path" % world},
            "e2e": {"value": e2e_value, "unit": "GCUPS", "h2d_bytes_per_step": st["h2d_bytes"] // args.steps,
                    "d2h_bytes_per_step": st["d2h_bytes"] // args.steps, "ms_per_step": e2e_s * 1e3,
                    "host_pack_ms_per_step": st["host_ms_pack"] / args.steps, "identical_to_resident_run": same},
            "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks}
    if cpu is not None:
        line["parity"] = parity
    if clock_extra:
        line["clocks"]["window"] = "timed region + %d further untimed passes of the same step (the timed steps are shorter than the 200 ms sampling tick)" % clock_extra
    if strong is not None:
        line["strong"] = strong
    print(json.dumps(line), flush=True)
    del ctypes


if __name__ == "__main__":
    main()
