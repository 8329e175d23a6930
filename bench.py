#!/usr/bin/env python
"""bench.py -- SV records audited/s of the audt breakpoint-refinement hot path on N B200s (one process per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A "step" is one pass of the hot path over one batch: every SV record of the workload (BASELINE.json configs[1]:
one 64 Mb contig, 30x 15 kb HiFi-like reads, 10k DEL/INS records => ~15k refine_* windows) is refined once.
At N > 1 every rank owns its own region shard of the same shape (a different contig, different seed): weak
scaling, no data-path collective -- torch.distributed is only used for the barrier and the
max-over-ranks of the device time.

  value   records/s with reads and windows already resident in HBM (svtk_refine_windows_device + svtk_sync),
          device time from CUDA events on the ctx stream, L2 flushed between timed steps
  e2e     records/s through the host-buffer C ABI: svtk_load_reads (pinned host SoA -> HBM) +
          svtk_refine_windows (host windows in, host results out) every step
  roofline  the fused CIGAR-walk + consensus kernel: algorithmic bytes (4 B per visited CIGAR word
          + 16 B per (window, read) pair + 4 B per signature + 28 B per window) / its CUDA-event launch duration
          measured live inside the timed region (svtk_get_timing), against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline  the CPU oracle port (oracle/svtk_oracle.c, 1 thread) on the same windows -- a checker timed
          beside the product, never on the product path
  --impl reference   the UNMODIFIED reference CLI (oracle/_ref/svtrek audt -t <all cores>) on a bounded sample of
          the same workload written out as BAM+VCF (falls back to the oracle port if oracle/_ref is missing)
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "SV records audited/s"
UNIT = "records/s"
SEED0 = 20261104


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c4", "c5"])
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the workload (debug only; the default is the full config)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--ref-sample-mb", type=float, default=3.0, help="--impl reference: contig length of the bounded sample")
    ap.add_argument("--ref-threads", type=int, default=0)
    ap.add_argument("--no-bam", action="store_true", help="skip the file-to-file leg (svtk_load_bam + svtk_process_vcf on the reference arm's sample)")
    ap.add_argument("--poa-sets", type=int, default=4000, help="INS read sets per GPU of the POA (K3) leg: a bounded sample of configs[3]'s 20k records; 0 skips the leg")
    ap.add_argument("--poa-steps", type=int, default=2)
    return ap.parse_args()


WORKLOAD_DESC = {
    "c1": "BASELINE.json configs[0]: synthetic 1 Mb contig, 500 DEL/INS, 30x 15 kb HiFi-like reads",
    "c2": "BASELINE.json configs[1]: synthetic chr20-sized 64 Mb contig, 30x 15 kb HiFi-like reads, 10k DEL/INS records",
    "c4": "BASELINE.json configs[3]: insertion-heavy, 20k INS at 60x ONT-like error profile",
    "c5": "BASELINE.json configs[4]: stress, 200x depth over tandem-repeat SV clusters",
}


def make_workload(name, rank, scale):
    from svtrek_b200 import simreads as S
    seed = SEED0 + sorted(S.CONFIGS).index(name) + 1000 * rank
    rs, truth = S.make_config(name, seed=seed, scale=scale)
    w = S.windows_from_truth(truth)
    n_records = int(np.unique(w["record"]).shape[0])      # records that print a line (DEL with end-pos <= 50 print nothing)
    return rs, truth, w, n_records


# ------------------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._pump, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.th.join(timeout=2)
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "power_w_max": float(max(power)),
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy)"
        except Exception:
            pass
    return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


class NvmlSampler:
    """Same readings as the nvidia-smi clocks line, taken through NVML every 5 ms: the timed region of this path
    lasts milliseconds, too short for `nvidia-smi -lms`.  Falls back to ClockSampler when pynvml is missing."""

    def __init__(self, index):
        self.index = index
        self.ok = False
        self.sm, self.reasons = [], set()
        self.smax = None
        self.power = []
        self._stop = threading.Event()
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            h = None
            uuid = None
            try:
                import torch
                uuid = "GPU-" + str(torch.cuda.get_device_properties(index).uuid)
                h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if hasattr(uuid, "encode") else uuid)
            except Exception:
                h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.h = h
            self.smax = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:
            self.fallback = ClockSampler(index)

    def start(self):
        if not self.ok:
            return self.fallback.start()
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()

    def _run(self):
        nv = self.nv
        bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for k, b in bits.items():
                    if r & b:
                        self.reasons.add(k)
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            except Exception:
                pass
            time.sleep(0.005)

    def stop(self):
        if not self.ok:
            return self.fallback.stop()
        self._stop.set()
        self.th.join(timeout=2)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.smax, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": self.smax, "power_w_max": float(max(self.power)) if self.power else None,
                "samples": len(self.sm), "reasons": sorted(self.reasons), "source": "nvml, 5 ms period, over the device-timed and e2e loops"}


def traffic_from_profiles():
    """dram bytes per launch of the dominant kernel, from the committed `ncu --set full` capture (profiles/*.json)."""
    best = None
    pdir = os.path.join(ROOT, "profiles")
    if os.path.isdir(pdir):
        for f in sorted(os.listdir(pdir)):
            if f.endswith("_traffic.json"):
                try:
                    best = json.load(open(os.path.join(pdir, f)))
                except Exception:
                    pass
    return best


def pinned_i32(arr):
    import torch
    a = np.ascontiguousarray(arr)
    if a.dtype.itemsize == 8:
        t = torch.from_numpy(a.view(np.int64))
    elif a.dtype.itemsize == 4:
        t = torch.from_numpy(a.view(np.int32))
    else:
        t = torch.from_numpy(a.view(np.uint8))
    return t.pin_memory()


def cpu_port_baseline(rs, w, n_records, budget_s=10.0):
    """The CPU oracle port (1 thread) on the same windows: checker timed beside the product."""
    from oracle import binding as B
    orc = B.OracleReads(rs)
    t0 = time.perf_counter()
    reps = 0
    while True:
        B.refine_batch(orc, w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
        reps += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or reps >= 50:
            break
    return {"value": n_records * reps / dt, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"oracle/svtk_oracle.c orc_refine_batch, 1 thread, the whole workload ({len(w['chrom'])} windows, {n_records} records) x {reps} passes in {dt:.2f} s; reads already decoded (no BGZF inflate)"}


# ------------------------------------------------------------------------------------------------------------
POA_DESC = "BASELINE.json configs[3] sample: INS read sets, inserted length log-uniform 50 bp - 10 kb, 60x depth x 0.8 carriers, 8 % ONT-like errors"


def poa_workload(rank, n_sets):
    from svtrek_b200 import simpoa as P
    sets = P.ins_workload(SEED0 + 77 + 1000 * rank, n_sets, depth=60, carrier=0.8, err=0.08, len_lo=50, len_hi=10000)
    return sets, P.pack(sets)


def poa_leg(a, rank, n_sets, steps):
    """K3: svtk_poa_consensus (host buffers in, consensus out) on this rank's read sets.  GCUPS = band cells the kernel
    itself counted (5-state convex cell updates) / CUDA-event time of k_poa_sets; e2e = the same cells / wall
    time of the C-ABI call (H2D of the reads, kernel, D2H of the consensus)."""
    sets, (so, qo, b) = poa_workload(rank, n_sets)
    a.poa_consensus(so, qo, b)                       # warm-up: scratch allocation, first-touch
    kms, wall, st = 0.0, 0.0, None
    for _ in range(steps):
        t0 = time.perf_counter()
        cons, st = a.poa_consensus(so, qo, b)
        wall += time.perf_counter() - t0
        kms += st["kernel_ms"]
    return sets, cons, dict(cells=st["cells"], cells_full=st["cells_full"], n_sets=n_sets, n_reads=int(len(qo) - 1), n_bases=int(len(b)),
                            kernel_ms=kms / steps, wall_ms=wall * 1e3 / steps, workers=st["workers"], relaunched=st["relaunched"],
                            phase_cycles=st["phase_cycles"])


def poa_cpu_baseline(sets, cons, budget_s=10.0):
    """abPOA 1.5.4 itself (oracle/_ref/libabpoa_ref.so, the reference's submodule compiled unmodified), 1 thread, on a
    bounded sample of the same read sets; every sampled consensus is also compared with the GPU's."""
    from oracle import binding as B
    if not B.have_abpoa():
        return None
    rng = np.random.default_rng(1)
    order = rng.permutation(len(sets))
    t0 = time.perf_counter()
    cells, n = 0, 0
    for i in order:
        r = B.ref_poa_steps(sets[i])
        assert np.array_equal(r["cons"], cons[i]), f"POA consensus of set {i} differs from abPOA"
        cells += int(r["cells"].sum()); n += 1
        if time.perf_counter() - t0 >= budget_s:
            break
    dt = time.perf_counter() - t0
    return {"value": cells / dt / 1e9, "unit": "GCUPS", "cores": 1, "kind": "reference", "sets_per_s": n / dt,
            "sample": f"abpoa_msa default path (abPOA 1.5.4 compiled from the reference's submodule, runtime SIMD dispatch), 1 thread, {n} randomly "
                      f"chosen read sets of the GPU leg's batch in {dt:.1f} s; cells = abPOA's own dp_beg/dp_end; consensus compared with the GPU's: identical"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from svtrek_b200 import api

    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path is CUDA only (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    K, W = args.steps, max(args.warmup, 3)

    rs, truth, w, n_records = make_workload(args.workload, rank, args.scale)
    n_win = int(w["chrom"].shape[0])
    stream = torch.cuda.Stream(device=dev)
    a = api.Auditor(local, cuda_stream=stream.cuda_stream)

    # host side: pinned SoA (the e2e leg's inputs) ; device side: resident copies (the kernel leg's inputs)
    h_tid = np.ascontiguousarray(rs.tid_off, dtype=np.int64)
    hp = {k: pinned_i32(np.ascontiguousarray(v)) for k, v in dict(pos=rs.pos.astype(np.uint32), end=rs.end.astype(np.uint32),
          n_cigar=rs.n_cigar.astype(np.uint32), cigar_off=rs.cigar_off.astype(np.uint64), cigar=rs.cigar.astype(np.uint32)).items()}
    hw = {k: pinned_i32(np.ascontiguousarray(w[k], dtype=(np.int32 if k == "chrom" else np.uint8 if k == "kind" else np.uint32)))
          for k in ("chrom", "start", "end", "pos", "kind")}
    h_res = torch.empty(n_win, dtype=torch.int32).pin_memory()
    n_reads, n_words = int(rs.pos.shape[0]), int(rs.cigar.shape[0])
    reads_bytes = 8 * h_tid.shape[0] + 20 * n_reads + 4 * n_words
    win_bytes = 17 * n_win

    def load_reads():
        a.load_reads_ptr(h_tid.shape[0] - 1, n_reads, n_words, h_tid.ctypes.data, hp["pos"].data_ptr(), hp["end"].data_ptr(),
                         hp["n_cigar"].data_ptr(), hp["cigar_off"].data_ptr(), hp["cigar"].data_ptr())

    load_reads()
    with torch.cuda.stream(stream):
        dw = {k: v.to(dev, non_blocking=True) for k, v in hw.items()}
        d_res = torch.empty(n_win, dtype=torch.int32, device=dev)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
    stream.synchronize()
    wargs = (n_win, dw["chrom"].data_ptr(), dw["start"].data_ptr(), dw["end"].data_ptr(), dw["pos"].data_ptr(), dw["kind"].data_ptr())

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # work counters (instrumented instantiations, outside the timed region): the product path's two kernels, and the
    # reference-shaped walker (one CIGAR walk per (window, read) pair) as an independent cross-check of the results
    walker = a.measure_windows(*wargs, d_res.data_ptr())
    walker_res = d_res.cpu().numpy().copy()
    stats = a.measure_index(*wargs, d_res.data_ptr())
    ref_res = d_res.cpu().numpy().copy()
    assert np.array_equal(walker_res, ref_res), "indexed path and per-pair walker disagree"

    def step_device():
        a.index_reads()                      # the whole job every step: one pass over every CIGAR word ...
        a.refine_windows_device(*wargs, d_res.data_ptr())   # ... then every window (asynchronous on the ctx stream)

    with torch.cuda.stream(stream):
        for _ in range(W):
            flush.zero_()
            step_device()
            a.sync()
    overflow_work = int(stats["n_overflow"]) != 0 or int(walker["n_overflow"]) != 0
    sampler = NvmlSampler(local)
    barrier()
    sampler.start()
    # per-kernel CUDA events live inside the timed region, on every prof_stride-th step: an event pair serialises the stream
    # around its kernel (B200, C2: a step is 92.8 us with both pairs on every step, 77.5 us without any -- tools/step_overhead.py)
    prof_stride = max(1, min(4, K // 4))
    a.set_profiling(prof_stride)
    launches0 = a.launch_count
    evs = []
    t_wall0 = time.perf_counter()
    with torch.cuda.stream(stream):
        for _ in range(K):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            step_device()
            if overflow_work:
                a.sync()                     # windows beyond the fast path are finished by svtk_sync: that work belongs to the step
            e1.record(stream)                # closes the step's device work; the host-side wait comes after it
            a.sync()
            evs.append((e0, e1))
    barrier()
    t_wall = time.perf_counter() - t_wall0
    dev_ms = float(sum(e0.elapsed_time(e1) for e0, e1 in evs))
    launches = a.launch_count - launches0
    tim = a.timing()
    a.set_profiling(False)
    assert np.array_equal(d_res.cpu().numpy(), ref_res), "timed steps changed the result"

    # ---- e2e: host buffers through the C ABI, H2D of reads + windows and D2H of the results inside the timed region
    for _ in range(3):
        load_reads(); a.refine_windows_ptr(n_win, *[hw[k].data_ptr() for k in ("chrom", "start", "end", "pos", "kind")], h_res.data_ptr())
    barrier()
    e2e_s = 0.0
    for _ in range(K):
        with torch.cuda.stream(stream):
            flush.zero_()
        stream.synchronize()
        t0 = time.perf_counter()
        load_reads()
        a.refine_windows_ptr(n_win, *[hw[k].data_ptr() for k in ("chrom", "start", "end", "pos", "kind")], h_res.data_ptr())
        e2e_s += time.perf_counter() - t0
    barrier()
    clocks = sampler.stop()
    assert np.array_equal(h_res.numpy(), ref_res), "e2e result differs from the device-resident result"

    # ---- file-to-file leg on its own context, like the C driver's process would have (before the POA leg claims most of HBM)
    bam_leg = None
    if world == 1 and not args.no_bam:
        with api.Auditor(local) as a_bam:
            bam_leg = bam_e2e_leg(a_bam, args)

    # ---- K3: POA consensus leg (GCUPS half of the metric)
    poa = None
    if args.poa_sets > 0:
        poa_sets_l, poa_cons, poa = poa_leg(a, rank, args.poa_sets, max(1, args.poa_steps))
        pr = torch.tensor([poa["kernel_ms"], poa["wall_ms"]], dtype=torch.float64, device=dev)
        pt = torch.tensor([float(poa["cells"]), float(poa["cells_full"]), float(poa["n_sets"])], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(pr, op=dist.ReduceOp.MAX)
            dist.all_reduce(pt, op=dist.ReduceOp.SUM)
        poa_kms, poa_wall = [float(x) for x in pr.cpu()]
        poa_cells, poa_full, poa_nsets = [float(x) for x in pt.cpu()]

    # ---- max over ranks, whole-job aggregate
    red = torch.tensor([dev_ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(n_records), float(launches), float(reads_bytes + win_bytes), float(4 * n_win)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    dev_ms_max, e2e_ms_max = [float(x) for x in red.cpu()]
    rec_all, launches_all, h2d_all, d2h_all = [float(x) for x in tot.cpu()]

    if rank == 0:
        peak, peak_src = measured_peak()
        idx_ms = tim["index_kernel_ms"] / max(1, tim["index_kernel_launches"])
        ref_ms = tim["refine_kernel_ms"] / max(1, tim["refine_kernel_launches"])
        kern = {"k_index_reads": (idx_ms, stats["index_bytes"]), "k_refine_indexed": (ref_ms, stats["refine_bytes"])}
        dom = max(kern, key=lambda k: kern[k][0])
        oth = [k for k in kern if k != dom][0]
        k_ms = kern[dom][0]
        achieved = kern[dom][1] / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
        tr = traffic_from_profiles()
        out = {
            "metric": METRIC, "value": rec_all * K / (dev_ms_max * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dev_ms_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32",
            "data": "synthetic", "impl": "ours",
            "config": {"workload": WORKLOAD_DESC[args.workload] + (f" (scale {args.scale})" if args.scale != 1.0 else ""),
                       "records_per_gpu": n_records, "windows_per_gpu": n_win, "reads_per_gpu": n_reads, "cigar_words_per_gpu": n_words,
                       "sharding": "one region shard (own contig) per GPU, no collective" if world > 1 else "single GPU",
                       "l2": "256 MB buffer written between timed steps (L2 flush); CIGAR array is %.0f MB" % (4 * n_words / 1e6),
                       "kernel_timing": "CUDA event pairs around k_index_reads and k_refine_warp on every %d%s timed step (%d and %d launches timed); "
                                        "all %d steps are inside the step events that give `value`" % (
                                            prof_stride, "th" if prof_stride > 3 else ("st" if prof_stride == 1 else "nd" if prof_stride == 2 else "rd"),
                                            tim["index_kernel_launches"], tim["refine_kernel_launches"], K),
                       "params": "wider 20000, median 10000, narrow 2000, range 500, interval 5, min-count 3 (params.h defaults)"},
            "clocks": clocks,
            "e2e": {"value": rec_all * K / (e2e_ms_max * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d_all), "d2h_bytes_per_step": int(d2h_all),
                    "ms_per_step": e2e_ms_max / K,
                    "path": "svtk_load_reads(pinned host SoA) + svtk_refine_windows(host windows -> host results), wall clock, every step"},
            "gpu_launches": int(launches_all),
            "kernel_share": {"index_kernel_ms_per_step": idx_ms, "refine_kernel_ms_per_step": ref_ms, "step_ms": dev_ms / K,
                             "share": (idx_ms + ref_ms) / (dev_ms / K) if dev_ms else None, "wall_ms_per_step_incl_flush": t_wall * 1e3 / K},
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": (tr or {}).get(dom), "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": kern[dom][1], "kernel_ms": k_ms,
                         "bytes_rule": "k_index_reads: 4 B/CIGAR word + 24 B/read + 16 B/event + 16 B/chunk; "
                                       "k_refine_indexed: 16 B/event scanned + 16 B/segment + 28 B/window (DESIGN.md)",
                         "work": stats},
            "roofline_other": {"kernel": oth, "achieved": kern[oth][1] / (kern[oth][0] * 1e-3) / 1e9 if kern[oth][0] > 0 else 0.0, "unit": "GB/s",
                               "peak": peak, "frac": (kern[oth][1] / (kern[oth][0] * 1e-3) / 1e9 / peak) if kern[oth][0] > 0 and peak else None,
                               "note": "the step's two kernels take about the same time; `roofline` is whichever ran longer in this run",
                               "algorithmic_bytes_per_launch": kern[oth][1], "kernel_ms": kern[oth][0], "traffic": (tr or {}).get(oth)},
            "reference_equivalent": {"note": "bytes the reference's per-(window, read) walks touch (SURVEY.md 8d) / time of both kernels",
                                     "bytes_per_step": walker["algorithmic_bytes"], "GBps": walker["algorithmic_bytes"] / ((idx_ms + ref_ms) * 1e-3) / 1e9 if (idx_ms + ref_ms) > 0 else 0.0,
                                     "work": {k: walker[k] for k in ("n_windows", "n_pairs", "n_words_visited", "n_signatures", "n_overflow")}},
            "poa_gcups": None,
        }
        if poa is not None:
            cyc = [float(x) for x in poa["phase_cycles"]]
            out["poa_gcups"] = {
                "value": poa_cells / (poa_kms * 1e-3) / 1e9, "unit": "GCUPS (band cells: 5-state convex cell updates actually computed)",
                "full_matrix_equivalent_gcups": poa_full / (poa_kms * 1e-3) / 1e9,
                "e2e": {"value": poa_cells / (poa_wall * 1e-3) / 1e9, "unit": "GCUPS", "path": "svtk_poa_consensus: host read sets in, host consensus out, wall clock"},
                "read_sets_per_s": poa_nsets / (poa_kms * 1e-3), "kernel_ms_per_step": poa_kms, "steps": max(1, args.poa_steps),
                "workload": POA_DESC + f"; {poa['n_sets']} read sets / {poa['n_reads']} reads / {poa['n_bases'] / 1e6:.0f} Mbases per GPU",
                "kernel": "k_poa_sets (one 32-thread CTA per read set, persistent workers)", "resident_workers": poa["workers"],
                "relaunched_sets": poa["relaunched"],
                "phase_share": dict(zip(("row_dp", "backtrack", "graph_fusion", "topological_sort"), [c / max(sum(cyc), 1.0) for c in cyc])),
                "bound": "integer ALU / instruction latency (no tensor cores: not a contraction)"}
            if world == 1 and not args.no_cpu:
                out["poa_gcups"]["cpu_baseline"] = poa_cpu_baseline(poa_sets_l, poa_cons)
        if world == 1 and not args.no_bam:
            out["bam_e2e"] = bam_leg
            out["e2e"]["note"] = ("inputs are decoded reads in pinned host arrays (the C ABI's svtk_reads); the reference CLI also inflates and parses "
                                  "the BAM -- the file-to-file number on the reference arm's own sample is bam_e2e.value")
            out["e2e"]["file_to_file_records_per_s"] = out["bam_e2e"].get("value")
        if world == 1 and not args.no_cpu:
            out["cpu_baseline"] = cpu_port_baseline(rs, w, n_records)
        emit(out)
    a.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def reference_sample(args):
    """The bounded sample both arms' file-level legs run on: a contig of --ref-sample-mb Mb with the workload's coverage,
    read length, error rate and SV density (same seed => the same BAM + VCF on both arms)."""
    from svtrek_b200 import simreads as S
    cfg = dict(S.CONFIGS[args.workload])
    L = int(args.ref_sample_mb * 1e6)
    full_L = float(sum(cfg["contig_lens"]))
    cfg["n_sv"] = max(10, int(round(cfg["n_sv"] * L / full_L)))
    cfg["contig_lens"] = [L]
    seed = SEED0 + sorted(S.CONFIGS).index(args.workload)
    if args.workload == "c5":
        return S._make_c5(seed, cfg)
    return S.simulate(seed, **cfg)


def bam_e2e_leg(a, args, steps=5):
    """File to file, like the reference CLI: svtk_load_bam (BGZF inflate + record extraction on the GPU) + svtk_process_vcf
    (VCF in, the reference's stdout lines out) on the sample the reference arm times: same reads, same VCF; the BAM + BAI are
    written by svtrek_b200.simreads.write_bam (HTSlib's block layout, zlib level 6; the reference CLI prints the same lines
    from it as from the HTSlib-written file of the same reads -- tests/test_bam.py)."""
    from svtrek_b200 import simreads as S
    rs, truth = reference_sample(args)
    cores = args.ref_threads or os.cpu_count() or 1
    with tempfile.TemporaryDirectory(prefix="svtk_bam_") as d:
        bam, vcf, out = (os.path.join(d, x) for x in ("s.bam", "s.vcf", "out.txt"))
        S.write_bam(bam, rs, truth=truth)
        S.write_vcf(vcf, S.vcf_lines(truth, end_first=True), pad=3 * cores + 8)      # the same file the reference arm reads
        times, loads, n = [], [], 0
        for i in range(2 + steps):
            t0 = time.perf_counter()
            st = a.load_bam(bam)
            n = a.process_vcf(vcf, out)
            dt = time.perf_counter() - t0
            if i >= 2:
                times.append(dt); loads.append(st)
        got = a.copy_reads()
        same = bool(np.array_equal(got["pos"], rs.pos) and np.array_equal(got["end"], rs.end) and np.array_equal(got["cigar"], rs.cigar))
        # beyond the reference's output: inserted sequences of the refined INS breakpoints -> POA consensus
        w = S.windows_from_truth(truth)
        refined = a.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
        t0 = time.perf_counter()
        sets = a.ins_sequences(w["chrom"], w["start"], w["end"], w["pos"], w["kind"], refined, tol=5)
        t1 = time.perf_counter()
        cons, pst = a.poa_consensus(sets["set_off"], sets["seq_off"], sets["bases"]) if len(sets["window"]) else ([], {"cells": 0, "kernel_ms": 0.0})
        t2 = time.perf_counter()
        L = truth.length[w["record"][sets["window"]]] if len(sets["window"]) else np.zeros(0)
        near = int(sum(abs(len(c) - int(l)) <= max(2, int(l) // 50) for c, l in zip(cons, L)))
        ins = {"read_sets": int(len(sets["window"])), "sequences": int(len(sets["read"])), "bases": int(len(sets["bases"])),
               "extract_ms": (t1 - t0) * 1e3, "poa_ms": (t2 - t1) * 1e3, "poa_kernel_ms": float(pst["kernel_ms"]), "poa_cells": int(pst["cells"]),
               "consensus_within_2pct_of_planted_length": near,
               "path": "svtk_ins_sequences (SEQ of the I >= 50 ops within 5 bases of the refined position) + svtk_poa_consensus; not part of the records/s above (the reference prints no sequence)"}
        # the same files through the C driver as a process: process start -> last line (CUDA context creation included)
        cli = os.path.join(ROOT, "svtrek_b200", "svtrek_gpu")
        cli_ms = None
        if os.path.exists(cli):
            ts = []
            for _ in range(3):
                t0 = time.perf_counter()
                subprocess.run([cli, "audt", "-b", bam, "-v", vcf], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=True)
                ts.append((time.perf_counter() - t0) * 1e3)
            cli_ms = min(ts)
        tot = float(sum(times))
        best = min(loads, key=lambda r: r["total_ms"])
        return {"value": n * len(times) / tot, "svtrek_gpu_process_ms": cli_ms, "unit": UNIT, "ms_per_step": tot * 1e3 / len(times), "steps": len(times), "records_per_step": int(n),
                "path": "svtk_load_bam(BAM + BAI files) + svtk_process_vcf(VCF file -> output file), wall clock, every step; page cache warm",
                "sample": f"the reference arm's sample: {args.ref_sample_mb} Mb contig, {rs.n_reads} reads, {os.path.getsize(bam) / 1e6:.0f} MB BAM "
                          f"(random SEQ/QUAL)",
                "loaded_reads_identical_to_written": same, "ins_consensus": ins,
                "load_bam": {"file_MB": best["file_bytes"] / 1e6, "inflated_MB": best["inflated_bytes"] / 1e6, "bgzf_blocks": best["n_blocks"],
                             "host_scan_ms": best["host_scan_ms"], "upload_ms": best["upload_ms"], "inflate_kernel_ms": best["inflate_ms"],
                             "record_kernels_ms": best["parse_ms"], "total_ms": best["total_ms"],
                             "inflate_GBps": best["inflated_bytes"] / max(best["inflate_ms"], 1e-6) / 1e6}}


# ------------------------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference arm: the UNMODIFIED reference CLI on the host cores (oracle/_ref/svtrek, compiled from
    /root/reference by oracle/build_ref.sh; the built files travel with the repo snapshot).  Each step is one
    `svtrek audt -b sample.bam -v sample.vcf -t <cores>` run over a bounded sample of the workload: a contig of
    --ref-sample-mb Mb with the workload's coverage, read length, error rate and SV density."""
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
    if rank != 0:
        return
    from oracle import binding as B
    from svtrek_b200 import simreads as S
    K, W = args.steps, args.warmup
    cores = args.ref_threads or os.cpu_count() or 1
    rs, truth = reference_sample(args)
    w = S.windows_from_truth(truth)
    n_records = int(np.unique(w["record"]).shape[0])
    base = {"metric": METRIC, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic", "impl": "reference",
            "config": {"workload": WORKLOAD_DESC[args.workload], "sample": f"{args.ref_sample_mb} Mb contig at the workload's depth/read length/SV density: "
                       f"{rs.n_reads} reads, {n_records} records"}}
    if B.have_ref() and os.path.exists(os.path.join(B.REF_DIR, "test_view")):
        with tempfile.TemporaryDirectory(prefix="svtk_ref_") as d:
            sam, bam, vcf = os.path.join(d, "s.sam"), os.path.join(d, "s.bam"), os.path.join(d, "s.vcf")
            S.write_sam(sam, rs)
            B.make_bam(sam, bam)
            os.unlink(sam)
            S.write_vcf(vcf, S.vcf_lines(truth, end_first=True), pad=3 * cores + 8)     # the reference drops queued lines at EOF
            bam_mb = os.path.getsize(bam) / 1e6
            cmd = [os.path.join(B.REF_DIR, "svtrek"), "audt", "-b", bam, "-v", vcf, "-t", str(cores)]
            times, printed = [], 0
            for i in range(W + K):
                t0 = time.perf_counter()
                p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True)
                dt = time.perf_counter() - t0
                if i >= W:
                    times.append(dt)
                    printed = p.stdout.count(b"(DEL) ") + p.stdout.count(b"(INS) ") + p.stdout.count(b"(INV) ")
            tot = float(sum(times))
            value = printed * K / tot
            base.update({"value": value, "ms_per_step": tot * 1e3 / K,
                         "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                                          "sample": f"oracle/_ref/svtrek audt -t {cores} (unmodified reference + its HTSlib, zlib inflate) on a {bam_mb:.0f} MB BAM "
                                                    f"(random SEQ/QUAL), {printed} records printed per run, page cache warm"},
                         "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    else:
        # the reference did not compile / did not travel: the oracle port stands in (1 thread, decoded reads)
        orc = B.OracleReads(rs)
        times = []
        for i in range(W + K):
            t0 = time.perf_counter()
            B.refine_batch(orc, w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
            if i >= W:
                times.append(time.perf_counter() - t0)
        tot = float(sum(times))
        value = n_records * K / tot
        base.update({"value": value, "ms_per_step": tot * 1e3 / K,
                     "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port",
                                      "sample": "oracle/_ref missing: oracle/svtk_oracle.c orc_refine_batch on decoded reads, 1 thread"},
                     "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    emit(base)


_JSON_FD = None


def emit(obj):
    """The ONE JSON line of the contract, on the process's real stdout."""
    line = (json.dumps(obj) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(line.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, line)


def main():
    global _JSON_FD
    args = parse_args()
    # Libraries write banners to fd 1 (NCCL prints "NCCL version ..." there when NCCL_DEBUG=VERSION is set on the box, child
    # processes of the reference arm inherit it): everything that is not the JSON line goes to stderr.
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
