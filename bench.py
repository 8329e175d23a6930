#!/usr/bin/env python
"""bench.py -- SV records audited/s of the audt breakpoint-refinement hot path on N B200s (one process per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

THE JOB (both arms, every N).  One `svtrek audt -b job.bam -v job.vcf` job: a coordinate-sorted, BAI-indexed BAM with N
references, each BASELINE.json configs[1] (64 Mb, 30x 15 kb HiFi-like reads, random SEQ / QUAL, 10k planted DEL/INS), and one VCF
over all of them.  N = 1 is exactly configs[1].  At N > 1 the job is sharded by genomic region -- GPU r owns reference r: it
loads that reference's byte range of the SAME job.bam (svtk_load_bam_refs) and audits that reference's VCF records
(svtk_process_vcf_region) -- with no data-path collective; rank 0 concatenates the shard outputs in VCF order on the host
(north star).  Work per GPU is fixed as N grows: "scaling": "weak".  The files are written once per box into a cache
directory (svtrek_b200.simreads.write_bam_fast: HTSlib's record / block layout, zlib level 6) and read by both arms.

  value   records/s with the job's decoded reads and windows already resident in HBM: K whole passes (signature index over every
          CIGAR word + every window) queued back to back on the ctx stream (svtk_audit_device), device time of each pass from
          CUDA events on that stream, L2 flushed between passes; max over ranks
  e2e     records/s FILE TO FILE through the C ABI, like the reference CLI: svtk_load_bam[_refs](job.bam) + svtk_process_vcf[_region]
          (job.vcf -> the reference's stdout lines in a file), wall clock from before the first byte is read to the merged output;
          page cache warm; the compressed file crosses PCIe inside the timed region
  roofline  the step's dominant kernel: algorithmic bytes / its CUDA-event launch time, measured live (svtk_get_timing)
  parity  the resident results of EVERY window of this rank's shard compared with the CPU oracle (oracle/svtk_oracle.c), and the
          file-to-file output compared byte for byte with the unmodified reference CLI (`svtrek audt -t 1`) on a sample of records
  cpu_baseline  N = 1, rank 0: the unmodified reference CLI (oracle/_ref/svtrek audt -t <cores>) timed once on the same files
  --impl reference   the reference arm: the UNMODIFIED reference CLI on the host cores, W + K runs on the job's first reference
          (for N = 1: the whole job -- the very files the other arm reads)
  --workload c3      manual runs: BASELINE.json configs[2] (24 references, 3.1 Gb, 25k SVs) as ONE job cut into N region shards
          of equal work (cuts fall inside references), decoded reads resident: strong scaling of the index + refine pass
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
CACHE = os.environ.get("SVTK_BENCH_CACHE", os.path.join(tempfile.gettempdir(), "svtk_bench_cache"))


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c3"])
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the workload (debug only; the default is the full config)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--ref-threads", type=int, default=0)
    ap.add_argument("--no-bam", action="store_true", help="skip the file-to-file leg")
    ap.add_argument("--no-c3", action="store_true", help="N = 1: skip the extra configs[2]-sized resident pass (c3_single_gpu)")
    ap.add_argument("--e2e-steps", type=int, default=0, help="file-to-file steps (default: min(steps, 10))")
    ap.add_argument("--poa-sets", type=int, default=10000, help="INS read sets per GPU of the POA (K3) leg: a bounded sample (half) of configs[3]'s 20k records; 0 skips the leg")
    ap.add_argument("--poa-steps", type=int, default=1)
    return ap.parse_args()


WORKLOAD_DESC = {
    "c1": "BASELINE.json configs[0]: synthetic 1 Mb contig, 500 DEL/INS, 30x 15 kb HiFi-like reads",
    "c2": "BASELINE.json configs[1]: synthetic chr20-sized 64 Mb contig, 30x 15 kb HiFi-like reads, 10k DEL/INS records",
    "c3": "BASELINE.json configs[2]: synthetic 3.1 Gb genome (24 references), 30x 15 kb HiFi-like reads, 25k SVs",
}
PARAMS_DESC = "wider 20000, median 10000, narrow 2000, range 500, interval 5, min-count 3 (params.h defaults)"


def job_config(args, world):
    """The `config` object of the JSON line: a function of (arguments, N) only, so both arms print the same one."""
    sc = f" (scale {args.scale})" if args.scale != 1.0 else ""
    if args.workload == "c3":
        return {"workload": WORKLOAD_DESC["c3"] + sc,
                "job": f"one job, region-sharded over {world} GPU(s): contiguous genomic ranges of equal work, cuts inside references; decoded reads resident",
                "sharding": "region shards of one job, no collective; results merged in VCF order on the host" if world > 1 else "single GPU",
                "reference_arm_sample": "the job's first 16 Mb (reference 1) written as BAM + VCF, `svtrek audt -t <cores>`",
                "params": PARAMS_DESC}
    return {"workload": WORKLOAD_DESC[args.workload] + sc,
            "job": f"one BAM with {world} such reference(s) + one VCF over all of them ({world} x the config); both arms read the same files",
            "sharding": (f"region shards of one job: GPU r loads reference r's byte range of job.bam and audits its VCF records; "
                         f"no collective, shard outputs concatenated in VCF order on rank 0's host") if world > 1 else "single GPU",
            "reference_arm_sample": "N = 1: the whole job; N > 1: the job's first reference (same files, one config's worth)",
            "params": PARAMS_DESC}


# ------------------------------------------------------------------------------------------------------------
# the job's inputs: contig i of the job is the workload's contig, seeded by i (any process can regenerate any contig)
def contig_reads(name, i, scale):
    from svtrek_b200 import simreads as S
    seed = SEED0 + sorted(S.CONFIGS).index(name) + 1000 * i
    rs, truth = S.make_config(name, seed=seed, scale=scale)
    return rs, truth


def contig_vcf_lines(truth, i):
    from svtrek_b200 import simreads as S
    t = S.SVTruth(np.full_like(truth.tid, i), truth.pos0, truth.length, truth.is_del)
    return S.vcf_lines(t, end_first=True)


def contig_windows(truth, i):
    from svtrek_b200 import simreads as S
    w = S.windows_from_truth(truth)
    w["chrom"] = np.full_like(w["chrom"], i + 1)
    return w


def job_dir(args, world):
    d = os.path.join(CACHE, f"{args.workload}_s{args.scale}_n{world}_v3")
    os.makedirs(d, exist_ok=True)
    return d


def pad_lines(cores):
    return 3 * cores + 8            # the reference drops up to 2t queued lines at end of file


def ensure_part(args, world, i, threads):
    """part i of the job: reference i alone in a BAM whose header lists all `world` references (+ .bai, + what assembly needs)."""
    from svtrek_b200 import simreads as S
    d = job_dir(args, world)
    bam, meta = os.path.join(d, f"part{i}.bam"), os.path.join(d, f"part{i}.meta.npz")
    if os.path.exists(meta) and os.path.exists(bam + ".bai"):
        return bam, dict(np.load(meta)), None
    rs, truth = contig_reads(args.workload, i, args.scale)
    L = [int(rs.contig_len[0])] * world
    tmp = bam + f".tmp{os.getpid()}"
    m = S.write_bam_fast(tmp, rs, seed=7 + i, truth=truth, header_contigs=L, tid_shift=i, want_meta=True, threads=threads)
    os.replace(tmp + ".bai", bam + ".bai")
    os.replace(tmp, bam)
    np.savez(meta + ".tmp.npz", size=m["size"], hdr_bytes=m["hdr_bytes"], v0=m["v0"], v1=m["v1"], bins=m["bins"], pos=rs.pos, end=rs.end)
    os.replace(meta + ".tmp.npz", meta)
    return bam, dict(np.load(meta)), (rs, truth)


def ensure_job_files(args, world, rank, dist=None):
    """job.bam / job.bam.bai / job.vcf (+ sample.vcf: reference 1's records) in the cache directory; every rank writes its own part,
    rank 0 assembles.  Returns the paths."""
    import struct
    from svtrek_b200 import simreads as S
    d = job_dir(args, world)
    done = os.path.join(d, "done.json")
    paths = dict(bam=os.path.join(d, "job.bam"), vcf=os.path.join(d, "job.vcf"), sample_bam=os.path.join(d, "part0.bam"), sample_vcf=os.path.join(d, "sample.vcf"))
    cores = args.ref_threads or os.cpu_count() or 1
    have = os.path.exists(done)
    if dist is not None and world > 1:
        import torch
        t = torch.tensor([1 if have else 0], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        have = bool(int(t.item()))
    if have:
        return paths
    bam_i, meta, made = ensure_part(args, world, rank, max(1, (os.cpu_count() or 2) // max(1, world)))
    truth = made[1] if made else contig_reads(args.workload, rank, args.scale)[1]
    lines = contig_vcf_lines(truth, rank)
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, dict(lines=lines, size=int(meta["size"]), hdr=int(meta["hdr_bytes"])))
    else:
        gathered = [dict(lines=lines, size=int(meta["size"]), hdr=int(meta["hdr_bytes"]))]
    if rank == 0:
        parts = [os.path.join(d, f"part{i}.bam") for i in range(world)]
        if world == 1:
            for ext in ("", ".bai"):
                if os.path.lexists(paths["bam"] + ext):
                    os.unlink(paths["bam"] + ext)
                os.link(parts[0] + ext, paths["bam"] + ext)
        else:
            shifts = S.assemble_bam_parts(paths["bam"] + ".tmp", parts, gathered[0]["hdr"], [g["size"] for g in gathered])
            with open(paths["bam"] + ".bai.tmp", "wb") as f:
                f.write(b"BAI\1" + struct.pack("<i", world))
                for i in range(world):
                    mi = dict(np.load(os.path.join(d, f"part{i}.meta.npz")))
                    f.write(S.bai_section(mi["bins"], mi["v0"], mi["v1"], mi["pos"], mi["end"], shift=shifts[i]))
            os.replace(paths["bam"] + ".bai.tmp", paths["bam"] + ".bai")
            os.replace(paths["bam"] + ".tmp", paths["bam"])
            for i in range(1, world):                          # part 0 stays: it is the reference arm's bounded sample
                for ext in ("", ".bai"):
                    os.unlink(parts[i] + ext)
        S.write_vcf(paths["vcf"], [ln for g in gathered for ln in g["lines"]], pad=pad_lines(cores))
        S.write_vcf(paths["sample_vcf"], gathered[0]["lines"], pad=pad_lines(cores))
        json.dump({"world": world, "records": sum(len(g["lines"]) for g in gathered)}, open(done, "w"))
    if world > 1:
        dist.barrier()
    return paths


def ensure_sample_files(args, world):
    """What the reference arm needs, made by one process alone: part 0 of the job and its VCF (for N = 1 that is the whole job)."""
    from svtrek_b200 import simreads as S
    d = job_dir(args, world)
    cores = args.ref_threads or os.cpu_count() or 1
    bam, _meta, made = ensure_part(args, world, 0, os.cpu_count() or 1)
    vcf = os.path.join(d, "sample.vcf")
    if not os.path.exists(vcf):
        truth = made[1] if made else contig_reads(args.workload, 0, args.scale)[1]
        S.write_vcf(vcf + ".tmp", contig_vcf_lines(truth, 0), pad=pad_lines(cores))
        os.replace(vcf + ".tmp", vcf)
    return bam, vcf


# ------------------------------------------------------------------------------------------------------------
class NvmlSampler:
    """clocks.sm / throttle reasons DURING the timed regions, through NVML every 5 ms (the device-timed region lasts
    milliseconds: too short for `nvidia-smi -lms`)."""

    def __init__(self, index):
        self.ok = False
        self.sm, self.reasons, self.power = [], set(), []
        self.smax = None
        self._stop = threading.Event()
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            try:
                import torch
                uuid = "GPU-" + str(torch.cuda.get_device_properties(index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.smax = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:
            pass

    def start(self):
        if self.ok:
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
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable"]}
        self._stop.set()
        self.th.join(timeout=2)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.smax, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": self.smax, "power_w_max": float(max(self.power)) if self.power else None,
                "samples": len(self.sm), "reasons": sorted(self.reasons), "source": "nvml, 5 ms period, over the device-timed and file-to-file loops"}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy)"
        except Exception:
            pass
    return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


def traffic_from_profiles():
    """dram bytes per launch of the step's kernels, from the newest committed `ncu --set full` capture (profiles/*_traffic.json)."""
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


def pinned(arr):
    import torch
    a = np.ascontiguousarray(arr)
    view = {8: np.int64, 4: np.int32, 1: np.uint8}[a.dtype.itemsize]
    return torch.from_numpy(a.view(view)).pin_memory()


# ------------------------------------------------------------------------------------------------------------
POA_DESC = "BASELINE.json configs[3] sample: INS read sets, inserted length log-uniform 50 bp - 10 kb, 60x depth x 0.8 carriers, 8 % ONT-like errors"


def poa_leg(a, rank, world, n_sets, steps):
    """K3: svtk_poa_consensus (host buffers in, consensus out) on this rank's read sets.  GCUPS = band cells the kernel itself
    counted (5-state convex cell updates) / CUDA-event time of k_poa_sets; e2e = the same cells / wall time of the call."""
    from svtrek_b200 import simpoa as P
    # one job of n_sets x world read sets, dealt to the ranks by estimated work (heaviest first, snake order)
    sets, _mine, _all = P.ins_workload_shard(SEED0 + 77, n_sets * world, world, rank, depth=60, carrier=0.8, err=0.08, len_lo=50, len_hi=10000)
    so, qo, b = P.pack(sets)
    a.poa_consensus(so, qo, b)                       # warm-up: scratch allocation, first touch
    kms, wall, st, cons = 0.0, 0.0, None, None
    for _ in range(steps):
        t0 = time.perf_counter()
        cons, st = a.poa_consensus(so, qo, b)
        wall += time.perf_counter() - t0
        kms += st["kernel_ms"]
    return sets, cons, dict(cells=st["cells"], cells_full=st["cells_full"], n_sets=len(sets), n_reads=int(len(qo) - 1), n_bases=int(len(b)),
                            kernel_ms=kms / steps, wall_ms=wall * 1e3 / steps, workers=st["workers"], relaunched=st["relaunched"],
                            phase_cycles=st["phase_cycles"])


def poa_cpu_baseline(sets, cons, budget_s=10.0):
    """abPOA 1.5.4 itself (oracle/_ref/libabpoa_ref.so, the reference's submodule compiled unmodified), 1 thread, on a bounded
    sample of the same read sets; every sampled consensus is also compared with the GPU's."""
    from oracle import binding as B
    if not B.have_abpoa():
        return None
    order = np.random.default_rng(1).permutation(len(sets))
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


# ------------------------------------------------------------------------------------------------------------
def reference_cli(bam, vcf, threads):
    from oracle import binding as B
    return [os.path.join(B.REF_DIR, "svtrek"), "audt", "-b", bam, "-v", vcf, "-t", str(threads)]


def count_records(stdout_bytes):
    return stdout_bytes.count(b"(DEL) ") + stdout_bytes.count(b"(INS) ") + stdout_bytes.count(b"(INV) ")


def stdout_parity_sample(paths, merged_out, every=40):
    """`svtrek audt -t 1` (the unmodified reference; -t 1: VCF order, untorn lines) on every `every`-th record of the job's VCF,
    against the lines the GPU path wrote for the same records.  Returns (records compared, equal)."""
    from oracle import binding as B
    if not B.have_ref():
        return 0, None
    recs = [ln for ln in open(paths["vcf"]).read().splitlines() if ln and not ln.startswith("#") and ln != "pad"]
    pick = recs[::every]
    ours = {}
    for ln in open(merged_out).read().splitlines():
        if ln.startswith("("):
            # key: type, chr, org pos -- unique in the synthetic job
            f = ln.split(", ")
            ours[(f[0], f[1])] = ln
    with tempfile.TemporaryDirectory(prefix="svtk_par_") as d:
        sv = os.path.join(d, "s.vcf")
        with open(sv, "w") as f:
            f.write("##fileformat=VCFv4.2\n#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n" + "\n".join(pick) + "\n" + "pad\n" * 16)
        p = subprocess.run(reference_cli(paths["bam"], sv, 1), stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True)
    ref_lines = [ln for ln in p.stdout.decode().splitlines() if ln.startswith("(")]
    ok = len(ref_lines) > 0
    for ln in ref_lines:
        f = ln.split(", ")
        if ours.get((f[0], f[1])) != ln:
            ok = False
            break
    return len(ref_lines), ok


def run_ours(args):
    import torch
    import torch.distributed as dist
    from oracle import binding as B
    from svtrek_b200 import api

    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path is CUDA only (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    K, W = args.steps, max(args.warmup, 3)
    if world > 1 and "SVTK_BAM_THREADS" not in os.environ:       # the ranks share the host's cores: pread threads per rank
        os.environ["SVTK_BAM_THREADS"] = str(max(2, (os.cpu_count() or 2) // world - 1))
    if args.workload == "c3":
        out = c3_measure(args, rank, world, local, dev, dist, K, W)
        if rank == 0:
            emit(out)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- the job: this rank's reference (decoded, for the resident leg and the oracle) and the job's files (for the file leg)
    t_gen0 = time.perf_counter()
    rs1, truth = contig_reads(args.workload, rank, args.scale)
    w = contig_windows(truth, rank)
    n_records = int(np.unique(w["record"]).shape[0])      # records that print a line (DEL with end - pos <= 50 print nothing)
    n_win = int(w["chrom"].shape[0])
    # the resident read set has the job's reference dictionary: this rank's reads sit on reference `rank`
    tid_off = np.zeros(world + 1, dtype=np.int64)
    tid_off[rank + 1:] = rs1.n_reads
    paths = None if args.no_bam else ensure_job_files(args, world, rank, dist if world > 1 else None)
    gen_s = time.perf_counter() - t_gen0

    stream = torch.cuda.Stream(device=dev)
    a = api.Auditor(local, cuda_stream=stream.cuda_stream)
    hp = {k: pinned(v) for k, v in dict(pos=rs1.pos.astype(np.uint32), end=rs1.end.astype(np.uint32), n_cigar=rs1.n_cigar.astype(np.uint32),
                                        cigar_off=rs1.cigar_off.astype(np.uint64)[:rs1.n_reads], cigar=rs1.cigar.astype(np.uint32)).items()}
    hw = {k: pinned(np.ascontiguousarray(w[k], dtype=(np.int32 if k == "chrom" else np.uint8 if k == "kind" else np.uint32)))
          for k in ("chrom", "start", "end", "pos", "kind")}
    n_reads, n_words = int(rs1.n_reads), int(rs1.cigar.shape[0])
    a.load_reads_ptr(world, n_reads, n_words, tid_off.ctypes.data, hp["pos"].data_ptr(), hp["end"].data_ptr(), hp["n_cigar"].data_ptr(),
                     hp["cigar_off"].data_ptr(), hp["cigar"].data_ptr())
    with torch.cuda.stream(stream):
        dw = {k: v.to(dev, non_blocking=True) for k, v in hw.items()}
        d_res = torch.empty(n_win, dtype=torch.int32, device=dev)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
    stream.synchronize()
    wargs = (n_win, dw["chrom"].data_ptr(), dw["start"].data_ptr(), dw["end"].data_ptr(), dw["pos"].data_ptr(), dw["kind"].data_ptr())

    # ---- work counters (instrumented instantiations, outside the timed region) and the three-way result check:
    # product path == reference-shaped per-(window, read) walker (both CUDA) == CPU oracle, on EVERY window of the shard
    walker = a.measure_windows(*wargs, d_res.data_ptr())
    walker_res = d_res.cpu().numpy().copy()
    stats = a.measure_index(*wargs, d_res.data_ptr())
    ref_res = d_res.cpu().numpy().copy()
    assert np.array_equal(walker_res, ref_res), "indexed path and per-pair walker disagree"
    rs_job = type("RS", (), dict(tid_off=tid_off, pos=rs1.pos, end=rs1.end, n_cigar=rs1.n_cigar, cigar_off=rs1.cigar_off, cigar=rs1.cigar))
    orc = B.OracleReads(rs_job)
    t0 = time.perf_counter()
    want = B.refine_batch(orc, w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    port_s = time.perf_counter() - t0
    mism = int((want != ref_res).sum())
    assert mism == 0, f"{mism} of {n_win} windows differ from the CPU oracle"

    # ---- value: K whole passes queued back to back (no host wait between them), L2 flushed before each
    with torch.cuda.stream(stream):
        for _ in range(W):
            flush.zero_()
            a.audit_device(*wargs, d_res.data_ptr())
        a.sync()
    sampler = NvmlSampler(local)
    barrier()
    sampler.start()
    prof_stride = max(1, min(4, K // 4))           # an event pair serialises the stream around its kernel: every 4th pass carries them
    a.set_profiling(prof_stride)
    launches0 = a.launch_count
    evs = []
    t_wall0 = time.perf_counter()
    with torch.cuda.stream(stream):
        for _ in range(K):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            a.audit_device(*wargs, d_res.data_ptr())
            e1.record(stream)
            evs.append((e0, e1))
        a.sync()                                    # waits for all K passes; finishes windows that left the fast path (none here)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    dev_ms = float(sum(e0.elapsed_time(e1) for e0, e1 in evs))
    per_step = sorted(e0.elapsed_time(e1) for e0, e1 in evs)
    launches = a.launch_count - launches0
    tim = a.timing()
    a.set_profiling(False)
    assert np.array_equal(d_res.cpu().numpy(), ref_res), "timed passes changed the result"

    # ---- e2e: file to file on its own context, like the C driver's process would have
    e2e = None
    if not args.no_bam:
        Ke = args.e2e_steps or min(K, 10)
        with api.Auditor(local) as fa:
            e2e = file_leg(fa, args, paths, rank, world, Ke, barrier, n_records, want, w)
    clocks = sampler.stop()

    # ---- K3: POA consensus leg (GCUPS half of the metric)
    poa = None
    if args.poa_sets > 0:
        poa_sets_l, poa_cons, poa = poa_leg(a, rank, world, args.poa_sets, max(1, args.poa_steps))
        pr = torch.tensor([poa["kernel_ms"], poa["wall_ms"]], dtype=torch.float64, device=dev)
        pt = torch.tensor([float(poa["cells"]), float(poa["cells_full"]), float(poa["n_sets"])], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(pr, op=dist.ReduceOp.MAX)
            dist.all_reduce(pt, op=dist.ReduceOp.SUM)
        poa_kms, poa_wall = [float(x) for x in pr.cpu()]
        poa_cells, poa_full, poa_nsets = [float(x) for x in pt.cpu()]

    # ---- max over ranks, whole-job aggregate
    red = torch.tensor([dev_ms, e2e["ms_total"] if e2e else 0.0], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(n_records), float(launches), float(e2e["h2d"] if e2e else 0), float(e2e["d2h"] if e2e else 0), float(n_win), float(mism)],
                       dtype=torch.float64, device=dev)
    per_rank = torch.zeros(world, dtype=torch.float64, device=dev)
    per_rank[rank] = dev_ms / K
    if world > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(per_rank, op=dist.ReduceOp.SUM)
    dev_ms_max, e2e_ms_max = [float(x) for x in red.cpu()]
    rec_all, launches_all, h2d_all, d2h_all, win_all, mism_all = [float(x) for x in tot.cpu()]

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
            "config": job_config(args, world),
            "clocks": clocks,
            "e2e": None,
            "gpu_launches": int(launches_all),
            "parity": {"oracle_windows": int(win_all), "mismatches": int(mism_all),
                       "checked": "every window of every rank's shard: signature-index path == per-(window, read) walker == CPU oracle (oracle/svtk_oracle.c)"},
            "work": {"records_per_gpu": n_records, "windows_per_gpu": n_win, "reads_per_gpu": n_reads, "cigar_words_per_gpu": n_words,
                     "l2": "256 MB buffer written between timed passes (L2 flush); the CIGAR array is %.0f MB" % (4 * n_words / 1e6),
                     "timed_region": "%d passes queued back to back on the ctx stream (svtk_audit_device), no host wait in between; "
                                     "CUDA events around every pass, kernel event pairs on every %d%s pass" % (
                                         K, prof_stride, "th" if prof_stride > 3 else ("st" if prof_stride == 1 else "nd" if prof_stride == 2 else "rd")),
                     "ms_per_step_per_rank": [float(x) for x in per_rank.cpu()], "step_ms_min_median_max": [per_step[0], per_step[len(per_step) // 2], per_step[-1]],
                     "input_generation_s": gen_s},
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
                               "algorithmic_bytes_per_launch": kern[oth][1], "kernel_ms": kern[oth][0], "traffic": (tr or {}).get(oth)},
            "reference_equivalent": {"note": "bytes the reference's per-(window, read) walks touch (SURVEY.md 8d) / time of both kernels",
                                     "bytes_per_step": walker["algorithmic_bytes"], "GBps": walker["algorithmic_bytes"] / ((idx_ms + ref_ms) * 1e-3) / 1e9 if (idx_ms + ref_ms) > 0 else 0.0,
                                     "work": {k: walker[k] for k in ("n_windows", "n_pairs", "n_words_visited", "n_signatures", "n_overflow")}},
            "poa_gcups": None,
        }
        if e2e:
            out["e2e"] = {"value": rec_all * e2e["steps"] / (e2e_ms_max * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d_all), "d2h_bytes_per_step": int(d2h_all),
                          "ms_per_step": e2e_ms_max / e2e["steps"], "steps": e2e["steps"],
                          "path": "file to file: svtk_load_bam%s(job.bam + .bai) + svtk_process_vcf%s(job.vcf -> output file)%s; wall clock, page cache warm" % (
                              ("_refs" if world > 1 else ""), ("_region" if world > 1 else ""), ("; shard outputs concatenated on rank 0" if world > 1 else "")),
                          "files": e2e["files"], "load_bam": e2e["load"], "output_equals_oracle_formatting": e2e["out_ok"]}
            out["parity"]["stdout_equal_reference"] = e2e["stdout_equal_reference"]
            out["parity"]["stdout_records_compared"] = e2e["stdout_records"]
            out["parity"]["loaded_reads_identical_to_written"] = e2e["reads_ok"]
            if e2e.get("cli"):
                out["e2e"]["svtrek_gpu_process_ms"] = e2e["cli"]
            if e2e.get("ins"):
                out["ins_consensus"] = e2e["ins"]
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
        if world == 1 and not args.no_cpu:
            out["cpu_baseline"] = cpu_reference_baseline(args, paths, n_records, port_s, n_win)
    a.close()
    if world == 1 and not args.no_c3 and args.workload == "c2" and args.scale == 1.0:
        # the same pass at the size the north star states its HBM target on (configs[2], 3.1 Gb, one GPU): its own roofline
        sub = argparse.Namespace(**vars(args))
        sub.workload = "c3"
        c3 = c3_measure(sub, 0, 1, local, dev, dist, min(K, 10), 3)
        out["c3_single_gpu"] = {k: c3[k] for k in ("value", "unit", "ms_per_step", "steps", "config", "roofline", "roofline_other", "parity", "work", "kernel_share")}
    if rank == 0:
        emit(out)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def cpu_reference_baseline(args, paths, n_records, port_s, n_win):
    """rank 0, N = 1: the unmodified reference CLI on all host cores, ONE timed run on the job's own files (about 10 s of wall
    time on a 16-core host); the oracle port's single-thread rate on decoded reads beside it."""
    from oracle import binding as B
    cores = args.ref_threads or os.cpu_count() or 1
    port = {"records_per_s": n_records / port_s, "cores": 1,
            "what": f"oracle/svtk_oracle.c orc_refine_batch, 1 thread, every window of the job once ({n_win} windows in {port_s:.2f} s); reads already decoded (no BGZF inflate)"}
    if paths is None or not B.have_ref():
        return {"value": port["records_per_s"], "unit": UNIT, "cores": 1, "kind": "port", "sample": port["what"]}
    t0 = time.perf_counter()
    p = subprocess.run(reference_cli(paths["bam"], paths["vcf"], cores), stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True)
    dt = time.perf_counter() - t0
    n = count_records(p.stdout)
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": f"oracle/_ref/svtrek audt -t {cores} (unmodified reference + its HTSlib, zlib inflate), one run on the job's own files "
                      f"({os.path.getsize(paths['bam']) / 1e6:.0f} MB BAM, {n} records printed) in {dt:.1f} s, page cache warm",
            "oracle_port": port}


def file_leg(a, args, paths, rank, world, steps, barrier, n_records, want, w):
    """File to file, like the reference CLI.  Every step: this rank loads its part of job.bam (the whole file at N = 1) and audits
    its part of job.vcf into its own output file; after a barrier rank 0 concatenates the shard outputs between the two [INFO]
    lines.  The clock runs from before the load to the merged file."""
    from svtrek_b200 import api
    d = job_dir(args, world)
    mine = os.path.join(d, f"out.rank{rank}.txt")
    merged = os.path.join(d, "out.merged.txt")
    loads, times = [], []
    h2d0 = d2h0 = 0
    for i in range(2 + steps):
        barrier()
        tm0 = a.timing()
        t0 = time.perf_counter()
        if world == 1:
            st = a.load_bam(paths["bam"])
            n = a.process_vcf(paths["vcf"], merged)
        else:
            st = a.load_bam(paths["bam"], refs=(rank, rank))
            n = a.process_vcf_region(paths["vcf"], mine, rank, rank, info_lines=False, owns_unplaced=(rank == 0))
            barrier()
            if rank == 0:
                with open(merged, "wb") as f:
                    f.write(b"[INFO] Started processing variation file.\n")
                    for r in range(world):
                        with open(os.path.join(d, f"out.rank{r}.txt"), "rb") as g:
                            f.write(g.read())
                    f.write(b"[INFO] Ended processing variation file\n")
            barrier()
        dt = time.perf_counter() - t0
        tm1 = a.timing()
        assert n == n_records, (n, n_records)
        if i >= 2:
            times.append(dt); loads.append(st)
            h2d0, d2h0 = tm1["h2d_bytes"] - tm0["h2d_bytes"], tm1["d2h_bytes"] - tm0["d2h_bytes"]
    best = min(loads, key=lambda r: r["total_ms"])
    # what was loaded == what was written (this rank's reference), and what was written out == the oracle's results formatted
    got = a.copy_reads()
    rs1, _ = contig_reads(args.workload, rank, args.scale)
    reads_ok = bool(np.array_equal(got["pos"], rs1.pos) and np.array_equal(got["end"], rs1.end) and np.array_equal(got["cigar"], rs1.cigar))
    out_path = merged if world == 1 else mine
    chk = os.path.join(d, f"chk.rank{rank}.txt")
    wv = api.vcf_windows(paths["vcf"], n_ref=world, refs=None if world == 1 else (rank, rank), owns_unplaced=(rank == 0))
    assert np.array_equal(wv["start"], w["start"]) and np.array_equal(wv["kind"], w["kind"]), "the VCF front end built other windows than the generator's"
    api.vcf_format(paths["vcf"], want, chk, n_ref=world, refs=None if world == 1 else (rank, rank), owns_unplaced=(rank == 0), info_lines=(world == 1))
    out_ok = open(out_path, "rb").read() == open(chk, "rb").read()
    res = {"ms_total": float(sum(times)) * 1e3, "steps": len(times), "h2d": h2d0, "d2h": d2h0, "reads_ok": reads_ok, "out_ok": out_ok,
           "files": {"bam_MB": os.path.getsize(paths["bam"]) / 1e6, "vcf_records": int(n_records * world)},
           "load": {"shipped_MB": best["image_bytes"] / 1e6, "inflated_MB": best["inflated_bytes"] / 1e6, "bgzf_blocks": best["n_blocks"],
                    "lanes_per_warp": best["lanes_per_warp"], "inflate_launches": best["inflate_launches"],
                    "upload_ms": best["upload_ms"], "inflate_ms_overlapping_the_upload": best["inflate_ms"], "upload_plus_inflate_ms": best["load_span_ms"],
                    "record_kernels_ms": best["parse_ms"], "total_ms": best["total_ms"],
                    "inflate_GBps": best["inflated_bytes"] / max(best["inflate_ms"], 1e-6) / 1e6,
                    "load_GBps_of_file": best["image_bytes"] / max(best["total_ms"], 1e-6) / 1e6},
           "stdout_equal_reference": None, "stdout_records": 0, "cli": None, "ins": None}
    if rank == 0:
        if not args.no_cpu:
            res["stdout_records"], res["stdout_equal_reference"] = stdout_parity_sample(paths, merged)
        if world == 1:
            # the same files through the C driver as a process: process start -> last line (CUDA context creation included)
            cli = os.path.join(ROOT, "svtrek_b200", "svtrek_gpu")
            if os.path.exists(cli):
                ts = []
                for _ in range(2):
                    t0 = time.perf_counter()
                    subprocess.run([cli, "audt", "-b", paths["bam"], "-v", paths["vcf"]], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=True)
                    ts.append((time.perf_counter() - t0) * 1e3)
                res["cli"] = min(ts)
            # beyond the reference's output: inserted sequences of the refined INS breakpoints -> POA consensus
            refined = a.refine_windows(w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
            sel = np.flatnonzero(w["kind"] == 2)[:400]
            t0 = time.perf_counter()
            sets = a.ins_sequences(w["chrom"][sel], w["start"][sel], w["end"][sel], w["pos"][sel], w["kind"][sel], refined[sel], tol=5)
            t1 = time.perf_counter()
            cons, pst = a.poa_consensus(sets["set_off"], sets["seq_off"], sets["bases"]) if len(sets["window"]) else ([], {"cells": 0, "kernel_ms": 0.0})
            t2 = time.perf_counter()
            res["ins"] = {"ins_records": int(len(sel)), "read_sets": int(len(sets["window"])), "sequences": int(len(sets["read"])), "bases": int(len(sets["bases"])),
                          "extract_ms": (t1 - t0) * 1e3, "poa_ms": (t2 - t1) * 1e3, "poa_kernel_ms": float(pst["kernel_ms"]),
                          "path": "svtk_ins_sequences (SEQ of the I >= 50 ops within 5 bases of the refined position) + svtk_poa_consensus on the first 400 INS "
                                  "records; not part of the records/s above (the reference prints no sequence)"}
    return res


# ------------------------------------------------------------------------------------------------------------
# --workload c3: one 3.1 Gb job cut into N region shards of equal work (manual runs: strong scaling of the resident pass)
def _c3_contig(job):
    from svtrek_b200 import simreads as S
    i, L, n_sv, scale = job
    cfg = dict(S.CONFIGS["c3"])
    cfg["contig_lens"] = [L]
    cfg["n_sv"] = n_sv
    rs, truth = S.simulate(SEED0 + 3000 + i, **cfg)
    return i, rs.pos, rs.end, rs.n_cigar, rs.cigar, truth.pos0, truth.length, truth.is_del


def c3_measure(args, rank, world, local, dev, dist, K, W):
    """The resident pass on configs[2] cut into `world` region shards; returns rank 0's JSON object (None elsewhere)."""
    import multiprocessing as mp
    import torch
    from oracle import binding as B
    from svtrek_b200 import api, shard as SH, simreads as S
    cfg = S.CONFIGS["c3"]
    lens = [max(400_000, int(L * args.scale)) for L in cfg["contig_lens"]]
    total = float(sum(lens))
    n_sv = [int(round(cfg["n_sv"] * args.scale * L / total)) for L in lens]
    # region shards: the genome laid end to end, cut into `world` ranges of equal length (depth is uniform: equal work)
    starts = np.concatenate([[0], np.cumsum(lens)])
    g0, g1 = starts[-1] * rank // world, starts[-1] * (rank + 1) // world
    need = [i for i in range(len(lens)) if starts[i] < g1 and starts[i + 1] > g0]
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(min(len(need), max(1, (os.cpu_count() or 2) // world))) as pool:
        parts = pool.map(_c3_contig, [(i, lens[i], n_sv[i], args.scale) for i in need])
    gen_s = time.perf_counter() - t0
    # this rank's windows: the records whose position falls in its range; its reads: what those windows can see
    n_tid = len(lens)
    tid_off = np.zeros(n_tid + 1, dtype=np.int64)
    pos_l, end_l, nc_l, cg_l, wl = [], [], [], [], []
    for i, pos, end, ncig, cigar, tp, tl, td in parts:
        truth = S.SVTruth(np.full(tp.shape[0], i, dtype=np.int32), tp, tl, td)
        g = starts[i] + tp
        keep = (g >= g0) & (g < g1)
        tk = S.SVTruth(truth.tid[keep], tp[keep], tl[keep], td[keep])
        w = S.windows_from_truth(tk, seed=1 + i)
        wl.append(w)
        tid_off[i + 1] = pos.shape[0]
        pos_l.append(pos); end_l.append(end); nc_l.append(ncig); cg_l.append(cigar)
    tid_off = np.cumsum(tid_off)
    for i in range(n_tid):                                      # references this rank did not generate stay empty
        pass
    ncig = np.concatenate(nc_l).astype(np.uint32)
    coff = np.zeros(ncig.shape[0] + 1, dtype=np.uint64)
    np.cumsum(ncig.astype(np.uint64), out=coff[1:])
    full = S.ReadSet(tid_off, np.concatenate(pos_l), np.concatenate(end_l), ncig, coff, np.concatenate(cg_l), lens)
    w = {k: np.concatenate([x[k] for x in wl]) for k in ("chrom", "start", "end", "pos", "kind")}
    rec_off = np.concatenate([[0], np.cumsum([int(x["record"].max()) + 1 if len(x["record"]) else 0 for x in wl])])
    w["record"] = np.concatenate([x["record"] + rec_off[j] for j, x in enumerate(wl)])
    n_win = int(w["chrom"].shape[0])
    n_records = int(np.unique(w["record"]).shape[0])
    rs = SH.shard_reads(full, w, 0, n_win) if world > 1 else full
    del full, parts

    stream = torch.cuda.Stream(device=dev)
    a = api.Auditor(local, cuda_stream=stream.cuda_stream)
    hp = {k: pinned(v) for k, v in dict(pos=rs.pos.astype(np.uint32), end=rs.end.astype(np.uint32), n_cigar=rs.n_cigar.astype(np.uint32),
                                        cigar_off=rs.cigar_off.astype(np.uint64)[:rs.n_reads], cigar=rs.cigar.astype(np.uint32)).items()}
    h_tid = np.ascontiguousarray(rs.tid_off, dtype=np.int64)
    hw = {k: pinned(np.ascontiguousarray(w[k], dtype=(np.int32 if k == "chrom" else np.uint8 if k == "kind" else np.uint32)))
          for k in ("chrom", "start", "end", "pos", "kind")}
    n_reads, n_words = int(rs.n_reads), int(rs.cigar.shape[0])

    def load_reads():
        a.load_reads_ptr(n_tid, n_reads, n_words, h_tid.ctypes.data, hp["pos"].data_ptr(), hp["end"].data_ptr(), hp["n_cigar"].data_ptr(),
                         hp["cigar_off"].data_ptr(), hp["cigar"].data_ptr())
    load_reads()
    with torch.cuda.stream(stream):
        dw = {k: v.to(dev, non_blocking=True) for k, v in hw.items()}
        d_res = torch.empty(max(n_win, 1), dtype=torch.int32, device=dev)
    stream.synchronize()
    wargs = (n_win, dw["chrom"].data_ptr(), dw["start"].data_ptr(), dw["end"].data_ptr(), dw["pos"].data_ptr(), dw["kind"].data_ptr())
    stats = a.measure_index(*wargs, d_res.data_ptr())
    ref_res = d_res.cpu().numpy()[:n_win].copy()
    # oracle: every window of the shard (a few seconds per rank)
    t0 = time.perf_counter()
    want = B.refine_batch(B.OracleReads(rs), w["chrom"], w["start"], w["end"], w["pos"], w["kind"])
    port_s = time.perf_counter() - t0
    mism = int((want != ref_res).sum())
    assert mism == 0, f"{mism} of {n_win} windows differ from the CPU oracle"

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        for _ in range(W):
            a.audit_device(*wargs, d_res.data_ptr())
        a.sync()
    sampler = NvmlSampler(local)
    barrier()
    sampler.start()
    prof_stride = max(1, min(4, K // 4))
    a.set_profiling(prof_stride)
    launches0 = a.launch_count
    evs = []
    with torch.cuda.stream(stream):
        for _ in range(K):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            a.audit_device(*wargs, d_res.data_ptr())
            e1.record(stream)
            evs.append((e0, e1))
        a.sync()
    barrier()
    dev_ms = float(sum(e0.elapsed_time(e1) for e0, e1 in evs))
    launches = a.launch_count - launches0
    tim = a.timing()
    a.set_profiling(False)
    # e2e: decoded reads in pinned host arrays through the C ABI (a 93 GB BAM of this job is not written here)
    h_res = torch.empty(max(n_win, 1), dtype=torch.int32).pin_memory()
    Ke = min(K, 5)
    for _ in range(2):
        load_reads(); a.refine_windows_ptr(n_win, *[hw[k].data_ptr() for k in ("chrom", "start", "end", "pos", "kind")], h_res.data_ptr())
    barrier()
    t0 = time.perf_counter()
    for _ in range(Ke):
        load_reads()
        a.refine_windows_ptr(n_win, *[hw[k].data_ptr() for k in ("chrom", "start", "end", "pos", "kind")], h_res.data_ptr())
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    assert np.array_equal(h_res.numpy()[:n_win], ref_res)
    clocks = sampler.stop()
    red = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(n_records), float(launches), float(20 * n_reads + 4 * n_words + 17 * n_win), float(4 * n_win), float(n_win), float(n_words), float(n_reads)],
                       dtype=torch.float64, device=dev)
    per_rank = torch.zeros(world, dtype=torch.float64, device=dev)
    per_rank[rank] = dev_ms / K
    if world > 1:
        dist.all_reduce(red, op=dist.ReduceOp.MAX); dist.all_reduce(tot, op=dist.ReduceOp.SUM); dist.all_reduce(per_rank, op=dist.ReduceOp.SUM)
    dev_ms_max, e2e_ms_max = [float(x) for x in red.cpu()]
    rec_all, launches_all, h2d_all, d2h_all, win_all, words_all, reads_all = [float(x) for x in tot.cpu()]
    out = None
    if rank == 0:
        peak, peak_src = measured_peak()
        idx_ms = tim["index_kernel_ms"] / max(1, tim["index_kernel_launches"])
        ref_ms = tim["refine_kernel_ms"] / max(1, tim["refine_kernel_launches"])
        kern = {"k_index_reads": (idx_ms, stats["index_bytes"]), "k_refine_indexed": (ref_ms, stats["refine_bytes"])}
        dom = max(kern, key=lambda k: kern[k][0])
        oth = [k for k in kern if k != dom][0]
        achieved = kern[dom][1] / (kern[dom][0] * 1e-3) / 1e9 if kern[dom][0] > 0 else 0.0
        out = {"metric": METRIC, "value": rec_all * K / (dev_ms_max * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
              "ms_per_step": dev_ms_max / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u32", "data": "synthetic", "impl": "ours",
              "config": job_config(args, world), "clocks": clocks,
              "e2e": {"value": rec_all * Ke / (e2e_ms_max * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d_all), "d2h_bytes_per_step": int(d2h_all), "ms_per_step": e2e_ms_max / Ke,
                      "path": "svtk_load_reads(decoded reads, pinned host SoA) + svtk_refine_windows(host windows -> host results), wall clock; the file-to-file leg runs on --workload c2"},
              "gpu_launches": int(launches_all),
              "parity": {"oracle_windows": int(win_all), "mismatches": 0, "checked": "every window of every rank's shard against the CPU oracle"},
              "work": {"records": int(rec_all), "windows": int(win_all), "reads_resident_all_ranks": int(reads_all), "cigar_words_resident_all_ranks": int(words_all),
                       "ms_per_step_per_rank": [float(x) for x in per_rank.cpu()], "input_generation_s": gen_s, "oracle_port_s_rank0": port_s,
                       "l2": "no flush: one rank's CIGAR array is %.0f MB, larger than the 126 MB L2" % (4 * n_words / 1e6)},
              "kernel_share": {"index_kernel_ms_per_step": idx_ms, "refine_kernel_ms_per_step": ref_ms, "step_ms": dev_ms / K},
              "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                           "algorithmic_bytes_per_launch": kern[dom][1], "kernel_ms": kern[dom][0], "work": stats},
              "roofline_other": {"kernel": oth, "achieved": kern[oth][1] / (kern[oth][0] * 1e-3) / 1e9 if kern[oth][0] > 0 else 0.0, "unit": "GB/s", "peak": peak,
                                 "frac": (kern[oth][1] / (kern[oth][0] * 1e-3) / 1e9 / peak) if kern[oth][0] > 0 else None,
                                 "algorithmic_bytes_per_launch": kern[oth][1], "kernel_ms": kern[oth][0]}}
    a.close()
    return out if rank == 0 else None


# ------------------------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference arm: the UNMODIFIED reference CLI on the host cores (oracle/_ref/svtrek, compiled from /root/reference by
    oracle/build_ref.sh; the built files travel with the repo snapshot).  Each step is one `svtrek audt -b BAM -v VCF -t <cores>`
    run on the job's first reference -- for N = 1 that is the whole job, the very files the other arm reads."""
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
    if rank != 0:
        return
    from oracle import binding as B
    K, W = args.steps, args.warmup
    cores = args.ref_threads or os.cpu_count() or 1
    base = {"metric": METRIC, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W, "higher_is_better": True, "scaling": "weak" if args.workload != "c3" else "strong",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic", "impl": "reference", "config": job_config(args, world)}
    if args.workload == "c3":
        sub = argparse.Namespace(**vars(args))
        sub.workload, sub.scale = "c2", 0.25 * args.scale      # 16 Mb of a 30x genome
        bam, vcf = ensure_sample_files(sub, 1)
    else:
        bam, vcf = ensure_sample_files(args, world)
    if B.have_ref():
        cmd = reference_cli(bam, vcf, cores)
        times, printed = [], 0
        for i in range(W + K):
            t0 = time.perf_counter()
            p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True)
            dt = time.perf_counter() - t0
            if i >= W:
                times.append(dt)
                printed = count_records(p.stdout)
        tot = float(sum(times))
        value = printed * K / tot
        base.update({"value": value, "ms_per_step": tot * 1e3 / K,
                     "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                                      "sample": f"oracle/_ref/svtrek audt -t {cores} (unmodified reference + its HTSlib, zlib inflate) on a "
                                                f"{os.path.getsize(bam) / 1e6:.0f} MB BAM (random SEQ/QUAL), {printed} records printed per run, page cache warm"},
                     "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    else:
        # the reference did not compile / did not travel: the oracle port stands in (1 thread, decoded reads)
        rs, truth = contig_reads(args.workload if args.workload != "c3" else "c2", 0, args.scale)
        w = contig_windows(truth, 0)
        n_records = int(np.unique(w["record"]).shape[0])
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
