#!/usr/bin/env python3
"""bench.py — the `match_trie` hot path on B200: LR segments/s against 25 k barcodes (ED <= 2).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step = one pass of the hot path over one batch of synthetic segments (BASELINE.json configs[1]: 1 M segments x 25 k
16-mer barcodes, -mr 2, -bl 16, seeds of SURVEY.md §8(d)).  With N > 1 (torchrun, one rank per GPU) every rank owns
its own batch of the same shape (reads shard, the barcode table is replicated; no data-path collective) -> weak scaling.

  value      whole-job segments/s with the batch already resident in HBM (sct_stage once, then K x sct_run_staged),
             device time from CUDA events on the library's launching stream, max over ranks; L2 flushed between steps.
  e2e        the same metric through the public host call (Matcher.match -> sct_match): HOST numpy buffers in, HOST
             result arrays out, every step; pinned staging, H2D and D2H are inside the timed region (wall clock between
             barrier + synchronize on both sides, max over ranks).
  roofline   the dominant kernel (match_seed_kernel) against the INT32 ALU issue rate: algorithmic lane-ops
             (14 per Myers column update x 2B x sum(len)) / kernel time (CUDA events) / the LOP3 issue
             peak measured live on this GPU (sct_measure_int32_peak).  HBM figures are reported as the minor counter.
  cpu_baseline / --impl reference   the CPU port of the reference's trie algorithm (oracle/trie_oracle.c, OpenMP over
             the reference's own prefix tasks) on the host cores, on a bounded sample of the same workload.  The
             reference itself is a Python script under /root/reference, which does not exist on the GPU box.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "lr_segments_per_sec_vs_25k_barcodes_ed2"
UNIT = "segments/s"
N_BARCODES = 25_000
N_SEGMENTS = 1_000_000
MAX_ERROR = 2
THREADS_FLAG = 16          # the reference's default -t (fixes k = 2; irrelevant for the clean ACGT bench data)
OPS_PER_COLUMN = 14        # LOP3/IADD3-fused Hyyro column update
CPU_SAMPLE = 1_000_000     # segments per CPU-arm step / cpu_baseline sample (the whole per-GPU workload)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--segments", type=int, default=N_SEGMENTS, help=argparse.SUPPRESS)
    ap.add_argument("--barcodes", type=int, default=N_BARCODES, help=argparse.SUPPRESS)
    ap.add_argument("--max-error", type=int, default=MAX_ERROR, help=argparse.SUPPRESS)
    ap.add_argument("--no-cpu", action="store_true", help=argparse.SUPPRESS)
    return ap.parse_args()


def workload_name(a):
    return (f"match_trie {a.segments:,} segments x {a.barcodes:,} barcodes, -mr {a.max_error}, -bl 16 per GPU "
            f"(BASELINE configs[1]); synthetic, seeds 1002/2002+rank")


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: an in-process NVML polling thread (every 2 ms —
    the timed region of this workload is tens of milliseconds, too short for `nvidia-smi -lms`); falls back to the
    nvidia-smi recipe of B200_PROFILING.md when NVML cannot be loaded."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.sm, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        self._smi = None
        self._nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._h = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(index))
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self._nv = pynvml
        except Exception:
            self._nv = None

    @staticmethod
    def _physical_index(i):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [x for x in vis.split(",") if x != ""]
            if i < len(ids) and ids[i].isdigit():
                return int(ids[i])
        return i

    def _poll(self):
        nv = self._nv
        bits = (("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40),
                ("sw_power_cap", 0x4))          # nvmlClocksEventReason* bit values
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                for name, bit in bits:
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.002)

    def start(self):
        if self._nv is not None:
            self._thread = threading.Thread(target=self._poll, daemon=True)
            self._thread.start()
            return
        try:
            self._f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            self._smi = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=self._f,
                                         stderr=subprocess.DEVNULL)
        except OSError:
            self._smi = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "how": None}
        if self._thread is not None:
            self._stop.set()
            self._thread.join(timeout=2)
            if self.sm:
                out.update(sm_mhz=float(np.median(self.sm)), sm_max_mhz=self.max_mhz, reasons=sorted(self.reasons),
                           samples=len(self.sm), how="NVML polled every 2 ms inside the timed region")
            return out
        if self._smi is None:
            return out
        self._smi.terminate()
        try:
            self._smi.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self._smi.kill()
        self._f.flush()
        self._f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self._f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.unlink(self._f.name)
        except OSError:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm),
                       how="nvidia-smi -lms 100 during the timed region")
        return out


def cpu_arm(barcodes, n_sample, max_error, seed, repeats=1):
    """Time the CPU port of the reference's algorithm (trie build + error-bounded DFS per prefix task) on a bounded
    sample.  -> (segments/s, cores, seconds)"""
    from oracle import oracle
    from sctagger_b200 import synth
    cores = os.cpu_count() or 1
    sb, off = synth.make_segments(barcodes, n_sample, seed)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        oracle.match(barcodes, sb, off, max_error, threads=cores, workers=cores, algo="trie")   # reference: -t = host cores
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return n_sample / best, cores, best


def run_reference(a, rank):
    """--impl reference: the CPU arm alone, rank 0 only (other ranks exit 0 without work).  -> the JSON line"""
    if rank != 0:
        return None
    from sctagger_b200 import synth
    bc = synth.make_barcodes(a.barcodes, synth.SEEDS["cfg2"])
    n = min(CPU_SAMPLE, a.segments)
    for i in range(a.warmup):
        cpu_arm(bc, n, a.max_error, 2002 + i)
    t = 0.0
    cores = os.cpu_count() or 1
    for i in range(a.steps):
        _, cores, dt = cpu_arm(bc, n, a.max_error, 3002 + i)
        t += dt
    v = n * a.steps / t
    sample = f"{n:,} of the workload's {a.segments:,} segments per step x {a.barcodes:,} barcodes, -t {cores}"
    return json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * t / a.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": workload_name(a), "cpu_sample": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "CPU port (oracle/trie_oracle.c) of the reference's trie build + error-bounded DFS, one OpenMP thread "
                "per core over the reference's own prefix tasks; the reference itself is a Python script that is not "
                "present on the GPU box (its measured throughput in the build container is in BASELINE.md)",
    })


def main():
    # everything but the one JSON line goes to STDERR, also what libraries write to file descriptor 1 themselves
    # (NCCL prints its version banner there when NCCL_DEBUG is set on the box)
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        line = _main()
    finally:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    if line is not None:
        print(line, flush=True)


def _main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference":
        return run_reference(a, rank)

    import torch
    import torch.distributed as dist
    from sctagger_b200 import _lib, synth

    if not torch.cuda.is_available() or _lib.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: sctagger_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    bc = synth.make_barcodes(a.barcodes, synth.SEEDS["cfg2"])
    sb, off = synth.make_segments(bc, a.segments, 2002 + rank)
    n = a.segments
    total_len = int(off[-1])
    m = _lib.Matcher(bc, a.max_error, THREADS_FLAG, mem_gb=16.0, device=local)

    # ---- INT32 issue peak of this GPU (roofline denominator), measured live -----------------------------------
    peak_lop3 = _lib.measure_int32_peak(local, 0, 5)
    peak_mix = _lib.measure_int32_peak(local, 2, 5)

    # ---- value: batch resident in HBM ---------------------------------------------------------------------------
    m.stage(sb, off)
    for _ in range(a.warmup):
        m.flush_l2()
        m.run_staged()
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    dev_ms = kern_ms = 0.0
    launches = 0
    for _ in range(a.steps):
        m.flush_l2()
        p, k = m.run_staged()
        dev_ms += p
        kern_ms += k
        launches += m.stats()["launches"]
    barrier()
    clocks = sampler.stop()
    stats = m.stats()
    res_staged = m.fetch_staged()
    dev_ms_max = max_over_ranks(dev_ms)
    value = world * n * a.steps / (dev_ms_max * 1e-3)

    # ---- e2e: host buffers in, host results out, every step -----------------------------------------------------
    for _ in range(min(a.warmup, 2)):
        m.match(sb, off)
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        res = m.match(sb, off)
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    st_e2e = m.stats()
    e2e_value = world * n * a.steps / e2e_s
    same = res.equals(res_staged)
    rows = int((res.min_ed >= 0).sum())
    checksum = int((res.min_ed.astype(np.int64) + 1).sum() * 1_000_003 + res.tie_ids.astype(np.int64).sum())

    # ---- roofline of the dominant kernel -------------------------------------------------------------------------
    work = OPS_PER_COLUMN * 2.0 * a.barcodes * total_len          # lane-ops per launch
    kern_s = kern_ms * 1e-3 / a.steps
    achieved = work / kern_s
    all_rows = sum_over_ranks(rows)
    seeded = stats["seed_candidates"] > 0
    kernel_name = (f"match_seed_kernel<{a.max_error},true>" if seeded else f"match_fast_kernel<{a.max_error},true>")
    prof = _ncu_profile(f"r01_ncu_match_seed_K{a.max_error}.txt") if seeded else {}
    executed = None
    if seeded and prof.get("alu_pct") and prof.get("ms"):
        # ALU-pipe lane-ops the kernel actually issues, per verified candidate, from the committed ncu capture
        # (pipe utilisation x measured LOP3 peak x kernel time / candidates); scaled by this run's candidates and time.
        cand = stats["seed_candidates"]
        per_cand = prof["alu_pct"] / 100.0 * 18.5e12 * prof["ms"] * 1e-3 / (prof.get("cand_per_launch") or cand)
        executed = {"candidates_per_segment": cand / n, "alu_lane_ops_per_candidate": per_cand,
                    "alu_T_lane_ops_per_s": cand * per_cand / kern_s / 1e12,
                    "frac_of_peak": cand * per_cand / kern_s / peak_lop3,
                    "ncu_alu_pipe_pct": prof["alu_pct"], "ncu_issue_ipc": prof.get("ipc"),
                    "source": f"profiles/r01_ncu_match_seed_K{a.max_error}.txt"}
    m.close()

    out = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dev_ms_max / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32", "data": "synthetic",
            "config": {"workload": workload_name(a), "segments_per_gpu": n, "barcodes": a.barcodes,
                       "max_error": a.max_error, "barcode_length": 16, "mean_segment_length": total_len / n,
                       "parallelism": f"reads sharded over {world} GPU(s), barcode table replicated, no collective",
                       "l2": "flushed between timed steps (256 MiB memset)"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": st_e2e["h2d_bytes"],
                    "d2h_bytes_per_step": st_e2e["d2h_bytes"], "ms_per_step": 1e3 * e2e_s / a.steps,
                    "api": "sctagger_b200._lib.Matcher.match -> sct_match (host numpy in/out)"},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "int32_alu", "kernel": kernel_name,
                         "achieved": achieved / 1e12, "peak": peak_lop3 / 1e12, "unit": "T lane-op/s",
                         "frac": achieved / peak_lop3, "traffic": prof.get("dram_bytes"),
                         "note": "achieved = ALGORITHMIC lane-ops (14 per Myers column update x 2B x sum(len), no credit "
                                 "taken for skipped work) / kernel time; the seeded kernel verifies only the pigeonhole "
                                 "candidates, so frac reads far above 1 - kernel quality is `executed` (ALU-pipe "
                                 "utilisation of the instructions actually issued)",
                         "executed": executed,
                         "peak_source": "measured live: sct_measure_int32_peak (independent LOP3 chains, all SMs)",
                         "peak_lop3_plus_imad": peak_mix / 1e12,
                         "algorithmic_ops_per_launch": work, "kernel_ms": kern_ms / a.steps,
                         "kernel_share_of_step": kern_ms / dev_ms,
                         "hbm_minor": {"algorithmic_GBps": (total_len + 24.0 * n) / kern_s / 1e9,
                                       "peak_GBps": _peaks().get("hbm_gbs")}},
            "check": {"e2e_equals_staged": bool(same), "rows_rank0": rows, "rows_all": all_rows,
                      "checksum_rank0": checksum, "fast_segments": stats["fast_segments"],
                      "generic_segments": stats["generic_segments"]},
        }
        if not a.no_cpu and world == 1:
            v, cores, dt = cpu_arm(bc, min(CPU_SAMPLE, n), a.max_error, 3002)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{min(CPU_SAMPLE, n):,} segments of the same generator x {a.barcodes:,} "
                                              f"barcodes, {dt:.1f} s, oracle/trie_oracle.c with -t {cores}"}
        out = json.dumps(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def _ncu_profile(name):
    """Numbers of the committed ncu capture of the dominant kernel (profiles/<name>), launch 0."""
    out = {}
    try:
        with open(os.path.join(ROOT, "profiles", name)) as f:
            lines = f.read().split("--- launch ID 1")[0].splitlines()
    except OSError:
        return out
    rd = wr = None
    for ln in lines:
        c = ln.split()
        if not c:
            continue
        try:
            v = float(c[-1].replace(",", ""))
        except ValueError:
            continue
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(c[1] if len(c) > 2 else "", 1.0)
        if c[0] == "gpu__time_duration.sum":
            out["ms"] = v
        elif c[0] == "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active":
            out["alu_pct"] = v
        elif c[0] == "smsp__issue_active.avg.per_cycle_active":
            out["ipc"] = v
        elif c[0] == "dram__bytes_read.sum":
            rd = v * scale
        elif c[0] == "dram__bytes_write.sum":
            wr = v * scale
        elif c[0] == "#" and "candidates_per_launch" in ln:
            out["cand_per_launch"] = v
    if rd is not None and wr is not None:
        out["dram_bytes"] = rd + wr
    return out


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except OSError:
        return {}


if __name__ == "__main__":
    main()
