#!/usr/bin/env python3
"""bench.py -- decompressed GB/s of the B200 bzip2 path on BASELINE.json's headline workload
(config 3: 8 GB mixed-entropy synthetic, bzip2 -9, full decode with CRC verify).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over the whole file: magic finder + Huffman/MTF + inverse BWT + RLE1 + CRC.
With N ranks the compressed stream is cut into N byte ranges; each rank owns the blocks whose magic starts in its
range and decodes them on its own GPU (no collective: blocks are independent).  torch.distributed is used only for
the barrier and the max-over-ranks of the timings.

  value  device-resident: compressed bytes and decoded bytes stay in HBM, timed with CUDA events on the library's
         stream around exactly K steps.
  e2e    the same through the C ABI with pinned HOST buffers: H2D of the compressed bytes and D2H of the decoded
         bytes happen inside every timed step.
  --impl reference  the reference's own ParallelBZ2Reader (compiled unmodified into oracle/_ref) on the box's host
         cores, on a bounded sample (one copy of the unique base, see benchdata.py) of the same workload.

Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import benchdata  # noqa: E402

METRIC = "decompressed GB/s (bzip2 -9, 8 GB synthetic)"
UNIT = "GB/s"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--kind", default=os.environ.get("IBZ2_BENCH_KIND", "mixed"), choices=list(benchdata.GENERATORS))
    ap.add_argument("--level", type=int, default=int(os.environ.get("IBZ2_BENCH_LEVEL", "9")))
    ap.add_argument("--base-bytes", type=int, default=int(os.environ.get("IBZ2_BENCH_BASE", str(1_000_000_000))))
    ap.add_argument("--copies", type=int, default=int(os.environ.get("IBZ2_BENCH_COPIES", "8")))
    ap.add_argument("--batch-blocks", type=int, default=int(os.environ.get("IBZ2_BATCH_BLOCKS", "0")))
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def get_workload(args, rank: int, barrier):
    seed = {"mixed": 3, "base64": 1, "longruns": 4}[args.kind]
    base = (args.base_bytes + benchdata.SEGMENT - 1) // benchdata.SEGMENT * benchdata.SEGMENT
    key = f"{args.kind}-l{args.level}-base{base}-x{args.copies}-seed{seed}"

    def build():
        return benchdata.build_single_stream(args.kind, seed, base, args.copies, args.level, log=log)

    w = None
    if rank == 0:
        t0 = time.time()
        w = benchdata.cached(build, key, log=log)
        log(f"[bench] workload ready in {time.time() - t0:.1f}s: {w.meta}")
    barrier()
    if rank != 0:
        w = benchdata.cached(build, key, log=None)
    return w, key


def workload_config(w, args) -> dict:
    return {
        "workload": f"config 3: {w.raw_size / 1e9:.2f} GB {args.kind} synthetic (alternating 1 MiB text-like / random "
                    f"segments), bzip2 -{args.level}, ONE stream of {w.blocks} blocks = the blocks of a "
                    f"{w.raw_base.size / 1e9:.2f} GB unique base repeated x{w.copies} (bit-spliced, every copy at its own "
                    f"bit alignment and address); full decode + per-block CRC verify",
        "raw_bytes": w.raw_size, "compressed_bytes": int(w.compressed.size), "blocks": w.blocks,
        "l2": "inputs and outputs larger than L2 (no flush needed)",
    }


# ---------------------------------------------------------------------------------------------------------------
# the reference on the host cores (oracle/_ref = the unmodified reference behind a C shim; test infrastructure)
# ---------------------------------------------------------------------------------------------------------------

def time_reference(path: Path, raw_bytes: int, cores: int, steps: int, warmup: int):
    import _oracle as orc

    if not orc.ref_available():
        return None
    times = []
    for i in range(warmup + steps):
        r = orc.ref_decode_file(str(path), parallelization=cores, serial_class=False, want_data=False)
        if r.rc != 0:
            raise RuntimeError(f"reference failed: {r.error}")
        if i >= warmup:
            times.append(r.seconds)
    return times


def time_oracle_port(sample: np.ndarray, steps: int, warmup: int):
    import _oracle as orc

    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        r = orc.parallel_reader(sample, want_data=False)
        if r.status != 0:
            raise RuntimeError(f"oracle failed with status {r.status}")
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return times


def cpu_reference_run(w, steps: int, warmup: int):
    """(GB/s, cpu_baseline dict) of the reference's ParallelBZ2Reader on one copy of the unique base."""
    cores = host_cores()
    sample_raw = int(w.raw_base.size)
    sample_desc = (f"one copy of the unique base as its own bzip2 stream: {sample_raw / 1e9:.2f} GB raw, "
                   f"{w.sample.size / 1e9:.2f} GB compressed, read from /dev/shm page cache")
    times = None
    kind = "reference"
    if w.sample_path is not None and Path(w.sample_path).exists():
        times = time_reference(Path(w.sample_path), sample_raw, cores, steps, warmup)
    if times is None:
        kind = "port"
        cores = 1
        sample_desc = "oracle port (oracle/bz2_oracle.c), single thread; compiled reference not available: " + sample_desc
        times = time_oracle_port(w.sample, steps, warmup)
    sec = sum(times) / len(times)
    gbs = sample_raw / sec / 1e9
    return gbs, sec, {"value": round(gbs, 4), "unit": UNIT, "cores": cores, "kind": kind, "sample": sample_desc,
                      "impl": "ParallelBZ2Reader(StandardFileReader(path), cores).read(-1, nullptr, SIZE_MAX), "
                              "built -O3 -DNDEBUG from the unmodified reference headers (oracle/ref_shim.cpp)"
                      if kind == "reference" else "oracle/bz2_oracle.c"}


# ---------------------------------------------------------------------------------------------------------------
# clocks during the timed region (recipe of /opt/skills/guides/B200_PROFILING.md)
# ---------------------------------------------------------------------------------------------------------------

class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.tmp = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(gpu_index)], stdout=self.tmp, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.tmp.flush()
        rows = [r.split(",") for r in Path(self.tmp.name).read_text().splitlines() if r.count(",") >= 7]
        os.unlink(self.tmp.name)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except ValueError:
                continue
            for name, v in zip(names, r[4:8]):
                if v.strip().lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------------------

STAGES = ["find_ms", "huffman_ms", "bwt_prepare_ms", "bwt_traverse_ms", "rle_scan_ms", "rle_write_ms", "crc_ms"]
# timings that cover ONE kernel launch per wave (CUDA events on the library's stream around that launch): the roofline
# is quoted for the slowest of these
KERNELS = {"find_ms": "K1 findMagicsKernel", "huffman_bits_ms": "K2b huffmanBitsKernel (+ its 2 ms resolve pass)",
           "bwt_prepare_ms": "K4a bwtPrepareKernel", "bwt_walk_ms": "K4b bwtWalkKernel", "rle_scan_ms": "K5a rleScanKernel",
           "rle_write_ms": "K5b rleWriteKernel", "crc_ms": "K6 crcKernel"}


# DRAM bytes (read + write) of one launch from the ncu capture of this same command on one B200
# (profiles/r1_v25_launches.csv: dram__bytes_read.sum + dram__bytes_write.sum, summed over the launches of a step);
# only valid for the default workload at N=1
NCU_TRAFFIC_8GB = {"find_ms": 4.9e9, "huffman_bits_ms": 16.5e9, "bwt_prepare_ms": 49.7e9, "bwt_walk_ms": 107.4e9,
                   "rle_scan_ms": 20.2e9, "rle_write_ms": 24.3e9, "crc_ms": 8.0e9}


def algorithmic_bytes(stage: str, c: int, n: int, u: int) -> int:
    """per-step algorithmic HBM bytes of each kernel (DESIGN.md "Kernels": SURVEY.md section 8(d) per-block figures
    x the blocks one step processes): C compressed bytes, N BWT symbols, U decoded bytes"""
    return {"find_ms": c, "huffman_ms": c + n, "huffman_bits_ms": c + 2 * n, "bwt_prepare_ms": 5 * n,
            "bwt_traverse_ms": 5 * n, "bwt_walk_ms": 5 * n,
            "rle_scan_ms": 5 * n, "rle_write_ms": 5 * n + u, "crc_ms": u}[stage]


def run_ours(args):
    import torch
    import torch.distributed as dist

    from indexed_bzip2_b200 import _capi

    rank, world, local_rank = dist_env()
    if world != args.gpus:
        log(f"[bench] WORLD_SIZE={world} but --gpus {args.gpus}; using WORLD_SIZE")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(value: float, op) -> float:
        if world == 1:
            return value
        t = torch.tensor([value], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    RMAX = dist.ReduceOp.MAX if world > 1 else None
    RSUM = dist.ReduceOp.SUM if world > 1 else None

    w, key = get_workload(args, rank, barrier)
    comp = w.compressed
    C = int(comp.size)

    # this rank's share: blocks whose magic starts in [first_byte, last_byte); the window runs past last_byte so that
    # the last owned block is complete (a block never exceeds ~1.1 MB compressed)
    from indexed_bzip2_b200.partition import rank_share

    share = rank_share(C, rank, world)
    window = comp[share.first_byte:share.end_byte]
    first_bit, last_bit = share.first_bit, share.last_bit
    out_cap = int(w.raw_size / world * 1.1) + (64 << 20)

    ctx = _capi.Context(local_rank)
    if args.batch_blocks:
        ctx.set_batch_blocks(args.batch_blocks)

    # ---- device-resident ("value") ----
    d_in = ctx.device_alloc(window.size)
    d_out = ctx.device_alloc(out_cap)
    ctx.h2d(d_in, window)
    max_blocks = w.blocks + 16

    def step_resident():
        return ctx.decode_buffer(d_in, d_out, out_cap, nbytes=window.size, data_device=True, out_device=True,
                                 first_bit=first_bit, last_bit=last_bit, max_blocks=max_blocks)

    for _ in range(args.warmup):
        results, total = step_resident()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    stage_ms = {s: 0.0 for s in STAGES + ["huffman_bits_ms", "bwt_walk_ms"]}
    launches = 0
    ctx.timer_start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        results, total = step_resident()
        t = ctx.timings()
        for s in stage_ms:
            stage_ms[s] += t[s]
        launches += t["launches"]
    dev_ms = ctx.timer_stop()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    barrier()
    clocks = sampler.stop() if sampler else None

    bad = int((results["status"] != 0).sum())
    if bad:
        raise RuntimeError(f"rank {rank}: {bad} blocks failed, first status {results['status'][results['status'] != 0][0]}")
    n_symbols = int(results["n_symbols"].sum())
    my_c = int((results["enc_size_bits"].sum() + 7) // 8)
    ms_dev = reduce(dev_ms, RMAX) / args.steps
    ms_wall = reduce(wall_ms, RMAX) / args.steps
    total_bytes = reduce(float(total), RSUM)
    total_launches = reduce(float(launches), RSUM)

    # ---- end to end through the C ABI with pinned host buffers ----
    pin_in = ctx.pinned(window.size)
    pin_out = ctx.pinned(out_cap)
    pin_in.array[:] = window

    def step_e2e():
        return ctx.decode_buffer(pin_in.array, pin_out.array, first_bit=first_bit, last_bit=last_bit, max_blocks=max_blocks)

    for _ in range(max(1, min(args.warmup, 2))):
        step_e2e()
    barrier()
    ctx.timer_start()
    t0 = time.perf_counter()
    e2e_steps = args.steps
    for _ in range(e2e_steps):
        res_e, total_e = step_e2e()
    e2e_timings = {k: round(float(v), 3) for k, v in ctx.timings().items()}   # of the last step: what was exposed
    e2e_dev_ms = ctx.timer_stop()
    e2e_wall_ms = (time.perf_counter() - t0) * 1e3
    barrier()
    e2e_ms = reduce(max(e2e_dev_ms, e2e_wall_ms), RMAX) / e2e_steps
    h2d_bytes = reduce(float(window.size), RSUM)
    d2h_bytes = reduce(float(total_e), RSUM)

    # ---- parity of what was just timed: bytes == the raw data, every block CRC == header CRC ----
    verified = None
    if not args.no_verify:
        if world > 1:
            totals = [torch.zeros(1, dtype=torch.float64, device="cuda") for _ in range(world)]
            dist.all_gather(totals, torch.tensor([float(total_e)], dtype=torch.float64, device="cuda"))
            offset = int(sum(t.item() for t in totals[:rank]))
        else:
            offset = 0
        ok = bool((res_e["calculated_crc"] == res_e["expected_crc"]).all()) and int(total_e) == int(total)
        out = pin_out.array[:total_e]
        base = w.raw_base
        pos = 0
        while ok and pos < total_e:
            b = (offset + pos) % base.size
            n = min(base.size - b, total_e - pos)
            ok = bool(np.array_equal(out[pos:pos + n], base[b:b + n]))
            pos += n
        host = np.empty(min(int(total), 256 << 20), dtype=np.uint8)  # and a slice of the device-resident output
        ctx.d2h(host, d_out)
        ok = ok and bool(np.array_equal(host, out[:host.size]))
        verified = bool(reduce(1.0 if ok else 0.0, dist.ReduceOp.MIN if world > 1 else None) > 0.5)
        if world == 1 and int(total_bytes) != w.raw_size:
            verified = False

    line = None
    if rank == 0:
        value = total_bytes / (ms_dev * 1e-3) / 1e9
        peaks_path = ROOT / "MEASURED_PEAKS.json"
        if peaks_path.exists():
            peak, peak_src = float(json.loads(peaks_path.read_text())["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy)"
        else:
            peak, peak_src = 6650.0, "fallback of B200_PROFILING.md"
        per_step = {s: stage_ms[s] / args.steps for s in stage_ms}
        dominant = max(KERNELS, key=lambda k: per_step[k])
        waves = max(1, -(-len(results) // (args.batch_blocks or 148 * 64)))
        alg = algorithmic_bytes(dominant, my_c, n_symbols, int(total))
        achieved = alg / (per_step[dominant] * 1e-3) / 1e9
        roofline = {"bound": "hbm", "kernel": KERNELS[dominant], "achieved": round(achieved, 2), "peak": peak,
                    "unit": "GB/s", "frac": round(achieved / peak, 5),
                    "traffic": (NCU_TRAFFIC_8GB[dominant] if (world == 1 and args.kind == "mixed" and args.copies == 8
                                                             and args.base_bytes == 1_000_000_000 and args.level == 9) else None),
                    "traffic_source": "profiles/r1_v25_launches.csv (ncu, same command, launches of one step)",
                    "peak_source": peak_src,
                    "launches_per_step": waves, "ms_per_launch": round(per_step[dominant] / waves, 4),
                    "algorithmic_bytes_per_step": alg, "rank": 0,
                    "whole_path": {"achieved": round((my_c + int(total)) / (ms_dev * 1e-3) / 1e9, 2),
                                   "frac": round((my_c + int(total)) / (ms_dev * 1e-3) / 1e9 / peak, 5),
                                   "bytes": "C + U of rank 0 per step"},
                    "stage_ms_per_step": {k.replace("_ms", ""): round(v, 3) for k, v in per_step.items()}}
        line = {
            "metric": METRIC, "value": round(value, 4), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_dev, 3), "wall_ms_per_step": round(ms_wall, 3),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": dict(workload_config(w, args), parallelism=f"block ranges over {world} GPU(s), no collective"),
            "e2e": {"value": round(d2h_bytes / (e2e_ms * 1e-3) / 1e9, 4), "unit": UNIT,
                    "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": int(d2h_bytes),
                    "ms_per_step": round(e2e_ms, 3), "api": "ibz2_decode_buffer, pinned host in/out",
                    "last_step_timings_rank0": e2e_timings},
            "gpu_launches": int(total_launches), "clocks": clocks, "roofline": roofline, "verified": verified,
        }
        if world == 1 and not args.no_cpu_baseline:
            try:
                _, _, cpu = cpu_reference_run(w, steps=1, warmup=0)
                line["cpu_baseline"] = cpu
            except Exception as e:  # the baseline is reported, never allowed to sink the bench line
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": host_cores(), "kind": "reference",
                                        "sample": f"failed: {e}"}
    pin_in.free()
    pin_out.free()
    ctx.device_free(d_in)
    ctx.device_free(d_out)
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line), flush=True)


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    w, _ = get_workload(args, 0, lambda: None)
    gbs, sec, cpu = cpu_reference_run(w, steps=args.steps, warmup=min(args.warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbs, 4), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": round(sec * 1e3, 3), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": dict(workload_config(w, args), parallelism=f"{cpu['cores']} host threads",
                       step=f"bounded sample per step: {cpu['sample']}"),
        "cpu_baseline": cpu,
        "e2e": {"value": round(gbs, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
