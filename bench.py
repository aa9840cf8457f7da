#!/usr/bin/env python3
"""bench.py -- decompressed GB/s of the B200 bzip2 path on BASELINE.json's headline workload
(config 3: 8 GB mixed-entropy synthetic, bzip2 -9, full decode with CRC verify) and, with --config, on the other four
configurations of BASELINE.json:

    --config 1   100 MB base64-random, bzip2 -9 (the reference's own CPU-runnable case, ParallelBZ2Reader P=1)
    --config 2   BlockFinder-only scan of a 4 GiB bzip2 bit stream: GB/s scanned, offsets equal to the reference's finder
    --config 3   8 GB mixed-entropy, bzip2 -9, full decode + CRC                                  (default, the headline)
    --config 4   8 GB long runs, bzip2 -1 (100 kB blocks that decode to ~5 MB each)
    --config 5   importIndex, then 10 000 x (seek + 1 MiB read) at random offsets of an 8 GB pbzip2-style multistream file

    python bench.py [--config C] [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
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
    ap.add_argument("--config", type=int, default=int(os.environ.get("IBZ2_BENCH_CONFIG", "3")), choices=[1, 2, 3, 4, 5])
    ap.add_argument("--kind", default=os.environ.get("IBZ2_BENCH_KIND"), choices=list(benchdata.GENERATORS))
    ap.add_argument("--level", type=int, default=int(os.environ.get("IBZ2_BENCH_LEVEL", "0")))
    ap.add_argument("--base-bytes", type=int, default=int(os.environ.get("IBZ2_BENCH_BASE", "0")))
    ap.add_argument("--copies", type=int, default=int(os.environ.get("IBZ2_BENCH_COPIES", "0")))
    ap.add_argument("--reads", type=int, default=10_000, help="config 5: number of seek + read operations")
    ap.add_argument("--batch-blocks", type=int, default=int(os.environ.get("IBZ2_BATCH_BLOCKS", "0")))
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    # what each configuration decodes (the decode configurations share one code path); explicit flags override
    preset = {1: ("base64", 9, 100_000_000, 1), 2: ("mixed", 9, 1_000_000_000, 8), 3: ("mixed", 9, 1_000_000_000, 8),
              4: ("longruns", 1, 1_000_000_000, 8), 5: ("mixed", 9, 1_000_000_000, 8)}[args.config]
    args.kind = args.kind or preset[0]
    args.level = args.level or preset[1]
    args.base_bytes = args.base_bytes or preset[2]
    args.copies = args.copies or preset[3]
    return args


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
    if args.config == 5:
        key += "-multistream"

    def build():
        if args.config == 5:
            return benchdata.build_multistream_repeated(args.kind, seed, base, args.copies, 900_000, args.level, log=log)
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


KIND_TEXT = {"mixed": "mixed synthetic (alternating 1 MiB text-like / random segments)",
             "base64": "base64 of random bytes (76-column lines)",
             "longruns": "long runs of one byte value (log-uniform lengths 1..1e5)"}
L2_BYTES = 126 << 20


def workload_config(w, args) -> dict:
    repeated = (f" = the blocks of a {w.raw_base.size / 1e9:.2f} GB unique base repeated x{w.copies} (bit-spliced, every "
                f"copy at its own bit alignment and address)") if w.copies > 1 else ""
    small = int(w.compressed.size) < 4 * L2_BYTES
    return {
        "workload": f"config {args.config}: {w.raw_size / 1e9:.2f} GB {KIND_TEXT[args.kind]}, bzip2 -{args.level}, ONE "
                    f"stream of {w.blocks} blocks{repeated}; full decode + per-block CRC and stream CRC verify",
        "raw_bytes": w.raw_size, "compressed_bytes": int(w.compressed.size), "blocks": w.blocks,
        "l2": ("L2 flushed between timed steps (a 512 MB buffer is overwritten), every step timed on its own" if small
               else "inputs and outputs larger than L2 (no flush needed)"),
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


def _compress_slice(job):
    import bz2

    data, level = job
    return bz2.compress(data, level)


def system_bzip2_run(w, level: int, sample_bytes: int = 100 << 20):
    """`bzip2 -dc` (the system's libbz2 CLI, one core) on a bounded sample: the first
    sample_bytes of the raw data, compressed here in 10 MB streams and concatenated (bzip2 reads multi-stream files)."""
    import shutil
    from concurrent.futures import ProcessPoolExecutor

    exe = shutil.which("bzip2")
    if exe is None:
        return {"value": None, "unit": UNIT, "cores": 1, "kind": "system bzip2", "sample": "bzip2 not installed"}
    raw = w.raw_base[:min(sample_bytes, w.raw_base.size)]
    piece = 10_000_000
    with ProcessPoolExecutor(max_workers=host_cores()) as pool:
        parts = list(pool.map(_compress_slice, [(raw[i:i + piece].tobytes(), level) for i in range(0, raw.size, piece)]))
    path = Path(os.environ.get("IBZ2_BENCH_CACHE", "/dev/shm/ibz2_bench")) / "system_bzip2_sample.bz2"
    path.parent.mkdir(parents=True, exist_ok=True)
    path.write_bytes(b"".join(parts))
    t0 = time.perf_counter()
    rc = subprocess.run([exe, "-dc", str(path)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL).returncode
    sec = time.perf_counter() - t0
    path.unlink()
    return {"value": round(raw.size / sec / 1e9, 4) if rc == 0 else None, "unit": UNIT, "cores": 1, "kind": "system bzip2",
            "sample": f"`bzip2 -dc` of the first {raw.size / 1e6:.0f} MB of the raw data (10 MB streams, /dev/shm), "
                      f"{sec:.2f} s wall", "impl": subprocess.run([exe, "--version"], capture_output=True, text=True,
                                                                    stdin=subprocess.DEVNULL).stderr.split("\n")[0]}


def reader_api_run(w):
    """What a user of the drop-in surface gets: `ibzip2 -d -c -P 0 file > /dev/null` (our CLI: file -> mmap -> HBM ->
    ParallelBZ2Reader -> write()) on the whole file, wall clock including process start, CUDA context and upload; next to
    it the reference's CLI on the bounded sample."""
    import _oracle as orc

    out = {"unit": UNIT}
    cli = ROOT / "indexed_bzip2_b200" / "bin" / "ibzip2"
    if w.path is None or not cli.exists():
        return dict(out, value=None, note="no cached file or CLI not built")
    best = None
    for _ in range(2):
        t0 = time.perf_counter()
        rc = subprocess.run([str(cli), "-d", "-c", "-P", "0", str(w.path)], stdout=subprocess.DEVNULL,
                            stderr=subprocess.DEVNULL).returncode
        sec = time.perf_counter() - t0
        if rc != 0:
            return dict(out, value=None, note=f"ibzip2 exited with {rc}")
        best = sec if best is None else min(best, sec)
    out.update(value=round(w.raw_size / best / 1e9, 3), seconds=round(best, 3),
               command="ibzip2 -d -c -P 0 <file> > /dev/null (ours), best of 2, wall clock of the whole process")
    if orc.ref_available() and orc.REF_CLI.exists() and w.sample_path is not None:
        t0 = time.perf_counter()
        rc = subprocess.run([str(orc.REF_CLI), "-d", "-c", "-P", "0", str(w.sample_path)], stdout=subprocess.DEVNULL,
                            stderr=subprocess.DEVNULL).returncode
        sec = time.perf_counter() - t0
        if rc == 0:
            out["reference_cli"] = {"value": round(w.raw_base.size / sec / 1e9, 4), "unit": UNIT, "cores": host_cores(),
                                    "seconds": round(sec, 3), "sample": "the reference's ibzip2 -d -c -P 0 on one copy of the base"}
    return out


def reference_single_thread_runs(w):
    """config 1 as BASELINE.json words it: the reference with parallelization=1, in process (ParallelBZ2Reader(file, 1):
    the decode runs deferred on the caller, core/BlockFetcher.hpp:183-184) and through its CLI, where -P 1 selects the
    serial BZ2Reader (tools/ibzip2.cpp:374-379)."""
    import _oracle as orc

    out = {}
    if not orc.ref_available() or w.sample_path is None:
        return out
    r = orc.ref_decode_file(str(w.sample_path), parallelization=1, serial_class=False, want_data=False)
    if r.rc == 0:
        out["ParallelBZ2Reader_P1"] = {"value": round(w.raw_base.size / r.seconds / 1e9, 5), "unit": UNIT, "cores": 1,
                                       "seconds": round(r.seconds, 3)}
    if orc.REF_CLI.exists():
        t0 = time.perf_counter()
        rc = subprocess.run([str(orc.REF_CLI), "-d", "-c", "-P", "1", str(w.sample_path)], stdout=subprocess.DEVNULL,
                            stderr=subprocess.DEVNULL).returncode
        sec = time.perf_counter() - t0
        if rc == 0:
            out["ibzip2_cli_P1"] = {"value": round(w.raw_base.size / sec / 1e9, 5), "unit": UNIT, "cores": 1,
                                    "seconds": round(sec, 3), "note": "-P 1 = the serial BZ2Reader class"}
    return out


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
SUB_STAGES = ["huffman_bits_ms", "mtf_chunk_ms", "bwt_walk_ms"]
# timings that cover ONE kernel launch per wave (CUDA events on the library's stream around that launch): the roofline
# is quoted for the slowest of these
KERNELS = {"find_ms": "K1 findMagicsKernel", "huffman_bits_ms": "K2b huffmanBitsKernel + resolveSymbolsKernel",
           "mtf_chunk_ms": "K2c mtfChunkKernel",
           "bwt_prepare_ms": "K4a bwtPrepareKernel", "bwt_walk_ms": "K4b bwtWalkKernel", "rle_scan_ms": "K5a rleScanKernel",
           "rle_write_ms": "K5b rleWriteKernel", "crc_ms": "K6 crcKernel"}


# DRAM bytes (read + write) per step of each timed stage, summed over its launches in the committed ncu launch list of
# this same command on one B200 (`--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum`); only
# valid for the default workload at N=1, and only as fresh as the list
TRAFFIC_CSV = ROOT / "profiles" / "r2_launches.csv"
TRAFFIC_CSV_CONFIG1 = ROOT / "profiles" / "r2_launches_cfg1.csv"
TRAFFIC_KERNELS = {"find_ms": ["findMagicsKernel"],
                   "huffman_bits_ms": ["huffmanBitsKernel", "huffmanGroupsKernel", "huffmanPipelineKernel", "resolveSymbolsKernel"],
                   "mtf_chunk_ms": ["mtfChunkKernel"],
                   "bwt_prepare_ms": ["bwtPrepareKernel"], "bwt_walk_ms": ["bwtWalkKernel"], "rle_scan_ms": ["rleScanKernel"],
                   "rle_write_ms": ["rleWriteKernel"], "crc_ms": ["crcKernel"]}


def k2b_kernel_name(blocks: int, sms: int = 148) -> str:
    """which of the two Huffman bit kernels a launch of `blocks` blocks takes (launchHuffmanBits, csrc/huffman_bits.cu)"""
    forced = os.environ.get("IBZ2_K2B_KERNEL", "")[:1]
    per_sm = int(os.environ.get("IBZ2_K2B_GROUPS_PER_SM", "2"))
    groups = forced in ("g", "p") or (forced != "l" and blocks <= sms * per_sm)
    first = forced == "g"
    return (("K2b huffmanGroupsKernel (a CTA per block)" if first else "K2b huffmanPipelineKernel (a CTA per block, pipelined)")
            if groups else "K2b huffmanBitsKernel (a lane per block)") \
        + " + resolveSymbolsKernel"


def ncu_traffic(stage: str, csv_path: Path = TRAFFIC_CSV):
    """(bytes per step, source) from the launch list: the launches between two findMagicsKernel launches are one step"""
    import csv

    TRAFFIC_CSV = csv_path
    if not TRAFFIC_CSV.exists():
        return None, "no launch list committed"
    rows = list(csv.reader(TRAFFIC_CSV.read_text().splitlines()))
    try:
        head = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    except StopIteration:
        return None, "launch list unreadable"
    h = rows[head]
    ik, im, iv, iid = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
    per_launch: dict[int, tuple[str, float]] = {}
    for r in rows[head + 1:]:
        if len(r) > iv and r[im] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            name, total = per_launch.get(int(r[iid]), (r[ik], 0.0))
            per_launch[int(r[iid])] = (name, total + float(r[iv].replace(",", "")))
    ids = sorted(per_launch)
    starts = [i for i in ids if "findMagicsKernel" in per_launch[i][0]]
    if len(starts) < 2:
        return None, "launch list holds less than one whole step"
    step = [per_launch[i] for i in ids if starts[0] <= i < starts[1]]
    total = sum(v for name, v in step if any(k in name for k in TRAFFIC_KERNELS[stage]))
    return total, f"{TRAFFIC_CSV.relative_to(ROOT)} (ncu, same command, the launches of one step)"


def algorithmic_bytes(stage: str, c: int, n: int, u: int, s: int = 0) -> int:
    """per-step algorithmic HBM bytes of each kernel (DESIGN.md "Kernels": SURVEY.md section 8(d) per-block figures
    x the blocks one step processes): C compressed bytes, N BWT symbols, U decoded bytes, S Huffman-coded symbols.
    SURVEY 8(d) gives the whole Huffman/MTF stage C + N; K2c on its own reads 2 bytes and writes 1 per coded symbol."""
    return {"find_ms": c, "huffman_ms": c + n, "huffman_bits_ms": c + n, "mtf_chunk_ms": 3 * s, "bwt_prepare_ms": 5 * n,
            "bwt_traverse_ms": 5 * n, "bwt_walk_ms": 5 * n,
            "rle_scan_ms": 5 * n, "rle_write_ms": 5 * n + u, "crc_ms": u}[stage]


def multi_api_run(w, devices: int, steps: int) -> dict:
    from indexed_bzip2_b200 import _capi

    comp = w.compressed
    with _capi.MultiContext(list(range(devices))) as multi:
        pin_in = multi.pinned(comp.size)
        pin_in[:] = comp
        out = multi.pinned(w.raw_size + (64 << 20))
        results, total = multi.decode_buffer(pin_in, out, max_blocks=w.blocks + 16)   # warm-up: contexts, scratch
        wall = []
        device_ms = []
        for _ in range(steps):
            t0 = time.perf_counter()
            results, total = multi.decode_buffer(pin_in, out, max_blocks=w.blocks + 16)
            wall.append((time.perf_counter() - t0) * 1e3)
            t = multi.timings()
            device_ms.append(max(d["total_ms"] for d in t["devices"]))
        ok = bool((results["status"] == 0).all()) and total == w.raw_size and w.verify(out[:total]) \
            and t["stream_crc_mismatches"] == 0 and bool(results[-1]["stream_flags"] & 4)
        ms = sum(wall) / len(wall)
        line = {"value": round(w.raw_size / (ms * 1e-3) / 1e9, 3), "unit": UNIT, "ms_per_step": round(ms, 3),
                "device_ms_max": round(sum(device_ms) / len(device_ms), 3), "devices": devices, "verified": ok,
                "api": "ibz2_multi_decode_buffer: host buffer in, ONE pinned host buffer out, one process; wall clock of "
                       "the call (its slowest device's CUDA-event time beside it)",
                "last_call": {k: v for k, v in t.items() if k != "devices"}}
        multi.free_pinned()
    return line


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
    stage_ms = {s: 0.0 for s in STAGES + SUB_STAGES}
    launches = 0
    # a workload that L2 could hold between two steps: flush (outside the timed regions) and time every step on its own
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda") if C < 4 * L2_BYTES else None
    dev_ms = 0.0
    if flush is None:
        ctx.timer_start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        if flush is not None:
            flush.fill_(1)
            torch.cuda.synchronize()
            ctx.timer_start()
        results, total = step_resident()
        if flush is not None:
            dev_ms += ctx.timer_stop()
        t = ctx.timings()
        for s in stage_ms:
            stage_ms[s] += t[s]
        launches += t["launches"]
    if flush is None:
        dev_ms = ctx.timer_stop()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3 if flush is None else dev_ms
    barrier()
    clocks = sampler.stop() if sampler else None

    bad = int((results["status"] != 0).sum())
    if bad:
        raise RuntimeError(f"rank {rank}: {bad} blocks failed, first status {results['status'][results['status'] != 0][0]}")
    n_symbols = int(results["n_symbols"].sum())
    n_coded = int(results["n_coded"].sum())
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
    e2e_steps = args.steps
    e2e_dev_ms = e2e_wall_ms = 0.0
    for _ in range(e2e_steps):
        if flush is not None:
            flush.fill_(1)
            torch.cuda.synchronize()
        ctx.timer_start()
        t0 = time.perf_counter()
        res_e, total_e = step_e2e()
        e2e_dev_ms += ctx.timer_stop()
        e2e_wall_ms += (time.perf_counter() - t0) * 1e3
    e2e_timings = {k: round(float(v), 3) for k, v in ctx.timings().items()}   # of the last step: what was exposed
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
        # the stream CRC (BZ2Reader.hpp:406-416, 481-484).  One rank: the call folded the whole stream itself.  Several
        # ranks: a share starts in the middle of the stream, so every rank folds its own block CRCs from zero and the
        # partial folds combine in rank order: fold is linear, n more blocks rotate what came before n times.
        if world == 1:
            last = res_e[-1]
            ok = ok and bool(last["stream_flags"] & 4) and int(last["stream_crc_calculated"]) == int(last["stream_crc_expected"])
        else:
            part = 0
            for c in res_e["calculated_crc"].tolist():
                part = (((part << 1) | (part >> 31)) & 0xFFFFFFFF) ^ c
            mine = torch.tensor([float(part), float(len(res_e)), float(res_e[-1]["stream_crc_expected"]),
                                 float(res_e[-1]["stream_flags"])], dtype=torch.float64, device="cuda")
            parts = [torch.zeros(4, dtype=torch.float64, device="cuda") for _ in range(world)]
            dist.all_gather(parts, mine)
            crc = 0
            for t in parts:
                rot = int(t[1].item()) % 32
                crc = (((crc << rot) | (crc >> (32 - rot))) & 0xFFFFFFFF) ^ int(t[0].item())
            ok = ok and (int(parts[-1][3].item()) & 2) != 0 and crc == int(parts[-1][2].item())
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
        alg = algorithmic_bytes(dominant, my_c, n_symbols, int(total), n_coded)
        default_workload = (world == 1 and args.config == 3 and args.kind == "mixed" and args.copies == 8
                            and args.base_bytes == 1_000_000_000 and args.level == 9)
        config1_workload = (world == 1 and args.config == 1 and args.kind == "base64" and args.copies == 1
                            and args.base_bytes == 100_000_000 and args.level == 9)
        traffic, traffic_src = (ncu_traffic(dominant) if default_workload
                                else ncu_traffic(dominant, TRAFFIC_CSV_CONFIG1) if config1_workload
                                else (None, "no capture for this workload"))
        achieved = alg / (per_step[dominant] * 1e-3) / 1e9
        kernel_name = KERNELS[dominant] if dominant != "huffman_bits_ms" else k2b_kernel_name(-(-len(results) // waves))
        roofline = {"bound": "hbm", "kernel": kernel_name, "achieved": round(achieved, 2), "peak": peak,
                    "unit": "GB/s", "frac": round(achieved / peak, 5),
                    "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": peak_src,
                    "launches_per_step": waves, "ms_per_launch": round(per_step[dominant] / waves, 4),
                    "algorithmic_bytes_per_step": alg, "rank": 0,
                    "symbols": {"C": my_c, "N": n_symbols, "S": n_coded, "U": int(total)},
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
            try:
                line["cpu_baseline"]["system_bzip2"] = system_bzip2_run(w, args.level)
                if args.config == 1:
                    line["cpu_baseline"]["single_thread"] = reference_single_thread_runs(w)
            except Exception as e:
                line["cpu_baseline"]["system_bzip2"] = {"value": None, "sample": f"failed: {e}"}
    pin_in.free()
    pin_out.free()
    ctx.device_free(d_in)
    ctx.device_free(d_out)
    ctx.close()
    if line is not None and world == 1 and not args.no_cpu_baseline:
        # the drop-in surface as a user runs it, once this process has given its pinned memory and HBM back
        try:
            line["reader_api"] = reader_api_run(w)
        except Exception as e:
            line["reader_api"] = {"value": None, "note": f"failed: {e}"}
    if world > 1:
        # the same job through the product's own multi-GPU call (ibz2_multi_decode_buffer: ONE process, one host thread
        # and one context per GPU, one pinned output buffer); the other ranks have released their GPUs and wait
        # (a CPU-side barrier: an NCCL barrier would keep a kernel spinning on the waiting ranks' GPUs, which are the
        # GPUs rank 0's call is using -- measured: 508 ms per call with it, 325 ms without)
        dist.barrier()
        torch.cuda.synchronize()
        waiting = dist.new_group(backend="gloo")
        if rank == 0 and os.environ.get("IBZ2_BENCH_MULTI", "1") != "0":
            try:
                line["multi_api"] = multi_api_run(w, world, args.steps)
            except Exception as e:
                line["multi_api"] = {"value": None, "note": f"failed: {e}"}
        dist.barrier(group=waiting)
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line), flush=True)



def measured_peak():
    peaks_path = ROOT / "MEASURED_PEAKS.json"
    if peaks_path.exists():
        return float(json.loads(peaks_path.read_text())["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy)"
    return 6650.0, "fallback of B200_PROFILING.md"


# ---------------------------------------------------------------------------------------------------------------
# config 2: BlockFinder only
# ---------------------------------------------------------------------------------------------------------------

def run_finder(args, reference_arm: bool = False):
    """GB/s scanned by K1 (both 48-bit magics at every bit offset) over the first 4 GiB of the config-3 bit stream,
    resident in HBM; offsets compared with the reference's finder (BitStringFinder.hpp:158-320 behind
    ParallelBitStringFinder, as `ibzip2 -l` drives it: one pass per magic, tools/ibzip2.cpp:96-143) on the host cores."""
    import _oracle as orc

    rank, world, local_rank = dist_env()
    if rank != 0:
        return
    w, _ = get_workload(args, 0, lambda: None)
    nbytes = min(int(w.compressed.size), 4 << 30)
    data = w.compressed[:nbytes]
    cores = host_cores()
    metric = "GB/s scanned (BlockFinder only, 4 GiB bzip2 bit stream)"
    config = {"workload": f"config 2: both block magics at every bit offset of the first {nbytes / 2**30:.2f} GiB of the "
                          f"config-3 stream ({w.blocks} blocks in the whole file)", "scanned_bytes": nbytes,
              "l2": "input larger than L2 (no flush needed)"}

    def reference_scan(sample: np.ndarray):
        """seconds for both magics (two passes, like `ibzip2 -l`), offsets of both"""
        t0 = time.perf_counter()
        blocks = orc.ref_find_magic(sample, benchdata.MAGIC_BLOCK, parallelization=cores, capacity=1 << 22)
        eos = orc.ref_find_magic(sample, benchdata.MAGIC_EOS, parallelization=cores, capacity=1 << 22)
        return time.perf_counter() - t0, blocks, eos

    if reference_arm:
        if not orc.ref_available():
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built"}), flush=True)
            return
        times = []
        for i in range(min(args.warmup, 1) + args.steps):
            sec, _, _ = reference_scan(data)
            if i >= min(args.warmup, 1):
                times.append(sec)
        sec = sum(times) / len(times)
        gbs = nbytes / sec / 1e9
        cpu = {"value": round(gbs, 4), "unit": UNIT, "cores": cores, "kind": "reference",
               "sample": f"the whole {nbytes / 2**30:.2f} GiB, in memory, one pass per magic (what `ibzip2 -l` does)",
               "impl": "ParallelBitStringFinder<48> over BufferViewFileReader (oracle/ref_shim.cpp ref_find_magics)"}
        print(json.dumps({"impl": "reference", "metric": metric, "value": round(gbs, 4), "unit": UNIT, "n_gpus": args.gpus,
                          "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": round(sec * 1e3, 3),
                          "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8",
                          "data": "synthetic", "config": dict(config, parallelism=f"{cores} host threads"),
                          "cpu_baseline": cpu, "gpu_launches": 0,
                          "e2e": {"value": round(gbs, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}),
              flush=True)
        return

    import torch
    from indexed_bzip2_b200 import _capi

    torch.cuda.set_device(local_rank)
    ctx = _capi.Context(local_rank)
    d_in = ctx.device_alloc(nbytes)
    ctx.h2d(d_in, data)
    for _ in range(args.warmup):
        found = ctx.find_magics(d_in, nbytes, device=True)
    sampler = ClockSampler(local_rank)
    kernel_ms = 0.0
    launches = 0
    ctx.timer_start()
    for _ in range(args.steps):
        found = ctx.find_magics(d_in, nbytes, device=True)
        t = ctx.timings()
        kernel_ms += t["find_ms"]
        launches += t["launches"]
    call_ms = ctx.timer_stop() / args.steps
    clocks = sampler.stop()
    kernel_ms /= args.steps
    # host input: the upload is part of the call
    pin = ctx.pinned(nbytes)
    pin.array[:] = data
    ctx.find_magics(pin.array)
    ctx.timer_start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        found_e = ctx.find_magics(pin.array)
    e2e_ms = max(ctx.timer_stop(), (time.perf_counter() - t0) * 1e3) / args.steps
    pin.free()

    verified = None
    cpu = None
    if not args.no_verify or not args.no_cpu_baseline:
        if orc.ref_available():
            sec, blocks, eos = reference_scan(data)
            both = np.concatenate([blocks.astype(np.uint64), eos.astype(np.uint64) | np.uint64(1 << 63)])
            want = both[np.argsort(both & np.uint64((1 << 63) - 1), kind="stable")]
            verified = bool(np.array_equal(found, want) and np.array_equal(found, found_e))
            cpu = {"value": round(nbytes / sec / 1e9, 4), "unit": UNIT, "cores": cores, "kind": "reference",
                   "sample": f"the whole {nbytes / 2**30:.2f} GiB, in memory, one pass per magic (what `ibzip2 -l` does), "
                             f"{sec:.2f} s", "offsets_found": int(want.size)}
        else:
            t0 = time.perf_counter()
            sample = data[:256 << 20]
            want = orc.find_magics(sample)
            sec = time.perf_counter() - t0
            mine = found & np.uint64((1 << 63) - 1)
            verified = bool(np.array_equal(mine[mine + np.uint64(48) <= np.uint64(sample.size * 8)], want))
            cpu = {"value": round(sample.size / sec / 1e9, 4), "unit": UNIT, "cores": 1, "kind": "port",
                   "sample": "first 256 MiB, oracle/bz2_oracle.c"}
    peak, peak_src = measured_peak()
    achieved = nbytes / (kernel_ms * 1e-3) / 1e9
    line = {"metric": metric, "value": round(achieved, 2), "unit": UNIT, "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(kernel_ms, 4), "call_ms_per_step": round(call_ms, 4),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": dict(config, parallelism="one GPU"),
            "e2e": {"value": round(nbytes / (e2e_ms * 1e-3) / 1e9, 3), "unit": UNIT, "h2d_bytes_per_step": nbytes,
                    "d2h_bytes_per_step": int(found.size) * 8, "ms_per_step": round(e2e_ms, 3),
                    "api": "ibz2_find_magics, pinned host input"},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": KERNELS["find_ms"], "achieved": round(achieved, 2), "peak": peak,
                         "unit": "GB/s", "frac": round(achieved / peak, 5), "traffic": None,
                         "peak_source": peak_src, "launches_per_step": 1, "ms_per_launch": round(kernel_ms, 4),
                         "algorithmic_bytes_per_step": nbytes},
            "verified": verified, "offsets_found": int(found.size)}
    if cpu is not None:
        line["cpu_baseline"] = cpu
    ctx.device_free(d_in)
    ctx.close()
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------
# config 5: random access through the reader API
# ---------------------------------------------------------------------------------------------------------------

def run_random_access(args, reference_arm: bool = False):
    """importIndex, then --reads x (seek to a random offset + read 1 MiB) through the drop-in reader
    (ParallelBZ2Reader.hpp:365-378 setBlockOffsets, :278-333 seek, :167-269 read) on an 8 GB pbzip2-style file."""
    import _oracle as orc

    rank, world, local_rank = dist_env()
    if rank != 0:
        return
    w, _ = get_workload(args, 0, lambda: None)
    path = str(w.path) if w.path is not None else None
    if path is None:
        raise RuntimeError("config 5 needs the workload cache (/dev/shm) for the file the readers open")
    cores = host_cores()
    read_size = 1 << 20
    raw_size = w.raw_size
    base = w.raw_base
    metric = "random 1 MiB reads/s (importIndex + seek + read, 8 GB pbzip2-style multistream)"
    config = {"workload": f"config 5: {raw_size / 1e9:.2f} GB mixed synthetic as {w.blocks} independent bzip2 -{args.level} "
                          f"streams of 900 kB (pbzip2 style; a {base.size / 1e9:.2f} GB unique base x{w.copies}), block map "
                          f"imported, then seek + 1 MiB read at uniformly random offsets",
              "raw_bytes": raw_size, "compressed_bytes": int(w.compressed.size), "streams": w.blocks, "read_bytes": read_size}
    rng = np.random.default_rng(5)
    positions = rng.integers(0, raw_size - read_size, args.reads).astype(np.uint64)

    import indexed_bzip2_b200 as ibz2

    def expected(pos: int) -> np.ndarray:
        b = pos % base.size
        if b + read_size <= base.size:
            return base[b:b + read_size]
        return np.concatenate([base[b:], base[:read_size - (base.size - b)]])

    # the index: one sequential pass of our reader (also the reader-level sequential throughput)
    index_path = Path(path).with_suffix(".index.json")
    if index_path.exists():
        index = {int(k): int(v) for k, v in json.loads(index_path.read_text()).items()}
        seq = None
    else:
        if reference_arm and not _gpu_present():
            raise RuntimeError("the index is produced by the GPU arm; run it first")
        first = ibz2.open(path, parallelization=0)
        t0 = time.perf_counter()
        first.seek(0, 2)
        seq = time.perf_counter() - t0
        index = first.block_offsets()
        first.close()
        index_path.write_text(json.dumps(index))
    pairs = np.array(sorted(index.items()), dtype=np.uint64).reshape(-1)

    def reference_reads(count: int):
        out = np.zeros(read_size, dtype=np.uint8)
        checksum = C_u64(0)
        seconds = C_double(0)
        rc = orc.ref().ref_random_reads(path.encode(), cores, pairs.ctypes.data, pairs.size // 2, positions.ctypes.data,
                                        count, read_size, out.ctypes.data, byref(checksum), byref(seconds))
        if rc != 0:
            raise RuntimeError(orc.ref().ref_last_error().decode())
        want = sum(int(expected(int(p)).sum(dtype=np.uint64)) for p in positions[:count].tolist())
        return seconds.value, checksum.value == want

    from ctypes import byref, c_double as C_double, c_uint64 as C_u64

    if reference_arm:
        if not orc.ref_available():
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built"}), flush=True)
            return
        count = min(args.reads, 100 * max(1, args.steps))
        sec, ok = reference_reads(count)
        rps = count / sec
        cpu = {"value": round(rps, 3), "unit": "reads/s", "cores": cores, "kind": "reference",
               "sample": f"the first {count} of the {args.reads} positions, index imported with setBlockOffsets, {sec:.1f} s",
               "mean_ms_per_read": round(sec / count * 1e3, 3), "verified": ok}
        print(json.dumps({"impl": "reference", "metric": metric, "value": round(rps, 3), "unit": "reads/s", "n_gpus": args.gpus,
                          "steps": count, "warmup": 0, "ms_per_step": round(sec / count * 1e3, 3), "higher_is_better": True,
                          "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                          "config": dict(config, parallelism=f"{cores} host threads"), "cpu_baseline": cpu, "gpu_launches": 0,
                          "e2e": {"value": round(rps, 3), "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}),
              flush=True)
        return

    reader = ibz2.open(path, parallelization=0)
    reader.set_block_offsets(index)
    for p in positions[:min(20, args.reads)].tolist():   # warm-up: scratch memory, the compressed file in HBM
        reader.seek(int(p))
        reader.read(read_size)
    sampler = ClockSampler(local_rank)
    lat = np.zeros(args.reads)
    ok = True
    t_all = time.perf_counter()
    for i, p in enumerate(positions.tolist()):
        t0 = time.perf_counter()
        reader.seek(int(p))
        data = reader.read(read_size)
        lat[i] = time.perf_counter() - t0
        if i % 16 == 0 or not ok:   # every 16th read is compared with the raw data (the comparison is not timed)
            ok = ok and len(data) == read_size and bool(np.array_equal(np.frombuffer(data, dtype=np.uint8), expected(int(p))))
    wall = time.perf_counter() - t_all
    clocks = sampler.stop()
    reader.close()
    busy = float(lat.sum())
    line = {"metric": metric, "value": round(args.reads / busy, 3), "unit": "reads/s", "n_gpus": 1, "steps": args.reads,
            "warmup": min(20, args.reads), "ms_per_step": round(busy / args.reads * 1e3, 3), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": dict(config, parallelism="one GPU, one reader object"),
            "latency_ms": {"p50": round(float(np.percentile(lat, 50)) * 1e3, 3), "p90": round(float(np.percentile(lat, 90)) * 1e3, 3),
                           "p99": round(float(np.percentile(lat, 99)) * 1e3, 3), "max": round(float(lat.max()) * 1e3, 3)},
            "delivered_GBps": round(args.reads * read_size / busy / 1e9, 4),
            "e2e": {"value": round(args.reads / busy, 3), "unit": "reads/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": read_size, "api": "indexed_bzip2_b200.open(path).set_block_offsets / seek / read: "
                    "the compressed file is uploaded once when the reader opens; every read brings its blocks back over PCIe"},
            "gpu_launches": None, "clocks": clocks, "verified": ok, "wall_s": round(wall, 2),
            "roofline": {"bound": "latency", "note": "a read decodes two or three blocks; what bounds it is the sequential "
                         "Huffman chain of one block (K2b), not bandwidth", "frac": None}}
    if seq is not None:
        line["reader_sequential"] = {"value": round(raw_size / seq / 1e9, 3), "unit": UNIT, "seconds": round(seq, 2),
                                     "api": "open(path).seek(0, SEEK_END): the first pass that builds the block map"}
    if not args.no_cpu_baseline and orc.ref_available():
        try:
            count = min(args.reads, 100)
            sec, ref_ok = reference_reads(count)
            line["cpu_baseline"] = {"value": round(count / sec, 3), "unit": "reads/s", "cores": cores, "kind": "reference",
                                    "sample": f"the first {count} positions, index imported with setBlockOffsets, {sec:.1f} s",
                                    "mean_ms_per_read": round(sec / count * 1e3, 3), "verified": ref_ok}
        except Exception as e:
            line["cpu_baseline"] = {"value": None, "kind": "reference", "sample": f"failed: {e}"}
    print(json.dumps(line), flush=True)


def _gpu_present() -> bool:
    try:
        from indexed_bzip2_b200 import _capi

        return _capi.load_library().ibz2_device_count() > 0
    except Exception:
        return False


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    if args.config == 2:
        return run_finder(args, reference_arm=True)
    if args.config == 5:
        return run_random_access(args, reference_arm=True)
    w, _ = get_workload(args, 0, lambda: None)
    gbs, sec, cpu = cpu_reference_run(w, steps=args.steps, warmup=min(args.warmup, 1))
    if args.config == 1:
        cpu["single_thread"] = reference_single_thread_runs(w)
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
    elif a.config == 2:
        run_finder(a)
    elif a.config == 5:
        run_random_access(a)
    else:
        run_ours(a)
