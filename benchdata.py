"""Synthetic bzip2 workloads for bench.py and the full-size GPU tests (BASELINE.json configs 1, 3, 4, 5).

Bench / test infrastructure, not product code.  Everything is seeded and generated on the machine that runs it
(there is no network and 8 GB fixtures cannot travel).  Compressing goes through CPython's bz2 (libbz2 1.0.8, the
same encoder as the `bzip2` CLI the reference's tests shell out to, testBZ2Reader.cpp:305) in a process pool; the
independent per-chunk streams are then cut at their block magics and re-joined bit by bit into ONE stream
(`single_stream`) -- possible because a bzip2 block is self-contained and the stream CRC only depends on the block
CRCs (BZ2Reader.hpp:481-484) -- or concatenated as they are (`multistream`, what pbzip2 writes).

An 8 GB single stream is built from a unique base (default 1e9 raw bytes) whose blocks are repeated: every copy
lands at a different bit alignment and at different addresses, so nothing is shared between copies on the GPU.
"""
from __future__ import annotations

import base64
import bz2
import hashlib
import json
import os
import time
from concurrent.futures import ProcessPoolExecutor, ThreadPoolExecutor
from dataclasses import dataclass
from pathlib import Path

import numpy as np

MAGIC_BLOCK = 0x314159265359
MAGIC_EOS = 0x177245385090
SEGMENT = 1 << 20


# ---------------------------------------------------------------------------------------------------------------
# raw data generators (deterministic per 1 MiB segment, so any chunking gives the same bytes)
# ---------------------------------------------------------------------------------------------------------------

_vocab_cache: dict[int, tuple[np.ndarray, np.ndarray, np.ndarray]] = {}


def _vocab(seed: int):
    if seed not in _vocab_cache:
        rng = np.random.default_rng([seed, 0xC0FFEE])
        lens = rng.integers(2, 10, 4096)
        words = [rng.integers(97, 123, int(n), dtype=np.uint8).tobytes() + b" " for n in lens]
        flat = np.frombuffer(b"".join(words), dtype=np.uint8)
        wl = (lens + 1).astype(np.int64)
        starts = np.concatenate([[0], np.cumsum(wl)[:-1]])
        _vocab_cache[seed] = (flat, starts, wl)
    return _vocab_cache[seed]


def text_segment(seed: int, k: int, n: int = SEGMENT) -> np.ndarray:
    """Zipf(1.3) draws from a 4096-word lowercase vocabulary, space separated, newline every 80 bytes."""
    flat, starts, wl = _vocab(seed)
    rng = np.random.default_rng([seed, k])
    ids = np.minimum(rng.zipf(1.3, n // 3 + 16), 4095)
    lens = wl[ids]
    ends = np.cumsum(lens)
    m = int(np.searchsorted(ends, n)) + 1
    ids, lens, ends = ids[:m], lens[:m], ends[:m]
    idx = np.repeat(starts[ids] - (ends - lens), lens) + np.arange(int(ends[-1]))
    out = flat[idx[:n]].copy()
    out[79::80] = 10
    return out


def mixed_chunk(seed: int, first_segment: int, nsegments: int) -> np.ndarray:
    """config 3: alternating 1 MiB text-like / uniform random segments"""
    parts = []
    for k in range(first_segment, first_segment + nsegments):
        if k % 2 == 0:
            parts.append(text_segment(seed, k))
        else:
            parts.append(np.random.default_rng([seed, k]).integers(0, 256, SEGMENT, dtype=np.uint8))
    return np.concatenate(parts)


def base64_chunk(seed: int, first_segment: int, nsegments: int) -> np.ndarray:
    """config 1: base64 of random bytes, 76-column lines (base64.encodebytes)"""
    parts = []
    for k in range(first_segment, first_segment + nsegments):
        raw = np.random.default_rng([seed, k]).bytes(SEGMENT * 3 // 4)
        parts.append(np.frombuffer(base64.encodebytes(raw)[:SEGMENT], dtype=np.uint8))
    return np.concatenate(parts)


def longrun_chunk(seed: int, first_segment: int, nsegments: int) -> np.ndarray:
    """config 4: runs of one byte value with log-uniform lengths 1..1e5 (incl. the RLE1 edge lengths)"""
    parts = []
    for k in range(first_segment, first_segment + nsegments):
        rng = np.random.default_rng([seed, k])
        lengths = np.concatenate([[3, 4, 5, 255, 256, 258, 259, 260, 1, 2],
                                  np.exp(rng.uniform(0, np.log(1e5), 4096)).astype(np.int64)])
        ends = np.cumsum(lengths)
        m = int(np.searchsorted(ends, SEGMENT)) + 1
        values = rng.integers(0, 256, m, dtype=np.uint8)
        parts.append(np.repeat(values, lengths[:m])[:SEGMENT])
    return np.concatenate(parts)


GENERATORS = {"mixed": mixed_chunk, "base64": base64_chunk, "longruns": longrun_chunk}


# ---------------------------------------------------------------------------------------------------------------
# bit-level helpers
# ---------------------------------------------------------------------------------------------------------------

def find_magics_py(data: bytes) -> list[tuple[int, bool]]:
    """(bit offset, is_eos) of both 48-bit magics, ascending.  Same idea as the reference's finder
    (core/BitStringFinder.hpp:158-238): 8 shifted byte-aligned searches plus head / tail bit checks."""
    found = []
    n = len(data)
    for magic, is_eos in ((MAGIC_BLOCK, False), (MAGIC_EOS, True)):
        for s in range(8):
            mid = ((magic >> s) & ((1 << 40) - 1)).to_bytes(5, "big")
            head = magic >> (40 + s)
            head_mask = (1 << (8 - s)) - 1
            tail = magic & ((1 << s) - 1)
            q = data.find(mid, 1)
            while q != -1:
                p = q - 1
                ok = (data[p] & head_mask) == head
                if ok and s > 0:
                    ok = q + 5 < n and (data[q + 5] >> (8 - s)) == tail
                if ok:
                    found.append((p * 8 + s, is_eos))
                q = data.find(mid, q + 1)
    found.sort()
    return found


def read_bits(data: bytes, pos: int, nbits: int) -> int:
    first, last = pos >> 3, (pos + nbits + 7) >> 3
    v = int.from_bytes(data[first:last], "big")
    return (v >> ((last - first) * 8 - (pos & 7) - nbits)) & ((1 << nbits) - 1)


def extract_bits(data: np.ndarray, start: int, end: int) -> np.ndarray:
    """bits [start, end) of a uint8 array, re-aligned to bit 0 (last byte zero padded)"""
    s = start & 7
    a = data[start >> 3:(end + 7) >> 3]
    nbytes = (end - start + 7) >> 3
    if s == 0:
        out = a[:nbytes].copy()
    else:
        nxt = np.zeros(a.size, dtype=np.uint8)
        nxt[:-1] = a[1:]
        out = ((a << s) | (nxt >> (8 - s)))[:nbytes]
    spare = nbytes * 8 - (end - start)
    if spare:
        out[-1] &= (0xFF << spare) & 0xFF
    return out


def shifted(payload: np.ndarray, s: int) -> np.ndarray:
    """payload moved right by s bits (0..7): len + 1 bytes"""
    out = np.zeros(payload.size + 1, dtype=np.uint8)
    if s == 0:
        out[:-1] = payload
    else:
        out[:-1] = payload >> s
        out[1:] |= payload << (8 - s)
    return out


class BitAssembler:
    """Appends bit strings (given re-aligned to bit 0) to a preallocated byte buffer."""

    def __init__(self, capacity_bytes: int):
        self.buf = np.zeros(capacity_bytes, dtype=np.uint8)
        self.nbits = 0

    def append(self, payload: np.ndarray, nbits: int):
        o, s = self.nbits >> 3, self.nbits & 7
        sh = shifted(payload, s)
        nb = (s + nbits + 7) >> 3
        self.buf[o] |= sh[0]
        self.buf[o + 1:o + nb] = sh[1:nb]
        self.nbits += nbits

    def append_int(self, value: int, nbits: int):
        nbytes = (nbits + 7) >> 3
        payload = np.frombuffer((value << (nbytes * 8 - nbits)).to_bytes(nbytes, "big"), dtype=np.uint8)
        self.append(payload, nbits)

    def finish(self) -> np.ndarray:
        return self.buf[:(self.nbits + 7) >> 3]


# ---------------------------------------------------------------------------------------------------------------
# pool workers
# ---------------------------------------------------------------------------------------------------------------

def _compress_chunk(args):
    kind, seed, first_segment, nsegments, level, want_blocks = args
    raw = GENERATORS[kind](seed, first_segment, nsegments)
    comp = bz2.compress(raw.tobytes(), level)
    digest = hashlib.sha256(raw).digest()
    if not want_blocks:
        return comp, None, 0, [], raw.size, digest
    magics = find_magics_py(comp)
    assert magics and magics[-1][1] and not any(e for _, e in magics[:-1]), "unexpected magic layout in a chunk"
    crcs = [read_bits(comp, p + 48, 32) for p, _ in magics[:-1]]
    start, end = magics[0][0], magics[-1][0]
    payload = extract_bits(np.frombuffer(comp, dtype=np.uint8), start, end)
    return None, payload, end - start, crcs, raw.size, digest


def _raw_chunk(args):
    kind, seed, first_segment, nsegments = args
    return GENERATORS[kind](seed, first_segment, nsegments)


# ---------------------------------------------------------------------------------------------------------------
# workloads
# ---------------------------------------------------------------------------------------------------------------

@dataclass
class Workload:
    name: str
    compressed: np.ndarray      # the .bz2 bytes
    raw_base: np.ndarray        # unique raw bytes; the decoded file is raw_base repeated `copies` times
    copies: int
    blocks: int                 # data blocks in the file
    path: Path | None           # where `compressed` lives on disk (for the CPU reference), if cached
    meta: dict
    sample: np.ndarray | None = None   # a valid .bz2 of ONE copy of raw_base (bounded CPU-baseline sample)
    sample_path: Path | None = None

    @property
    def raw_size(self) -> int:
        return int(self.raw_base.size) * self.copies

    def verify(self, out: np.ndarray) -> bool:
        """decoded bytes == raw_base repeated"""
        n = self.raw_base.size
        if out.size != n * self.copies:
            return False
        return all(np.array_equal(out[i * n:(i + 1) * n], self.raw_base) for i in range(self.copies))


def default_workers() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def _chunk_jobs(kind, seed, nsegments_total, segs_per_chunk):
    jobs, first = [], 0
    while first < nsegments_total:
        n = min(segs_per_chunk, nsegments_total - first)
        jobs.append((kind, seed, first, n))
        first += n
    return jobs


def build_single_stream(kind: str, seed: int, base_bytes: int, copies: int, level: int = 9,
                        workers: int | None = None, segs_per_chunk: int = 32, log=None) -> Workload:
    """One bzip2 stream whose blocks are the blocks of `base_bytes` of generated data, repeated `copies` times."""
    workers = workers or default_workers()
    nseg = (base_bytes + SEGMENT - 1) // SEGMENT
    jobs = _chunk_jobs(kind, seed, nseg, segs_per_chunk)
    t0 = time.time()
    with ProcessPoolExecutor(max_workers=workers) as pool:
        results = list(pool.map(_compress_chunk, [j + (level, True) for j in jobs]))
        raws = list(pool.map(_raw_chunk, jobs))
    raw_base = np.concatenate(raws)
    del raws
    t1 = time.time()
    # the base payload: all blocks of all chunks, bit-contiguous from bit 0
    total_bits = sum(r[2] for r in results)
    base = BitAssembler((total_bits + 7) // 8 + 8)
    crcs: list[int] = []
    for _, payload, nbits, chunk_crcs, _, _ in results:
        base.append(payload, nbits)
        crcs.extend(chunk_crcs)
    payload = base.finish()
    del results
    out = _assemble(payload, total_bits, crcs, copies, level)
    sample = _assemble(payload, total_bits, crcs, 1, level) if copies > 1 else out
    t2 = time.time()
    meta = {"kind": kind, "seed": seed, "level": level, "layout": "single stream, blocks of a %d-byte unique base x %d" % (raw_base.size, copies),
            "raw_bytes": int(raw_base.size) * copies, "compressed_bytes": int(out.size), "blocks": len(crcs) * copies,
            "gen_compress_s": round(t1 - t0, 2), "splice_s": round(t2 - t1, 2), "workers": workers}
    if log:
        log(f"[benchdata] {meta}")
    return Workload(f"{kind}-l{level}", out, raw_base, copies, len(crcs) * copies, None, meta, sample=sample)


def _assemble(payload: np.ndarray, total_bits: int, crcs: list[int], copies: int, level: int) -> np.ndarray:
    """'BZh<level>' + `copies` x payload (bit-contiguous) + EOS magic + stream CRC, padded to a byte."""
    out_bits = 32 + copies * total_bits + 48 + 32
    out = np.zeros((out_bits + 7) // 8, dtype=np.uint8)
    out[:4] = np.frombuffer(b"BZh" + str(level).encode(), dtype=np.uint8)

    def place(c):
        pos = 32 + c * total_bits
        o, s = pos >> 3, pos & 7
        sh = shifted(payload, s)
        nb = (s + total_bits + 7) >> 3
        out[o + 1:o + nb - 1] = sh[1:nb - 1]
        return o, sh[0], o + nb - 1, sh[nb - 1] if nb > 1 else 0

    # copies are placed by threads (numpy shifts release the GIL); the shared seam bytes are or-ed afterwards
    with ThreadPoolExecutor(max_workers=min(copies, 8)) as tp:
        seams = list(tp.map(place, range(copies)))
    for o0, b0, o1, b1 in seams:
        out[o0] |= b0
        if o1 != o0:
            out[o1] |= b1
    stream_crc = 0
    for _ in range(copies):
        for c in crcs:
            stream_crc = (((stream_crc << 1) | (stream_crc >> 31)) & 0xFFFFFFFF) ^ c
    tail = BitAssembler(16)
    tail.nbits = (32 + copies * total_bits) & 7
    tail.append_int(MAGIC_EOS, 48)
    tail.append_int(stream_crc, 32)
    tb = tail.finish()
    o = (32 + copies * total_bits) >> 3
    out[o] |= tb[0]
    out[o + 1:o + tb.size] = tb[1:]
    return out


def build_multistream(kind: str, seed: int, total_bytes: int, piece_bytes: int = 900_000, level: int = 9,
                      workers: int | None = None, log=None) -> Workload:
    """config 5: independent streams of piece_bytes raw bytes each, concatenated (pbzip2 style)."""
    workers = workers or default_workers()
    nseg = (total_bytes + SEGMENT - 1) // SEGMENT
    with ProcessPoolExecutor(max_workers=workers) as pool:
        raws = list(pool.map(_raw_chunk, _chunk_jobs(kind, seed, nseg, 32)))
        raw = np.concatenate(raws)[:total_bytes]
        del raws
        pieces = [raw[i:i + piece_bytes].tobytes() for i in range(0, total_bytes, piece_bytes)]
        comp = list(pool.map(_compress_piece, [(p, level) for p in pieces], chunksize=8))
    out = np.frombuffer(b"".join(comp), dtype=np.uint8)
    meta = {"kind": kind, "seed": seed, "level": level, "layout": "multistream, %d-byte pieces" % piece_bytes,
            "raw_bytes": int(raw.size), "compressed_bytes": int(out.size), "streams": len(comp)}
    if log:
        log(f"[benchdata] {meta}")
    return Workload(f"{kind}-multistream-l{level}", out, raw, 1, len(comp), None, meta)


def build_multistream_repeated(kind: str, seed: int, base_bytes: int, copies: int, piece_bytes: int = 900_000,
                               level: int = 9, log=None) -> Workload:
    """config 5 at size: the streams of a unique base, the whole file repeated `copies` times (streams are byte
    aligned and independent, so the copies are plain concatenation)"""
    w = build_multistream(kind, seed, base_bytes, piece_bytes, level, log=log)
    comp = np.tile(w.compressed, copies) if copies > 1 else w.compressed
    meta = dict(w.meta, layout=w.meta["layout"] + f", unique base x {copies}", raw_bytes=int(w.raw_base.size) * copies,
                compressed_bytes=int(comp.size), streams=w.blocks * copies)
    return Workload(w.name, comp, w.raw_base, copies, w.blocks * copies, None, meta, sample=w.compressed)


def _compress_piece(args):
    data, level = args
    return bz2.compress(data, level)


def cached(workload_fn, key: str, cache_dir: str | None = None, log=None) -> Workload:
    """Keeps the generated files under /dev/shm (or cache_dir) so that the reference arm and ours, run back to back
    on one box, generate once."""
    root = Path(cache_dir or os.environ.get("IBZ2_BENCH_CACHE", "/dev/shm/ibz2_bench"))
    d = root / key
    meta_path = d / "meta.json"
    if meta_path.exists():
        meta = json.loads(meta_path.read_text())
        comp = np.fromfile(d / "data.bz2", dtype=np.uint8)
        raw = np.fromfile(d / "raw_base.bin", dtype=np.uint8)
        if comp.size == meta["compressed_bytes"] and raw.size * meta["copies"] == meta["raw_bytes"]:
            if log:
                log(f"[benchdata] cache hit {d}")
            sample = np.fromfile(d / "sample.bz2", dtype=np.uint8) if (d / "sample.bz2").exists() else comp
            return Workload(meta["name"], comp, raw, meta["copies"], meta["blocks"], d / "data.bz2", meta,
                            sample=sample, sample_path=d / "sample.bz2" if (d / "sample.bz2").exists() else d / "data.bz2")
    w = workload_fn()
    try:
        d.mkdir(parents=True, exist_ok=True)
        w.compressed.tofile(d / "data.bz2")
        w.raw_base.tofile(d / "raw_base.bin")
        meta = dict(w.meta, name=w.name, copies=w.copies, blocks=w.blocks)
        meta_path.write_text(json.dumps(meta))
        w.path = d / "data.bz2"
        w.sample_path = w.path
        if w.sample is not None and w.sample is not w.compressed:
            w.sample.tofile(d / "sample.bz2")
            w.sample_path = d / "sample.bz2"
    except OSError as e:  # cache is best effort
        if log:
            log(f"[benchdata] cache write failed: {e}")
    return w
