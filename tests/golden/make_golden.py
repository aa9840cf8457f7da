#!/usr/bin/env python3
"""Regenerates tests/golden/: inputs + the outputs of the compiled, UNMODIFIED reference for them.

Run in the build container (needs /root/reference and `make -C oracle ref`):
    python tests/golden/make_golden.py

What lands in tests/golden/:
  cases/<name>.bz2      the compressed inputs: (i) the reference's own 11 single-block fixtures
                        (src/tests/data/*.bz2, data files, copied verbatim), (ii) generated multi-block,
                        multi-stream, long-run, periodic and damaged files (seeded; python bz2 = libbz2 1.0.8)
  golden.json           per case: sha256 + size of the decoded data, blockOffsets() of ParallelBZ2Reader (P=4)
                        and of BZ2Reader, the `ibzip2 -l` magic list, error text / exit status of the CLI,
                        all produced by oracle/_ref (reference @ 30c04e66)
Nothing here is read from /root/reference at test time.
"""
from __future__ import annotations

import base64
import bz2
import hashlib
import json
import subprocess
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT / "tests"))
sys.path.insert(0, str(HERE))
import _oracle  # noqa: E402
import craft  # noqa: E402  (tests/golden/craft.py: hand-written streams libbz2 never produces)

REF_DATA = Path("/root/reference/src/tests/data")
CASES = HERE / "cases"


def text_like(rng: np.random.Generator, n: int) -> bytes:
    vocab = ["".join(chr(97 + c) for c in rng.integers(0, 26, rng.integers(2, 10))) for _ in range(4096)]
    ranks = np.minimum(rng.zipf(1.3, n // 3), 4095)
    words = " ".join(vocab[r] for r in ranks)
    raw = words.encode()[:n]
    out = bytearray(raw)
    out[79::80] = b"\n" * len(out[79::80])
    return bytes(out)


def long_runs(rng: np.random.Generator, n: int) -> bytes:
    parts = []
    total = 0
    special = [3, 4, 5, 255, 256, 258, 259, 260, 1, 2]
    i = 0
    while total < n:
        length = special[i] if i < len(special) else int(np.exp(rng.uniform(0, np.log(1e5))))
        i += 1
        parts.append(bytes([int(rng.integers(0, 256))]) * length)
        total += length
    return b"".join(parts)[:n]


def generated_cases() -> dict[str, bytes]:
    rng = np.random.default_rng(20261104)
    ref = {p.stem: p.read_bytes() for p in REF_DATA.glob("*.bz2")}
    cases: dict[str, bytes] = {}

    # Appendix B multi-stream example: data, tiny, EMPTY stream, data
    cases["multistream-b64-1B-empty-b64"] = (ref["base64-32KiB"] + ref["1B"] + ref["empty"] + ref["base64-32KiB"])
    # multi-block, level 1 (100 kB blocks): text-like
    txt = text_like(rng, 450_000)
    cases["text-450k-l1"] = bz2.compress(txt, 1)
    # multi-block mixed entropy (text + random), level 2
    mixed = b"".join(text_like(rng, 65536) if i % 2 == 0 else rng.bytes(65536) for i in range(6))
    cases["mixed-384k-l2"] = bz2.compress(mixed, 2)
    # base64 random, level 1, 3 blocks (config 1 miniature)
    b64 = base64.encodebytes(rng.bytes(225_000))[:300_000]
    cases["base64-300k-l1"] = bz2.compress(b64, 1)
    # long runs (config 4 miniature): RLE1 expansion, RUNA/RUNB heavy
    runs = long_runs(rng, 3_000_000)
    cases["longruns-3M-l1"] = bz2.compress(runs, 1)
    # periodic data (testPythonWrappers.py:476-500 regression sizes, Appendix C-15)
    for size in [1, 2, 4, 128, 1000, 1024, 100_000, 200_000, 512 * 1024 + 2]:
        cases[f"A-x{size}-l1"] = bz2.compress(b"A" * size, 1)
    cases["AB-x150000-l1"] = bz2.compress(b"AB" * 150_000, 1)
    cases["stripes-257-l3"] = bz2.compress((b"A" * 257 + b"B" * 257) * 400, 3)
    cases["abracadabra-x3"] = bz2.compress(b"abracadabra" * 3, 9)
    # all 256 byte values, every RLE1 run length 1..300 of a few bytes
    edge = b"".join(bytes([b]) * k for k, b in zip(range(1, 301), [0, 255, 4, 65] * 75))
    cases["rle1-runlengths-1..300"] = bz2.compress(bytes(range(256)) * 3 + edge, 9)
    # pbzip2-style: many small independent streams, different levels
    pieces = [text_like(rng, 20_000), rng.bytes(5000), b"", b"x", long_runs(rng, 50_000)]
    cases["pbzip2-style-5-streams"] = b"".join(bz2.compress(p, lvl) for p, lvl in zip(pieces, [9, 1, 5, 3, 2]))
    # damaged inputs
    good = cases["text-450k-l1"]
    cases["err-trailing-garbage"] = ref["base64-32KiB"] + b"\x00garbage after the end of stream"
    cases["err-truncated"] = good[: len(good) // 2]
    flipped = bytearray(good)
    flipped[len(flipped) // 3] ^= 0x10
    cases["err-bitflip"] = bytes(flipped)
    cases["err-not-bzip2"] = b"this is not a bzip2 file at all, not even a little bit\n" * 3
    cases["err-gap-then-valid"] = ref["1B"] + b"\x00\x00\x00" + ref["1B"]
    hdr = bytearray(ref["dolorem-ipsum.txt"])
    hdr[3] = ord("1")  # BZh1 header in front of a block produced at level 9 (Appendix C-1: level digit ignored)
    cases["level-digit-ignored"] = bytes(hdr)

    # ---- round 2 (appended, so that the generator state of the cases above is unchanged) ----
    rng2 = np.random.default_rng(20261105)
    # 2, 3, 4, 5 and 6 Huffman groups: libbz2 picks the count from the number of MTF symbols (< 200, 600, 1200, 2400)
    for groups, size in [(2, 150), (3, 450), (4, 900), (5, 1800), (6, 4000)]:
        cases[f"groups-{groups}"] = bz2.compress(rng2.bytes(size), 9)
    # codes of up to 20 bits (bzip2.hpp:97; libbz2's encoder stops at 17): three hand-made tables, two of them with
    # lengths 1, 2, 3, ..., 19, 20, 20 in opposite symbol orders, so that the 18..20-bit codes really occur
    vals = np.arange(19, dtype=np.uint8) + 65
    prob = 0.5 ** np.arange(1, 20)
    prob[-1] *= 2
    skew = bytes(rng2.choice(vals, 6000, p=prob / prob.sum())) + bytes(vals) * 3
    packed = craft.rle1(skew)
    last, _ = craft.bwt(packed)
    used = sorted(set(packed))
    symbols = craft.mtf_rle2(last, used)
    alphabet = len(used) + 2
    by_freq = sorted(range(alphabet), key=lambda x: (-symbols.count(x), x))
    tables = [craft.skewed_lengths(alphabet, by_freq), craft.skewed_lengths(alphabet, by_freq[::-1]),
              craft.skewed_lengths(alphabet, by_freq, longest=12)]
    cases["long-codes-20bit"] = craft.stream([dict(data=skew, tables=tables,
                                                   selectors_of=lambda g: (0, 0, 2, 1, 0, 2)[g % 6])])
    # an EMPTY symbol map: the alphabet is RUNA, RUNB and the end-of-block symbol would be 1 = RUNB, which
    # bzip2.hpp:723 takes for a run symbol first, so the block never ends (selector out of range)
    cases["err-empty-symbol-map-runb"] = craft.stream([dict(data=b"x", tables=[[1, 1], [1, 1]], raw_symbols=[1])])
    cases["err-empty-symbol-map-runs"] = craft.stream([dict(data=b"x", tables=[[1, 1], [1, 1]],
                                                            raw_symbols=[0] * 30 + [1])])
    # wrong STREAM CRC (the 32 bits behind the end-of-stream magic) with every block CRC right: BZ2Reader.hpp:406-416
    # throws, ParallelBZ2Reader does not look at it
    bad = bytearray(cases["text-450k-l1"])
    bad[-3] ^= 0x04
    cases["err-stream-crc"] = bytes(bad)
    second = bytearray(bz2.compress(text_like(rng2, 30_000), 1))
    second[-3] ^= 0x40
    cases["err-stream-crc-multistream"] = ref["base64-32KiB"] + bytes(second) + ref["1B"]
    # as many selectors as the 15-bit field holds (bzip2.hpp:574-637 reads them all; libbz2 writes 18002 at most), with
    # every move-to-front distance among them; and a long list whose selector 19000 has one 1 bit too many
    sel_text = text_like(rng2, 5_000)
    cases["selectors-32767"] = craft.stream([dict(data=sel_text, tables=None, total_selectors=32767,
                                                  selectors_of=lambda g: (g * 7 + g // 5) % 2)])
    sel_tables6 = [[8] * 254 + [9] * 4 for _ in range(6)]   # 254 / 256 + 4 / 512 = 1
    every_byte = bytes(range(256)) * 40
    cases["selectors-20001-six-tables"] = craft.stream([dict(data=every_byte, tables=sel_tables6, total_selectors=20001,
                                                             selectors_of=lambda g: (g * g + g // 3) % 6)])
    cases["err-selector-unary-late"] = craft.stream([dict(data=every_byte, tables=sel_tables6, total_selectors=20001,
                                                          selectors_of=lambda g: (g * g + g // 3) % 6,
                                                          bad_selector=19000)])
    # one stream per header check of bzip2.hpp:479-685 and :794-798 that no other case reaches: what the reference
    # throws, and with which words, is part of the interface
    small = text_like(rng2, 3_000)
    cases["err-origptr"] = craft.stream([dict(data=small, tables=None, fields={"orig": 0xFFFFF0})])
    cases["err-group-count-1"] = craft.stream([dict(data=small, tables=None, fields={"ngroups": 1})])
    cases["err-group-count-7"] = craft.stream([dict(data=small, tables=None, fields={"ngroups": 7})])
    cases["err-selector-count-0"] = craft.stream([dict(data=small, tables=None, fields={"nselectors": 0})])
    cases["err-code-length-0"] = craft.stream([dict(data=small, tables=None, fields={"first_length": 0})])
    cases["err-code-length-21"] = craft.stream([dict(data=small, tables=None, fields={"first_length": 21})])
    # a stream header with something else where the first block's magic should be (bzip2.hpp:502-507)
    cases["err-bad-magic-after-header"] = b"BZh9" + bytes([0xAB, 0xCD, 0xEF, 0x01, 0x23, 0x45, 0x67, 0x89] * 4)
    # what readBlockData itself rejects (bzip2.hpp:748-756, 794-798) and what the table construction rejects (:680-683)
    cases["err-origptr-error"] = craft.stream([dict(data=small, tables=None, fields={"orig": 5000})])
    alphabet_small = len(set(craft.rle1(small))) + 2
    cases["err-code-lengths-oversubscribed"] = craft.stream([dict(data=small, tables=[[1] * alphabet_small] * 2,
                                                                  fields={"skip_data": True})])
    # one byte value, then RUNB twenty-one times: a run of 2^22 - 2 bytes, more than a block holds
    cases["err-run-longer-than-block"] = craft.stream([dict(data=b"x", tables=[[1, 2, 2], [1, 2, 2]],
                                                            raw_symbols=[1] * 21 + [2], fields={"used": [120]})])
    return cases


def reference_cli(args: list[str]) -> tuple[int, str, bytes]:
    p = subprocess.run([str(_oracle.REF_CLI)] + args, capture_output=True)
    return p.returncode, p.stderr.decode(errors="replace"), p.stdout


def main() -> None:
    assert _oracle.ref_available() and _oracle.REF_CLI.exists(), "run `make -C oracle ref` first"
    CASES.mkdir(parents=True, exist_ok=True)
    cases: dict[str, bytes] = {}
    for p in sorted(REF_DATA.glob("*.bz2")):
        cases["ref-" + p.stem] = p.read_bytes()
    cases.update(generated_cases())

    golden = {}
    for name, blob in sorted(cases.items()):
        path = CASES / f"{name}.bz2"
        path.write_bytes(blob)
        entry: dict = {"compressed_size": len(blob), "compressed_sha256": hashlib.sha256(blob).hexdigest()}

        par = _oracle.ref_decode_file(str(path), parallelization=4, serial_class=False)
        entry["parallel"] = {"rc": par.rc, "error": par.error, "size": len(par.data),
                             "sha256": hashlib.sha256(par.data).hexdigest(), "offsets": par.offsets}
        ser = _oracle.ref_decode_file(str(path), parallelization=1, serial_class=True)
        entry["serial"] = {"rc": ser.rc, "error": ser.error, "size": len(ser.data),
                           "sha256": hashlib.sha256(ser.data).hexdigest(), "offsets": ser.offsets}

        rc, err, out = reference_cli(["-l", "-", str(path)])
        entry["finder"] = [int(x) for x in out.split()] if rc == 0 else None
        rc, err, out = reference_cli(["-d", "-c", "-P", "4", str(path)])
        entry["cli_parallel"] = {"rc": rc, "stderr": err.strip(), "stdout_size": len(out),
                                 "stdout_sha256": hashlib.sha256(out).hexdigest()}
        rc, err, out = reference_cli(["-d", "-c", "-P", "1", str(path)])
        entry["cli_serial"] = {"rc": rc, "stderr": err.strip(), "stdout_size": len(out),
                               "stdout_sha256": hashlib.sha256(out).hexdigest()}
        # independent cross-check (libbz2) where the file is valid
        try:
            if name.startswith("err-stream-crc"):
                # CPython's bz2.decompress silently drops an invalid stream that follows a valid one: no cross-check
                raise ValueError("stream CRC damaged on purpose")
            raw = bz2.decompress(blob)
            entry["libbz2_sha256"] = hashlib.sha256(raw).hexdigest()
        except Exception as e:  # noqa: BLE001
            entry["libbz2_sha256"] = None
            entry["libbz2_error"] = type(e).__name__
        golden[name] = entry
        print(f"{name:40s} C={len(blob):8d} par rc={par.rc} U={len(par.data):8d} entries={len(par.offsets):3d} "
              f"ser rc={ser.rc} cliP4 rc={entry['cli_parallel']['rc']} {entry['cli_parallel']['stderr'][:60]!r}")

    (HERE / "golden.json").write_text(json.dumps(golden, indent=1, sort_keys=True) + "\n")
    total = sum(len(b) for b in cases.values())
    print(f"{len(cases)} cases, {total} compressed bytes")


if __name__ == "__main__":
    main()
