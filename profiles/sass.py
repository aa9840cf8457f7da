#!/usr/bin/env python3
"""Writes profiles/sass/<kernel>.sass: the SASS of the hot kernels as they are in the shipped library
(cuobjdump -sass indexed_bzip2_b200/libibz2_cuda.so), one file per kernel, encodings stripped, every backward
branch marked as the end of a loop.  With an `ncu --set full --import-source on` capture of the kernel the
listing also carries, per instruction, how often it was executed and its share of the warp-stall samples, and the
loop that holds most of the samples is marked HOT.

    python profiles/sass.py [capture.ncu-rep ...]
"""
from __future__ import annotations

import csv
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "indexed_bzip2_b200" / "libibz2_cuda.so"
OUT = ROOT / "profiles" / "sass"
KERNELS = ["findMagicsKernel", "huffmanHeaderKernel", "huffmanBitsKernelILi4E", "huffmanGroupsKernel", "huffmanPipelineKernel", "resolveSymbolsKernel", "mtfChunkKernel",
           "mtfExpandKernel", "bwtPrepareKernel", "bwtWalkKernel", "bwtRankKernel", "bwtPlaceKernel", "rleScanKernel",
           "rleWriteKernel", "crcKernel"]


def functions() -> dict[str, list[str]]:
    text = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    out: dict[str, list[str]] = {}
    name = None
    for line in text.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = m.group(1)
            out[name] = []
        elif name is not None and re.match(r"\s*/\*[0-9a-f]{4,}\*/", line):
            out[name].append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", line).rstrip())
    return out


def stall_profile(rep: Path) -> dict[str, dict[int, tuple[int, float]]]:
    """kernel name -> { instruction offset -> (executed, share of samples) }"""
    text = subprocess.run(["ncu", "-i", str(rep), "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(text.splitlines()))
    result: dict[str, dict[int, tuple[int, float]]] = {}
    kernel = None
    i = 0
    while i < len(rows):
        r = rows[i]
        if r and r[0] == "Kernel Name":
            kernel = r[1]
        if r and r[0] == "Address" and kernel is not None:
            h = r
            ia, ie, isamp = h.index("Address"), h.index("Instructions Executed"), h.index("# Samples")
            body = []
            i += 1
            while i < len(rows) and len(rows[i]) == len(h) and rows[i][0] != "Address":
                body.append(rows[i])
                i += 1
            total = sum(int(b[isamp]) for b in body) or 1
            base = int(body[0][ia], 16)
            result[kernel] = {int(b[ia], 16) - base: (int(b[ie]), int(b[isamp]) / total) for b in body}
            continue
        i += 1
    return result


def main() -> None:
    OUT.mkdir(parents=True, exist_ok=True)
    profiles: dict[str, dict[int, tuple[int, float]]] = {}
    for rep in sys.argv[1:]:
        profiles.update(stall_profile(Path(rep)))
    for mangled, lines in functions().items():
        short = next((k for k in KERNELS if k in mangled), None)
        if short is None:
            continue
        pretty = short.replace("ILi4E", "_lanes4")
        prof = next((p for k, p in profiles.items() if pretty.split("_")[0] in k), None)
        target = OUT / f"{pretty}.sass"
        if prof is None and target.exists() and "share of the stall samples" in target.read_text():
            # an annotated listing from an earlier capture is worth more than a bare one: leave it
            print(f"{pretty}: kept (annotated by an earlier capture)")
            continue
        # loops: a branch to a lower address closes one
        loops = []
        for line in lines:
            m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(?:@!?U?P\d\s+)?BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?0x([0-9a-f]+)", line)
            if m and int(m.group(2), 16) < int(m.group(1), 16):
                loops.append((int(m.group(2), 16), int(m.group(1), 16)))
        hot = None
        if prof and loops:
            def weight(loop):
                return sum(s for a, (_, s) in prof.items() if loop[0] <= a <= loop[1]) / (loop[1] - loop[0] + 16)
            hot = max(loops, key=lambda l: sum(s for a, (_, s) in prof.items() if l[0] <= a <= l[1]) if weight(l) > 0 else 0)
            inner = [l for l in loops if hot[0] <= l[0] and l[1] <= hot[1] and l != hot
                     and sum(s for a, (_, s) in prof.items() if l[0] <= a <= l[1]) > 0.6 * sum(s for a, (_, s) in prof.items() if hot[0] <= a <= hot[1])]
            if inner:
                hot = min(inner, key=lambda l: l[1] - l[0])
        text = [f"// {mangled}", f"// from {LIB.relative_to(ROOT)} (sm_100a), {len(lines)} instructions, loops: "
                + ", ".join(f"{a:#06x}..{b:#06x}" for a, b in loops)]
        if hot:
            share = sum(s for a, (_, s) in prof.items() if hot[0] <= a <= hot[1])
            text.append(f"// HOT loop {hot[0]:#06x}..{hot[1]:#06x}: {100 * share:.1f} % of the warp-stall samples of the capture")
        if prof:
            text.append("// columns behind the instruction: times executed (warp level), share of the stall samples")
        for line in lines:
            m = re.match(r"\s*/\*([0-9a-f]+)\*/", line)
            addr = int(m.group(1), 16)
            mark = ""
            if any(addr == a for a, _ in loops):
                mark += "   // <- loop head"
            if any(addr == b for _, b in loops):
                mark += "   // <- loop end"
            if hot and addr == hot[0]:
                mark += " HOT"
            if prof and addr in prof:
                mark = f"   // {prof[addr][0]:>12d} {100 * prof[addr][1]:5.2f}%" + mark.replace("   //", "")
            text.append(line.strip() + mark)
        (OUT / f"{pretty}.sass").write_text("\n".join(text) + "\n")
        print(f"{pretty}: {len(lines)} instructions, {len(loops)} loops" + (f", hot {hot[0]:#x}..{hot[1]:#x}" if hot else ""))


if __name__ == "__main__":
    main()
