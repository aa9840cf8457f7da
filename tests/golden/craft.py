"""Hand-made bzip2 streams for the cases libbz2's encoder never produces (test infrastructure).

libbz2 caps its code lengths at 17 bits and always fills the symbol map, so codes of 18..20 bits
(bzip2.hpp:97 allows 20), chosen numbers of Huffman groups on long blocks and an empty symbol map
(bzip2.hpp:526-571 followed by :723) have to be written by hand.  This is a plain-Python bzip2
block WRITER following the format as the reference reads it (bzip2.hpp:479-807): RLE1, naive BWT,
MTF + RUNA/RUNB, canonical codes from caller-chosen lengths.  Only tests/golden/make_golden.py
uses it; what the streams decode to is decided by the compiled reference, not by this file.
"""
from __future__ import annotations

MAGIC_BLOCK = 0x314159265359
MAGIC_EOS = 0x177245385090


class BitWriter:
    def __init__(self) -> None:
        self.bits: list[int] = []

    def put(self, value: int, n: int) -> None:
        for i in range(n - 1, -1, -1):
            self.bits.append((value >> i) & 1)

    def tobytes(self) -> bytes:
        bits = self.bits + [0] * (-len(self.bits) % 8)
        out = bytearray()
        for i in range(0, len(bits), 8):
            b = 0
            for bit in bits[i:i + 8]:
                b = (b << 1) | bit
            out.append(b)
        return bytes(out)


def crc32_bzip2(data: bytes) -> int:
    crc = 0xFFFFFFFF
    for byte in data:
        crc ^= byte << 24
        for _ in range(8):
            crc = ((crc << 1) ^ 0x04C11DB7) & 0xFFFFFFFF if crc & 0x80000000 else (crc << 1) & 0xFFFFFFFF
    return crc ^ 0xFFFFFFFF


def rle1(data: bytes) -> bytes:
    out = bytearray()
    i = 0
    while i < len(data):
        j = i
        while j < len(data) and data[j] == data[i] and j - i < 255:
            j += 1
        run = j - i
        if run >= 4:
            out += bytes([data[i]]) * 4 + bytes([run - 4])
        else:
            out += bytes([data[i]]) * run
        i = j
    return bytes(out)


def bwt(data: bytes) -> tuple[bytes, int]:
    n = len(data)
    doubled = data + data
    order = sorted(range(n), key=lambda i: doubled[i:i + n])
    return bytes(doubled[i + n - 1] for i in order), order.index(0)


def mtf_rle2(last: bytes, used: list[int]) -> list[int]:
    """symbols 0 = RUNA, 1 = RUNB, k + 1 = MTF index k > 0, len(used) + 1 = end of block"""
    mtf = list(used)
    symbols: list[int] = []
    zeros = 0

    def flush() -> None:
        nonlocal zeros
        while zeros > 0:   # bijective base 2
            symbols.append((zeros - 1) & 1)
            zeros = (zeros - 1) >> 1

    for byte in last:
        k = mtf.index(byte)
        if k == 0:
            zeros += 1
            continue
        flush()
        mtf.insert(0, mtf.pop(k))
        symbols.append(k + 1)
    flush()
    symbols.append(len(used) + 1)
    return symbols


def canonical_codes(lengths: list[int]) -> list[int]:
    codes = [0] * len(lengths)
    code = 0
    for length in range(min(lengths), max(lengths) + 1):
        for s, l in enumerate(lengths):
            if l == length:
                codes[s] = code
                code += 1
        code <<= 1
    return codes


def skewed_lengths(alphabet: int, order: list[int] | None = None, longest: int = 20) -> list[int]:
    """A complete prefix code whose longest codes have `longest` bits: 1, 2, 3, ... for the first symbols of `order`
    (default: symbol order), the remaining symbols share what is left as evenly as a complete code allows."""
    order = list(range(alphabet)) if order is None else order
    lengths = [0] * alphabet
    chain = max(0, min(alphabet - 2, longest - 1))   # symbols with lengths 1..chain
    while chain > 0 and chain + (alphabet - chain - 1).bit_length() > longest:
        chain -= 1
    rest = alphabet - chain                    # they share a subtree that hangs at depth `chain`
    for i in range(chain):
        lengths[order[i]] = i + 1
    # complete code for `rest` leaves below depth `chain`: depths d or d + 1
    depth = 0
    while (1 << (depth + 1)) <= rest:
        depth += 1
    deeper = 2 * (rest - (1 << depth))        # leaves at depth + 1
    for i in range(rest):
        lengths[order[chain + i]] = chain + depth + (1 if i >= rest - deeper else 0)
    assert max(lengths) <= 20 and abs(sum(2.0 ** -l for l in lengths) - 1.0) < 1e-12, lengths
    return lengths


def block_bits(w: BitWriter, data: bytes, tables: list[list[int]] | None, selectors_of=None,
               raw_symbols: list[int] | None = None, total_selectors: int | None = None,
               bad_selector: int | None = None, fields: dict | None = None) -> int:
    """Appends one block; returns its CRC.  tables: code lengths per group over the alphabet (len(used) + 2 symbols),
    None = two tables of balanced codes.  selectors_of(group index) -> table.  raw_symbols: write these symbols
    under an EMPTY symbol map (alphabet RUNA, RUNB) instead of what `data` transforms to.  total_selectors: write
    this many selectors (at most 32767, the 15-bit field), more than the block's groups need.  bad_selector: selector
    number whose unary code has as many ones as there are tables (one too many).  fields: header fields written
    with the given value instead of the right one: "orig" (24 bits), "ngroups" (3), "nselectors" (15),
    "first_length" (the 5-bit start length of table 0); "used": with raw_symbols, the byte values of the symbol map
    (default none); "skip_data": nothing behind the code lengths (for tables no code can be made of)."""
    fields = fields or {}
    crc = crc32_bzip2(data)
    packed = rle1(data)
    last, orig = bwt(packed)
    used = sorted(set(packed))
    symbols = mtf_rle2(last, used)
    empty_symbol_map = raw_symbols is not None
    if empty_symbol_map:
        used, symbols, orig = list(fields.get("used", [])), list(raw_symbols), 0
        empty_symbol_map = not used
    alphabet = len(used) + 2
    if tables is None:
        tables = [skewed_lengths(alphabet, longest=0), skewed_lengths(alphabet, longest=0)]   # balanced
    assert 2 <= len(tables) <= 6 and all(len(t) == alphabet for t in tables)
    ngroups = (len(symbols) + 49) // 50
    nselectors = ngroups if total_selectors is None else total_selectors
    assert ngroups <= nselectors <= 32767
    selectors = [(selectors_of(g) if selectors_of else g % len(tables)) for g in range(nselectors)]

    w.put(MAGIC_BLOCK, 48)
    w.put(crc, 32)
    w.put(0, 1)
    w.put(fields.get("orig", orig), 24)
    if empty_symbol_map:
        w.put(0, 16)
    else:
        groups_used = 0
        for b in used:
            groups_used |= 1 << (15 - (b >> 4))
        w.put(groups_used, 16)
        for g in range(16):
            if groups_used & (1 << (15 - g)):
                mask = 0
                for b in used:
                    if b >> 4 == g:
                        mask |= 1 << (15 - (b & 15))
                w.put(mask, 16)
    w.put(fields.get("ngroups", len(tables)), 3)
    w.put(fields.get("nselectors", len(selectors)), 15)
    order = list(range(len(tables)))
    for i, s in enumerate(selectors):
        k = order.index(s)
        order.insert(0, order.pop(k))
        if i == bad_selector:
            k = len(tables)
        w.put((1 << (k + 1)) - 2, k + 1)   # k ones and a zero
    for number, lengths in enumerate(tables):
        cur = lengths[0]
        w.put(fields.get("first_length", cur) if number == 0 else cur, 5)
        for l in lengths:
            while cur != l:
                w.put(2 if cur < l else 3, 2)
                cur += 1 if cur < l else -1
            w.put(0, 1)
    if fields.get("skip_data"):
        return crc
    codes = [canonical_codes(t) for t in tables]
    for i, s in enumerate(symbols):
        t = selectors[i // 50]
        w.put(codes[t][s], tables[t][s])
    return crc


def stream(blocks: list[dict], level: int = 9) -> bytes:
    """blocks: keyword arguments of block_bits (data=..., tables=..., selectors_of=..., empty_symbol_map=...)"""
    w = BitWriter()
    for ch in b"BZh" + bytes([48 + level]):
        w.put(ch, 8)
    combined = 0
    for blk in blocks:
        crc = block_bits(w, **blk)
        combined = (((combined << 1) | (combined >> 31)) & 0xFFFFFFFF) ^ crc
    w.put(MAGIC_EOS, 48)
    w.put(combined, 32)
    return w.tobytes()
