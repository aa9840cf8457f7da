"""Pins oracle/bz2_oracle.c (our CPU restatement) to the reference: its own known-answer vectors and the golden
outputs of the compiled, unmodified reference committed in tests/golden/golden.json. CPU only."""
import bz2
import hashlib
import json

import numpy as np
import pytest

import _oracle as orc

GOLDEN = json.loads((orc.ROOT / "tests" / "golden" / "golden.json").read_text())
VALID = sorted(k for k, v in GOLDEN.items() if v["parallel"]["rc"] == 0)
ALL = sorted(GOLDEN)

# reference error text -> status of the restatement
ERROR_STATUS = {"err-bitflip": 13, "err-truncated": 14, "err-empty-symbol-map-runb": 9, "err-empty-symbol-map-runs": 9,
                "err-stream-crc": 17, "err-stream-crc-multistream": 17, "err-selector-unary-late": 6,
                "err-origptr": 3, "err-group-count-1": 4, "err-group-count-7": 4, "err-selector-count-0": 5,
                "err-code-length-0": 7, "err-code-length-21": 7, "err-origptr-error": 12,
                "err-code-lengths-oversubscribed": 8, "err-run-longer-than-block": 10}


def case(name):
    return (orc.ROOT / "tests" / "golden" / "cases" / f"{name}.bz2").read_bytes()


# --- known-answer vectors held by the reference's own tests -----------------------------------------------------

BLOCK_MAGIC_BYTES = [0x31, 0x41, 0x59, 0x26, 0x53, 0x59]


@pytest.mark.parametrize("data,expected", [
    ([0x11, 0x41, 0x59, 0x26, 0x53, 0x59], []),
    ([0x31, 0x41, 0x59, 0x26, 0x53, 0x58], []),
    ([0x31, 0x41, 0x59, 0x26, 0x53, 0x59], [0]),
    ([0x31, 0x41, 0x59, 0x26, 0x53, 0x59, 0, 0], [0]),
    ([0, 0x31, 0x41, 0x59, 0x26, 0x53, 0x59, 0, 0], [8]),
    ([0, 0, 0x31, 0x41, 0x59, 0x26, 0x53, 0x59, 0, 0], [16]),
    ([0, 0, 0, 0x31, 0x41, 0x59, 0x26, 0x53, 0x59, 0, 0], [24]),
    ([0, 0, 0, 0, 0x31, 0x41, 0x59, 0x26, 0x53, 0x59, 0, 0], [32]),
    ([0x18, 0xA0, 0xAC, 0x93, 0x29, 0xAC, 0x80], [1]),
    ([0x00, 0x62, 0x82, 0xB2, 0x4C, 0xA6, 0xB2], [7]),
])
def test_finder_kat(data, expected):
    """src/tests/core/testBitStringFinder.cpp:117-128"""
    got = orc.find_magic(bytes(data), orc.MAGIC_BLOCK)
    assert got.tolist() == expected


@pytest.mark.parametrize("gap", [1, 100, 123, 1024, 2000])
def test_finder_second_match(gap):
    """src/tests/core/testBitStringFinder.cpp:131-146"""
    buf = [0, 0, 0, 0] + BLOCK_MAGIC_BYTES + [0, 0]
    data = bytes(buf) + bytes(gap) + bytes(BLOCK_MAGIC_BYTES)
    assert orc.find_magic(data, orc.MAGIC_BLOCK).tolist() == [32, (len(buf) + gap) * 8]


RUN_TABLE = """A B AA BA AB BB AAA BAA ABA BBA AAB BAB ABB BBB AAAA BAAA ABAA BBAA AABA BABA ABBA BBBA AAAB BAAB ABAB
BBAB AABB BABB ABBB BBBB AAAAA BAAAA ABAAA BBAAA AABAA BABAA ABBAA BBBAA AAABA BAABA ABABA BBABA AABBA BABBA ABBBA
BBBBA AAAAB BAAAB ABAAB BBAAB AABAB BABAB ABBAB BBBAB AAABB BAABB ABABB BBABB AABBB BABBB ABBBB BBBBB AAAAAA""".split()


def test_run_ab_table():
    """src/tests/indexed_bzip2/testRunAB.cpp:12-77: the i-th string encodes run length i (bijective base 2)."""
    assert len(RUN_TABLE) == 63
    for i, s in enumerate(RUN_TABLE, start=1):
        assert orc.run_length(s) == i


def test_crc_check_value():
    """CRC-32/BZIP2 catalogue check value; table construction src/indexed_bzip2/bzip2.hpp:59-91."""
    assert orc.crc32(b"123456789") == 0xFC891918
    assert orc.crc32(b"") == 0


def test_first_block_offset_is_32():
    """src/tests/indexed_bzip2/testBZ2Reader.cpp:284-285"""
    r = orc.serial_reader(case("ref-base64-32KiB"))
    assert r.offsets[0] == (32, 0)


def test_1B_index():
    """src/tests/rapidgzip/testParallelGzipReader.cpp:1133: 1B.bz2 -> {32:0, 296:1} through the gzip-architecture
    route (chunk index without the EOS entry); ParallelBZ2Reader adds the EOS block at bit 211."""
    r = orc.parallel_reader(case("ref-1B"))
    assert dict(r.offsets) == {32: 0, 211: 1, 296: 1}


# --- golden outputs of the compiled reference -------------------------------------------------------------------

@pytest.mark.parametrize("name", ALL)
def test_finder_matches_reference_cli(name):
    g = GOLDEN[name]
    assert orc.find_magics(case(name)).tolist() == g["finder"]


@pytest.mark.parametrize("name", ALL)
def test_parallel_reader_matches_reference(name):
    g = GOLDEN[name]["parallel"]
    r = orc.parallel_reader(case(name))
    if g["rc"] != 0:
        assert r.status == ERROR_STATUS[name]
        return
    assert r.status == 0
    assert len(r.data) == g["size"]
    assert hashlib.sha256(r.data).hexdigest() == g["sha256"]
    assert [list(x) for x in r.offsets] == g["offsets"]
    assert r.warn_trailing == ("Trailing garbage" in GOLDEN[name]["cli_parallel"]["stderr"])


@pytest.mark.parametrize("name", ALL)
def test_serial_reader_matches_reference(name):
    g = GOLDEN[name]["serial"]
    r = orc.serial_reader(case(name))
    if g["rc"] != 0:
        assert r.status != 0
        if name in ERROR_STATUS:
            assert r.status == ERROR_STATUS[name]
        return
    assert r.status == 0
    assert hashlib.sha256(r.data).hexdigest() == g["sha256"]
    assert [list(x) for x in r.offsets] == g["offsets"]


@pytest.mark.parametrize("name", VALID)
def test_block_stages_consistent(name):
    """Stage-level outputs of the restatement are self-consistent and agree with libbz2 on the payload."""
    data = case(name)
    g = GOLDEN[name]
    blocks = [o for o in orc.find_magic(data, orc.MAGIC_BLOCK).tolist()]
    if g["libbz2_sha256"] is not None:
        assert hashlib.sha256(orc.parallel_reader(data).data).hexdigest() == g["libbz2_sha256"]
    for off in blocks[:3]:
        b = orc.decode_block(data, off)
        if b.status != 0:
            continue
        n = b.info.n
        assert np.bincount(b.lbytes, minlength=256).tolist() == b.hist.tolist()
        assert (b.tvec & 0xFF).astype(np.uint8).tolist() == b.lbytes.tolist()
        t = (b.tvec >> 8).astype(np.int64)
        assert sorted(t.tolist()) == list(range(n))             # a permutation
        assert np.all(np.diff(b.lbytes[t].astype(np.int64)) >= 0)  # L[T[k]] = F[k] is sorted
        assert orc.crc32(b.out.tobytes()) == b.info.calc_crc == b.info.header_crc


def test_status_table_matches_header():
    import re
    hdr = (orc.ROOT / "oracle" / "bz2_oracle.h").read_text()
    vals = {m.group(1): int(m.group(2)) for m in re.finditer(r"ORC_(?:ERR_)?(\w+) = (\d+)", hdr)}
    for code, nm in orc.STATUS_NAMES.items():
        assert vals[nm] == code
