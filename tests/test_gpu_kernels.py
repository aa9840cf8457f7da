"""GPU parity tests: every kernel stage of the sm_100a path, called through the C ABI (include/ibz2_cuda.h), against
the CPU oracle (oracle/bz2_oracle.c, pinned to the reference by tests/test_oracle.py) and the golden outputs of the
compiled reference (tests/golden/golden.json).  Bit-exact: all of this is integer / byte work.

Run on the B200 box: python -m pytest tests -m gpu
"""
import hashlib
import os
import json

import numpy as np
import pytest

import _oracle as orc
from indexed_bzip2_b200 import _capi

pytestmark = pytest.mark.gpu

GOLDEN = json.loads((orc.ROOT / "tests" / "golden" / "golden.json").read_text())
VALID = sorted(k for k, v in GOLDEN.items() if v["parallel"]["rc"] == 0)
ALL = sorted(GOLDEN)
EOS_FLAG = np.uint64(1 << 63)
BLOCK_MAGIC_BYTES = [0x31, 0x41, 0x59, 0x26, 0x53, 0x59]
EOS_MAGIC_BYTES = [0x17, 0x72, 0x45, 0x38, 0x50, 0x90]


def case(name):
    return (orc.ROOT / "tests" / "golden" / "cases" / f"{name}.bz2").read_bytes()


# K2b has two kernels, a lane per block for launches with many blocks and a CTA per block for launches with few
# (csrc/huffman_bits.cu); the launches of this file are all small, so each kernel is forced in turn
@pytest.fixture(scope="module", params=["lanes", "groups", "pipeline"])
def ctx(request):
    os.environ["IBZ2_K2B_KERNEL"] = request.param
    c = _capi.Context(0)
    # these tests compare block by block with the oracle's decode_block, which knows nothing about streams: the library's
    # stream CRC verdict (IBZ2_ST_STREAM_CRC on the last block of a stream) is switched on where it is the subject
    c.set_verify_stream_crc(False)
    yield c
    c.close()
    os.environ.pop("IBZ2_K2B_KERNEL", None)


def split_found(found):
    offsets = (found & ~EOS_FLAG).astype(np.uint64)
    is_eos = (found & EOS_FLAG) != 0
    return offsets, is_eos


def data_blocks(ctx, data):
    offsets, is_eos = split_found(ctx.find_magics(data))
    return offsets[~is_eos]


# ---------------------------------------------------------------------------------------------------------------
# K1 finder
# ---------------------------------------------------------------------------------------------------------------

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
def test_finder_kat(ctx, data, expected):
    """src/tests/core/testBitStringFinder.cpp:117-128"""
    got = ctx.find_magics(bytes(data), which=_capi.FIND_BLOCK)
    assert got.tolist() == expected


@pytest.mark.parametrize("gap", [1, 100, 123, 1024, 2000, 4095, 4096, 4097, 28 * 1024, 4 << 20])
def test_finder_second_match(ctx, gap):
    """src/tests/core/testBitStringFinder.cpp:131-146, testParallelBitStringFinder.cpp:228-249"""
    buf = [0, 0, 0, 0] + BLOCK_MAGIC_BYTES + [0, 0]
    data = bytes(buf) + bytes(gap) + bytes(BLOCK_MAGIC_BYTES)
    assert ctx.find_magics(data, which=_capi.FIND_BLOCK).tolist() == [32, (len(buf) + gap) * 8]


def plant(buf: np.ndarray, bitpos: int, magic: int):
    """write the 48-bit magic at an arbitrary bit position of a uint8 buffer (MSB-first)"""
    for i in range(48):
        bit = (magic >> (47 - i)) & 1
        p = bitpos + i
        byte, sh = p >> 3, 7 - (p & 7)
        buf[byte] = (int(buf[byte]) & ~(1 << sh)) | (bit << sh)


def test_finder_planted_every_alignment(ctx):
    """both magics at every bit alignment and straddling 4 B / 16 B / 128 B / 512 B / 4 KiB / CTA-tile seams"""
    rng = np.random.default_rng(2)
    n = 3 * 1024 * 1024 + 7
    buf = rng.integers(0, 256, n, dtype=np.uint8)
    positions = []
    cursor = 64
    for seam in [4, 16, 128, 512, 4096, 16384, 65536, 1 << 20]:
        for back in range(0, 56, 1):
            base = (cursor // seam + 1) * seam * 8  # bit index of a seam
            p = base - back
            positions.append(p)
            cursor = p // 8 + 40
    for s in range(64):
        positions.append((cursor + 50) * 8 + s)
        cursor += 20
    positions.append((n - 6) * 8)  # flush with the end of the buffer
    positions = sorted(set(positions))
    kept = []
    last_end = -1
    for i, p in enumerate(positions):
        if p <= last_end or p + 48 > n * 8:
            continue
        plant(buf, p, _capi.MAGIC_BLOCK if i % 2 == 0 else _capi.MAGIC_EOS)
        kept.append(p)
        last_end = p + 48
    expected = orc.find_magics(buf)
    assert set(kept) <= set(expected.tolist())
    found = ctx.find_magics(buf)
    offsets, is_eos = split_found(found)
    assert offsets.tolist() == expected.tolist()
    assert offsets[is_eos].tolist() == orc.find_magic(buf, orc.MAGIC_EOS).tolist()
    only_block = ctx.find_magics(buf, which=_capi.FIND_BLOCK)
    assert only_block.tolist() == orc.find_magic(buf, orc.MAGIC_BLOCK).tolist()


@pytest.mark.parametrize("n", list(range(0, 41)) + [63, 64, 65, 127, 129, 255, 257])
def test_finder_ragged_sizes(ctx, n):
    """tiny and ragged buffers: magic flush with the end, sizes that are not multiples of 4 or 16"""
    buf = np.zeros(n, dtype=np.uint8)
    if n >= 6:
        buf[n - 6:] = EOS_MAGIC_BYTES
    if n >= 14:
        plant(buf, 3, _capi.MAGIC_BLOCK)
    found = ctx.find_magics(buf)
    offsets, _ = split_found(found)
    assert offsets.tolist() == orc.find_magics(buf).tolist()


@pytest.mark.parametrize("name", ALL)
def test_finder_golden(ctx, name):
    """`ibzip2 -l` of the compiled reference (tools/ibzip2.cpp:96-143), committed in golden.json"""
    data = case(name)
    offsets, is_eos = split_found(ctx.find_magics(data))
    assert offsets.tolist() == GOLDEN[name]["finder"]
    assert offsets[is_eos].tolist() == orc.find_magic(data, orc.MAGIC_EOS).tolist()


def test_finder_device_resident_and_base_offset(ctx):
    data = np.frombuffer(case("text-450k-l1"), dtype=np.uint8)
    dptr = ctx.device_alloc(data.size)
    try:
        ctx.h2d(dptr, data)
        found = ctx.find_magics(dptr, data.size, device=True, base_bit_offset=1000)
        offsets, _ = split_found(found)
        assert (offsets - np.uint64(1000)).tolist() == GOLDEN["text-450k-l1"]["finder"]
    finally:
        ctx.device_free(dptr)


# ---------------------------------------------------------------------------------------------------------------
# K2..K6 stage by stage
# ---------------------------------------------------------------------------------------------------------------

def candidate_offsets(name):
    return GOLDEN[name]["finder"]


@pytest.mark.parametrize("name", ALL)
def test_stage_parity(ctx, name):
    """every candidate block of every golden file: header fields, L, packed T vector, pre-RLE1 bytes, output, CRC"""
    data = case(name)
    for off in candidate_offsets(name):
        want = orc.decode_block(data, off)
        got = ctx.debug_decode_block(data, off, out_capacity=max(int(want.info.out_size), 1) + 16)
        res = got["result"]
        label = f"{name}@{off}"
        assert bool(res.is_eos) == bool(want.info.is_eos), label
        if want.info.is_eos:
            assert res.status == want.info.status, label
            assert res.enc_size_bits == want.info.enc_size_bits, label
            assert bool(res.is_eof) == bool(want.info.at_eof), label
            assert res.expected_crc == want.info.header_crc, label
            continue
        if want.info.status not in (0, 13):
            assert res.status == want.info.status, f"{label}: status {_capi.STATUS_NAMES.get(res.status)} vs oracle {orc.STATUS_NAMES.get(want.info.status)}"
            continue
        # stage by stage, so that the first differing stage is the one reported
        assert res.expected_crc == want.info.header_crc, label
        assert res.n_symbols == want.info.n, f"{label}: N (status {_capi.STATUS_NAMES.get(res.status)})"
        assert res.orig_ptr == want.info.orig_ptr, label
        assert res.enc_size_bits == want.info.enc_size_bits, label
        assert np.array_equal(got["lbytes"], want.lbytes), f"{label}: L differs"
        # ours packs (T[k] << 8) | L[T[k]], the reference (T[k] << 8) | L[k] (see bwt_prepare.cu)
        t = want.tvec >> 8
        assert np.array_equal(got["tvec"], (t << 8) | want.lbytes[t]), f"{label}: T vector differs"
        assert np.array_equal(got["prerle"], want.prerle), f"{label}: pre-RLE1 bytes differ at {np.flatnonzero(got['prerle'] != want.prerle)[:8]}"
        assert res.out_size == want.info.out_size, label
        assert np.array_equal(got["out"], want.out), f"{label}: output differs at {np.flatnonzero(got['out'] != want.out)[:8]}"
        assert res.calculated_crc == want.info.calc_crc, f"{label}: CRC {res.calculated_crc:08x} vs {want.info.calc_crc:08x}"
        assert res.status == want.info.status, f"{label}: status {_capi.STATUS_NAMES.get(res.status)} vs oracle {orc.STATUS_NAMES.get(want.info.status)}"


# ---------------------------------------------------------------------------------------------------------------
# whole files through ibz2_decode_blocks
# ---------------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("name", VALID)
def test_decode_file_matches_reference(ctx, name):
    data = case(name)
    blocks = data_blocks(ctx, data)
    size = GOLDEN[name]["parallel"]["size"]
    out = np.zeros(size + 64, dtype=np.uint8)
    results, total = ctx.decode_blocks(data, blocks, out)
    if blocks.size == 0:
        assert total == 0 == size
        return
    # err-gap-then-valid style files contain blocks the reader never reaches; only compare what the reference emits
    if name.startswith("err-"):
        return
    assert total == size
    assert (results["status"] == 0).all(), results["status"]
    assert hashlib.sha256(out[:total].tobytes()).hexdigest() == GOLDEN[name]["parallel"]["sha256"]
    assert (results["calculated_crc"] == results["expected_crc"]).all()
    # blocks are laid out back to back
    assert results["out_offset"].tolist() == np.concatenate([[0], np.cumsum(results["out_size"])[:-1]]).tolist()
    # each data block ends where the next header starts (data block or EOS): the reference's block map keys
    ends = (results["enc_offset_bits"] + results["enc_size_bits"]).tolist()
    keys = [k for k, _ in GOLDEN[name]["parallel"]["offsets"]]
    finder = set(GOLDEN[name]["finder"])
    for e in ends:
        assert e in keys or e in finder


def test_decode_multi_wave_and_device_buffers(ctx):
    name = "text-450k-l1"
    data = np.frombuffer(case(name), dtype=np.uint8)
    blocks = data_blocks(ctx, data)
    assert blocks.size >= 4
    size = GOLDEN[name]["parallel"]["size"]
    din = ctx.device_alloc(data.size)
    dout = ctx.device_alloc(size)
    try:
        ctx.set_batch_blocks(2)  # forces several waves with a carried output offset
        ctx.h2d(din, data)
        results, total = ctx.decode_blocks(din, blocks, dout, size, nbytes=data.size, data_device=True, out_device=True)
        assert total == size and (results["status"] == 0).all()
        host = np.zeros(size, dtype=np.uint8)
        ctx.d2h(host, dout)
        assert hashlib.sha256(host.tobytes()).hexdigest() == GOLDEN[name]["parallel"]["sha256"]
        t = ctx.timings()
        assert t["launches"] >= 8 * 2 and t["huffman_ms"] > 0
    finally:
        ctx.set_batch_blocks(148 * 12)
        ctx.device_free(din)
        ctx.device_free(dout)


def test_decode_pinned_host_buffers(ctx):
    name = "mixed-384k-l2"
    raw = case(name)
    size = GOLDEN[name]["parallel"]["size"]
    pin_in = ctx.pinned(len(raw))
    pin_out = ctx.pinned(size)
    try:
        pin_in.array[:] = np.frombuffer(raw, dtype=np.uint8)
        blocks = data_blocks(ctx, pin_in.array)
        results, total = ctx.decode_blocks(pin_in.array, blocks, pin_out.array)
        assert total == size and (results["status"] == 0).all()
        assert hashlib.sha256(pin_out.array.tobytes()).hexdigest() == GOLDEN[name]["parallel"]["sha256"]
    finally:
        pin_in.free()
        pin_out.free()


def test_output_capacity_and_size_only(ctx):
    name = "base64-300k-l1"
    data = case(name)
    blocks = data_blocks(ctx, data)
    size = GOLDEN[name]["parallel"]["size"]
    # size-only call: no output buffer
    results, total = ctx.decode_blocks(data, blocks, None)
    assert total == size and (results["status"] == 0).all()
    sizes = results["out_size"].copy()
    # buffer that holds all but the last block
    cap = int(sizes[:-1].sum()) + 5
    out = np.zeros(cap, dtype=np.uint8)
    results, total = ctx.decode_blocks(data, blocks, out)
    assert total == size
    assert results["status"][:-1].tolist() == [0] * (blocks.size - 1)
    assert results["status"][-1] == 15
    full = np.zeros(size, dtype=np.uint8)
    ctx.decode_blocks(data, blocks, full)
    assert np.array_equal(out[:cap - 5], full[:cap - 5])


def test_error_statuses_match_oracle(ctx):
    for name in [n for n in ALL if n.startswith("err-")]:
        data = case(name)
        offs = np.array(GOLDEN[name]["finder"], dtype=np.uint64)
        if offs.size == 0:
            continue
        out = np.zeros(4 << 20, dtype=np.uint8)
        results, _ = ctx.decode_blocks(data, offs, out)
        for off, r in zip(offs.tolist(), results):
            want = orc.decode_block(data, off, stages=False)
            assert r["status"] == want.info.status, f"{name}@{off}"


def reference_stream_crcs(data: bytes):
    """(bit offset of the stream's last data block, CRC field of its EOS block, fold of its block CRCs) per stream, from
    the oracle's block headers: BZ2Reader.hpp:406-416 / :481-484 restated on the host"""
    streams = []
    crc = 0
    last = None
    eos = set(orc.find_magic(data, orc.MAGIC_EOS).tolist())
    for offset in orc.find_magics(data).tolist():
        if offset in eos:
            header = _capi.read_block_header(data, offset)
            if last is not None:
                streams.append((last, header.expected_crc, crc))
            crc, last = 0, None
        else:
            block = orc.decode_block(data, offset, stages=False)
            crc = _capi.stream_crc_fold(crc, block.info.calc_crc)
            last = offset
    return streams


@pytest.mark.parametrize("name", VALID)
def test_stream_crc_fold(ctx, name):
    """north_star step 6: the block CRCs are combined into the stream CRC and compared with the end-of-stream block.
    Every stream of every golden file that the reference decodes gets checked; the two err-stream-crc files (every block
    CRC right, the stream CRC wrong) get IBZ2_ST_STREAM_CRC on the last block of the damaged stream, and only there."""
    data = case(name)
    if name == "err-gap-then-valid":
        pytest.skip("the finder lists a block behind a gap that belongs to no stream the readers reach")
    want = reference_stream_crcs(data)
    out = np.zeros(GOLDEN[name]["parallel"]["size"] + 64, dtype=np.uint8)
    ctx.set_verify_stream_crc(True)   # the library's default
    try:
        results, total = ctx.decode_buffer(data, out, max_blocks=4096)
    finally:
        ctx.set_verify_stream_crc(False)
    assert total == GOLDEN[name]["parallel"]["size"]
    closing = results[(results["stream_flags"] & 2) != 0]
    assert closing["enc_offset_bits"].tolist() == [w[0] for w in want]
    assert closing["stream_crc_expected"].tolist() == [w[1] for w in want]
    assert ((closing["stream_flags"] & 4) != 0).all()           # every stream starts inside the call: all checked
    assert closing["stream_crc_calculated"].tolist() == [w[2] for w in want]
    bad = [w[0] for w in want if w[1] != w[2]]
    assert results["enc_offset_bits"][results["status"] == 17].tolist() == bad
    assert (len(bad) > 0) == name.startswith("err-stream-crc")
    assert ((results["status"] == 0) | (results["status"] == 17)).all()
    # facts only, no verdict
    quiet, _ = ctx.decode_buffer(data, out, max_blocks=4096)
    assert (quiet["status"] == 0).all() and quiet["stream_crc_calculated"].tolist() == results["stream_crc_calculated"].tolist()
    # a call that begins in the middle of a stream cannot vouch for that stream
    if len(results) > 1 and (results["stream_flags"][1] & 1) == 0:
        ctx.set_verify_stream_crc(True)
        try:
            tail, _ = ctx.decode_blocks(data, results["enc_offset_bits"][1:].copy(), out)
        finally:
            ctx.set_verify_stream_crc(False)
        first_closing = tail[(tail["stream_flags"] & 2) != 0][0]
        assert (first_closing["stream_flags"] & 4) == 0 and first_closing["status"] == 0
    # the host-side fold over results laid end to end (what a multi-GPU caller does with its shares)
    again = quiet.copy()
    assert _capi.fold_stream_crcs(again, verify=True) == len(bad)
    assert again["status"].tolist() == results["status"].tolist()


def test_synthetic_corruptions_match_oracle(ctx):
    """flip bits in the header / tables / payload of a small block: the status must match the oracle's"""
    base = bytearray(case("ref-dolorem-ipsum.txt"))
    rng = np.random.default_rng(7)
    for trial in range(200):
        data = bytearray(base)
        pos = int(rng.integers(4 * 8, len(data) * 8))
        data[pos >> 3] ^= 1 << (7 - (pos & 7))
        want = orc.decode_block(bytes(data), 32, stages=False)
        out = np.zeros(1 << 20, dtype=np.uint8)
        results, _ = ctx.decode_blocks(bytes(data), [32], out)
        assert results["status"][0] == want.info.status, f"bit {pos}: {results['status'][0]} vs {want.info.status}"
        if want.info.status == 0:
            assert results["out_size"][0] == want.info.out_size


def test_truncations_match_oracle(ctx):
    base = case("ref-base64-32KiB")
    for cut in [10, 14, 20, 40, 100, 1000, len(base) - 11, len(base) - 10, len(base) - 5, len(base) - 1]:
        data = base[:cut]
        want = orc.decode_block(data, 32, stages=False)
        out = np.zeros(1 << 20, dtype=np.uint8)
        results, _ = ctx.decode_blocks(data, [32], out)
        assert results["status"][0] == want.info.status, f"cut {cut}"


# ---------------------------------------------------------------------------------------------------------------
# ibz2_decode_buffer with host buffers: one call, find + decode + copies (what bench.py times as e2e), also with the
# input uploaded in two parts (the path large inputs take)
# ---------------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("split", [0, 25, 50, 90])   # per cent of the input in the first part, 0 = one upload
@pytest.mark.parametrize("name", ALL)
def test_decode_buffer_host(ctx, monkeypatch, name, split):
    data = case(name)
    blocks = data_blocks(ctx, data)
    monkeypatch.setenv("IBZ2_SPLIT_INPUT_BYTES", "1" if split else str(1 << 62))
    monkeypatch.setenv("IBZ2_SPLIT_PERCENT", str(split or 50))
    want_results, want_total = ctx.decode_blocks(data, blocks, None)       # sizes and statuses, block by block
    cap = max(int(want_total), 1)
    want_out = np.zeros(cap, dtype=np.uint8)
    want_results, want_total = ctx.decode_blocks(data, blocks, want_out)
    out = np.zeros(cap, dtype=np.uint8)
    results, total = ctx.decode_buffer(np.frombuffer(data, dtype=np.uint8), out)
    assert len(results) == len(blocks) and total == want_total
    for field in ("status", "enc_offset_bits", "enc_size_bits", "expected_crc", "calculated_crc", "out_size"):
        assert results[field].tolist() == want_results[field].tolist(), field
    assert out[:total].tobytes() == want_out[:total].tobytes()
    if name in VALID and (results["status"] == 0).all() and total == GOLDEN[name]["parallel"]["size"] and total > 0:
        assert hashlib.sha256(out[:total].tobytes()).hexdigest() == GOLDEN[name]["parallel"]["sha256"]


# ---------------------------------------------------------------------------------------------------------------
# ibz2_multi_*: one file, several contexts (here: on the one device the test box has), ONE output buffer
# ---------------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("contexts", [2, 3])
def test_multi_device_decode_matches_single(ctx, contexts):
    """The multi-device call cuts the stream into byte ranges, decodes every range through its own context and places
    the bytes of all of them in one caller buffer: bytes, block list and stream CRC verdicts must be those of the
    single-context call (which earlier tests pin to the oracle and the reference)."""
    with _capi.MultiContext([0] * contexts) as multi:
        for name in ["text-450k-l1", "mixed-384k-l2", "multistream-b64-1B-empty-b64", "pbzip2-style-5-streams",
                     "longruns-3M-l1", "ref-1B", "ref-empty", "err-stream-crc", "err-stream-crc-multistream", "err-bitflip"]:
            data = case(name)
            size = GOLDEN[name]["parallel"]["size"] if GOLDEN[name]["parallel"]["rc"] == 0 else 1 << 20
            out_one = np.zeros(size + 64, dtype=np.uint8)
            ctx.set_verify_stream_crc(True)
            try:
                want, total_one = ctx.decode_buffer(data, out_one, max_blocks=4096)
            finally:
                ctx.set_verify_stream_crc(False)
            out = multi.pinned(size + 64)
            out[:] = 0xEE
            got, total = multi.decode_buffer(data, out, max_blocks=4096)
            label = f"{name} x{contexts}"
            assert total == total_one, label
            for field in ("enc_offset_bits", "enc_size_bits", "out_offset", "out_size", "expected_crc", "calculated_crc",
                          "status", "n_symbols", "stream_flags", "stream_crc_expected", "stream_crc_calculated"):
                assert got[field].tolist() == want[field].tolist(), f"{label}: {field}"
            ok = want["status"] == 0
            if ok.all():
                assert np.array_equal(out[:total], out_one[:total]), label
                if GOLDEN[name]["parallel"]["rc"] == 0 and not name.startswith("err-"):
                    assert hashlib.sha256(out[:total].tobytes()).hexdigest() == GOLDEN[name]["parallel"]["sha256"], label
            t = multi.timings()
            assert t["stream_crc_mismatches"] == int((want["status"] == 17).sum()), label
            assert len(t["devices"]) == contexts
            multi.free_pinned()
        # the caller's buffer is too small: the blocks that do not fit say so, the ones in front of them are delivered
        data = case("text-450k-l1")
        small = multi.pinned(200_000)
        got, total = multi.decode_buffer(data, small, max_blocks=64)
        assert total == GOLDEN["text-450k-l1"]["parallel"]["size"]
        fits = got["out_offset"] + got["out_size"] <= 200_000
        assert (got["status"][fits] == 0).all() and (got["status"][~fits] == 15).all() and (~fits).any()
        full = orc.parallel_reader(data).data
        edge = int((got["out_offset"] + got["out_size"])[fits].max())
        assert small[:edge].tobytes() == full[:edge]
