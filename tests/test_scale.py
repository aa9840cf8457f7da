"""GPU tests of the BASELINE.json configurations 2, 4 and 5 at sizes between the golden files and the 8 GB bench runs,
through properties that do not need the oracle to decode the whole input: the finder against positions planted by the
test itself, decode -> bytes equal to what the generator compressed (every block CRC included), random access against
slices of the raw data.  Config 3 at full size is verified inside bench.py (`"verified"` in its JSON line).

Run on the B200 box: python -m pytest tests -m gpu
"""
import sys

import numpy as np
import pytest

import _oracle as orc

sys.path.insert(0, str(orc.ROOT))
import benchdata  # noqa: E402
import indexed_bzip2_b200 as ibz2  # noqa: E402
from indexed_bzip2_b200 import _capi  # noqa: E402

pytestmark = pytest.mark.gpu

EOS_FLAG = np.uint64(1 << 63)


@pytest.fixture(scope="module")
def ctx():
    c = _capi.Context(0)
    yield c
    c.close()


def plant(buf: np.ndarray, bit: int, magic: int):
    for k in range(48):
        b = (magic >> (47 - k)) & 1
        byte, sh = (bit + k) >> 3, 7 - ((bit + k) & 7)
        buf[byte] = (int(buf[byte]) & ~(1 << sh)) | (b << sh)


def test_config2_finder_scan_512mb(ctx):
    """config 2 (BlockFinder-only scan), 512 MiB of noise with both magics planted at every bit alignment, next to
    16 B / 512 B / 1 MiB / 64 MiB seams and flush with both ends; expected = what was planted (the chance of a 48-bit
    magic in 2^32 random bit positions is 2^-16 per magic, and the seed is fixed)"""
    rng = np.random.default_rng(2)
    n = 512 << 20
    buf = rng.integers(0, 256, n, dtype=np.uint8)
    positions = {0, (n - 6) * 8}
    for seam in (16, 512, 1 << 20, 64 << 20):
        for j, back in enumerate(range(0, 56)):
            positions.add((seam * (3 + 2 * j)) * 8 - back)
    for s in range(64):
        positions.add((100_000_007 + 97 * s) * 8 + s)
    kept, last_end = [], -1
    for i, p in enumerate(sorted(positions)):
        if p <= last_end or p + 48 > n * 8:
            continue
        plant(buf, p, _capi.MAGIC_BLOCK if i % 2 == 0 else _capi.MAGIC_EOS)
        kept.append((p, i % 2 == 1))
        last_end = p + 48
    found = ctx.find_magics(buf)
    offsets = (found & ~EOS_FLAG).tolist()
    is_eos = ((found & EOS_FLAG) != 0).tolist()
    assert list(zip(offsets, is_eos)) == kept


def test_config4_long_runs_level1(ctx):
    """config 4: long runs (RLE1 expands every block ~50x, RLE2 carries most of the symbols), bzip2 -1, 256 MB raw"""
    w = benchdata.build_single_stream("longruns", 4, 64 << 20, 4, 1)
    out = np.zeros(w.raw_size + 16, dtype=np.uint8)
    results, total = ctx.decode_buffer(w.compressed, out)
    assert total == w.raw_size and len(results) == w.blocks
    assert (results["status"] == 0).all()
    assert (results["calculated_crc"] == results["expected_crc"]).all()
    assert w.verify(out[:total])
    # blocks of 100 kB symbols that decode to several MB each: the point of the configuration
    assert int(results["out_size"].max()) > 8 * int(results["n_symbols"].max())


def test_config5_multistream_random_access(tmp_path):
    """config 5: pbzip2-style file (one stream per 900 kB), index from a first pass imported into a second reader, then
    seeks + 1 MiB reads at random offsets against the raw data"""
    w = benchdata.build_multistream("mixed", 3, 48 << 20, 900_000, 9)
    raw = w.raw_base
    path = tmp_path / "multistream.bz2"
    w.compressed.tofile(path)
    first = ibz2.open(str(path), parallelization=4)
    assert first.read(1 << 20) == raw[:1 << 20].tobytes()
    first.seek(0, 2)
    index = first.block_offsets()
    assert first.block_offsets_complete() and first.size() == raw.size
    first.close()
    # per stream at least one data block and one end-of-stream entry (900 000 raw bytes do not fit one 899 981-byte
    # block, so most streams have two data blocks), plus the end sentinel
    assert len(index) >= 2 * w.blocks + 1
    assert max(index.values()) == raw.size
    if orc.ref_available():
        ref = orc.ref_decode_file(str(path), parallelization=4, want_data=False)
        assert sorted(index.items()) == sorted((int(a), int(b)) for a, b in ref.offsets)

    second = ibz2.open(str(path), parallelization=4)
    second.set_block_offsets(index)
    assert second.block_offsets_complete()
    rng = np.random.default_rng(5)
    for _ in range(40):
        offset = int(rng.integers(0, raw.size - (1 << 20)))
        assert second.seek(offset) == offset
        assert second.read(1 << 20) == raw[offset:offset + (1 << 20)].tobytes()
    assert second.seek(raw.size - 10) == raw.size - 10 and second.read() == raw[-10:].tobytes()
    second.close()


def test_sequential_read_with_prefetch_thread(tmp_path, monkeypatch):
    """Consecutive reads: once a batch has been decoded on demand, the next one is decoded by a second host thread while the
    caller consumes the current one (host/BlockSource.hpp startPrefetch; the reference's FetchNextAdaptive,
    core/Prefetcher.hpp:159-176).  Here the threshold is lowered so that a 64 MB file already gets there; bytes, block map
    and a seek back into an evicted region must be what they are without it."""
    w = benchdata.build_single_stream("mixed", 7, 64 << 20, 1, 9)
    path = tmp_path / "stream.bz2"
    w.compressed.tofile(path)
    raw = w.raw_base.tobytes()
    maps = []
    for min_blocks in ("1", "0"):
        monkeypatch.setenv("IBZ2_PREFETCH_MIN_BLOCKS", min_blocks)
        f = ibz2.open(str(path), parallelization=4)
        got = bytearray()
        while True:
            piece = f.read(3_000_000)
            if not piece:
                break
            got += piece
        assert bytes(got) == raw
        maps.append(f.block_offsets())
        f.seek(1_000_000)
        assert f.read(100) == raw[1_000_000:1_000_100]
        f.close()
    assert maps[0] == maps[1] and len(maps[0]) == w.blocks + 2
