"""Drop-in surface tests: the C++ readers (indexed_bzip2_b200/host), the Python module and the ibzip2 CLI against the
golden outputs of the compiled, unmodified reference (tests/golden/golden.json: decoded bytes, blockOffsets() of
ParallelBZ2Reader and BZ2Reader, `ibzip2 -l`, CLI exit status / stderr) and against the CPU oracle.

The GPU tests read like the reference's own testParallelBZ2Reader.cpp / testBZ2Reader.cpp / testPythonWrappers.py:
decode, compare bytes; seek around, compare slices; export the index, import it into a fresh reader, read again.
"""
import ctypes as C
import hashlib
import io
import json
import os
import subprocess

import numpy as np
import pytest

import _oracle as orc
import indexed_bzip2_b200 as ibz2

ROOT = orc.ROOT
GOLDEN = json.loads((ROOT / "tests" / "golden" / "golden.json").read_text())
ALL = sorted(GOLDEN)
VALID = sorted(k for k, v in GOLDEN.items() if v["parallel"]["rc"] == 0)
CLI = ROOT / "indexed_bzip2_b200" / "bin" / "ibzip2"


def path(name):
    return str(ROOT / "tests" / "golden" / "cases" / f"{name}.bz2")


def sha(data):
    return hashlib.sha256(data).hexdigest()


# ---------------------------------------------------------------------------------------------------------------
# host logic without a GPU
# ---------------------------------------------------------------------------------------------------------------

def simulate_index(triples, finish=True, probe=0):
    lib = ibz2._load()
    lib.ibz2r_index_simulate.restype = C.c_longlong
    lib.ibz2r_index_simulate.argtypes = [C.POINTER(C.c_uint64), C.c_size_t, C.c_int, C.POINTER(C.c_uint64), C.c_size_t,
                                         C.c_uint64, C.POINTER(C.c_uint64)]
    flat = (C.c_uint64 * max(1, 3 * len(triples)))(*[x for t in triples for x in t])
    pairs = (C.c_uint64 * 4096)()
    located = (C.c_uint64 * 5)()
    count = lib.ibz2r_index_simulate(flat, len(triples), 1 if finish else 0, pairs, 2048, probe, located)
    assert count >= 0
    return [[pairs[2 * i], pairs[2 * i + 1]] for i in range(count)], list(located)


@pytest.mark.parametrize("name", VALID)
def test_block_index_reproduces_reference_maps(name):
    """Feeding the blocks of the reference's own ParallelBZ2Reader map back through BlockIndex (append per block and
    EOS, then finish) must give that map again, sentinel included (core/BlockMap.hpp:69-93,174-198)."""
    want = GOLDEN[name]["parallel"]["offsets"]
    entries = want[:-1] if len(want) > 1 else []
    triples = [(bits, want[i + 1][0] - bits, want[i + 1][1] - nbytes) for i, (bits, nbytes) in enumerate(entries)]
    got, _ = simulate_index(triples)
    assert got == want
    unfinished, located = simulate_index(triples, finish=False)
    assert unfinished == entries
    assert located[4] == sum(1 for t in triples if t[2] > 0)   # dataBlockCount() while the map is still growing


def test_block_index_locate_prefers_the_block_after_an_eos():
    # base64-32KiB | 1B | empty | base64-32KiB (Appendix B of SURVEY.md)
    triples = [(32, 199677, 32768), (199709, 115, 0), (199824, 179, 1), (200003, 117, 0), (200232, 199677, 32768),
               (399909, 83, 0)]
    offsets, located = simulate_index(triples, probe=32768)
    assert offsets == [[32, 0], [199709, 32768], [199824, 32768], [200003, 32769], [200232, 32769], [399909, 65537],
                       [399992, 65537]]
    assert located[:4] == [1, 199824, 32768, 1]   # the 1-byte block, not the EOS entry that starts at the same byte
    _, located = simulate_index(triples, probe=65537)
    assert located[3] == 0                        # nothing holds the byte one past the end
    assert simulate_index([], finish=True)[0] == [[0, 0]]   # empty file (BlockMap.hpp:184-187)


def test_host_library_exports_the_reader_entry_points():
    lib = ibz2._load()
    for name in ("ibz2r_open_path", "ibz2r_open_fd", "ibz2r_open_memory", "ibz2r_destroy", "ibz2r_read", "ibz2r_read_to_fd",
                 "ibz2r_seek", "ibz2r_tell", "ibz2r_size", "ibz2r_tell_compressed", "ibz2r_eof", "ibz2r_closed",
                 "ibz2r_close", "ibz2r_block_offsets_complete", "ibz2r_block_offsets", "ibz2r_set_block_offsets",
                 "ibz2r_join_threads", "ibz2r_cli", "ibz2r_last_error", "ibz2r_last_error_kind"):
        assert hasattr(lib, name), name
    assert ibz2.__version__ == "1.7.0"
    with pytest.raises(TypeError):
        ibz2.open(path("ref-1B"), parallelization="4")


def test_reference_module_name_is_importable():
    """python/indexed_bzip2/pyproject.toml:14,51-57: the module is `indexed_bzip2`, the console script `ibzip2` -> cli"""
    import indexed_bzip2

    assert indexed_bzip2.open is ibz2.open and indexed_bzip2.IndexedBzip2File is ibz2.IndexedBzip2File
    assert indexed_bzip2.cli is ibz2.cli and indexed_bzip2.__version__ == "1.7.0"
    text = (ROOT / "pyproject.toml").read_text()
    assert 'ibzip2 = "indexed_bzip2_b200:cli"' in text


def test_cli_version_and_help_need_no_gpu():
    out = subprocess.run([str(CLI), "--version"], capture_output=True, text=True)
    assert out.returncode == 0
    assert out.stdout == "ibzip2, CLI to the indexed and seekable bzip2 decoding library indexed-bzip2 version 1.7.0.\n"
    assert subprocess.run([str(CLI), "-h"], capture_output=True).returncode == 0
    out = subprocess.run([str(CLI)], capture_output=True, text=True, stdin=subprocess.DEVNULL)
    assert out.returncode == 1


# ---------------------------------------------------------------------------------------------------------------
# on the B200
# ---------------------------------------------------------------------------------------------------------------

STREAM_CRC_CASES = [n for n in ALL if n.startswith("err-stream-crc")]


@pytest.mark.gpu
@pytest.mark.parametrize("name", ALL)
def test_parallel_reader_matches_reference(name):
    want = GOLDEN[name]["parallel"]
    # the reference's ParallelBZ2Reader never looks at the stream CRC (only BZ2Reader does, BZ2Reader.hpp:406-416);
    # ours does by default -- see test_parallel_reader_checks_the_stream_crc -- and is compared here with the check off
    f = ibz2.open(path(name), parallelization=4, verify_stream_crc=name not in STREAM_CRC_CASES)
    if want["rc"] != 0:
        with pytest.raises(Exception) as info:
            f.read()
        if want["error"] != "std::exception":
            assert want["error"] in str(info.value)
        else:
            assert "end of file" in str(info.value).lower()
        return
    data = f.read()
    assert len(data) == want["size"] and sha(data) == want["sha256"]
    assert f.block_offsets_complete()
    assert [list(kv) for kv in sorted(f.block_offsets().items())] == want["offsets"]
    assert f.size() == want["size"] and f.tell() == want["size"]
    oracle = orc.parallel_reader(open(path(name), "rb").read())
    assert data == oracle.data


@pytest.mark.gpu
@pytest.mark.parametrize("name", ALL)
def test_serial_reader_matches_reference(name):
    want = GOLDEN[name]["serial"]
    f = ibz2.open(path(name), parallelization=1)
    if want["rc"] != 0:
        with pytest.raises(Exception) as info:
            f.read()
        if want["error"] != "std::exception":
            assert want["error"] in str(info.value)
        return
    data = f.read()
    assert len(data) == want["size"] and sha(data) == want["sha256"]
    assert [list(kv) for kv in sorted(f.block_offsets().items())] == want["offsets"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", STREAM_CRC_CASES)
def test_parallel_reader_checks_the_stream_crc(name):
    """Deliberate deviation: the parallel reader folds the block CRCs per stream while it builds the
    block map and throws what the reference's SERIAL reader throws (BZ2Reader.hpp:406-416), after it has handed out
    every byte in front of the end-of-stream block -- the same bytes the serial reference writes before it throws."""
    want = GOLDEN[name]["serial"]
    assert want["rc"] != 0 and "Stream CRC" in want["error"]
    everything = ibz2.open(path(name), parallelization=4, verify_stream_crc=False).read()
    assert sha(everything) == GOLDEN[name]["parallel"]["sha256"]
    delivered = GOLDEN[name]["cli_serial"]["stdout_size"]   # what the serial reference has written when it throws
    # unbuffered, in small pieces: a read that crosses the end of the damaged stream throws (like the reference's serial
    # reader, the exception takes that call's buffer with it), every read before it has delivered its bytes
    f = ibz2.IndexedBzip2FileRaw(path(name), 4)
    got = bytearray()
    piece = bytearray(4096)
    with pytest.raises(RuntimeError) as info:
        while True:
            n = f.readinto(piece)
            if n == 0:
                break
            got += piece[:n]
    assert str(info.value) == want["error"]
    assert delivered - len(piece) <= len(got) <= delivered and bytes(got) == everything[:len(got)]
    # the check happens once, while the block map is built (BZ2Reader.hpp:409): afterwards the file reads like any other
    rest = bytearray()
    while True:
        n = f.readinto(piece)
        if n == 0:
            break
        rest += piece[:n]
    assert bytes(rest) == everything[f.tell() - len(rest):]
    assert f.tell() == len(everything)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["text-450k-l1", "multistream-b64-1B-empty-b64", "pbzip2-style-5-streams", "longruns-3M-l1"])
@pytest.mark.parametrize("parallelization", [1, 4])
def test_seek_and_partial_reads(name, parallelization):
    raw = orc.parallel_reader(open(path(name), "rb").read()).data
    rng = np.random.default_rng(5)
    f = ibz2.open(path(name), parallelization=parallelization)
    assert f.seekable() and f.readable() and f.tell() == 0
    assert f.read(10) == raw[:10] and f.tell() == 10
    for _ in range(25):
        offset = int(rng.integers(0, len(raw)))
        size = int(rng.integers(1, 70000))
        assert f.seek(offset) == offset
        assert f.read(size) == raw[offset:offset + size]
    assert f.seek(0, io.SEEK_END) == len(raw) and f.read(1) == b""
    assert f.seek(-5, io.SEEK_END) == len(raw) - 5 and f.read() == raw[-5:]
    assert f.seek(3) == 3 and f.seek(4, io.SEEK_CUR) == 7 and f.read(3) == raw[7:10]
    assert f.seek(len(raw) + 1000) == len(raw)
    f.close()
    assert f.closed


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["text-450k-l1", "pbzip2-style-5-streams", "multistream-b64-1B-empty-b64", "longruns-3M-l1"])
def test_parallel_reader_over_several_devices(name, monkeypatch):
    """ParallelBZ2Reader( ..., devices ) and `ibzip2 --gpus N`: every batch is cut into one part per device (here: two and
    three contexts on the one GPU of the box, and parts of a single block), bytes and block map stay what they are."""
    monkeypatch.setenv("IBZ2_MIN_BLOCKS_PER_DEVICE", "1")
    want = GOLDEN[name]["parallel"]
    for devices in ([0, 0], [0, 0, 0]):
        f = ibz2.open(path(name), parallelization=4, gpus=devices)
        data = f.read()
        assert sha(data) == want["sha256"] and [list(kv) for kv in sorted(f.block_offsets().items())] == want["offsets"]
        rng = np.random.default_rng(11)
        for _ in range(10):
            offset = int(rng.integers(0, max(1, len(data))))
            f.seek(offset)
            assert f.read(50_000) == data[offset:offset + 50_000]
        f.close()
    env = dict(os.environ, IBZ2_MIN_BLOCKS_PER_DEVICE="1")
    out = subprocess.run([str(CLI), "-d", "-c", "-P", "4", "--gpus", "1", path(name)], capture_output=True, env=env)
    assert out.returncode == 0 and sha(out.stdout) == want["sha256"]


@pytest.mark.gpu
def test_index_export_import_roundtrip(tmp_path):
    name = "pbzip2-style-5-streams"
    raw = orc.parallel_reader(open(path(name), "rb").read()).data
    first = ibz2.open(path(name), parallelization=4)
    offsets = first.block_offsets()
    index_file = tmp_path / "index.csv"
    first.export_index(index_file)
    assert index_file.read_text() == "".join(f"{b},{n}\n" for b, n in GOLDEN[name]["parallel"]["offsets"])

    for flavour in (GOLDEN[name]["parallel"]["offsets"], GOLDEN[name]["serial"]["offsets"]):
        second = ibz2.open(path(name), parallelization=4)
        second.set_block_offsets({b: n for b, n in flavour})   # either map flavour is accepted (ParallelBZ2Reader.hpp:463-472)
        assert second.block_offsets_complete()
        assert second.size() == len(raw)
        rng = np.random.default_rng(7)
        for _ in range(20):
            offset = int(rng.integers(0, len(raw)))
            second.seek(offset)
            assert second.read(4096) == raw[offset:offset + 4096]
    third = ibz2.open(path(name), parallelization=4)
    third.import_index(index_file)
    assert third.block_offsets() == offsets
    third.seek(len(raw) // 2)
    assert third.read() == raw[len(raw) // 2:]
    with pytest.raises(ValueError):
        ibz2.open(path(name), parallelization=4).set_block_offsets({})
    with pytest.raises(ValueError):
        ibz2.open(path(name), parallelization=4).set_block_offsets({32: 0})


@pytest.mark.gpu
def test_python_sources_fd_fileobject_bytesio():
    name = "text-450k-l1"
    raw = orc.parallel_reader(open(path(name), "rb").read()).data
    with open(path(name), "rb") as fobj:
        assert ibz2.open(fobj, parallelization=2).read() == raw                 # object with fileno()
        assert ibz2.open(fobj.fileno(), parallelization=2).read() == raw        # plain descriptor
    memory = io.BytesIO(open(path(name), "rb").read())
    memory.seek(17)
    f = ibz2.open(memory, parallelization=3)                                    # read/seek/tell/seekable object
    assert memory.tell() == 17
    assert f.read(100) == raw[:100] and f.readline() == raw[100:raw.index(b"\n", 100) + 1]
    raw_file = ibz2.IndexedBzip2FileRaw(path(name), 4)
    buffer = bytearray(1000)
    assert raw_file.readinto(buffer) == 1000 and bytes(buffer) == raw[:1000]
    assert raw_file.bz2reader.tell_compressed() == 32
    raw_file.join_threads()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ALL)
@pytest.mark.parametrize("flavour,parallelism", [("cli_parallel", "4"), ("cli_serial", "1")])
def test_cli_decompress_matches_reference(name, flavour, parallelism):
    want = GOLDEN[name][flavour]
    env = dict(os.environ)
    if name in STREAM_CRC_CASES and flavour == "cli_parallel":
        # with its stream CRC check on (the default) the parallel reader behaves like the reference's serial one
        out = subprocess.run([str(CLI), "-d", "-c", "-P", parallelism, path(name)], capture_output=True)
        serial = GOLDEN[name]["cli_serial"]
        assert out.returncode == serial["rc"] and out.stderr.decode().strip() == serial["stderr"]
        assert len(out.stdout) == serial["stdout_size"] and sha(out.stdout) == serial["stdout_sha256"]
        env["IBZ2_VERIFY_STREAM_CRC"] = "0"   # ... and like the reference's parallel one with the check off
    out = subprocess.run([str(CLI), "-d", "-c", "-P", parallelism, path(name)], capture_output=True, env=env)
    assert out.returncode == want["rc"]
    assert out.stderr.decode().strip() == want["stderr"]
    if want["rc"] == 0 or flavour == "cli_parallel":
        assert len(out.stdout) == want["stdout_size"] and sha(out.stdout) == want["stdout_sha256"]
    else:
        # the serial reference writes the bytes of a block before it finds the CRC wrong; this reader does not
        assert len(out.stdout) <= want["stdout_size"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["multistream-b64-1B-empty-b64", "err-gap-then-valid", "text-450k-l1", "ref-empty"])
def test_cli_offset_lists(name, tmp_path):
    out = subprocess.run([str(CLI), "-l", "-", path(name)], capture_output=True, text=True)
    assert out.returncode == 0
    assert [int(x) for x in out.stdout.split()] == GOLDEN[name]["finder"]
    listing = tmp_path / "offsets.txt"
    out = subprocess.run([str(CLI), "-L", str(listing), "-P", "4", path(name)], capture_output=True, text=True)
    assert out.returncode == 0
    assert listing.read_text() == "".join(f"{b},{n}\n" for b, n in GOLDEN[name]["parallel"]["offsets"])
    out = subprocess.run([str(CLI), "-L", str(listing), path(name)], capture_output=True, text=True)
    assert out.returncode == 1 and "already exists" in out.stderr


@pytest.mark.gpu
def test_cli_output_file_rules(tmp_path):
    source = tmp_path / "sample.BZ2"
    source.write_bytes(open(path("text-450k-l1"), "rb").read())
    raw = orc.parallel_reader(source.read_bytes()).data
    out = subprocess.run([str(CLI), "-d", "-t", "-v", str(source)], capture_output=True, text=True)
    entries = len(GOLDEN["text-450k-l1"]["parallel"]["offsets"])   # data blocks + EOS + end sentinel
    assert out.returncode == 0 and f"Found {entries} blocks" in out.stdout
    assert (tmp_path / "sample").read_bytes() == raw                   # '.bz2' stripped, case-insensitively
    out = subprocess.run([str(CLI), "-d", str(source)], capture_output=True, text=True)
    assert out.returncode == 1 and "already exists" in out.stderr
    assert subprocess.run([str(CLI), "-d", "-f", str(source)], capture_output=True).returncode == 0
    piped = subprocess.run([str(CLI), "-d", "-c"], input=source.read_bytes(), capture_output=True)
    assert piped.returncode == 0 and piped.stdout == raw               # standard input -> standard output


# ---------------------------------------------------------------------------------------------------------------
# GZIDX (rapidgzip/IndexFileFormat.hpp:37-64): written here, read by the compiled reference, and back
# ---------------------------------------------------------------------------------------------------------------

REF_GZIDX = ROOT / "oracle" / "_ref" / "libref_gzidx.so"


def ref_gzidx():
    lib = C.CDLL(str(REF_GZIDX))
    lib.refgz_last_error.restype = C.c_char_p
    lib.refgz_reads_with_index.restype = C.c_int
    lib.refgz_reads_with_index.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p,
                                           C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.refgz_export_index.restype = C.c_int
    lib.refgz_export_index.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
    return lib


def gzidx_write(offsets, compressed_size, target):
    lib = ibz2._load()
    lib.ibz2r_gzidx_write.restype = C.c_int
    lib.ibz2r_gzidx_write.argtypes = [C.POINTER(C.c_uint64), C.c_size_t, C.c_uint64, C.c_char_p]
    flat = (C.c_uint64 * (2 * len(offsets)))(*[x for pair in offsets for x in pair])
    assert lib.ibz2r_gzidx_write(flat, len(offsets), compressed_size, os.fsencode(str(target))) == 0


def gzidx_read(source):
    lib = ibz2._load()
    lib.ibz2r_gzidx_read.restype = C.c_longlong
    lib.ibz2r_gzidx_read.argtypes = [C.c_char_p, C.POINTER(C.c_uint64), C.c_size_t]
    pairs = (C.c_uint64 * 8192)()
    count = lib.ibz2r_gzidx_read(os.fsencode(str(source)), pairs, 4096)
    assert count >= 0
    return [[pairs[2 * i], pairs[2 * i + 1]] for i in range(count)]


@pytest.mark.parametrize("name", [n for n in VALID if GOLDEN[n]["parallel"]["size"] > 0 and not n.startswith("err-")])
def test_gzidx_written_here_is_read_by_the_reference(name, tmp_path):
    """The block map of the reference's ParallelBZ2Reader (golden), written as GZIDX by our writer: one checkpoint per
    data block and the end of the file, layout byte for byte as IndexFileFormat.hpp:37-64 describes it; the reference's
    gzip-architecture reader imports it and returns the right bytes at random offsets (needs oracle/_ref/libref_gzidx.so)."""
    want = GOLDEN[name]["parallel"]
    offsets = want["offsets"]
    compressed_size = GOLDEN[name]["compressed_size"]
    index = tmp_path / "index.gzidx"
    gzidx_write(offsets, compressed_size, index)
    blob = index.read_bytes()
    data_blocks = [o for i, o in enumerate(offsets) if i + 1 == len(offsets) or offsets[i + 1][1] != o[1]]
    assert blob[:7] == b"GZIDX\x01\x00" and len(blob) == 35 + 18 * len(data_blocks)
    assert int.from_bytes(blob[7:15], "little") == compressed_size
    assert int.from_bytes(blob[15:23], "little") == want["size"]
    assert int.from_bytes(blob[27:31], "little") == 32768 and int.from_bytes(blob[31:35], "little") == len(data_blocks)
    for i, (bits, nbytes) in enumerate(data_blocks):
        entry = blob[35 + 18 * i:35 + 18 * (i + 1)]
        assert int.from_bytes(entry[:8], "little") * 8 - entry[16] == bits and entry[16] < 8 and entry[17] == 0
        assert int.from_bytes(entry[8:16], "little") == nbytes
    assert gzidx_read(index) == data_blocks
    if not REF_GZIDX.exists():
        pytest.skip("oracle/_ref/libref_gzidx.so not built (needs zlib)")
    raw = orc.parallel_reader(open(path(name), "rb").read()).data
    rng = np.random.default_rng(3)
    size = min(4096, len(raw))
    positions = np.sort(rng.integers(0, len(raw) - size + 1, 6)).astype(np.uint64)
    out = np.zeros(size, dtype=np.uint8)
    checksum, decoded = C.c_uint64(0), C.c_uint64(0)
    lib = ref_gzidx()
    rc = lib.refgz_reads_with_index(os.fsencode(path(name)), os.fsencode(str(index)), 2, positions.ctypes.data, len(positions),
                                    size, out.ctypes.data, C.byref(checksum), C.byref(decoded))
    assert rc == 0, lib.refgz_last_error()
    assert decoded.value == len(raw)
    assert out.tobytes() == raw[int(positions[0]):int(positions[0]) + size]
    assert checksum.value == sum(sum(raw[int(p):int(p) + size]) for p in positions)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["text-450k-l1", "pbzip2-style-5-streams", "multistream-b64-1B-empty-b64", "ref-1B"])
def test_gzidx_export_import_through_the_readers(name, tmp_path):
    raw = orc.parallel_reader(open(path(name), "rb").read()).data
    first = ibz2.open(path(name), parallelization=4)
    index = tmp_path / "index.gzidx"
    first.export_index(index, format="gzidx")
    want = GOLDEN[name]["parallel"]["offsets"]
    assert gzidx_read(index) == [o for i, o in enumerate(want) if i + 1 == len(want) or want[i + 1][1] != o[1]]
    for parallelization in (4, 1):
        second = ibz2.open(path(name), parallelization=parallelization)
        second.import_index(index)
        assert second.block_offsets_complete() and second.size() == len(raw)
        rng = np.random.default_rng(9)
        for _ in range(8):
            offset = int(rng.integers(0, len(raw)))
            second.seek(offset)
            assert second.read(3000) == raw[offset:offset + 3000]
    buffer = io.BytesIO()
    first.export_index(buffer, format="gzidx")
    assert buffer.getvalue() == index.read_bytes()
    if REF_GZIDX.exists():
        # the reference's own index of the same file: chunk starts only.  One block per chunk: importable; else refused.
        theirs = tmp_path / "theirs.gzidx"
        assert ref_gzidx().refgz_export_index(os.fsencode(path(name)), os.fsencode(str(theirs)), 2) == 0
        third = ibz2.open(path(name), parallelization=4)
        blocks = sum(1 for i, o in enumerate(want[:-1]) if want[i + 1][1] != o[1])
        if len(gzidx_read(theirs)) == blocks + 1:
            third.import_index(theirs)
            third.seek(len(raw) // 3)
            assert third.read(1000) == raw[len(raw) // 3:len(raw) // 3 + 1000]
        else:
            with pytest.raises(ValueError, match="does not list every bzip2 block"):
                third.import_index(theirs)


# ---------------------------------------------------------------------------------------------------------------
# Bzip2Chunk::decodeChunk (rapidgzip/chunkdecoding/Bzip2Chunk.hpp:200-268) on the GPU, against the chunk map of the
# compiled reference's gzip-architecture reader
# ---------------------------------------------------------------------------------------------------------------

def decode_chunk(name, chunk_offset, until_offset, max_size=(1 << 64) - 1, out_capacity=1 << 22):
    lib = ibz2._load()
    lib.ibz2r_decode_chunk.restype = C.c_int
    lib.ibz2r_decode_chunk.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64),
                                       C.POINTER(C.c_uint64), C.c_size_t, C.POINTER(C.c_uint64), C.c_size_t,
                                       C.POINTER(C.c_uint64), C.c_void_p, C.c_uint64]
    info, counts = (C.c_uint64 * 4)(), (C.c_uint64 * 2)()
    boundaries, footers = (C.c_uint64 * 2048)(), (C.c_uint64 * 2048)()
    out = np.zeros(out_capacity, dtype=np.uint8)
    rc = lib.ibz2r_decode_chunk(os.fsencode(path(name)), chunk_offset, until_offset, max_size, info, boundaries, 1024, footers,
                                1024, counts, out.ctypes.data, out_capacity)
    if rc != 0:
        raise RuntimeError((lib.ibz2r_last_error(None) or b"").decode())
    return {"encoded_offset": info[0], "encoded_size": info[1], "decoded_size": info[2], "stopped": bool(info[3]),
            "boundaries": [(boundaries[2 * i], boundaries[2 * i + 1]) for i in range(counts[0])],
            "footers": [(footers[2 * i], footers[2 * i + 1]) for i in range(counts[1])],
            "data": out[:min(info[2], out_capacity)].tobytes()}


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["text-450k-l1", "pbzip2-style-5-streams", "multistream-b64-1B-empty-b64", "mixed-384k-l2"])
def test_bzip2_chunk_adapter_matches_the_reference_chunk_map(name):
    if not REF_GZIDX.exists():
        pytest.skip("oracle/_ref/libref_gzidx.so not built (needs zlib)")
    lib = ref_gzidx()
    lib.refgz_chunk_offsets.restype = C.c_int
    lib.refgz_chunk_offsets.argtypes = [C.c_char_p, C.c_int, C.c_uint64, C.POINTER(C.c_uint64), C.c_uint64, C.POINTER(C.c_uint64)]
    pairs, count = (C.c_uint64 * 4096)(), C.c_uint64(0)
    # 8 KiB chunks: the reference cuts these small files into several chunks (a chunk holds at least one block)
    assert lib.refgz_chunk_offsets(os.fsencode(path(name)), 2, 8192, pairs, 2048, C.byref(count)) == 0, lib.refgz_last_error()
    chunks = [(pairs[2 * i], pairs[2 * i + 1]) for i in range(count.value)]
    raw = orc.parallel_reader(open(path(name), "rb").read()).data
    assert chunks[-1][1] == len(raw) and len(chunks) >= 2
    want_map = GOLDEN[name]["parallel"]["offsets"]
    magics = orc.find_magics(open(path(name), "rb").read())
    for (start, decoded_start), (end, decoded_end) in zip(chunks, chunks[1:]):
        got = decode_chunk(name, start, end)
        assert got["encoded_offset"] == start and got["encoded_offset"] + got["encoded_size"] == end
        if decoded_start == decoded_end:
            # a chunk that starts at an end-of-stream block and holds nothing else (the reference's map has such entries)
            assert got["decoded_size"] == 0 and got["boundaries"] == [] and len(got["footers"]) >= 1
            continue
        assert got["decoded_size"] == decoded_end - decoded_start and got["data"] == raw[decoded_start:decoded_end]
        assert not got["stopped"]
        # every block boundary inside the chunk is one of the reference's ParallelBZ2Reader block-map entries
        # every block inside the chunk is a boundary, the end-of-stream blocks too -- also the one of an empty stream, which
        # the block map of ParallelBZ2Reader leaves out -- as far as the chunk holds data in front of it
        # (Bzip2Chunk.hpp:119-121); the decoded offset of a magic is that of the next block-map entry
        inside = []
        for magic in (int(m) for m in magics if start < m < end):
            decoded_at = next(d for b, d in want_map if b >= magic)
            if decoded_at > decoded_start:
                inside.append((magic, decoded_at - decoded_start))
        assert got["boundaries"] == inside
        assert set(got["boundaries"]) >= {(b, d - decoded_start) for b, d in want_map[:-1] if start < b < end and d > decoded_start}
        assert all(f[1] <= got["decoded_size"] for f in got["footers"])
    # the whole file as one chunk, cut short by the size limit; and a range without any block
    whole = decode_chunk(name, chunks[0][0], chunks[-1][0])
    assert whole["decoded_size"] == len(raw) and whole["data"] == raw[:1 << 22]
    limited = decode_chunk(name, chunks[0][0], chunks[-1][0], max_size=1)
    assert limited["stopped"] and 0 < limited["decoded_size"] < len(raw) and limited["data"] == raw[:limited["decoded_size"]]
    with pytest.raises(RuntimeError, match="Failed to find any valid bzip2 block"):
        decode_chunk(name, chunks[-1][0], chunks[-1][0] + 1000)
