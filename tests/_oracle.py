"""ctypes bindings for the TEST-ONLY checkers: oracle/libbz2oracle.so (our C restatement) and, when it was
built (oracle/_ref/, needs /root/reference at build time), the compiled unmodified reference behind ref_shim.cpp.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE_DIR = ROOT / "oracle"
ORACLE_SO = ORACLE_DIR / "libbz2oracle.so"
REF_SO = ORACLE_DIR / "_ref" / "libref_bz2.so"
REF_CLI = ORACLE_DIR / "_ref" / "ibzip2"

MAGIC_BLOCK = 0x314159265359
MAGIC_EOS = 0x177245385090
DBUF_SIZE = 900000
SYMBOL_CAPACITY = 32767 * 50 + 1

STATUS_NAMES = {
    0: "OK", 1: "BAD_MAGIC", 2: "RANDOMISED", 3: "ORIGPTR_RANGE", 4: "GROUP_COUNT", 5: "SELECTOR_COUNT",
    6: "SELECTOR_UNARY", 7: "CODE_LENGTH", 8: "CODE_LENGTHS", 9: "SELECTOR_OVERRUN", 10: "DBUF_OVERFLOW",
    11: "INVALID_CODE", 12: "ORIGPTR", 13: "CRC", 14: "EOF", 15: "OUTPUT_CAPACITY", 16: "STREAM_HEADER",
    17: "STREAM_CRC", 18: "BLOCK_ORDER",
}


def build_oracle(force: bool = False) -> Path:
    src = ORACLE_DIR / "bz2_oracle.c"
    if force or not ORACLE_SO.exists() or ORACLE_SO.stat().st_mtime < src.stat().st_mtime:
        subprocess.check_call(["make", "-s", "-C", str(ORACLE_DIR), "oracle"])
    return ORACLE_SO


class BlockInfo(C.Structure):
    _fields_ = [
        ("status", C.c_int32), ("is_eos", C.c_int32), ("at_eof", C.c_int32),
        ("enc_offset_bits", C.c_uint64), ("enc_size_bits", C.c_uint64),
        ("header_crc", C.c_uint32), ("calc_crc", C.c_uint32), ("orig_ptr", C.c_uint32),
        ("symbol_count", C.c_uint32), ("symbol_to_byte", C.c_uint8 * 256),
        ("group_count", C.c_uint32), ("selectors_count", C.c_uint32),
        ("code_lengths", (C.c_uint8 * 258) * 6),
        ("nsymbols", C.c_uint32), ("n", C.c_uint32), ("out_size", C.c_uint64),
    ]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build_oracle()
        L = C.CDLL(str(ORACLE_SO))
        L.orc_crc32.restype = C.c_uint32
        L.orc_crc32.argtypes = [C.c_void_p, C.c_size_t]
        L.orc_crc32_update.restype = C.c_uint32
        L.orc_crc32_update.argtypes = [C.c_uint32, C.c_void_p, C.c_size_t]
        L.orc_stream_crc_fold.restype = C.c_uint32
        L.orc_stream_crc_fold.argtypes = [C.c_uint32, C.c_uint32]
        L.orc_find_magic.restype = C.c_size_t
        L.orc_find_magic.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p, C.c_size_t]
        L.orc_run_length.restype = C.c_uint32
        L.orc_run_length.argtypes = [C.c_void_p, C.c_size_t]
        L.orc_read_stream_header.restype = C.c_int
        L.orc_read_stream_header.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_uint64)]
        L.orc_decode_block.restype = C.c_int
        L.orc_decode_block.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64, C.POINTER(BlockInfo)] + [C.c_void_p] * 7 + [C.c_uint64]
        L.orc_parallel_reader.restype = C.c_int
        L.orc_parallel_reader.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64),
                                          C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_int)]
        L.orc_serial_reader.restype = C.c_int
        L.orc_serial_reader.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64),
                                        C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_int),
                                        C.POINTER(C.c_uint32)]
        _lib = L
    return _lib


def _buf(data) -> np.ndarray:
    a = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
    return np.ascontiguousarray(a)


def _ptr(a: np.ndarray | None):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def crc32(data) -> int:
    a = _buf(data)
    return lib().orc_crc32(_ptr(a), a.size)


def find_magic(data, magic: int) -> np.ndarray:
    a = _buf(data)
    n = lib().orc_find_magic(_ptr(a), a.size, magic, None, 0)
    out = np.zeros(max(n, 1), dtype=np.uint64)
    lib().orc_find_magic(_ptr(a), a.size, magic, _ptr(out), n)
    return out[:n]


def find_magics(data) -> np.ndarray:
    """Sorted offsets of both magics, like `ibzip2 -l` (src/tools/ibzip2.cpp:96-143)."""
    return np.sort(np.concatenate([find_magic(data, MAGIC_BLOCK), find_magic(data, MAGIC_EOS)]))


def run_length(ab: str) -> int:
    a = np.array([0 if c == "A" else 1 for c in ab], dtype=np.uint8)
    return lib().orc_run_length(_ptr(a), a.size) if a.size else 0


@dataclass
class DecodedBlock:
    info: BlockInfo
    selectors: np.ndarray
    symbols: np.ndarray
    lbytes: np.ndarray
    hist: np.ndarray
    tvec: np.ndarray
    prerle: np.ndarray
    out: np.ndarray

    @property
    def status(self) -> int:
        return self.info.status


def decode_block(data, bit_offset: int, out_cap: int = 64 << 20, stages: bool = True) -> DecodedBlock:
    a = _buf(data)
    info = BlockInfo()
    selectors = np.zeros(32768, dtype=np.uint8) if stages else None
    symbols = np.zeros(SYMBOL_CAPACITY, dtype=np.uint16) if stages else None
    lbytes = np.zeros(DBUF_SIZE, dtype=np.uint8) if stages else None
    hist = np.zeros(256, dtype=np.uint32) if stages else None
    tvec = np.zeros(DBUF_SIZE, dtype=np.uint32) if stages else None
    prerle = np.zeros(DBUF_SIZE, dtype=np.uint8) if stages else None
    out = np.zeros(out_cap, dtype=np.uint8)
    lib().orc_decode_block(_ptr(a), a.size, bit_offset, C.byref(info), _ptr(selectors), _ptr(symbols), _ptr(lbytes),
                           _ptr(hist), _ptr(tvec), _ptr(prerle), _ptr(out), out_cap)
    e = np.zeros(0, dtype=np.uint8)
    n = info.n
    return DecodedBlock(
        info=info,
        selectors=selectors[: info.selectors_count] if stages else e,
        symbols=symbols[: min(info.nsymbols, SYMBOL_CAPACITY)] if stages else e,
        lbytes=lbytes[:n] if stages else e,
        hist=hist if stages else e,
        tvec=tvec[:n] if stages else e,
        prerle=prerle[:n] if stages else e,
        out=out[: min(info.out_size, out_cap)],
    )


@dataclass
class ReaderResult:
    status: int
    data: bytes
    offsets: list[tuple[int, int]]
    warn_trailing: bool
    stream_crc: int = 0


def _reader(fn, data, want_data: bool, out_cap: int | None, serial: bool) -> ReaderResult:
    a = _buf(data)
    cap = out_cap if out_cap is not None else max(64 << 20, a.size * 60)
    out = np.zeros(cap, dtype=np.uint8) if want_data else None
    out_size = C.c_uint64(0)
    pairs_cap = a.size // 10 + 16
    offsets = np.zeros(2 * pairs_cap, dtype=np.uint64)
    count = C.c_size_t(0)
    warn = C.c_int(0)
    crc = C.c_uint32(0)
    if serial:
        st = fn(_ptr(a), a.size, _ptr(out), cap if want_data else 0, C.byref(out_size), _ptr(offsets), pairs_cap,
                C.byref(count), C.byref(warn), C.byref(crc))
    else:
        st = fn(_ptr(a), a.size, _ptr(out), cap if want_data else 0, C.byref(out_size), _ptr(offsets), pairs_cap,
                C.byref(count), C.byref(warn))
    n = min(count.value, pairs_cap)
    offs = [(int(offsets[2 * i]), int(offsets[2 * i + 1])) for i in range(n)]
    payload = out[: out_size.value].tobytes() if want_data else b""
    return ReaderResult(st, payload, offs, bool(warn.value), crc.value)


def parallel_reader(data, want_data: bool = True, out_cap: int | None = None) -> ReaderResult:
    return _reader(lib().orc_parallel_reader, data, want_data, out_cap, serial=False)


def serial_reader(data, want_data: bool = True, out_cap: int | None = None) -> ReaderResult:
    return _reader(lib().orc_serial_reader, data, want_data, out_cap, serial=True)


# ---------------------------------------------------------------------------------------------------------------
# The compiled, unmodified reference (only where oracle/_ref was built; it travels to the GPU box prebuilt)
# ---------------------------------------------------------------------------------------------------------------
_ref = None


def ref_available() -> bool:
    return REF_SO.exists()


def ref():
    global _ref
    if _ref is None:
        R = C.CDLL(str(REF_SO))
        R.ref_last_error.restype = C.c_char_p
        R.ref_decode_file.restype = C.c_int
        R.ref_decode_file.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64),
                                      C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_double)]
        R.ref_random_reads.restype = C.c_int
        R.ref_random_reads.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_uint64,
                                       C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_double)]
        R.ref_find_magics.restype = C.c_int
        R.ref_find_magics.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_uint64, C.c_void_p, C.c_uint64,
                                      C.POINTER(C.c_uint64), C.POINTER(C.c_double)]
        _ref = R
    return _ref


@dataclass
class RefResult:
    rc: int
    error: str
    data: bytes
    offsets: list[tuple[int, int]]
    seconds: float


def ref_decode_file(path: str, parallelization: int = 4, serial_class: bool = False, want_data: bool = True,
                    out_cap: int | None = None) -> RefResult:
    R = ref()
    size = os.path.getsize(path)
    cap = out_cap if out_cap is not None else max(64 << 20, size * 60)
    out = np.zeros(cap, dtype=np.uint8) if want_data else None
    out_size = C.c_uint64(0)
    pairs_cap = size // 10 + 16
    offsets = np.zeros(2 * pairs_cap, dtype=np.uint64)
    count = C.c_uint64(0)
    secs = C.c_double(0)
    rc = R.ref_decode_file(str(path).encode(), parallelization, int(serial_class), _ptr(out), cap, C.byref(out_size),
                           _ptr(offsets), pairs_cap, C.byref(count), C.byref(secs))
    err = R.ref_last_error().decode(errors="replace") if rc != 0 else ""
    n = min(count.value, pairs_cap) if rc == 0 else 0
    offs = [(int(offsets[2 * i]), int(offsets[2 * i + 1])) for i in range(n)]
    payload = out[: out_size.value].tobytes() if (want_data and rc == 0) else b""
    return RefResult(rc, err, payload, offs, secs.value)


def ref_find_magic(data, magic: int, parallelization: int = 1, buffer_bytes: int = 1 << 20,
                   capacity: int | None = None) -> np.ndarray:
    a = _buf(data)
    cap = capacity if capacity is not None else a.size // 6 + 16
    out = np.zeros(cap, dtype=np.uint64)
    count = C.c_uint64(0)
    rc = ref().ref_find_magics(_ptr(a), a.size, magic, parallelization, buffer_bytes, _ptr(out), cap, C.byref(count), None)
    if rc != 0:
        raise RuntimeError(ref().ref_last_error().decode())
    return out[: count.value]
