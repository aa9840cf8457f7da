"""ctypes binding of the C ABI in include/ibz2_cuda.h (libibz2_cuda.so, built in-tree by indexed_bzip2_b200.build).

This is the only door from Python into the decoder.  There is no CPU implementation behind it: a missing library
or a machine without a CUDA device raises, loudly.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "libibz2_cuda.so"

MAGIC_BLOCK = 0x314159265359
MAGIC_EOS = 0x177245385090
FIND_BLOCK = 1
FIND_EOS = 2
OFFSET_EOS_FLAG = 1 << 63
MEM_HOST = 0
MEM_DEVICE = 1

RC_NAMES = {0: "OK", 1: "INVALID_ARGUMENT", 2: "CUDA", 3: "NO_DEVICE", 4: "CAPACITY", 5: "OUT_OF_MEMORY"}
STATUS_NAMES = {
    0: "OK", 1: "BAD_MAGIC", 2: "RANDOMISED", 3: "ORIGPTR_RANGE", 4: "GROUP_COUNT", 5: "SELECTOR_COUNT",
    6: "SELECTOR_UNARY", 7: "CODE_LENGTH", 8: "CODE_LENGTHS", 9: "SELECTOR_OVERRUN", 10: "DBUF_OVERFLOW",
    11: "INVALID_CODE", 12: "ORIGPTR", 13: "CRC", 14: "EOF", 15: "OUTPUT_CAPACITY", 16: "STREAM_HEADER",
    17: "STREAM_CRC", 18: "BLOCK_ORDER",
}


class BlockResult(C.Structure):
    _fields_ = [
        ("enc_offset_bits", C.c_uint64), ("enc_size_bits", C.c_uint64), ("out_offset", C.c_uint64),
        ("out_size", C.c_uint64), ("expected_crc", C.c_uint32), ("calculated_crc", C.c_uint32),
        ("status", C.c_uint32), ("is_eos", C.c_uint32), ("is_eof", C.c_uint32), ("n_symbols", C.c_uint32),
        ("orig_ptr", C.c_uint32), ("stream_flags", C.c_uint32),
        ("stream_crc_expected", C.c_uint32), ("stream_crc_calculated", C.c_uint32),
        ("n_coded", C.c_uint32), ("reserved", C.c_uint32),
    ]


BLOCK_RESULT_DTYPE = np.dtype([
    ("enc_offset_bits", "<u8"), ("enc_size_bits", "<u8"), ("out_offset", "<u8"), ("out_size", "<u8"),
    ("expected_crc", "<u4"), ("calculated_crc", "<u4"), ("status", "<u4"), ("is_eos", "<u4"), ("is_eof", "<u4"),
    ("n_symbols", "<u4"), ("orig_ptr", "<u4"), ("stream_flags", "<u4"),
    ("stream_crc_expected", "<u4"), ("stream_crc_calculated", "<u4"), ("n_coded", "<u4"), ("reserved", "<u4"),
])
assert BLOCK_RESULT_DTYPE.itemsize == C.sizeof(BlockResult)


class Timings(C.Structure):
    _fields_ = [(n, C.c_float) for n in (
        "h2d_ms", "find_ms", "huffman_ms", "bwt_prepare_ms", "bwt_traverse_ms", "rle_scan_ms", "rle_write_ms",
        "crc_ms", "d2h_ms", "total_ms")] + [("launches", C.c_uint64), ("huffman_bits_ms", C.c_float),
                                             ("bwt_walk_ms", C.c_float), ("mtf_chunk_ms", C.c_float)]

    def as_dict(self) -> dict:
        return {n: getattr(self, n) for n, _ in self._fields_}


class MultiTimings(C.Structure):
    _fields_ = [("decode_ms", C.c_float), ("copy_ms", C.c_float), ("total_ms", C.c_float), ("reserved", C.c_float),
                ("stream_crc_mismatches", C.c_uint64)]


class Ibz2Error(RuntimeError):
    def __init__(self, rc: int, message: str):
        super().__init__(f"ibz2 error {RC_NAMES.get(rc, rc)}: {message}")
        self.rc = rc


EXPORTS = [
    "ibz2_abi_version", "ibz2_device_count", "ibz2_create", "ibz2_destroy", "ibz2_last_error", "ibz2_get_timings",
    "ibz2_set_batch_blocks", "ibz2_device_alloc", "ibz2_device_free", "ibz2_host_alloc_pinned",
    "ibz2_host_free_pinned", "ibz2_memcpy_h2d", "ibz2_memcpy_d2h", "ibz2_synchronize", "ibz2_find_magics",
    "ibz2_decode_blocks", "ibz2_decode_buffer", "ibz2_timer_start", "ibz2_timer_stop", "ibz2_debug_decode_block", "ibz2_read_stream_header", "ibz2_read_block_header",
    "ibz2_stream_crc_fold", "ibz2_set_verify_stream_crc", "ibz2_fold_stream_crcs",
    "ibz2_upload_pageable", "ibz2_memcpy_d2h_async", "ibz2_copy_synchronize", "ibz2_multi_create", "ibz2_multi_destroy", "ibz2_multi_last_error", "ibz2_multi_device_count",
    "ibz2_multi_host_alloc_pinned", "ibz2_multi_host_free_pinned", "ibz2_multi_decode_buffer", "ibz2_multi_get_timings",
    "ibz2_find_magics_noalloc",
]

_lib = None


def load_library() -> C.CDLL:
    """dlopen libibz2_cuda.so and declare every signature. Raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m indexed_bzip2_b200.build` "
            "(needs nvcc; there is no CPU fallback for the bzip2 decode path)")
    L = C.CDLL(str(LIB_PATH))
    vp, u64, u32, i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
    L.ibz2_abi_version.restype = i32
    L.ibz2_device_count.restype = i32
    L.ibz2_create.restype = i32
    L.ibz2_create.argtypes = [i32, C.POINTER(vp)]
    L.ibz2_destroy.restype = None
    L.ibz2_destroy.argtypes = [vp]
    L.ibz2_last_error.restype = C.c_char_p
    L.ibz2_last_error.argtypes = [vp]
    L.ibz2_get_timings.restype = i32
    L.ibz2_get_timings.argtypes = [vp, C.POINTER(Timings)]
    L.ibz2_set_batch_blocks.restype = i32
    L.ibz2_set_batch_blocks.argtypes = [vp, u32]
    L.ibz2_device_alloc.restype = i32
    L.ibz2_device_alloc.argtypes = [vp, u64, C.POINTER(vp)]
    L.ibz2_device_free.restype = i32
    L.ibz2_device_free.argtypes = [vp, vp]
    L.ibz2_host_alloc_pinned.restype = i32
    L.ibz2_host_alloc_pinned.argtypes = [vp, u64, C.POINTER(vp)]
    L.ibz2_host_free_pinned.restype = i32
    L.ibz2_host_free_pinned.argtypes = [vp, vp]
    L.ibz2_memcpy_h2d.restype = i32
    L.ibz2_memcpy_h2d.argtypes = [vp, vp, vp, u64]
    L.ibz2_memcpy_d2h.restype = i32
    L.ibz2_memcpy_d2h.argtypes = [vp, vp, vp, u64]
    L.ibz2_synchronize.restype = i32
    L.ibz2_synchronize.argtypes = [vp]
    L.ibz2_find_magics.restype = i32
    L.ibz2_find_magics.argtypes = [vp, vp, u64, i32, u64, u32, vp, u64, C.POINTER(u64)]
    L.ibz2_find_magics_noalloc.restype = i32
    L.ibz2_find_magics_noalloc.argtypes = [i32, vp, vp, u64, u64, u32, vp, u64, C.POINTER(u64)]
    L.ibz2_decode_blocks.restype = i32
    L.ibz2_decode_blocks.argtypes = [vp, vp, u64, i32, vp, u64, vp, u64, i32, vp, C.POINTER(u64)]
    L.ibz2_decode_buffer.restype = i32
    L.ibz2_decode_buffer.argtypes = [vp, vp, u64, i32, u64, u64, vp, u64, i32, vp, u64, C.POINTER(u64), C.POINTER(u64)]
    L.ibz2_timer_start.restype = i32
    L.ibz2_timer_start.argtypes = [vp]
    L.ibz2_timer_stop.restype = i32
    L.ibz2_timer_stop.argtypes = [vp, C.POINTER(C.c_float)]
    L.ibz2_debug_decode_block.restype = i32
    L.ibz2_debug_decode_block.argtypes = [vp, vp, u64, u64, C.POINTER(BlockResult), vp, vp, vp, vp, u64]
    L.ibz2_read_stream_header.restype = i32
    L.ibz2_read_stream_header.argtypes = [vp, u64, C.POINTER(u64)]
    L.ibz2_read_block_header.restype = i32
    L.ibz2_read_block_header.argtypes = [vp, u64, u64, C.POINTER(BlockResult)]
    L.ibz2_stream_crc_fold.restype = u32
    L.ibz2_stream_crc_fold.argtypes = [u32, u32]
    L.ibz2_set_verify_stream_crc.restype = i32
    L.ibz2_set_verify_stream_crc.argtypes = [vp, i32]
    L.ibz2_fold_stream_crcs.restype = u64
    L.ibz2_fold_stream_crcs.argtypes = [vp, u64, i32]
    L.ibz2_memcpy_d2h_async.restype = i32
    L.ibz2_memcpy_d2h_async.argtypes = [vp, vp, vp, u64]
    L.ibz2_copy_synchronize.restype = i32
    L.ibz2_copy_synchronize.argtypes = [vp]
    L.ibz2_upload_pageable.restype = i32
    L.ibz2_upload_pageable.argtypes = [vp, vp, vp, u64]
    L.ibz2_multi_create.restype = i32
    L.ibz2_multi_create.argtypes = [C.POINTER(i32), i32, C.POINTER(vp)]
    L.ibz2_multi_destroy.restype = None
    L.ibz2_multi_destroy.argtypes = [vp]
    L.ibz2_multi_last_error.restype = C.c_char_p
    L.ibz2_multi_last_error.argtypes = [vp]
    L.ibz2_multi_device_count.restype = i32
    L.ibz2_multi_device_count.argtypes = [vp]
    L.ibz2_multi_host_alloc_pinned.restype = i32
    L.ibz2_multi_host_alloc_pinned.argtypes = [vp, u64, C.POINTER(vp)]
    L.ibz2_multi_host_free_pinned.restype = i32
    L.ibz2_multi_host_free_pinned.argtypes = [vp, vp]
    L.ibz2_multi_decode_buffer.restype = i32
    L.ibz2_multi_decode_buffer.argtypes = [vp, vp, u64, vp, u64, vp, u64, C.POINTER(u64), C.POINTER(u64)]
    L.ibz2_multi_get_timings.restype = i32
    L.ibz2_multi_get_timings.argtypes = [vp, C.POINTER(MultiTimings), C.POINTER(Timings), i32]
    if L.ibz2_abi_version() != 3:
        raise ImportError("libibz2_cuda.so has an unexpected ABI version")
    _lib = L
    return L


def _ptr(a) -> int:
    """address of a numpy array / bytes-like (kept alive by the caller) or a raw device/host address"""
    if a is None:
        return 0
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return np.frombuffer(a, dtype=np.uint8).ctypes.data


def as_u8(data) -> np.ndarray:
    if isinstance(data, np.ndarray):
        return np.ascontiguousarray(data).view(np.uint8).reshape(-1)
    return np.frombuffer(data, dtype=np.uint8)


class PinnedBuffer:
    """Page-locked host memory owned by a Context (numpy view in .array)."""

    def __init__(self, ctx: "Context", nbytes: int):
        self._ctx = ctx
        p = C.c_void_p()
        ctx._check(ctx.lib.ibz2_host_alloc_pinned(ctx.handle, nbytes, C.byref(p)))
        self.address = p.value
        self.nbytes = nbytes
        self.array = np.ctypeslib.as_array((C.c_uint8 * max(nbytes, 1)).from_address(self.address))[:nbytes]

    def free(self):
        if self.address:
            self._ctx.lib.ibz2_host_free_pinned(self._ctx.handle, self.address)
            self.address = 0
            self.array = None


class Context:
    """One ibz2_ctx (= one GPU). Not thread-safe, like the reference readers (ParallelBZ2Reader.hpp:36-38)."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.ibz2_create(device, C.byref(h))
        if rc != 0:
            raise Ibz2Error(rc, (self.lib.ibz2_last_error(None) or b"").decode())
        self.handle = h
        self.device = device

    def close(self):
        if getattr(self, "handle", None):
            self.lib.ibz2_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc: int, allow=()):
        if rc != 0 and rc not in allow:
            raise Ibz2Error(rc, (self.lib.ibz2_last_error(self.handle) or b"").decode())
        return rc

    # ---- memory ----
    def device_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        self._check(self.lib.ibz2_device_alloc(self.handle, nbytes, C.byref(p)))
        return p.value

    def device_free(self, dptr: int):
        self._check(self.lib.ibz2_device_free(self.handle, dptr))

    def pinned(self, nbytes: int) -> PinnedBuffer:
        return PinnedBuffer(self, nbytes)

    def h2d(self, dptr: int, host, nbytes: int | None = None):
        a = as_u8(host)
        self._check(self.lib.ibz2_memcpy_h2d(self.handle, dptr, a.ctypes.data, a.size if nbytes is None else nbytes))

    def d2h(self, host: np.ndarray, dptr: int, nbytes: int | None = None):
        self._check(self.lib.ibz2_memcpy_d2h(self.handle, host.ctypes.data, dptr, host.nbytes if nbytes is None else nbytes))

    def synchronize(self):
        self._check(self.lib.ibz2_synchronize(self.handle))

    def set_batch_blocks(self, blocks: int):
        self._check(self.lib.ibz2_set_batch_blocks(self.handle, blocks))

    def set_verify_stream_crc(self, verify: bool):
        self._check(self.lib.ibz2_set_verify_stream_crc(self.handle, 1 if verify else 0))

    def timings(self) -> dict:
        t = Timings()
        self._check(self.lib.ibz2_get_timings(self.handle, C.byref(t)))
        return t.as_dict()

    # ---- K1 ----
    def find_magics(self, data, nbytes: int | None = None, *, device: bool = False, which: int = FIND_BLOCK | FIND_EOS,
                    base_bit_offset: int = 0, capacity: int = 1 << 16) -> np.ndarray:
        """Sorted bit offsets (u64, bit 63 set for EOS magics)."""
        if device:
            addr, n = int(data), int(nbytes)
        else:
            a = as_u8(data)
            addr, n = a.ctypes.data, a.size if nbytes is None else nbytes
        while True:
            out = np.empty(capacity, dtype=np.uint64)
            count = C.c_uint64()
            rc = self.lib.ibz2_find_magics(self.handle, addr, n, MEM_DEVICE if device else MEM_HOST, base_bit_offset,
                                           which, out.ctypes.data, capacity, C.byref(count))
            self._check(rc, allow=(4,))
            if rc == 0:
                return out[:count.value].copy()
            capacity = int(count.value)

    # ---- K2..K6 ----
    def decode_blocks(self, data, offsets, out=None, out_capacity: int = 0, *, nbytes: int | None = None,
                      data_device: bool = False, out_device: bool = False):
        """Returns (results structured array, total_out). `out` is a numpy array / PinnedBuffer.array / device address."""
        if data_device:
            daddr, n = int(data), int(nbytes)
        else:
            a = as_u8(data)
            daddr, n = a.ctypes.data, a.size if nbytes is None else nbytes
        offs = np.ascontiguousarray(offsets, dtype=np.uint64)
        results = np.zeros(offs.size, dtype=BLOCK_RESULT_DTYPE)
        total = C.c_uint64()
        if out is None:
            oaddr, cap = 0, 0
        elif out_device:
            oaddr, cap = int(out), int(out_capacity)
        else:
            oaddr, cap = out.ctypes.data, out.nbytes if not out_capacity else out_capacity
        self._check(self.lib.ibz2_decode_blocks(self.handle, daddr, n, MEM_DEVICE if data_device else MEM_HOST,
                                                offs.ctypes.data, offs.size, oaddr, cap,
                                                MEM_DEVICE if out_device else MEM_HOST,
                                                results.ctypes.data, C.byref(total)))
        return results, int(total.value)

    def decode_buffer(self, data, out=None, out_capacity: int = 0, *, nbytes: int | None = None,
                      data_device: bool = False, out_device: bool = False, first_bit: int = 0,
                      last_bit: int = (1 << 64) - 1, max_blocks: int | None = None):
        """find + decode in one call. Returns (results structured array, total_out)."""
        if data_device:
            daddr, n = int(data), int(nbytes)
        else:
            a = as_u8(data)
            daddr, n = a.ctypes.data, a.size if nbytes is None else nbytes
        if out is None:
            oaddr, cap = 0, 0
        elif out_device:
            oaddr, cap = int(out), int(out_capacity)
        else:
            oaddr, cap = out.ctypes.data, out.nbytes if not out_capacity else out_capacity
        capacity = max_blocks if max_blocks is not None else max(16, n // 64)  # a block takes > 100 bytes
        while True:
            results = np.zeros(capacity, dtype=BLOCK_RESULT_DTYPE)
            nblocks, total = C.c_uint64(), C.c_uint64()
            rc = self.lib.ibz2_decode_buffer(self.handle, daddr, n, MEM_DEVICE if data_device else MEM_HOST,
                                             first_bit, last_bit, oaddr, cap, MEM_DEVICE if out_device else MEM_HOST,
                                             results.ctypes.data, capacity, C.byref(nblocks), C.byref(total))
            self._check(rc, allow=(4,))
            if rc == 0:
                return results[:nblocks.value], int(total.value)
            capacity = int(nblocks.value)

    def timer_start(self):
        self._check(self.lib.ibz2_timer_start(self.handle))

    def timer_stop(self) -> float:
        ms = C.c_float()
        self._check(self.lib.ibz2_timer_stop(self.handle, C.byref(ms)))
        return float(ms.value)

    def debug_decode_block(self, data, bit_offset: int, out_capacity: int = 0):
        """Stage-level intermediates of one block: dict(result, lbytes, tvec, prerle, out)."""
        a = as_u8(data)
        res = BlockResult()
        lbytes = np.zeros(900000, dtype=np.uint8)
        tvec = np.zeros(900000, dtype=np.uint32)
        prerle = np.zeros(900000, dtype=np.uint8)
        out = np.zeros(max(out_capacity, 1), dtype=np.uint8)
        self._check(self.lib.ibz2_debug_decode_block(self.handle, a.ctypes.data, a.size, bit_offset, C.byref(res),
                                                     lbytes.ctypes.data, tvec.ctypes.data, prerle.ctypes.data,
                                                     out.ctypes.data if out_capacity else 0, out_capacity))
        n = res.n_symbols
        return {"result": res, "lbytes": lbytes[:n], "tvec": tvec[:n], "prerle": prerle[:n],
                "out": out[:res.out_size] if out_capacity else None}


# ---- host-side header probes (no GPU) ----
def read_stream_header(data, bit_pos: int = 0):
    """(level, new_bit_pos); level 0 = not a bzip2 stream header, -1 = file ended first (bzip2.hpp:114-142)."""
    a = as_u8(data)
    pos = C.c_uint64(bit_pos)
    level = load_library().ibz2_read_stream_header(a.ctypes.data, a.size, C.byref(pos))
    return level, int(pos.value)


def read_block_header(data, bit_offset: int) -> BlockResult:
    a = as_u8(data)
    res = BlockResult()
    load_library().ibz2_read_block_header(a.ctypes.data, a.size, bit_offset, C.byref(res))
    return res


def stream_crc_fold(stream_crc: int, block_crc: int) -> int:
    return int(load_library().ibz2_stream_crc_fold(stream_crc, block_crc))


class MultiContext:
    """ibz2_multi_*: one file decoded by several devices (or several contexts on one device) into ONE host buffer."""

    def __init__(self, devices):
        self.lib = load_library()
        self.devices = [int(d) for d in devices]
        handle = C.c_void_p()
        arr = (C.c_int * len(self.devices))(*self.devices)
        rc = self.lib.ibz2_multi_create(arr, len(self.devices), C.byref(handle))
        if rc != 0:
            raise Ibz2Error(rc, (self.lib.ibz2_multi_last_error(None) or b"").decode())
        self.handle = handle

    def close(self):
        if self.handle:
            self.lib.ibz2_multi_destroy(self.handle)
            self.handle = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc: int):
        if rc != 0:
            raise Ibz2Error(rc, (self.lib.ibz2_multi_last_error(self.handle) or b"").decode())

    def pinned(self, nbytes: int) -> np.ndarray:
        ptr = C.c_void_p()
        self._check(self.lib.ibz2_multi_host_alloc_pinned(self.handle, nbytes, C.byref(ptr)))
        self._pinned = getattr(self, "_pinned", []) + [ptr.value]
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(max(1, nbytes),))[:nbytes]

    def free_pinned(self):
        for ptr in getattr(self, "_pinned", []):
            self.lib.ibz2_multi_host_free_pinned(self.handle, ptr)
        self._pinned = []

    def decode_buffer(self, data, out: np.ndarray | None, max_blocks: int | None = None):
        a = as_u8(data)
        capacity = max_blocks if max_blocks is not None else max(1024, a.size // 4096)
        while True:
            results = np.zeros(capacity, dtype=BLOCK_RESULT_DTYPE)
            nblocks, total = C.c_uint64(), C.c_uint64()
            rc = self.lib.ibz2_multi_decode_buffer(self.handle, a.ctypes.data, a.size,
                                                   out.ctypes.data if out is not None else None,
                                                   out.size if out is not None else 0, results.ctypes.data, capacity,
                                                   C.byref(nblocks), C.byref(total))
            if rc == 4 and nblocks.value > capacity:
                capacity = int(nblocks.value)
                continue
            self._check(rc)
            return results[:nblocks.value], int(total.value)

    def timings(self) -> dict:
        t = MultiTimings()
        per = (Timings * len(self.devices))()
        self._check(self.lib.ibz2_multi_get_timings(self.handle, C.byref(t), per, len(self.devices)))
        return {"decode_ms": t.decode_ms, "copy_ms": t.copy_ms, "total_ms": t.total_ms,
                "stream_crc_mismatches": int(t.stream_crc_mismatches), "devices": [p.as_dict() for p in per]}


NOALLOC_MAX_MATCHES = 65536


def find_magics_noalloc(device: int, dptr: int, nbytes: int, *, which: int = FIND_BLOCK | FIND_EOS, base_bit_offset: int = 0,
                        capacity: int = NOALLOC_MAX_MATCHES, stream: int = 0) -> np.ndarray:
    """ibz2_find_magics_noalloc: K1 over DEVICE memory with the scratch in a static device array; no context, no
    allocation inside the call.  Sorted bit offsets like Context.find_magics."""
    lib = load_library()
    out = np.empty(max(capacity, 1), dtype=np.uint64)
    count = C.c_uint64()
    rc = lib.ibz2_find_magics_noalloc(device, stream or None, dptr or None, nbytes, base_bit_offset, which,
                                      out.ctypes.data if capacity else None, capacity, C.byref(count))
    if rc != 0:
        raise Ibz2Error(rc, (lib.ibz2_last_error(None) or b"").decode())
    return out[:count.value].copy()


def fold_stream_crcs(results: np.ndarray, verify: bool = True) -> int:
    """BZ2Reader.hpp:406-416 / :481-484 over an array of block results in file order (in place); returns the number of
    streams whose CRC did not match."""
    assert results.dtype == BLOCK_RESULT_DTYPE and results.flags.c_contiguous
    return int(load_library().ibz2_fold_stream_crcs(results.ctypes.data, len(results), 1 if verify else 0))
