"""indexed_bzip2_b200 -- the Python face of the B200 bzip2 readers.

Same surface as the reference's Cython module (python/indexed_bzip2/indexed_bzip2.pyx:69-366):

    open(filename, parallelization=1) -> IndexedBzip2File        (io.BufferedReader, 1 MiB buffer)
    IndexedBzip2FileRaw(filename, parallelization=1)             (io.RawIOBase)
    cli()                                                        (the ibzip2 command line on sys.argv)
    __version__

``filename`` may be a path, a file descriptor, an object with a valid ``fileno()`` or any object with
``read/seek/tell/seekable``.  ``parallelization == 1`` picks the serial reader class (BZ2Reader), anything else the
parallel one (ParallelBZ2Reader); both decode on the GPU through ``libibz2_cuda.so`` -- there is no CPU decoder, and
opening a file on a machine without a usable B200 raises.

The C++ classes live in ``host/``; they are bound here with ctypes through the flat entry points of
``host/reader_capi.cpp`` (the reference binds its classes with Cython; C++ exceptions are mapped to the Python
exception types Cython's ``except +`` produces).
"""
from __future__ import annotations

import ctypes as C
import io
import os
import sys
import tempfile
from pathlib import Path

__version__ = "1.7.0"

_PKG = Path(__file__).resolve().parent
_HOST_LIB = _PKG / "libibz2_host.so"
_lib = None

# enum ibz2r_error_kind (host/reader_capi.cpp) -> what Cython's "except +" raises for the C++ type
_ERRORS = {
    1: ValueError,     # std::invalid_argument
    2: ValueError,     # std::domain_error
    3: IndexError,     # std::out_of_range
    4: RuntimeError,   # std::runtime_error
    5: RuntimeError,   # std::logic_error
    6: RuntimeError,   # EndOfFileReached
    7: MemoryError,    # std::bad_alloc
    8: RuntimeError,
}


def _load():
    global _lib
    if _lib is not None:
        return _lib
    if not _HOST_LIB.exists():
        raise ImportError(
            f"{_HOST_LIB} is missing: build it with `python -m indexed_bzip2_b200.build` (needs nvcc and g++). "
            "There is no pure-Python or CPU decoder to fall back to.")
    lib = C.CDLL(str(_HOST_LIB))
    vp, sz, ll, i32 = C.c_void_p, C.c_size_t, C.c_longlong, C.c_int
    sigs = {
        "ibz2r_open_path": (vp, [C.c_char_p, sz]),
        "ibz2r_open_fd": (vp, [i32, sz]),
        "ibz2r_open_memory": (vp, [vp, sz, sz]),
        "ibz2r_destroy": (None, [vp]),
        "ibz2r_last_error": (C.c_char_p, [vp]),
        "ibz2r_last_error_kind": (i32, [vp]),
        "ibz2r_read": (ll, [vp, vp, sz]),
        "ibz2r_read_to_fd": (ll, [vp, i32, sz]),
        "ibz2r_seek": (ll, [vp, ll, i32]),
        "ibz2r_tell": (ll, [vp]),
        "ibz2r_size": (ll, [vp, i32]),
        "ibz2r_tell_compressed": (ll, [vp]),
        "ibz2r_eof": (i32, [vp]),
        "ibz2r_closed": (i32, [vp]),
        "ibz2r_close": (i32, [vp]),
        "ibz2r_block_offsets_complete": (i32, [vp]),
        "ibz2r_block_offsets": (ll, [vp, C.POINTER(C.c_uint64), sz, i32]),
        "ibz2r_set_block_offsets": (i32, [vp, C.POINTER(C.c_uint64), sz]),
        "ibz2r_join_threads": (i32, [vp]),
        "ibz2r_set_verify_stream_crc": (i32, [vp, i32]),
        "ibz2r_export_index": (i32, [vp, C.c_char_p, i32]),
        "ibz2r_import_index": (i32, [vp, C.c_char_p]),
        "ibz2r_set_open_device": (None, [i32]),
        "ibz2r_set_open_devices": (None, [C.POINTER(i32), i32]),
        "ibz2r_cli": (i32, [i32, C.POINTER(C.c_char_p)]),
    }
    for name, (restype, argtypes) in sigs.items():
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def _isFileObject(file):
    return all(callable(getattr(file, name, None)) for name in ("read", "seek", "tell", "seekable"))


def _hasValidFileno(file):
    if not hasattr(file, "fileno"):
        return False
    try:
        fileno = file.fileno()
        return isinstance(fileno, int) and fileno >= 0
    except Exception:
        return False


class _Reader:
    """What _IndexedBzip2File / _IndexedBzip2FileParallel are in the reference: a thin object around the C++ reader."""

    def __init__(self, file, parallelization, device=None, verify_stream_crc=True, gpus=None):
        if not isinstance(parallelization, int):
            raise TypeError(f"Parallelization argument must be an integer not '{parallelization}'!")
        lib = _load()
        self._lib = lib
        self._handle = None
        lib.ibz2r_set_open_device(-1 if device is None else int(device))   # None: $IBZ2_DEVICE, else device 0
        devices = list(range(gpus)) if isinstance(gpus, int) else [int(d) for d in (gpus or [])]
        lib.ibz2r_set_open_devices((C.c_int * max(1, len(devices)))(*devices), len(devices))
        if isinstance(file, int) and not isinstance(file, bool):
            handle = lib.ibz2r_open_fd(file, parallelization)
        elif _hasValidFileno(file):
            handle = lib.ibz2r_open_fd(file.fileno(), parallelization)
        elif _isFileObject(file):
            # the reference wraps the Python object in a C++ FileReader; the GPU readers keep the whole compressed
            # file in memory anyway, so it is read once, from its start, and its position is restored
            if not file.seekable():
                raise ValueError("Parallel BZ2 Reader will not work on non-seekable input like stdin (yet)!")
            position = file.tell()
            file.seek(0)
            data = file.read()
            file.seek(position)
            if not isinstance(data, bytes):
                data = bytes(data)   # bytearray / memoryview from a custom file object
            handle = lib.ibz2r_open_memory(C.cast(C.c_char_p(data), C.c_void_p), len(data), parallelization)
        elif isinstance(file, str):
            handle = lib.ibz2r_open_path(os.fsencode(file), parallelization)
        else:
            raise Exception("Expected file name string, file descriptor integer, or file-like object for BZ2Reader!")
        if not handle:
            self._raise(None)
        self._handle = handle
        self._parallel = parallelization != 1
        if not verify_stream_crc:
            lib.ibz2r_set_verify_stream_crc(handle, 0)

    def _raise(self, handle):
        kind = self._lib.ibz2r_last_error_kind(handle)
        message = (self._lib.ibz2r_last_error(handle) or b"").decode(errors="replace")
        raise _ERRORS.get(kind, RuntimeError)(message)

    def _checked(self, value):
        if value < 0:
            self._raise(self._handle)
        return value

    def _require(self):
        if not self._handle:
            raise Exception("Invalid file object!")
        return self._handle

    def __del__(self):
        try:
            if self._handle:
                self._lib.ibz2r_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    def close(self):
        if self._handle and not self._lib.ibz2r_closed(self._handle):
            self._checked(self._lib.ibz2r_close(self._handle))

    def closed(self):
        return not self._handle or bool(self._lib.ibz2r_closed(self._handle))

    def seekable(self):
        return bool(self._handle)

    def readinto(self, bytes_like):
        handle = self._require()
        view = memoryview(bytes_like).cast("B")
        if len(view) == 0:
            return 0
        buffer = (C.c_char * len(view)).from_buffer(view)
        return self._checked(self._lib.ibz2r_read(handle, C.addressof(buffer), len(view)))

    def read_to_fd(self, fd, size=-1):
        """Decoded bytes straight to a file descriptor (BZ2ReaderInterface::read( fd ))."""
        n = (1 << 64) - 1 if size is None or size < 0 else size
        return self._checked(self._lib.ibz2r_read_to_fd(self._require(), fd, n))

    def seek(self, offset, whence=io.SEEK_SET):
        return self._checked(self._lib.ibz2r_seek(self._require(), offset, whence))

    def tell(self):
        return self._checked(self._lib.ibz2r_tell(self._require()))

    def size(self):
        # std::optional::value_or( 0 ) like the reference: 0 until the block map is complete
        handle = self._require()
        value = self._lib.ibz2r_size(handle, 0)
        if value < 0 and self._lib.ibz2r_last_error_kind(handle) != 0:
            self._raise(handle)   # -1 is also what a caught exception returns
        return max(0, value)

    def tell_compressed(self):
        return self._checked(self._lib.ibz2r_tell_compressed(self._require()))

    def block_offsets_complete(self):
        return bool(self._lib.ibz2r_block_offsets_complete(self._require()))

    def _offsets(self, available_only):
        handle = self._require()
        count = self._checked(self._lib.ibz2r_block_offsets(handle, None, 0, available_only))
        pairs = (C.c_uint64 * (2 * max(count, 1)))()
        count = self._checked(self._lib.ibz2r_block_offsets(handle, pairs, count, 1))
        return {int(pairs[2 * i]): int(pairs[2 * i + 1]) for i in range(count)}

    def block_offsets(self):
        return self._offsets(0)

    def available_block_offsets(self):
        return self._offsets(1)

    def set_block_offsets(self, offsets):
        items = sorted(dict(offsets).items())
        pairs = (C.c_uint64 * (2 * max(len(items), 1)))()
        for i, (bits, nbytes) in enumerate(items):
            pairs[2 * i] = bits
            pairs[2 * i + 1] = nbytes
        self._checked(self._lib.ibz2r_set_block_offsets(self._require(), pairs, len(items)))

    def join_threads(self):
        self._lib.ibz2r_join_threads(self._require())


class IndexedBzip2FileRaw(io.RawIOBase):
    def __init__(self, filename, parallelization=1, **options):
        self.bz2reader = _Reader(filename, parallelization, **options)
        self.name = filename
        self.mode = "rb"
        self.readinto = self.bz2reader.readinto
        self.seek = self.bz2reader.seek
        self.tell = self.bz2reader.tell
        self.seekable = self.bz2reader.seekable
        if parallelization != 1:
            self.join_threads = self.bz2reader.join_threads
        # IOBase provides sane default implementations for read, readline, readlines, readall, ...

    def close(self):
        if self.closed:
            return
        super().close()
        self.bz2reader.close()

    def readable(self):
        return True


class IndexedBzip2File(io.BufferedReader):
    def __init__(self, filename, parallelization=1, **options):
        fobj = IndexedBzip2FileRaw(filename, parallelization, **options)
        self.bz2reader = fobj.bz2reader
        self.tell_compressed = self.bz2reader.tell_compressed
        self.block_offsets = self.bz2reader.block_offsets
        self.set_block_offsets = self.bz2reader.set_block_offsets
        self.block_offsets_complete = self.bz2reader.block_offsets_complete
        self.available_block_offsets = self.bz2reader.available_block_offsets
        self.size = self.bz2reader.size
        if parallelization != 1:
            self.join_threads = self.bz2reader.join_threads
        # Most of the calls like close, seekable, name, mode ... are forwarded to the given raw object
        # by BufferedReader or more specifically _BufferedIOMixin
        super().__init__(fobj, buffer_size=1024**2)

    def export_index(self, file, format="csv"):
        """The block map as the CLI's -L writes it ("csv": one "<encoded bits>,<decoded bytes>" line per entry) or as a
        GZIDX file ("gzidx": the indexed_gzip format rapidgzip reads and writes for bzip2 as well)."""
        if format not in ("csv", "gzidx"):
            raise ValueError("format must be 'csv' or 'gzidx'")
        reader = self.bz2reader
        if not hasattr(file, "write"):
            reader._checked(reader._lib.ibz2r_export_index(reader._require(), os.fsencode(str(file)), int(format == "gzidx")))
            return
        with tempfile.TemporaryDirectory() as tmp:
            path = os.path.join(tmp, "index")
            reader._checked(reader._lib.ibz2r_export_index(reader._require(), os.fsencode(path), int(format == "gzidx")))
            data = Path(path).read_bytes()
        file.write(data.decode() if isinstance(file, io.TextIOBase) else data)

    def import_index(self, file):
        """Either format (told apart by the GZIDX magic); a GZIDX file must list every block, as export_index's do."""
        reader = self.bz2reader
        if not hasattr(file, "read"):
            reader._checked(reader._lib.ibz2r_import_index(reader._require(), os.fsencode(str(file))))
            return
        data = file.read()
        with tempfile.TemporaryDirectory() as tmp:
            path = os.path.join(tmp, "index")
            Path(path).write_bytes(data.encode() if isinstance(data, str) else bytes(data))
            reader._checked(reader._lib.ibz2r_import_index(reader._require(), os.fsencode(path)))


def open(filename, parallelization=1, *, device=None, verify_stream_crc=True, gpus=None):
    """
    filename: can be a file path, a file descriptor, or a file object
              with suitable read, seekable, seek, and tell methods.
    device: CUDA device index (default: $IBZ2_DEVICE, else 0).  New next to the reference.
    verify_stream_crc: the parallel reader checks the CRC of every bzip2 stream while it builds the block map, which
              the reference only does in its serial reader; False restores the reference's behaviour.
    gpus: a number N (devices 0..N-1) or a list of device indices that share the decoding of every large batch of the
              parallel reader.  New next to the reference.
    """
    return IndexedBzip2File(filename, parallelization, device=device, verify_stream_crc=verify_stream_crc, gpus=gpus)


def cli(args=None):
    lib = _load()
    args = list(sys.argv if args is None else args)
    argv = (C.c_char_p * (len(args) + 1))(*[os.fsencode(a) for a in args], None)
    sys.stdout.flush()
    return lib.ibz2r_cli(len(args), argv)
