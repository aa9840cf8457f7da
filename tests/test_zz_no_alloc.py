"""ibz2_find_magics_noalloc (include/ibz2_cuda.h): K1 with its scratch in a static device array instead of cudaMalloc'ed
memory -- no counterpart in the reference.  Checked through the C ABI
against ibz2_find_magics (the existing entry point) and the golden finder offsets, bit-exact, on calls in a row (the
array is shared between calls and keeps the last call's contents).  Collected last on purpose:
a failure here must not hide the other tests.
"""
import json

import numpy as np
import pytest

import _oracle as orc
from indexed_bzip2_b200 import _capi

pytestmark = pytest.mark.gpu

GOLDEN = json.loads((orc.ROOT / "tests" / "golden" / "golden.json").read_text())


def case(name):
    return (orc.ROOT / "tests" / "golden" / "cases" / f"{name}.bz2").read_bytes()


@pytest.fixture()
def ctx():
    # the context only holds the INPUT in device memory and runs the existing entry point; the new one takes no context
    c = _capi.Context(0)
    yield c
    c.close()


def on_device(ctx, data: bytes):
    dptr = ctx.device_alloc(max(len(data), 16))
    ctx.h2d(dptr, np.frombuffer(data, dtype=np.uint8).copy())
    ctx.synchronize()
    return dptr


def test_noalloc_finder_matches_existing_entry_point_on_calls_in_a_row(ctx):
    # the second file has fewer matches than the first: stale slots of the first call lie behind the second call's count
    names = ["pbzip2-style-5-streams", "text-450k-l1", "pbzip2-style-5-streams", "mixed-384k-l2"]
    for name in names:
        data = case(name)
        dptr = on_device(ctx, data)
        try:
            want = ctx.find_magics(dptr, len(data), device=True)
            for _ in range(2):
                got = _capi.find_magics_noalloc(0, dptr, len(data))
                assert got.dtype == np.uint64 and got.tolist() == want.tolist(), name
            flagged = [int(o) & ~(1 << 63) for o in got]
            assert flagged == GOLDEN[name]["finder"], name
            # one magic only, and a base offset
            blocks = _capi.find_magics_noalloc(0, dptr, len(data), which=_capi.FIND_BLOCK, base_bit_offset=1000)
            assert blocks.tolist() == ctx.find_magics(dptr, len(data), device=True, which=_capi.FIND_BLOCK,
                                                      base_bit_offset=1000).tolist()
        finally:
            ctx.device_free(dptr)


def test_noalloc_finder_planted_magics_and_ragged_sizes(ctx):
    rng = np.random.default_rng(11)
    for n in [0, 5, 6, 7, 100, 4097, 1 << 20]:
        buf = rng.integers(0, 256, n, dtype=np.uint8)
        if n >= 64:
            for k, at in enumerate(rng.integers(0, n * 8 - 48, 40)):
                magic = _capi.MAGIC_BLOCK if k % 2 == 0 else _capi.MAGIC_EOS
                for b in range(48):
                    bit, p = (magic >> (47 - b)) & 1, int(at) + b
                    buf[p >> 3] = (int(buf[p >> 3]) & ~(0x80 >> (p & 7))) | (bit << (7 - (p & 7)))
        data = buf.tobytes()
        dptr = on_device(ctx, data)
        try:
            want = ctx.find_magics(dptr, n, device=True)
            assert _capi.find_magics_noalloc(0, dptr, n).tolist() == want.tolist(), n
            if n >= 6:
                assert sorted(int(o) & ~(1 << 63) for o in want) == sorted(int(o) for o in orc.find_magics(data)), n
            else:
                assert want.size == 0
        finally:
            ctx.device_free(dptr)


def test_noalloc_finder_refuses_what_it_cannot_hold(ctx):
    data = case("text-450k-l1")
    dptr = on_device(ctx, data)
    try:
        with pytest.raises(_capi.Ibz2Error, match="IBZ2_NOALLOC_MAX_MATCHES") as e:
            _capi.find_magics_noalloc(0, dptr, len(data), capacity=_capi.NOALLOC_MAX_MATCHES + 1)
        assert e.value.rc == 4
        # fewer slots than matches: the count is reported, the call says CAPACITY
        with pytest.raises(_capi.Ibz2Error, match="more matches than capacity") as e:
            _capi.find_magics_noalloc(0, dptr, len(data), capacity=2)
        assert e.value.rc == 4
        # host memory is refused before any kernel runs, and the device works afterwards
        host = np.frombuffer(data, dtype=np.uint8).copy()
        with pytest.raises(_capi.Ibz2Error, match="not device memory"):
            _capi.find_magics_noalloc(0, host.ctypes.data, host.size)
        with pytest.raises(_capi.Ibz2Error, match="device index"):
            _capi.find_magics_noalloc(64, dptr, len(data))
        assert _capi.find_magics_noalloc(0, dptr, len(data)).tolist() == ctx.find_magics(dptr, len(data), device=True).tolist()
    finally:
        ctx.device_free(dptr)
