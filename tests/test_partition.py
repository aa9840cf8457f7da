"""The N > 1 path on CPU: two gloo ranks shard a multi-block file the way bench.py / a multi-GPU reader do
(indexed_bzip2_b200/partition.py), decode their shares with the oracle standing in for the GPU, exchange nothing but
byte counts, and together reproduce the reference's output."""
import hashlib
import json
import os
import socket
import sys

import numpy as np
import pytest

import _oracle as orc
from indexed_bzip2_b200.partition import ALL_BITS, owned, rank_share

GOLDEN = json.loads((orc.ROOT / "tests" / "golden" / "golden.json").read_text())


def case(name):
    return (orc.ROOT / "tests" / "golden" / "cases" / f"{name}.bz2").read_bytes()


def block_magics(data):
    return [int(o) for o in orc.find_magic(data, 0x314159265359)]


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("name", ["text-450k-l1", "pbzip2-style-5-streams", "base64-300k-l1", "ref-1B"])
def test_shares_cover_every_block_exactly_once(name, world):
    data = case(name)
    magics = block_magics(data)
    seen = []
    for rank in range(world):
        share = rank_share(len(data), rank, world, slack=1 << 20)
        assert share.first_byte <= share.end_byte <= len(data)
        mine = owned(magics, share)
        seen += [o + share.first_byte * 8 for o in mine]
        for o in mine:   # an owned block lies completely inside the uploaded window
            info = orc.decode_block(data, o + share.first_byte * 8, stages=False)
            assert o + info.info.enc_size_bits <= (share.end_byte - share.first_byte) * 8
    assert seen == magics
    assert rank_share(len(data), world - 1, world).last_bit == ALL_BITS


def _worker(rank, world, port, name, queue):
    import torch
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        data = case(name)
        share = rank_share(len(data), rank, world, slack=1 << 20)
        window = data[share.first_byte:share.end_byte]
        parts = []
        for o in owned(block_magics(data), share):
            block = orc.decode_block(window, o)          # the rank only ever touches its own window
            assert block.info.status == 0 and block.info.calc_crc == block.info.header_crc
            parts.append(block.out)
        mine = b"".join(parts)
        counts = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(counts, torch.tensor([len(mine)], dtype=torch.int64))
        offset = int(sum(int(c) for c in counts[:rank]))
        dist.barrier()
        queue.put((rank, offset, mine))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["text-450k-l1", "pbzip2-style-5-streams"])
def test_two_gloo_ranks_reproduce_the_reference_output(name):
    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    world = 2
    procs = [ctx.Process(target=_worker, args=(r, world, port, name, queue)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted(queue.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    out = bytearray(GOLDEN[name]["parallel"]["size"])
    for _, offset, part in got:
        out[offset:offset + len(part)] = part
    assert hashlib.sha256(out).hexdigest() == GOLDEN[name]["parallel"]["sha256"]
