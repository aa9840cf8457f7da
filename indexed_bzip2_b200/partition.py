"""Sharding of one bzip2 file over the ranks of a multi-GPU job: no collective on the data path.

The compressed stream is cut into `world` byte ranges of equal length.  A rank owns the blocks whose 48-bit magic
STARTS inside its range, uploads its range plus some slack (so that its last block is complete: a block is at most
~1.1 MB compressed, 900000 symbols of at most 20 bits plus the header) and passes the range to ibz2_decode_buffer as
[first_bit, last_bit).  The ranks' outputs, concatenated in rank order, are the decoded file; only the byte counts
travel between ranks (one all_gather of a scalar) to place each rank's bytes in the shared output.
"""
from __future__ import annotations

from dataclasses import dataclass

SLACK_BYTES = 4 << 20   # > the largest compressed block (bzip2.hpp:414-416: 900000 symbols x 20 bits = 2.25 MB)
ALL_BITS = (1 << 64) - 1


@dataclass(frozen=True)
class Share:
    first_byte: int   # start of the window this rank uploads
    end_byte: int     # end of the window (exclusive), slack included
    first_bit: int    # ownership range relative to the window start ...
    last_bit: int     # ... blocks with first_bit <= magic offset < last_bit are this rank's


def rank_share(compressed_bytes: int, rank: int, world: int, slack: int = SLACK_BYTES) -> Share:
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    first = compressed_bytes * rank // world
    last = compressed_bytes * (rank + 1) // world
    end = min(compressed_bytes, last + slack)
    last_bit = (last - first) * 8 if rank + 1 < world else ALL_BITS
    return Share(first, end, 0, last_bit)


def owned(offsets_bits, share: Share):
    """The subset of absolute magic bit offsets that `share` owns, made relative to the window."""
    base = share.first_byte * 8
    return [o - base for o in offsets_bits if o >= base and share.first_bit <= o - base < share.last_bit]
