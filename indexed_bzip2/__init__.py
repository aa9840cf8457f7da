"""`import indexed_bzip2` for code written against the reference's Python module (python/indexed_bzip2/indexed_bzip2.pyx:
293-366): the same names, served by the B200 path of indexed_bzip2_b200.  Only this alias lives here."""
from indexed_bzip2_b200 import *  # noqa: F401,F403
from indexed_bzip2_b200 import (IndexedBzip2File, IndexedBzip2FileRaw, __version__, cli, open)  # noqa: F401

__all__ = ["IndexedBzip2File", "IndexedBzip2FileRaw", "__version__", "cli", "open"]
