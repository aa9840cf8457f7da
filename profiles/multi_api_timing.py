"""Per-call and per-device timings of ibz2_multi_decode_buffer on the config-3 workload, outside torchrun:
    python profiles/multi_api_timing.py [devices]        (profiles/r2_multi_api_n2_standalone.log is its output at 2)
"""
import sys, time, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import benchdata, bench
from indexed_bzip2_b200 import _capi

class A: pass
args = A(); args.kind='mixed'; args.level=9; args.base_bytes=1_000_000_000; args.copies=8; args.config=3
w, key = bench.get_workload(args, 0, lambda: None)
devices = int(sys.argv[1]) if len(sys.argv) > 1 else 2
comp = w.compressed
with _capi.MultiContext(list(range(devices))) as multi:
    pin_in = multi.pinned(comp.size); pin_in[:] = comp
    out = multi.pinned(w.raw_size + (64 << 20))
    for it in range(4):
        t0 = time.perf_counter()
        results, total = multi.decode_buffer(pin_in, out, max_blocks=w.blocks + 16)
        ms = (time.perf_counter() - t0) * 1e3
        t = multi.timings()
        print(it, round(ms, 1), {k: (round(v, 1) if isinstance(v, float) else v) for k, v in t.items() if k != 'devices'})
        for d in t['devices']:
            print('   ', {k: round(v, 1) for k, v in d.items()})
    print('verified', w.verify(out[:total]))
    multi.free_pinned()
