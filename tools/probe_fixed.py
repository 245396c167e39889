#!/usr/bin/env python
"""Developer probe: kernel time of one stats + filter launch against the batch size (fixed cost of a launch).
    python tools/probe_fixed.py [L | Lmin-Lmax]      (a range: ragged reads through the general loop)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g
q = g.load_package().hpgq
arg = sys.argv[1] if len(sys.argv) > 1 else "100"
ragged = "-" in arg
Lr = tuple(int(x) for x in arg.split("-")) if ragged else None
L = Lr[1] if ragged else int(arg)
par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_length=50, min_read_quality=20, max_N=5))
N = 1 << 23
with q.Ctx(0, max_reads=N, max_bytes=N * L, max_read_len=(L + 63) // 64 * 64, host_staging=False) as ctx:
    ctx.set_params(par)
    cfg = q.synth_cfg(Lr if ragged else L)
    for n in (1000, 100000, 1000000, 4000000, N):
        b = ctx.synth_device(cfg, 0, n, bytes_hint=None if not ragged else n * L)
        ms = []
        for i in range(12):
            ctx.submit_device(b, 0, uniform=not ragged)
            ctx.sync(0)
            ms.append(ctx.last_kernel_ms(0))
        ms = sorted(ms[3:])
        t = ms[len(ms) // 2]
        print(f"L={arg} n={n:8d} kernel {t*1000:8.1f} us  -> {(2 * b.n_bytes + 4 * n) / t / 1e6:8.1f} GB/s", flush=True)
        ctx.free_batch(b)
