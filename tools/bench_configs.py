#!/usr/bin/env python
"""Device-resident timings of the hot path on the BASELINE.json configs that are NOT the headline bench line
(bench.py measures configs[1]).  One JSON line per config: reads/s, algorithmic GB/s (SURVEY.md 8d bytes per
read) and the fraction of the measured HBM peak.  Developer / evidence tool, run under gpurun:
    python tools/bench_configs.py [--configs C1,C3,C4,C5] [--steps 5]
Sizes are per-batch samples of each config (the kernels stream batches; throughput does not depend on the
number of batches), inputs are generated on the device and are far larger than the 126 MB L2.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="C1,C3,C4,C4k,C4c,C5")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--variant", type=int, default=0)
    args = ap.parse_args()
    import __graft_entry__ as g
    q = g.load_package().hpgq
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    FILT = dict(min_read_length=50, min_read_quality=20, max_N=5)
    cfgs = {
        # name: (synth cfg, reads per batch, work, params kwargs, ctx kwargs, bytes-per-read fn)
        "C1": dict(L=100, n=1 << 20, work=q.WORK_STATS, par={}, ctx={}, bpr=lambda L, o: 2 * L + 4),
        "C3": dict(L=150, n=1 << 22, work=q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT,
                   par=dict(filter=FILT, edit=dict(left_length=10, min_left_quality=20, right_length=10,
                                                   min_right_quality=20)),
                   ctx={}, bpr=None, short_pct=2),
        "C4": dict(L=250, n=1 << 21, work=q.WORK_KMERS | q.WORK_CHAOS, par=dict(k2=11, cg_k=10),
                   ctx=dict(kmers=True, k2=11, cg_k=10), bpr=lambda L, o: 2 * L + 4, cap=1),
        "C4k": dict(L=250, n=1 << 21, work=q.WORK_KMERS, par=dict(k2=11), ctx=dict(kmers=True, k2=11),
                    bpr=lambda L, o: L + 4, cap=1),
        "C4c": dict(L=250, n=1 << 21, work=q.WORK_CHAOS, par=dict(cg_k=10), ctx=dict(cg_k=10),
                    bpr=lambda L, o: 2 * L + 4, cap=1),
        "C5": dict(L=(1000, 20000), n=1 << 16, work=q.WORK_STATS | q.WORK_FILTER,
                   par=dict(filter=dict(min_read_length=2000, min_read_quality=20, max_N=200)), ctx={}, bpr=None),
    }
    for name in args.configs.split(","):
        c = cfgs[name]
        L = c["L"]
        Lmax = L if isinstance(L, int) else L[1]
        scfg = q.synth_cfg(L, short_pct=c.get("short_pct", 0), cap_axis_runs=c.get("cap", 0))
        n = c["n"]
        cap_bytes = n * Lmax if isinstance(L, int) else n * (L[0] + L[1]) // 2 + (64 << 20)
        ctx = q.Ctx(0, max_reads=n, max_bytes=min(cap_bytes, (1 << 32) - 128), max_read_len=(Lmax + 63) // 64 * 64,
                    n_slots=1, host_staging=False, **c["ctx"])
        par = q.make_params(c["work"], kernel_variant=args.variant, **c["par"])
        ctx.set_params(par)
        b = ctx.synth_device(scfg, 0, n, bytes_hint=cap_bytes if not isinstance(L, int) else None)
        ms = []
        for i in range(args.warmup + args.steps):
            ctx.counters_reset()
            ctx.submit_device(b, 0)
            ctx.sync(0)
            if i >= args.warmup:
                ms.append(ctx.last_kernel_ms(0))
        # whole-submit time (kernels + compaction) with events on the slot's stream
        import torch
        st = torch.cuda.ExternalStream(ctx.slot_stream(0), device=torch.device("cuda", 0))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ctx.counters_reset()
        e0.record(st)
        for _ in range(args.steps):
            ctx.submit_device(b, 0)
            ctx.sync(0)
        e1.record(st)
        torch.cuda.synchronize()
        tot_ms = e0.elapsed_time(e1) / args.steps
        counters = ctx.counters()
        if name == "C3":
            bytes_total = b.n_bytes * 2 + n * 4 + n * 9  # read side; output side added below when known
            npass = int(counters[q.S["NUM_PASSED"]]) // args.steps
            # survivors are written trimmed: approximate the output side with the input bytes of all reads
            # (passed + failed partitions are both written): 2*L' + 8 per output read
            bytes_total += b.n_bytes * 2 + n * 8
        elif c["bpr"] is None:
            bytes_total = b.n_bytes * 2 + n * 5
        else:
            bytes_total = sum([c["bpr"](Lmax, None)]) * n if isinstance(L, int) else 0
        k_ms = float(np.mean(ms))
        line = {"config": name, "reads": n, "bases": b.n_bytes, "kernel_ms": k_ms, "submit_ms": tot_ms,
                "reads_per_s_kernel": n / (k_ms * 1e-3), "reads_per_s_total": n / (tot_ms * 1e-3),
                "algorithmic_bytes": bytes_total, "gbs_kernel": bytes_total / (k_ms * 1e-3) / 1e9,
                "gbs_total": bytes_total / (tot_ms * 1e-3) / 1e9, "peak_gbs": peak,
                "frac_total": bytes_total / (tot_ms * 1e-3) / 1e9 / peak, "launches": ctx.launch_count()}
        print(json.dumps(line), flush=True)
        ctx.free_batch(b)
        ctx.close()


if __name__ == "__main__":
    main()
