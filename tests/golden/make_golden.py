"""Writes the committed regression fixtures under tests/golden/.

The reference ships no inputs, golden vectors or tests and its per-read functions
are not importable, so these fixtures pin (a) the synthetic generator and (b) the oracle's counters on a
seeded batch; the hand-computed vectors and the reference-code comparison live in tests/test_oracle.py.
Run from the repo root:  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import __graft_entry__ as g  # noqa: E402
import oracle_binding as ob  # noqa: E402

q = g.load_package().hpgq

cases = []
for length, seed, cap, short_pct, shift, first, n in [
    (100, q.DEFAULT_SEED, 0, 0, 0, 0, 64), (150, q.DEFAULT_SEED + 1, 0, 2, 2, 1000, 64),
    (250, q.DEFAULT_SEED, 1, 0, 0, 5, 32), ([1000, 20000], q.DEFAULT_SEED, 0, 0, 0, 0, 8)]:
    cfg = q.synth_cfg(tuple(length) if isinstance(length, list) else length, seed=seed, cap_axis_runs=cap,
                      short_pct=short_pct, qual_shift=shift)
    seq, qual, idx = q.synth_host(cfg, first, n)
    cases.append(dict(length=length, seed=seed, cap=cap, short_pct=short_pct, qual_shift=shift, first=first, n=n,
                      head_len=80, seq_head=bytes(seq[:80]).decode(), qual_head=bytes(qual[:80]).decode(),
                      total=int(idx[-1]), seq_sum=int(seq.astype(np.uint64).sum()),
                      qual_sum=int(qual.astype(np.uint64).sum())))
with open(os.path.join(HERE, "synth_golden.json"), "w") as f:
    json.dump(dict(cases=cases), f, indent=1)

n = 5000
cfg = q.synth_cfg(100)
seq, qual, idx = q.synth_host(cfg, 0, n)
lay = ob.layout(q.Layout, 128, kmers=1, k2=7, cg_k=6)
par = q.make_params(q.WORK_STATS | q.WORK_FILTER | q.WORK_KMERS | q.WORK_CHAOS, k2=7, cg_k=6,
                    filter=dict(min_read_length=50, min_read_quality=20, max_N=5))
c, mask, _ = ob.process_batch(par, lay, seq, qual, idx)
sections = {}
for name in q.SEC_NAMES:
    sec = lay.section(c, name).astype(object)
    sections[name] = [int(sum(sec)), int(sum((i + 1) * int(v) for i, v in enumerate(sec)) % (1 << 61))]
with open(os.path.join(HERE, "oracle_golden.json"), "w") as f:
    json.dump(dict(n=n, n_passed=int(mask.sum()), sections=sections), f, indent=1)
print("wrote fixtures; passed", int(mask.sum()), "of", n)
