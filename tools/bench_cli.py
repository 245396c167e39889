#!/usr/bin/env python
"""End-to-end timing of the `hpg-fastq` program on a synthetic FASTQ file: FASTQ text in the page cache -> report /
output files, with the records cut on the GPU (default) and on the host (--host-parse).  Developer / evidence tool,
run under gpurun:  python tools/bench_cli.py [--reads 8000000] [--len 150]
One JSON line per (command, parser): wall seconds (best of --reps), reads/s and GB/s of FASTQ text.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def write_fastq(q, path, n, L, chunk=1 << 20):
    """Fixed-width records: '@SYN.%010d' / seq / '+' / qual, generated on the device and laid out with numpy."""
    cfg = q.synth_cfg(L)
    hdr_w = 5 + 10
    rec_w = hdr_w + 1 + L + 1 + 2 + L + 1
    with q.Ctx(0, max_reads=chunk, max_bytes=chunk * L, max_read_len=(L + 63) // 64 * 64, host_staging=False) as ctx, \
            open(path, "wb") as f:
        for first in range(0, n, chunk):
            m = min(chunk, n - first)
            b = ctx.synth_device(cfg, first, m)
            seq = ctx.dev_download(b.seq, m * L).reshape(m, L)
            qual = ctx.dev_download(b.qual, m * L).reshape(m, L)
            ctx.free_batch(b)
            rec = np.empty((m, rec_w), dtype=np.uint8)
            rec[:, :5] = np.frombuffer(b"@SYN.", dtype=np.uint8)
            ids = np.arange(first, first + m, dtype=np.int64)
            for k in range(10):
                rec[:, 5 + 9 - k] = (ids // 10 ** k % 10 + 48).astype(np.uint8)
            o = hdr_w
            rec[:, o] = 10
            rec[:, o + 1:o + 1 + L] = seq
            o += 1 + L
            rec[:, o] = 10
            rec[:, o + 1] = ord("+")
            rec[:, o + 2] = 10
            rec[:, o + 3:o + 3 + L] = qual
            rec[:, o + 3 + L] = 10
            rec.tofile(f)
    return n * rec_w


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=8_000_000)
    ap.add_argument("--len", type=int, default=150)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--batch-size", type=int, default=1 << 20)
    ap.add_argument("--dir", default=None)
    ap.add_argument("--commands", default="stats,filter")
    ap.add_argument("--parsers", default="device,host")
    ap.add_argument("--gpus", default="1", help="comma list: the program's --gpus N, one run per value")
    ap.add_argument("--scan", action="store_true", help="stats without --max-read-length: the program scans the file for the longest line")
    ap.add_argument("--env", action="append", default=[], help="A/B: run every case once more per NAME=VALUE[,NAME=VALUE]")
    args = ap.parse_args()
    import __graft_entry__ as g
    q = g.load_package().hpgq
    cli = os.path.join(g.PKG_DIR, "hpg-fastq")
    base = args.dir or tempfile.mkdtemp(prefix="hpgq_cli_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    path = os.path.join(base, "reads.fq")
    t0 = time.perf_counter()
    nbytes = write_fastq(q, path, args.reads, args.len)
    print(json.dumps({"file": path, "bytes": nbytes, "reads": args.reads, "gen_s": time.perf_counter() - t0}), flush=True)
    cases = [("stats", ["--read-length-range", "50,", "--read-quality-range", "20,", "--max-N", "5"]),
             ("filter", ["--read-length-range", "50,", "--read-quality-range", "20,", "--max-N", "5"])]
    envs = [("", dict(os.environ))]
    for spec in args.env:
        e = dict(os.environ)
        e.update(kv.split("=", 1) for kv in spec.split(","))
        envs.append((spec, e))
    parsers = [x for x in (("device", []), ("host", ["--host-parse"])) if x[0] in args.parsers.split(",")]
    combos = [(c, e, p, f, g, en, ev) for c, e in cases if c in args.commands.split(",") for p, f in parsers
              for g in [int(x) for x in args.gpus.split(",")] for en, ev in envs]
    for command, extra, parser, pflag, ngpu, env_name, env in combos:
        out = os.path.join(base, f"{command}_{parser}")
        os.makedirs(out, exist_ok=True)
        best, trace = None, []
        mrl = ["--max-read-length", str(args.len)] if command == "stats" and not args.scan else []
        argv = [cli, command, "-f", path, "-o", out, "--batch-size", str(args.batch_size)] + mrl + extra + pflag
        if ngpu != 1:
            argv += ["--gpus", str(ngpu)]
        for _ in range(args.reps):
            t0 = time.perf_counter()
            e0 = time.time()
            r = subprocess.run(argv, capture_output=True, text=True, cwd=out, env=env)
            dt = time.perf_counter() - t0
            if r.returncode != 0:
                print(json.dumps({"command": command, "parser": parser, "gpus": ngpu, "error": (r.stdout + r.stderr)[-600:]}), flush=True)
                best = None
                break
            best = dt if best is None else min(best, dt)
            trace = [l for l in r.stderr.splitlines() if l.startswith("[trace]")]  # HPGQ_TRACE=1
            if trace:
                trace = [f"[trace] spawned at epoch {e0:.3f}, returned at {time.time():.3f}"] + trace
        if best is not None:
            log = open(os.path.join(out, "hpg-fastq.log")).read().strip().splitlines()
            print(json.dumps({"command": command, "parser": parser, "gpus": ngpu, "env": env_name, "wall_s": best,
                              "reads_per_s": args.reads / best, "text_gbs": nbytes / best / 1e9, "batch_size": args.batch_size,
                              "log": [l for l in log if "reads on" in l or "phases" in l]}), flush=True)
            for l in trace:
                print(l, flush=True)
    if not args.dir:
        subprocess.run(["rm", "-rf", base])


if __name__ == "__main__":
    main()
