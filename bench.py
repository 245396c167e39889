#!/usr/bin/env python
"""bench.py -- headline benchmark of the per-read QC hot path (BASELINE.json configs[1]).

Workload (config "C2"): hpg-fastq `stats` with filters (--read-length-range 50, --read-quality-range 20,
--max-N 5) on 100 M synthetic 150 bp paired-end reads (= two 50 M-read mate streams) per GPU.
  value : reads/s with the reads already resident in HBM (CUDA events on the launching stream)
  e2e   : reads/s through the C ABI from pinned HOST batches (H2D copies, kernels, D2H of the masks and of
          the counter block inside the timed region, wall clock between device synchronisations)
  roofline : algorithmic bytes (305 B/read = 2L+4+1) / mean duration of the stats+filter kernel, against the
          measured HBM copy bandwidth in MEASURED_PEAKS.json
  cpu_baseline : the CPU oracle (port of the reference's path; the reference's own per-read functions live in
          a submodule that is absent from its tree) on the box's host cores, bounded sample
`--impl reference` times that CPU path alone (rank 0 only).
One process per GPU (torchrun for N > 1); reads are sharded by read index, no data-path collective; the only
collective is one sum-allreduce of the counter block per step.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

READ_LEN = 150
BYTES_PER_READ = 2 * READ_LEN + 4 + 1  # seq + qual + u32 offset + mask byte
FILTER = dict(min_read_length=50, min_read_quality=20, max_N=5)
METRIC = "reads/s (stats+filter, 150 bp PE, per-read QC hot path)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--reads", type=int, default=100_000_000, help="reads per GPU per step")
    ap.add_argument("--batch-reads", type=int, default=1 << 23)
    ap.add_argument("--e2e-batch-reads", type=int, default=1 << 21)
    ap.add_argument("--e2e-pool", type=int, default=8, help="distinct pinned host batches cycled by the e2e leg")
    ap.add_argument("--cpu-sample-reads", type=int, default=4_000_000)
    ap.add_argument("--configs", default="C1,C3,C4,C5", help="other BASELINE.json configs measured after the timed C2 region")
    ap.add_argument("--config-scale", type=float, default=1.0, help="fraction of each other config's named size")
    ap.add_argument("--no-configs", action="store_true")
    ap.add_argument("--verify-reads", type=int, default=10_000_000, help="reads checked against the oracle (rank 0)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML from a thread while the timed region runs."""

    def __init__(self, gpu, period=0.004):
        import threading
        self.sm, self.reasons, self.mx, self.err = [], set(), None, None
        self._stop = threading.Event()
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[gpu]) if vis and vis.split(",")[gpu].strip().isdigit() else gpu
            self.nv, self.h = nv, nv.nvmlDeviceGetHandleByIndex(phys)
            self.mx = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
        except Exception as e:  # no NVML: report it, do not fail the bench
            self.nv, self.err = None, f"nvml unavailable: {e}"
            return
        self.period = period
        self.t = threading.Thread(target=self._run, daemon=True)

    def start(self):
        """NVML was initialised in __init__ (tens of ms, and not the same on every rank): only the thread starts here,
        after the barrier that opens the timed region."""
        if self.nv is not None:
            self.t.start()

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception as e:
                self.err = str(e)
                break
            self._stop.wait(self.period)

    def stop(self):
        if self.nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": [self.err]}
        self._stop.set()
        self.t.join(timeout=2)
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.mx,
                "samples": len(self.sm), "reasons": sorted(self.reasons)}


def host_cores():
    """Cores this process may use.  Never OpenMP's default: torchrun exports OMP_NUM_THREADS=1."""
    return len(os.sched_getaffinity(0))


def cpu_sample(q, ctx, cfg, n_sample):
    if ctx is not None:
        b = ctx.synth_device(cfg, 0, n_sample)
        seq = ctx.dev_download(b.seq, b.n_bytes)
        qual = ctx.dev_download(b.qual, b.n_bytes)
        idx = ctx.dev_download(b.idx, (n_sample + 1) * 4).view(np.uint32)
        ctx.free_batch(b)
        return seq, qual, idx
    return q.synth_host(cfg, 0, n_sample)


def cpu_leg(q, ob, sample, steps, warmup, threads):
    """Times the CPU oracle (stats + filter of C2) on a bounded sample with an explicit thread count."""
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=FILTER)
    seq, qual, idx = sample
    lay = ob.layout(q.Layout, 160)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        ob.process_batch(par, lay, seq, qual, idx, threads=threads, finalize=True)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    dt = sum(times) / len(times)
    return (len(idx) - 1) / dt, dt


def cpu_baseline_block(q, ob, ctx, cfg, n_sample, steps=1, warmup=1):
    """cpu_baseline of the bench line: the oracle port on all host cores, the 2-thread figure that mirrors the
    reference's default --num-threads (src/stats_options.c:21), and -- once, on a C1-shaped sample -- the
    "reference-faithful merge": one serial consumer doing a hash insertion per counter per base
    (src/stats_fastq.c:338-382), which is where the reference itself spends its time."""
    cores = host_cores()
    sample = cpu_sample(q, ctx, cfg, n_sample)
    val, dt = cpu_leg(q, ob, sample, steps, warmup, cores)
    n2 = min(n_sample, 1_000_000)
    s2 = (sample[0][:int(sample[2][n2])], sample[1][:int(sample[2][n2])], sample[2][:n2 + 1])
    val2, _ = cpu_leg(q, ob, s2, 1, 0, 2)
    nf = 200_000
    fseq, fqual, fidx = q.synth_host(q.synth_cfg(100), 0, nf)
    lay = ob.layout(q.Layout, 128)
    t0 = time.perf_counter()
    ob.merge_hash_faithful(lay, fseq, fqual, fidx)
    tf = time.perf_counter() - t0
    return {"value": val, "unit": "reads/s", "cores": cores, "kind": "port",
            "sample": f"{n_sample} reads x {READ_LEN} bp of the same workload, {steps} pass(es) of {dt:.2f} s, "
                      "oracle/hpgq_oracle.c (unoptimised three-pass port; the reference's per-read functions are in a "
                      "submodule absent from its tree), OpenMP, batch-parallel like the reference's worker pool",
            "threads_2": {"value": val2, "unit": "reads/s", "cores": 2, "sample": f"{n2} reads",
                          "note": "the reference's default --num-threads (src/stats_options.c:21)"},
            "reference_faithful_merge": {"value": nf / tf, "unit": "reads/s", "cores": 1,
                                         "sample": f"{nf} reads x 100 bp (C1 shape), per-position counters only",
                                         "note": "serial consumer, one hash insertion per counter per base "
                                                 "(src/stats_fastq.c:338-382)"}}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import __graft_entry__ as g
    import oracle_binding as ob
    q = g.load_package().hpgq
    cfg = q.synth_cfg(READ_LEN)
    n = min(args.cpu_sample_reads, 2_000_000)  # host-generated here (3 s per million reads): keep it bounded
    cores = host_cores()
    sample = cpu_sample(q, None, cfg, n)
    val, dt = cpu_leg(q, ob, sample, max(args.steps, 1), min(args.warmup, 1), cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "reads/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8/int64", "data": "synthetic",
        "config": {"workload": "C2 stats+filter, 150 bp reads, CPU path", "sample_reads_per_step": n,
                   "filters": FILTER},
        "cpu_baseline": {"value": val, "unit": "reads/s", "cores": cores, "kind": "port",
                         "sample": f"{n} reads x {READ_LEN} bp per step, oracle/hpgq_oracle.c (unoptimised port, OpenMP, "
                                   "batch-parallel like the reference's worker pool) on every core of the affinity mask; "
                                   "the reference's per-read functions are in a submodule absent from its tree, so "
                                   "oracle/_ref holds only its chaos-game code"},
        "e2e": {"value": val, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


def bind_to_gpu_numa_node(local):
    """Pins this rank's threads to the CPUs of the NUMA node its GPU hangs off, BEFORE any pinned memory is
    allocated: first-touch then places the pinned batches next to the GPU's PCIe root, so that the H2D copies of
    eight ranks do not all stream out of one socket's DRAM (and across the inter-socket link)."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(local)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = set()
            for part in f.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        if allowed:
            os.sched_setaffinity(0, allowed)
            return node
    except Exception:
        pass
    return None


_REAL_STDOUT = None


def quiet_stdout():
    """Everything libraries print to fd 1 (NCCL's version banner, for one) goes to stderr: stdout carries exactly
    one JSON line, written by emit()."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT if _REAL_STDOUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


IGNORED_SCALARS = ("SLOW_READS", "BATCHES", "CG_LONG_RUNS", "CG_FIXED_READS", "CG_MOVED_WORDS")


def verify_against_oracle(q, ob, ctx, par, batches, want_reads, threads):
    """Outside the timed region: the first batches (>= want_reads reads) go through the GPU one at a time and, streamed
    chunk by chunk, through the CPU oracle; the whole counter block and every mask byte must be equal."""
    import ctypes as C
    ctx.counters_reset()
    lay = ctx.layout
    exp = np.zeros(lay.words, dtype=np.uint64)
    done = 0
    for b in batches:
        if done >= want_reads:
            break
        ctx.submit_device(b, 0, uniform=True)
        ctx.sync(0)
        views = q.Result()
        q._ck(ctx.lib.hpgq_batch_device_views(ctx.h, 0, C.byref(views)), "device_views")
        mask = ctx.dev_download(views.mask, b.n_reads)
        seq = ctx.dev_download(b.seq, b.n_bytes)
        qual = ctx.dev_download(b.qual, b.n_bytes)
        idx = ctx.dev_download(b.idx, (b.n_reads + 1) * 4).view(np.uint32)
        exp, emask, _ = ob.process_batch(par, lay, seq, qual, idx, counters=exp, threads=threads, finalize=False)
        if not np.array_equal(mask, emask):
            raise SystemExit(f"bench: pass/fail mask differs from the oracle in the batch starting at read {done}")
        done += b.n_reads
    ob.lib().oracle_finalize(C.byref(lay), exp.ctypes.data_as(C.c_void_p))
    got = ctx.counters()
    ign = {q.S[k] for k in IGNORED_SCALARS}
    bad = [int(i) for i in np.nonzero(got != exp)[0] if int(i) not in ign]
    if bad:
        raise SystemExit(f"bench: counter words {bad[:8]} differ from the oracle over the first {done} reads")
    return done


def h2d_probe(torch, dist, world, nbytes=512 << 20, reps=4):
    """Pinned cudaMemcpyAsync host -> device bandwidth with every rank copying at once, two copies in flight per GPU
    (like the e2e leg's two stream slots): the ceiling of the e2e leg, measured in the same run."""
    srcs = [torch.empty(nbytes, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
    dsts = [torch.empty(nbytes, dtype=torch.uint8, device="cuda") for _ in range(2)]
    streams = [torch.cuda.Stream() for _ in range(2)]
    for s_, d_ in zip(srcs, dsts):
        s_.fill_(7)
        d_.copy_(s_, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        for st, s_, d_ in zip(streams, srcs, dsts):
            with torch.cuda.stream(st):
                d_.copy_(s_, non_blocking=True)
    torch.cuda.synchronize()
    gbs = 2 * nbytes * reps / (time.perf_counter() - t0) / 1e9
    t = torch.tensor([gbs], device="cuda", dtype=torch.float64)
    if world > 1:
        g = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(g, t)
        return [float(x.item()) for x in g]
    return [gbs]


def mix64(a, b):
    """splitmix64 finaliser of (a, b): combines per-batch checksums with their global batch number."""
    M = (1 << 64) - 1
    z = (a * 0x9E3779B97F4A7C15 + (b + 1) * 0xD1B54A32D192ED03) & M
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
    return z ^ (z >> 31)


def config_table(q, scale):
    """BASELINE.json configs[0, 2, 3, 4] at their named size per GPU (scale < 1 shrinks them for quick runs)."""
    c3_edit = dict(left_length=10, min_left_quality=20, right_length=10, min_right_quality=20)
    return {
        "C1": dict(desc="stats, 1 M x 100 bp single-end", L=100, reads=max(int(1_000_000 * scale), 1), batch=1 << 20,
                   work=q.WORK_STATS, par={}, ctx={}, uniform=True, reps=20),
        "C3": dict(desc="edit (3'/5' quality trim) + filter + compacted output, 500 M x 150 bp over 8 GPUs = 62.5 M per GPU",
                   L=150, reads=int(62_500_000 * scale), batch=1 << 22, short_pct=2,
                   work=q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT, par=dict(filter=FILTER, edit=c3_edit), ctx={},
                   uniform=False, reps=1, checksum=True),
        "C4": dict(desc="k-mers (k = 5 by position, k = 11) + chaos game k = 10, 200 M x 250 bp", L=250,
                   reads=int(200_000_000 * scale), batch=1 << 21, cap=1, work=q.WORK_KMERS | q.WORK_CHAOS,
                   par=dict(k2=11, cg_k=10), ctx=dict(kmers=True, k2=11, cg_k=10), uniform=True, reps=1, halves=2),
        "C5": dict(desc="stats + filter, 1-20 kb reads, 10 Gbases", L=(1000, 20000), bases=int(10_000_000_000 * scale),
                   batch=1 << 16, work=q.WORK_STATS | q.WORK_FILTER,
                   par=dict(filter=dict(min_read_length=2000, min_read_quality=20, max_N=200)), ctx={}, uniform=False,
                   reps=1),
    }


def run_config(q, torch, dist, name, c, rank, world, local, peak):
    """One pass of the hot path over a config's reads, device-resident, batches cycled through two stream slots;
    CUDA events on slot 0's stream (slot 1 joined), barrier on both sides, max over ranks."""
    L = c["L"]
    fixed = isinstance(L, int)
    Lmax = L if fixed else L[1]
    mean = L if fixed else (L[0] + L[1]) // 2
    B = c["batch"]
    reads = c["reads"] if "reads" in c else max(c["bases"] // mean, 1)
    scfg = q.synth_cfg(L, short_pct=c.get("short_pct", 0), cap_axis_runs=c.get("cap", 0))
    cap_bytes = B * Lmax if fixed else B * mean + (96 << 20)
    ctx = q.Ctx(local, max_reads=B, max_bytes=min(cap_bytes, (1 << 32) - 128), max_read_len=(Lmax + 63) // 64 * 64,
                n_slots=2, host_staging=False, **c["ctx"])
    par = q.make_params(c["work"], **c["par"])
    ctx.set_params(par)
    streams = [torch.cuda.ExternalStream(ctx.slot_stream(i), device=torch.device("cuda", local)) for i in range(2)]
    halves = c.get("halves", 1)
    per_half = (reads + halves - 1) // halves
    first0 = rank * reads
    tot_ms, tot_reads, tot_bases, launches, n_batches = 0.0, 0, 0, 0, 0
    out_bytes, out_reads, job_ck, n_passed_ck = 0, 0, 0, 0
    kms = []
    for h in range(halves):
        lo, hi = h * per_half, min(reads, (h + 1) * per_half)
        batches = []
        f = lo
        while f < hi:
            nb = min(B, hi - f)
            batches.append((f, ctx.synth_device(scfg, first0 + f, nb, bytes_hint=None if fixed else cap_bytes)))
            f += nb

        def passes(reps):
            # batches of all passes are queued back to back on the two slots; a slot is awaited only to be reused
            busy = [False, False]
            k = 0
            for _ in range(reps):
                for _, b in batches:
                    sl = k & 1
                    k += 1
                    if busy[sl]:
                        ctx.sync(sl)
                        kms.append(ctx.last_kernel_ms(sl))
                    ctx.submit_device(b, sl, uniform=c["uniform"] and fixed)
                    busy[sl] = True
            for sl in (0, 1):
                if busy[sl]:
                    ctx.sync(sl)
                    kms.append(ctx.last_kernel_ms(sl))

        ctx.counters_reset()
        w = batches[:4]
        for i, (_, b) in enumerate(w):  # warm-up on the first batches
            ctx.submit_device(b, i & 1, uniform=c["uniform"] and fixed)
            ctx.sync(i & 1)
        ctx.counters_reset()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        l0 = ctx.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(streams[0])
        passes(c["reps"])
        j = torch.cuda.Event()
        j.record(streams[1])
        streams[0].wait_event(j)
        e1.record(streams[0])
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        tot_ms += e0.elapsed_time(e1) / c["reps"]
        launches += (ctx.launch_count() - l0) // c["reps"]
        tot_reads += sum(b.n_reads for _, b in batches)
        tot_bases += sum(b.n_bytes for _, b in batches)
        n_batches += len(batches)
        if c.get("checksum"):
            # untimed second pass: order-sensitive checksum of every batch's compacted output, keyed by the batch's
            # global number (= its place in the input order), so the sum over ranks is the checksum of the gathered job
            for f, b in batches:
                ctx.submit_device(b, 0)
                ctx.sync(0)
                r = ctx.checksum(0)
                gid = (first0 + f) // B
                ck = mix64(r["checksum"], gid)
                job_ck = (job_ck + ck) & ((1 << 64) - 1)
                n_passed_ck += r["n_passed"]
                out_bytes += r["passed_bytes"] + r["failed_bytes"]
                out_reads += b.n_reads
        for _, b in batches:
            ctx.free_batch(b)
    ctx.close()
    t = torch.tensor([tot_ms], device="cuda", dtype=torch.float64)
    agg = torch.tensor([tot_reads, tot_bases, out_bytes, out_reads], device="cuda", dtype=torch.float64)
    shard_ck = [job_ck]
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(agg)
        mine = torch.tensor([job_ck & 0xFFFFFFFF, job_ck >> 32], device="cuda", dtype=torch.int64)
        g = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(g, mine)
        shard_ck = [int(x[0].item()) | (int(x[1].item()) << 32) for x in g]
    ms = float(t.item())
    j_reads, j_bases, j_ob, j_or = (float(x) for x in agg.tolist()[:4])
    if name == "C1":
        alg = j_bases * 2 + j_reads * 4
    elif name == "C3":
        alg = j_bases * 2 + j_reads * (4 + 9) + 2 * j_ob + 8 * j_or
    elif name == "C4":
        alg = j_bases * 2 + j_reads * 4
    else:
        alg = j_bases * 2 + j_reads * 5
    out = {"workload": c["desc"], "reads_per_gpu": tot_reads, "bases_per_gpu": tot_bases, "batches_per_gpu": n_batches,
           "batch_reads": B, "ms": ms, "reads_per_s": j_reads / (ms * 1e-3), "bases_per_s": j_bases / (ms * 1e-3),
           "algorithmic_bytes": alg, "gbs": alg / (ms * 1e-3) / 1e9, "frac": alg / (ms * 1e-3) / 1e9 / (peak * world),
           "launches_per_pass": int(launches), "kernel_ms_sum": float(sum(kms)) / c["reps"],
           "resident_passes": halves}
    if c.get("checksum"):
        out["output_bytes_per_gpu"] = out_bytes * 2
        # Every batch's compacted output (passed reads, then failed reads, in input order) is hashed position by
        # position on the device and the hash is keyed by the batch's global number, so a shard's sum is the checksum
        # of its part of the output gathered in input order.  Shard 0 holds the same reads at every GPU count: its
        # value must not change with N; the job's value is the sum over shards (mod 2^64).
        out["checksum_shards"] = [f"{x:016x}" for x in shard_ck]
        out["checksum_job"] = f"{sum(shard_ck) & ((1 << 64) - 1):016x}"
        out["passed_fraction_rank0"] = n_passed_ck / max(out_reads, 1)
    return out


def main():
    args = parse()
    quiet_stdout()
    if args.impl == "reference":
        run_reference(args)
        return
    import torch
    import torch.distributed as dist

    import __graft_entry__ as g
    q = g.load_package().hpgq
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: there is no CPU fallback (use --impl reference for the CPU path)")
    torch.cuda.set_device(local)
    numa_node = bind_to_gpu_numa_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    n_reads = args.reads
    half = n_reads // 2
    B = args.batch_reads
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=FILTER)
    # ---- resident data: two mate streams, sharded by read index across ranks
    NSLOT = 4
    ctx = q.Ctx(local, max_reads=B, max_bytes=B * READ_LEN, max_read_len=160, n_slots=NSLOT, host_staging=False)
    ctx.set_params(par)
    mates = [q.synth_cfg(READ_LEN, seed=q.DEFAULT_SEED), q.synth_cfg(READ_LEN, seed=q.DEFAULT_SEED + 1, qual_shift=2)]
    batches = []
    for m, cfg in enumerate(mates):
        first = rank * half
        left = half if m == 0 else n_reads - half
        while left > 0:
            nb = min(B, left)
            batches.append(ctx.synth_device(cfg, first, nb))
            first += nb
            left -= nb
    total_reads = sum(b.n_reads for b in batches)
    streams = [torch.cuda.ExternalStream(ctx.slot_stream(i), device=torch.device("cuda", local)) for i in range(NSLOT)]
    stream = streams[0]

    # counter block as a torch tensor (for the one allreduce per step)
    dptr, nwords = ctx.counters_device_ptr()

    class _Holder:
        __cuda_array_interface__ = {"shape": (nwords,), "typestr": "<i8", "data": (dptr, False), "version": 3}

    ctr_t = torch.as_tensor(_Holder(), device=torch.device("cuda", local))

    kernel_ms = []

    def step(collect):
        # batches are queued back to back on the slots' streams; a slot is retired only when it is needed again
        ctx.counters_reset()
        inflight = [None] * NSLOT
        for i, b in enumerate(batches):
            sl = i % NSLOT
            if inflight[sl] is not None:
                ctx.sync(sl)
                if collect:
                    kernel_ms.append((ctx.last_kernel_ms(sl), inflight[sl]))
            ctx.submit_device(b, sl, uniform=True)
            inflight[sl] = b.n_reads
        for sl in range(NSLOT):
            if inflight[sl] is not None:
                ctx.sync(sl)
                if collect:
                    kernel_ms.append((ctx.last_kernel_ms(sl), inflight[sl]))
        if world > 1:
            with torch.cuda.stream(stream):
                dist.all_reduce(ctr_t)

    for _ in range(args.warmup):
        step(False)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    l0 = ctx.launch_count()
    ev0 = torch.cuda.Event(enable_timing=True)
    ev1 = torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        step(True)
    # the slots' streams join stream 0 before the closing event is recorded on it
    for st in streams[1:]:
        e = torch.cuda.Event()
        e.record(st)
        stream.wait_event(e)
    ev1.record(stream)
    barrier()
    clocks = sampler.stop()
    launches = ctx.launch_count() - l0
    ms_total = ev0.elapsed_time(ev1)
    t = torch.tensor([ms_total], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    value = total_reads * world / (ms_step * 1e-3)
    counters = ctx.counters()
    n_passed = int(counters[q.S["NUM_PASSED"]])
    verified = 0
    if rank == 0 and args.verify_reads > 0:
        import oracle_binding as ob
        verified = verify_against_oracle(q, ob, ctx, par, batches, args.verify_reads, host_cores())
    # roofline of the dominant kernel
    peak, peak_src = measured_peak()
    k_ms = sum(m for m, _ in kernel_ms)
    k_reads = sum(r for _, r in kernel_ms)
    achieved = (k_reads * BYTES_PER_READ) / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
            traffic = tj.get("hq_qc_tiled_kernel_bytes_per_launch")
            traffic_src = ("committed constant from profiles/traffic.json, NOT measured in this run: " + tj.get("source", ""))
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "hqt::qc_tiled_kernel<K=5,G=8,FULL=0,UNI=1>", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src,
                "frac_of_nominal_8000": achieved / 8000.0, "kernel_ms_mean": k_ms / max(len(kernel_ms), 1),
                "kernel_share_of_step": k_ms / (ms_step * args.steps) if ms_step > 0 else None,
                "algorithmic_bytes_per_read": BYTES_PER_READ}

    # ---- end to end from pinned host batches
    e2e = None
    if not args.no_e2e:
        EB = args.e2e_batch_reads
        pool = []
        for i in range(args.e2e_pool):
            cfg = mates[i % 2]
            b = ctx.synth_device(cfg, rank * half + (i // 2) * EB, EB)
            hs, ps = q.pinned_array(b.n_bytes + 64)
            hq, pq = q.pinned_array(b.n_bytes + 64)
            hi, pi = q.pinned_array((EB + 1) * 4 + 64)
            hs[:b.n_bytes] = ctx.dev_download(b.seq, b.n_bytes)
            hq[:b.n_bytes] = ctx.dev_download(b.qual, b.n_bytes)
            hi[:(EB + 1) * 4] = ctx.dev_download(b.idx, (EB + 1) * 4)
            ctx.free_batch(b)
            pool.append((EB, ps, pq, pi, b.n_bytes))
        for b in batches:
            ctx.free_batch(b)
        batches = []
        ectx = q.Ctx(local, max_reads=EB, max_bytes=EB * READ_LEN, max_read_len=160, n_slots=2, host_staging=True)
        ectx.set_params(par)
        n_sub = (n_reads + EB - 1) // EB

        # Batches are dealt by TICKET: a rank draws the next batch number when one of its two slots is free, so a GPU
        # whose PCIe uplink is shared (profiles/h2d_probe_n8_r1l.log: with eight ranks copying, GPUs 0-3 get 23.8 GB/s
        # each and GPUs 4-7 35.5) takes fewer batches instead of holding the others up. The job is still
        # world x n_sub batches per step; at N = 1 the counter is local.
        store = None
        if world > 1:
            try:
                store = dist.distributed_c10d._get_default_store()
            except Exception:
                store = None  # no shared counter: every rank takes its own n_sub batches (equal shares)
        total_sub = n_sub * world if store is not None else n_sub
        step_no = [0]

        def e2e_step():
            ectx.counters_reset()
            key = f"hpgq_e2e_ticket_{step_no[0]}"
            step_no[0] += 1
            local_next = [0]

            def draw():
                if store is None:
                    local_next[0] += 1
                    return local_next[0] - 1
                return store.add(key, 1) - 1

            h2d = d2h = 0
            done = 0
            busy = [False, False]
            i = 0
            while True:
                slot = i % 2
                if busy[slot]:
                    r = ectx.wait(slot)
                    busy[slot] = False
                    done += r["n_reads"]
                    d2h += r["n_reads"]
                t = draw()
                if t >= total_sub:
                    break
                nr, ps, pq, pi, nbytes = pool[t % len(pool)]
                ectx.submit_raw(nr, ps, pq, pi, slot)
                busy[slot] = True
                h2d += 2 * nbytes + (nr + 1) * 4
                i += 1
            for slot in ((i % 2), ((i + 1) % 2)):
                if busy[slot]:
                    r = ectx.wait(slot)
                    done += r["n_reads"]
                    d2h += r["n_reads"]
            c = ectx.counters()
            d2h += c.nbytes
            return done, h2d, d2h

        e2e_step()
        barrier()
        e_steps = max(1, min(args.steps, 3))
        t0 = time.perf_counter()
        for _ in range(e_steps):
            done, h2d, d2h = e2e_step()
        barrier()
        dt = (time.perf_counter() - t0) / e_steps
        tt = torch.tensor([dt], device="cuda", dtype=torch.float64)
        job = torch.tensor([done, h2d, d2h], device="cuda", dtype=torch.float64)
        per_rank_reads = [done]
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            g = [torch.zeros_like(job) for _ in range(world)]
            dist.all_gather(g, job)
            per_rank_reads = [int(x[0].item()) for x in g]
            job = torch.stack(g).sum(0)
        dt = float(tt.item())
        job_reads, job_h2d, job_d2h = (float(x) for x in job.tolist())
        e2e = {"value": job_reads / dt, "unit": "reads/s", "h2d_bytes_per_step": int(job_h2d / world),
               "d2h_bytes_per_step": int(job_d2h / world),
               "ms_per_step": dt * 1e3, "h2d_gbs_per_gpu": job_h2d / world / dt / 1e9, "batch_reads": EB,
               "pinned_pool_batches": len(pool), "numa_node_of_rank0": numa_node,
               "dealing": ("ticket per batch (shared counter); bytes per step are the mean over ranks"
                           if store is not None or world == 1 else "equal shares (no shared counter)"),
               "reads_per_rank_last_step": per_rank_reads,
               "timing": "wall clock between device synchronisations, max over ranks"}
        ectx.close()
        # the ceiling of this leg, measured in the same run: pinned H2D with every rank copying at once
        probe = h2d_probe(torch, dist, world)
        e2e["h2d_ceiling_gbs_per_gpu"] = probe
        e2e["h2d_ceiling_gbs"] = sum(probe)
        e2e["frac_of_h2d_ceiling"] = (job_h2d / dt / 1e9) / sum(probe) if sum(probe) > 0 else None
        e2e["bound"] = "host->device PCIe copies (not a kernel, not NCCL): see h2d_ceiling_gbs"
        for _, ps, pq, pi, _ in pool:
            q.free_pinned(ps)
            q.free_pinned(pq)
            q.free_pinned(pi)

    cpu = None
    if rank == 0 and not args.no_cpu:
        import oracle_binding as ob
        cpu = cpu_baseline_block(q, ob, ctx, mates[0], args.cpu_sample_reads)
    for b in batches:
        ctx.free_batch(b)
    batches = []
    ctx.close()
    # ---- the other BASELINE.json configs, after the timed C2 region (every rank takes part: barriers inside)
    configs = None
    if not args.no_configs:
        configs = {}
        table = config_table(q, args.config_scale)
        for name in [x for x in args.configs.split(",") if x]:
            configs[name] = run_config(q, torch, dist, name, table[name], rank, world, local, peak)
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "reads/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8/int64", "data": "synthetic",
            "config": {"workload": "C2: hpg-fastq stats + filter (min length 50, min mean Q20, max N 5) on "
                                   "synthetic 150 bp paired-end reads, two mate streams",
                       "reads_per_gpu": total_reads, "batch_reads": B, "read_len": READ_LEN, "filters": FILTER,
                       "l2_policy": f"inputs ({total_reads * (2 * READ_LEN + 4) / 1e9:.1f} GB per GPU, every batch read once per "
                                    "step) far larger than the 126 MB L2",
                       "sharding": "reads by index across ranks; one sum-allreduce of the counter block per step",
                       "passed_fraction": n_passed / max(total_reads * world if world > 1 else total_reads, 1)},
            "fastq_gbs": value * BYTES_PER_READ / 1e9,
            "roofline": roofline, "clocks": clocks, "gpu_launches": int(launches), "e2e": e2e, "cpu_baseline": cpu,
            "verified_reads": int(verified),
            "verified_how": "rank 0's first batches, one at a time through the GPU and chunk by chunk through the CPU "
                            "oracle after the timed region: every counter word and every mask byte equal",
            "configs": configs,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
