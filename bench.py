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


def cpu_leg(q, ob, ctx, cfg, n_sample, steps, warmup, threads=0):
    """Times the CPU oracle on a bounded sample (reads generated on the GPU when one is there)."""
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=FILTER)
    if ctx is not None:
        b = ctx.synth_device(cfg, 0, n_sample)
        seq = ctx.dev_download(b.seq, b.n_bytes)
        qual = ctx.dev_download(b.qual, b.n_bytes)
        idx = ctx.dev_download(b.idx, (n_sample + 1) * 4).view(np.uint32)
        ctx.free_batch(b)
    else:
        seq, qual, idx = q.synth_host(cfg, 0, n_sample)
    lay = ob.layout(q.Layout, 160)
    # torchrun exports OMP_NUM_THREADS=1: take the core count from the affinity mask, never from OpenMP's default
    cores = len(os.sched_getaffinity(0)) if threads <= 0 else threads
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        ob.process_batch(par, lay, seq, qual, idx, threads=cores, finalize=True)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    return n_sample / (sum(times) / len(times)), cores, sum(times) / len(times)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import __graft_entry__ as g
    import oracle_binding as ob
    q = g.load_package().hpgq
    cfg = q.synth_cfg(READ_LEN)
    n = min(args.cpu_sample_reads, 2_000_000)  # host-generated here: keep the generator cost bounded
    val, cores, dt = cpu_leg(q, ob, None, cfg, n, max(args.steps, 1), min(args.warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "reads/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8/int64", "data": "synthetic",
        "config": {"workload": "C2 stats+filter, 150 bp reads, CPU path", "sample_reads_per_step": n,
                   "filters": FILTER},
        "cpu_baseline": {"value": val, "unit": "reads/s", "cores": cores, "kind": "port",
                         "sample": f"{n} reads x {READ_LEN} bp per step, oracle/hpgq_oracle.c (OpenMP, batch-parallel "
                                   "like the reference's worker pool); the reference's per-read functions are in a "
                                   "submodule absent from its tree, so oracle/_ref holds only its chaos-game code"},
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
    # roofline of the dominant kernel
    peak, peak_src = measured_peak()
    k_ms = sum(m for m, _ in kernel_ms)
    k_reads = sum(r for _, r in kernel_ms)
    achieved = (k_reads * BYTES_PER_READ) / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get("hq_qc_tiled_kernel_bytes_per_launch")
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "hqt::qc_tiled_kernel<K=5,G=8,FULL=0>", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
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
        for _, ps, pq, pi, _ in pool:
            q.free_pinned(ps)
            q.free_pinned(pq)
            q.free_pinned(pi)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        import oracle_binding as ob
        val, cores, dtc = cpu_leg(q, ob, ctx, mates[0], args.cpu_sample_reads, 1, 1)
        cpu = {"value": val, "unit": "reads/s", "cores": cores, "kind": "port",
               "sample": f"{args.cpu_sample_reads} reads x {READ_LEN} bp of the same workload, one pass "
                         f"({dtc:.2f} s), oracle/hpgq_oracle.c with OpenMP on all host threads"}
    ctx.close()
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
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
