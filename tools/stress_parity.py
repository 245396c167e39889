#!/usr/bin/env python
"""Randomised parity stress on the GPU box: random batch shapes and option sets through the C ABI, every counter,
mask, trim record and compacted array compared with the oracle (test infrastructure, like tests/).  Developer tool:
    python tools/stress_parity.py [--cases 60] [--seed 1] [--text | --multi]
--text sends the batches as FASTQ text (random line ends, headers, '+' lines) through hpgq_batch_submit_text and
compares the records composed on the device (hpgq_batch_wait_text) as well.  --multi keeps one context over several
batches of changing shapes on two slots (host and device-resident submits mixed) and checks the accumulated counters.
Prints one line per case and exits non-zero at the first difference."""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

IGNORE = ("SLOW_READS", "BATCHES", "CG_LONG_RUNS", "CG_FIXED_READS", "CG_MOVED_WORDS", "BAD_BATCH")


def rand_batch(rng, n, lmin, lmax, p_n, lower, uniform, hiq):
    lens = np.full(n, lmax) if uniform else rng.integers(lmin, lmax + 1, size=n)
    idx = np.zeros(n + 1, dtype=np.uint32)
    idx[1:] = np.cumsum(lens)
    tot = int(idx[-1])
    seq = rng.choice(np.frombuffer(b"ACGTN", dtype=np.uint8), size=tot, p=[(1 - p_n) / 4] * 4 + [p_n]).astype(np.uint8)
    if lower > 0 and tot:
        m = rng.random(tot) < lower
        seq[m] |= 0x20
    qual = rng.integers(33, 75, size=tot).astype(np.uint8)
    if tot:
        # poly-A / poly-T / poly-G stretches (chaos-game axis runs) and stretches of odd quality bytes
        for _ in range(int(rng.integers(0, 6))):
            s = int(rng.integers(0, tot))
            seq[s:s + int(rng.integers(10, 90))] = rng.choice(np.frombuffer(b"ATG", dtype=np.uint8))
        if hiq:
            for _ in range(int(rng.integers(1, 5))):
                s = int(rng.integers(0, tot))
                qual[s:s + int(rng.integers(1, 40))] = int(rng.choice([20, 126, 200, 255]))
    return seq, qual, idx


def text_case(q, ob, rng, case, seed):
    """One random batch as FASTQ text: parser + QC + (sometimes) the output records composed on the device."""
    kind = rng.choice(["short", "short", "mid", "tiny", "long"])
    lmin, lmax, n = {"short": (int(rng.integers(0, 60)), int(rng.integers(60, 400)), int(rng.integers(1, 30000))),
                     "mid": (int(rng.integers(100, 640)), int(rng.integers(641, 1100)), int(rng.integers(1, 3000))),
                     "long": (int(rng.integers(200, 3000)), int(rng.integers(3000, 9000)), int(rng.integers(1, 300))),
                     "tiny": (0, int(rng.integers(1, 40)), int(rng.integers(1, 300)))}[kind]
    uniform = bool(rng.random() < 0.3)
    seq, qual, idx = rand_batch(rng, n, lmin, lmax, float(rng.choice([0.0, 0.01, 0.05])),
                                float(rng.choice([0.0, 0.0, 0.01])), uniform, False)
    eol = b"\r\n" if rng.random() < 0.3 else b"\n"
    final_eol = bool(rng.random() < 0.7)
    plus_names = bool(rng.random() < 0.3)
    hl = rng.integers(1, 60, size=n)
    hb = rng.choice(np.frombuffer(b"abcXYZ019:/_ @+\t#", dtype=np.uint8), size=int(hl.sum()) + 1).tobytes()
    ho = np.concatenate([[0], np.cumsum(hl)])
    hdr = lambda i: b"@" + hb[ho[i]:ho[i + 1]]
    sb, qb = seq.tobytes(), qual.tobytes()
    parts = []
    for i in range(n):
        h = hdr(i)
        parts.append(h + eol + sb[idx[i]:idx[i + 1]] + eol + (b"+" + h[1:] if plus_names else b"+") + eol + qb[idx[i]:idx[i + 1]] + eol)
    text = b"".join(parts)
    if not final_eol:
        text = text[:-len(eol)]
    flt = dict(min_read_length=int(rng.integers(0, max(lmax // 2, 1))), min_read_quality=int(rng.integers(0, 30)),
               max_N=int(rng.integers(0, max(lmax // 10, 2))))
    edit = dict(left_length=int(rng.integers(0, 40)), min_left_quality=int(rng.integers(10, 38)),
                right_length=int(rng.integers(0, 40)), min_right_quality=int(rng.integers(10, 38)))
    mode = rng.choice(["stats_filter", "edit_text", "filter_text"])
    if mode == "stats_filter":
        work, parkw = q.WORK_STATS | q.WORK_FILTER, dict(filter=flt)
    elif mode == "edit_text":
        work = q.WORK_EDIT | q.WORK_COMPACT | (q.WORK_FILTER if rng.random() < 0.7 else 0)
        parkw = dict(edit=edit, filter=flt if work & q.WORK_FILTER else None)
    else:
        work, parkw = q.WORK_FILTER | q.WORK_COMPACT, dict(filter=flt)
    par = q.make_params(int(work), **parkw)
    Lcap = (max(lmax, 1) + 63) // 64 * 64
    desc = (f"text case {case}: {kind} n={n} L={lmin}-{lmax}{' uniform' if uniform else ''} {mode} eol={eol!r} "
            f"final_eol={final_eol} plus_names={plus_names} bytes={len(text)}")
    with q.Ctx(0, max_reads=n + 4, max_bytes=max(int(idx[-1]), 16) + 64, max_read_len=Lcap) as ctx:
        ctx.set_params(par)
        info = ctx.submit_text(text, 0)
        if info.status != 0 or info.n_reads != n:
            print("FAIL " + desc + f" submit_text status {info.status} reads {info.n_reads} bad record {info.bad_record}")
            return False
        res = ctx.wait_text(0) if work & q.WORK_COMPACT else ctx.wait(0)
        got, lay = ctx.counters(), ctx.layout
    exp, emask, etrim = ob.process_batch(par, lay, seq, qual, idx)
    skip = {q.S[k] for k in IGNORE}
    bad = [int(b) for b in np.nonzero(got != exp)[0] if int(b) not in skip]
    ok = not bad
    if ok and work & q.WORK_COMPACT:
        cp = ob.compact(seq, qual, idx, emask if work & q.WORK_FILTER else None, etrim if work & q.WORK_EDIT else None)
        oseq, oqual, oidx, osrc = cp["out_seq"].tobytes(), cp["out_qual"].tobytes(), cp["out_idx"], cp["out_src"]
        rec = lambda a, b: b"".join(hdr(int(osrc[k])) + b"\n" + oseq[oidx[k]:oidx[k + 1]] + b"\n+\n" + oqual[oidx[k]:oidx[k + 1]] + b"\n"
                                    for k in range(a, b))
        npass = cp["n_passed"] if work & q.WORK_FILTER else n
        ok = res["n_passed"] == npass and res["passed_text"] == rec(0, npass) and res["failed_text"] == rec(npass, n)
    elif ok:
        ok = np.array_equal(res["mask"], emask)
    print(("ok   " if ok else "FAIL ") + desc + (f" first bad words {bad[:6]}" if bad else ""), flush=True)
    if not ok:
        with open(os.path.join(ROOT, "gpurun_out", f"stress_text_fail_{seed}_{case}.fq"), "wb") as f:
            f.write(text)
        print(f"     filter {flt if work & q.WORK_FILTER else None}, edit {edit if work & q.WORK_EDIT else None}")
    return ok


def multi_case(q, ob, rng, case):
    """One context, 3-7 batches of changing shapes over two slots (host and device-resident submits mixed), one option
    set: accumulated counters and every batch's mask / trim / compacted output."""
    import ctypes as C
    nb = int(rng.integers(3, 8))
    mode = rng.choice(["stats", "stats_filter", "filter_compact", "edit", "all_stats"])
    Lmax_ctx = int(rng.choice([120, 300, 700, 5000]))
    flt = dict(min_read_length=int(rng.integers(0, Lmax_ctx // 3)), min_read_quality=int(rng.integers(0, 30)),
               max_N=int(rng.integers(0, max(Lmax_ctx // 10, 2))))
    if rng.random() < 0.3:
        flt.update(max_out_of_quality=int(rng.integers(0, Lmax_ctx)), left_length=int(rng.integers(1, 30)),
                   min_left_quality=int(rng.integers(5, 30)), right_length=int(rng.integers(1, 30)),
                   min_right_quality=int(rng.integers(5, 30)))
    edit = dict(left_length=int(rng.integers(0, 40)), min_left_quality=int(rng.integers(10, 38)),
                right_length=int(rng.integers(0, 40)), min_right_quality=int(rng.integers(10, 38)))
    k2, ck = int(rng.integers(2, 12)), int(rng.integers(2, 11))
    ctxkw, parkw = {}, {}
    if mode == "stats":
        work = q.WORK_STATS
    elif mode == "stats_filter":
        work, parkw = q.WORK_STATS | q.WORK_FILTER, dict(filter=flt)
    elif mode == "filter_compact":
        work, parkw = q.WORK_FILTER | q.WORK_COMPACT, dict(filter=flt)
    elif mode == "edit":
        work = q.WORK_EDIT | q.WORK_COMPACT | (q.WORK_FILTER if rng.random() < 0.7 else 0)
        parkw = dict(edit=edit, filter=flt if work & q.WORK_FILTER else None)
    else:
        work = q.WORK_STATS | q.WORK_FILTER | q.WORK_KMERS | q.WORK_CHAOS
        parkw, ctxkw = dict(filter=flt, k2=k2, cg_k=ck), dict(kmers=True, k2=k2, cg_k=ck)
    par = q.make_params(int(work), **parkw)
    Lcap = (Lmax_ctx + 63) // 64 * 64
    max_reads, max_bytes = 20000, 20000 * min(Lmax_ctx, 400) + 4 * Lmax_ctx
    shapes = []
    for b in range(nb):
        lmax = int(rng.integers(1, Lmax_ctx + 1))
        lmin = int(rng.integers(0, lmax + 1))
        uniform = bool(rng.random() < 0.3)
        n = int(rng.integers(0, max_reads + 1))
        n = max(min(n, max_bytes // max(lmax, 1) - 1), 0)
        if work & (q.WORK_KMERS | q.WORK_CHAOS):
            n = min(n, 3000)
        shapes.append((n, lmin, lmax, uniform, bool(rng.random() < 0.5)))
    desc = f"multi case {case}: {mode} Lctx={Lmax_ctx} k2={k2} ck={ck} batches " + " ".join(
        f"[n={n} L={a}-{b}{'u' if u else ''}{' dev' if d else ''}]" for n, a, b, u, d in shapes)
    ok, why = True, ""
    with q.Ctx(0, max_reads=max_reads, max_bytes=max_bytes, max_read_len=Lcap, n_slots=2, **ctxkw) as ctx:
        ctx.set_params(par)
        lay = ctx.layout
        exp = np.zeros(lay.words, dtype=np.uint64)
        pending = [None, None]

        def retire(slot):
            nonlocal ok, why
            p = pending[slot]
            if p is None:
                return
            pending[slot] = None
            b, dev, seq, qual, idx, emask, etrim = p
            if dev is not None:
                ctx.sync(slot)
                views = q.Result()
                q._ck(ctx.lib.hpgq_batch_device_views(ctx.h, slot, C.byref(views)), "device_views")
                n = len(idx) - 1
                mask = ctx.dev_download(views.mask, n) if n and (work & q.WORK_FILTER) else emask
                ctx.free_batch(dev)
                if not np.array_equal(mask, emask):
                    ok, why = False, f"batch {b}: mask (device views)"
                return
            res = ctx.wait(slot)
            if not np.array_equal(res["mask"], emask):
                ok, why = False, f"batch {b}: mask differs at {np.nonzero(res['mask'] != emask)[0][:6]}"
            elif (work & q.WORK_EDIT) and not np.array_equal(res["trim"], etrim):
                ok, why = False, f"batch {b}: trim"
            elif work & q.WORK_COMPACT:
                cp = ob.compact(seq, qual, idx, emask if work & q.WORK_FILTER else None, etrim if work & q.WORK_EDIT else None)
                for k in ("out_idx", "out_src", "out_seq", "out_qual"):
                    if not np.array_equal(res[k], cp[k]):
                        ok, why = False, f"batch {b}: {k}"
                        break

        for b, (n, lmin, lmax, uniform, dev) in enumerate(shapes):
            slot = b & 1
            retire(slot)
            seq, qual, idx = rand_batch(rng, n, lmin, lmax, float(rng.choice([0.0, 0.01, 0.05])),
                                        float(rng.choice([0.0, 0.0, 0.01])), uniform, bool(rng.random() < 0.2))
            exp, emask, etrim = ob.process_batch(par, lay, seq, qual, idx, counters=exp, finalize=False)
            if dev and n > 0 and not (work & q.WORK_COMPACT):
                d = ctx.upload_batch(seq, qual, idx, max_len=lmax)
                uni = uniform and bool(rng.random() < 0.7)
                ctx.submit_device(d, slot, max_len_hint=None if (uni or rng.random() < 0.5) else 0, uniform=uni)
                pending[slot] = (b, d, seq, qual, idx, emask, etrim)
            else:
                ctx.submit(seq, qual, idx, slot)
                pending[slot] = (b, None, seq, qual, idx, emask, etrim)
        retire(0)
        retire(1)
        got = ctx.counters()
    ob.lib().oracle_finalize(C.byref(lay), exp.ctypes.data_as(C.c_void_p))
    skip = {q.S[k] for k in IGNORE}
    bad = [int(b) for b in np.nonzero(got != exp)[0] if int(b) not in skip]
    if bad and ok:
        names = {v: k for k, v in q.S.items()}
        ok, why = False, "counters: " + ", ".join(f"{names.get(b, b)} got {int(got[b])} exp {int(exp[b])}" for b in bad[:6])
    print(("ok   " if ok else "FAIL ") + desc + ("  -- " + why if why else ""), flush=True)
    if not ok:
        print(f"     filter {flt if work & q.WORK_FILTER else None}, edit {edit if work & q.WORK_EDIT else None}")
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=60)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--text", action="store_true")
    ap.add_argument("--multi", action="store_true")
    args = ap.parse_args()
    if args.multi:
        import __graft_entry__ as g
        import oracle_binding as ob
        q = g.load_package().hpgq
        ob.lib()
        rng = np.random.default_rng(2000 + args.seed)
        for case in range(args.cases):
            if not multi_case(q, ob, rng, case):
                return 1
        print(f"{args.cases} multi cases ok (seed {args.seed})")
        return 0
    if args.text:
        import __graft_entry__ as g
        import oracle_binding as ob
        q = g.load_package().hpgq
        ob.lib()
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        rng = np.random.default_rng(1000 + args.seed)
        for case in range(args.cases):
            if not text_case(q, ob, rng, case, args.seed):
                return 1
        print(f"{args.cases} text cases ok (seed {args.seed})")
        return 0
    import __graft_entry__ as g
    import oracle_binding as ob
    q = g.load_package().hpgq
    ob.lib()
    rng = np.random.default_rng(args.seed)
    skip = {q.S[k] for k in IGNORE}
    for case in range(args.cases):
        kind = rng.choice(["short", "short", "mid", "long", "tiny"])
        lmin, lmax, n = {"short": (int(rng.integers(1, 60)), int(rng.integers(60, 400)), int(rng.integers(1, 40000))),
                         "mid": (int(rng.integers(100, 640)), int(rng.integers(641, 1100)), int(rng.integers(1, 4000))),
                         "long": (int(rng.integers(200, 3000)), int(rng.integers(3000, 12000)), int(rng.integers(1, 600))),
                         "tiny": (0, int(rng.integers(1, 40)), int(rng.integers(0, 300)))}[kind]
        uniform = bool(rng.random() < 0.3) and lmax > 0
        mode = rng.choice(["stats", "stats_filter", "filter_compact", "edit", "kmers", "kmers_chaos", "all_stats"])
        flt = dict(min_read_length=int(rng.integers(0, max(lmax // 2, 1))), min_read_quality=int(rng.integers(0, 30)),
                   max_N=int(rng.integers(0, max(lmax // 10, 2))))
        if rng.random() < 0.3:
            flt.update(max_read_length=int(rng.integers(lmax // 2 + 1, 2 * lmax + 2)), max_read_quality=int(rng.integers(30, 45)))
        if rng.random() < 0.25:
            flt.update(max_out_of_quality=int(rng.integers(0, lmax + 1)), left_length=int(rng.integers(1, 30)),
                       min_left_quality=int(rng.integers(5, 30)), right_length=int(rng.integers(1, 30)),
                       min_right_quality=int(rng.integers(5, 30)))
        edit = dict(left_length=int(rng.integers(0, 40)), min_left_quality=int(rng.integers(10, 38)),
                    right_length=int(rng.integers(0, 40)), min_right_quality=int(rng.integers(10, 38)))
        k2, ck = int(rng.integers(2, 12)), int(rng.integers(2, 11))
        ctxkw, parkw = {}, {}
        if mode == "stats":
            work = q.WORK_STATS
        elif mode == "stats_filter":
            work, parkw = q.WORK_STATS | q.WORK_FILTER, dict(filter=flt)
        elif mode == "filter_compact":
            work, parkw = q.WORK_FILTER | q.WORK_COMPACT, dict(filter=flt)
        elif mode == "edit":
            work = q.WORK_EDIT | q.WORK_COMPACT | (q.WORK_FILTER if rng.random() < 0.7 else 0)
            parkw = dict(edit=edit, filter=flt if work & q.WORK_FILTER else None)
        elif mode == "kmers":
            work, parkw, ctxkw = q.WORK_KMERS, dict(k2=k2), dict(kmers=True, k2=k2)
        elif mode == "kmers_chaos":
            work, parkw, ctxkw = q.WORK_KMERS | q.WORK_CHAOS, dict(k2=k2, cg_k=ck), dict(kmers=True, k2=k2, cg_k=ck)
        else:
            work = q.WORK_STATS | q.WORK_FILTER | q.WORK_KMERS | q.WORK_CHAOS
            parkw, ctxkw = dict(filter=flt, k2=k2, cg_k=ck), dict(kmers=True, k2=k2, cg_k=ck)
        if (work & (q.WORK_KMERS | q.WORK_CHAOS)) and kind == "long":
            n = min(n, 80)
        seq, qual, idx = rand_batch(rng, n, lmin, lmax, float(rng.choice([0.0, 0.01, 0.05, 0.5])),
                                    float(rng.choice([0.0, 0.0, 0.01, 0.3])), uniform, bool(rng.random() < 0.3))
        if (work & q.WORK_COMPACT) and rng.random() < 0.4:
            parkw = dict(parkw, kernel_variant=2)  # decisions + compaction in one pass (hq_fz_kernel) where it applies
            mode += "/one_pass"
        par = q.make_params(int(work), **parkw)
        Lcap = (max(lmax, 1) + 63) // 64 * 64
        desc = f"case {case}: {kind} n={n} L={lmin}-{lmax}{' uniform' if uniform else ''} {mode} k2={k2} ck={ck}"
        if os.environ.get("STRESS_TRACE"):
            print("run  " + desc + f" filter {flt} edit {edit}", flush=True)
            np.savez(os.path.join(ROOT, "gpurun_out", "stress_last.npz"), seq=seq, qual=qual, idx=idx)
        with q.Ctx(0, max_reads=max(n, 1), max_bytes=max(int(idx[-1]), 16), max_read_len=Lcap, **ctxkw) as ctx:
            ctx.set_params(par)
            ctx.submit(seq, qual, idx, 0)
            res = ctx.wait(0)
            got, lay = ctx.counters(), ctx.layout
        exp, emask, etrim = ob.process_batch(par, lay, seq, qual, idx)
        bad = [int(b) for b in np.nonzero(got != exp)[0] if int(b) not in skip]
        ok = not bad and np.array_equal(res["mask"], emask)
        if ok and (work & q.WORK_EDIT):
            ok = np.array_equal(res["trim"], etrim)
        if ok and (work & q.WORK_COMPACT) and n > 0:  # an empty batch has no output arrays
            cp = ob.compact(seq, qual, idx, emask if work & q.WORK_FILTER else None, etrim if work & q.WORK_EDIT else None)
            ok = all(np.array_equal(res[k], cp[k]) for k in ("out_idx", "out_src", "out_seq", "out_qual"))
        print(("ok   " if ok else "FAIL ") + desc + (f" first bad words {bad[:6]}" if bad else ""), flush=True)
        if not ok:
            names = {v: k for k, v in q.S.items()}
            for b in bad[:10]:
                print(f"     word {b} ({names.get(b, 'table')}): got {int(got[b])} exp {int(exp[b])}")
            print(f"     mask differs at {np.nonzero(res['mask'] != emask)[0][:8]}, max quality byte {int(qual.max()) if len(qual) else 0}, "
                  f"filter {flt if work & q.WORK_FILTER else None}, edit {edit if work & q.WORK_EDIT else None}")
            for env in ({"HPGQ_NO_UNI": "1"}, {"HPGQ_NO_POSFUSE": "1"}, {"VARIANT": "1"}):
                os.environ.update({k: v for k, v in env.items() if k != "VARIANT"})
                par2 = q.make_params(int(work), **dict(parkw, kernel_variant=1 if "VARIANT" in env else 0))
                with q.Ctx(0, max_reads=max(n, 1), max_bytes=max(int(idx[-1]), 16), max_read_len=Lcap, **ctxkw) as ctx:
                    ctx.set_params(par2)
                    ctx.submit(seq, qual, idx, 0)
                    r2 = ctx.wait(0)
                    g2 = ctx.counters()
                b2 = [int(b) for b in np.nonzero(g2 != exp)[0] if int(b) not in skip]
                print(f"     with {env}: {len(b2)} bad words, mask equal {np.array_equal(r2['mask'], emask)}")
                for k in env:
                    os.environ.pop(k, None)
            np.savez(os.path.join(ROOT, "gpurun_out", f"stress_fail_{args.seed}_{case}.npz"), seq=seq, qual=qual, idx=idx)
            return 1
    print(f"{args.cases} cases ok (seed {args.seed})")
    return 0


if __name__ == "__main__":
    sys.exit(main())
