"""CPU tests of the C host (hpg-fastq_b200/host/): option surface, FASTQ reader / writer, report writer.
The report writer is checked against tests/report_oracle.py (a Python restatement of src/stats_report.c) on
counter blocks produced by the CPU oracle; the `hpg-fastq` program must fail loudly without a GPU."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import host_binding as hb
import report_oracle as ro
from test_oracle import MICRO, pack


def headers_for(n, tag=b"r"):
    return [b"@" + tag + str(i).encode() + b" len=" + str(i % 7).encode() for i in range(n)]


# ------------------------------------------------------------------ options
@pytest.mark.parametrize("text,exp", [("80,110", (80, 110)), ("20,", (20, -1)), (",40", (-1, 40)), ("35", (35, -1)),
                                      (None, (-1, -1))])
def test_parse_range_forms(text, exp):
    """Accepted forms of parse_range (reference src/commons_fastq.c:31-103)."""
    lo, hi = C.c_int(7), C.c_int(7)
    assert hb.lib().hq_parse_range(C.byref(lo), C.byref(hi), text.encode() if text else None, b"read length range") == 1
    assert (lo.value, hi.value) == exp


@pytest.mark.parametrize("text", ["110,80", "-3,5", "a,5", "5,b"])
def test_parse_range_rejects(text, capfd):
    lo, hi = C.c_int(0), C.c_int(0)
    assert hb.lib().hq_parse_range(C.byref(lo), C.byref(hi), text.encode(), b"read length range") == 0


def test_stats_options_to_filter_tuple(q, tmp_path):
    """12-int tuple in the order of src/stats_fastq.c:209-214; unset -> 0 / 100000 (:431-444)."""
    fq = tmp_path / "a.fq"
    fq.write_bytes(b"@r\nACGT\n+\nIIII\n")
    o = hb.Options("stats", ["-f", str(fq), "-o", str(tmp_path), "--read-length-range", "50,", "--read-quality-range",
                             "20,", "--max-N", "5", "--kmers", "--left-length", "10", "--left-quality-range", "15,45"],
                   fill=True)
    assert o.valid and o.runnable
    p = o.params(q.Params)
    f = p.filter
    assert p.work == q.WORK_STATS | q.WORK_FILTER | q.WORK_KMERS and p.phred == 33
    assert (f.min_read_length, f.max_read_length, f.min_read_quality, f.max_read_quality) == (50, 100000, 20, 100000)
    assert (f.max_out_of_quality, f.left_length, f.min_left_quality, f.max_left_quality) == (100000, 10, 15, 45)
    assert (f.right_length, f.min_right_quality, f.max_right_quality, f.max_N) == (0, 0, 100000, 5)
    o.close()


def test_edit_options_tuples(q, tmp_path):
    """edit: 9-int tuple from the window options, following filter has its windows disabled (src/edit_fastq.c:148-164)."""
    fq = tmp_path / "a.fq"
    fq.write_bytes(b"@r\nACGT\n+\nIIII\n")
    o = hb.Options("edit", ["-f", str(fq), "-o", str(tmp_path), "--left-length", "10", "--left-quality-range", "20,",
                            "--right-length", "8", "--right-quality-range", "25,60", "--read-length-range", "30,"])
    assert o.valid and o.runnable
    p = o.params(q.Params)
    assert p.work == q.WORK_EDIT | q.WORK_COMPACT | q.WORK_FILTER and p.phred == 33
    e, f = p.edit, p.filter
    assert (e.left_length, e.min_left_quality, e.max_left_quality) == (10, 20, 100000)
    assert (e.right_length, e.min_right_quality, e.max_right_quality) == (8, 25, 60)
    assert (f.left_length, f.min_left_quality, f.max_left_quality, f.right_length) == (0, 0, 100000, 0)
    assert (f.min_read_length, f.max_read_length) == (30, 100000)
    o.close()


def test_nothing_to_do_and_bad_options(tmp_path, capfd):
    fq = tmp_path / "a.fq"
    fq.write_bytes(b"@r\nACGT\n+\nIIII\n")
    o = hb.Options("filter", ["-f", str(fq)])
    assert o.valid and not o.runnable  # "Nothing to filter" (src/filter_options.c:191-194)
    o.close()
    o = hb.Options("edit", ["-f", str(fq), "--left-length", "5"])  # window option counts only with its range
    assert o.valid and not o.runnable
    o.close()
    assert not hb.Options("stats", ["-f", str(tmp_path / "missing.fq")]).valid
    assert not hb.Options("stats", ["-f", str(fq), "--quality-encoding", "phred99"]).valid
    assert not hb.Options("stats", ["-f", str(fq), "--no-such-option"]).valid
    assert not hb.Options("filter", ["-f", str(fq), "--kmers"]).valid  # --kmers is a stats option only


# ------------------------------------------------------------------ reader / writer
def synth_reads(q, n, length=(20, 90), seed=7):
    cfg = q.synth_cfg(length, seed=seed)
    return q.synth_host(cfg, 0, n)


@pytest.mark.parametrize("eol,final", [(b"\n", True), (b"\r\n", True), (b"\n", False)])
def test_reader_round_trip(q, tmp_path, eol, final):
    seq, qual, idx = synth_reads(q, 500)
    hdr = headers_for(500)
    path = str(tmp_path / "in.fq")
    hb.write_fastq(path, hdr, seq, qual, idx, eol=eol, final_eol=final)
    r = hb.Reader(path)
    mx, nlines = r.scan()
    assert nlines == 2000 and mx == int(np.max(np.diff(idx.astype(np.int64))))
    data = open(path, "rb").read()
    got_s, got_q, lens, heads = [], [], [], []
    nb = 0
    for s, qq, ix, ho, hl, ml in r.batches(cap_reads=64, cap_bytes=3000):
        nb += 1
        assert ix[0] == 0 and ml == int(np.max(np.diff(ix.astype(np.int64)))) and int(ix[-1]) <= 3000
        got_s.append(s)
        got_q.append(qq)
        lens.extend(np.diff(ix.astype(np.int64)).tolist())
        heads.extend(data[int(o):int(o) + int(l)] for o, l in zip(ho, hl))
    r.close()
    assert nb > 500 // 64  # some batches were cut by the byte capacity
    assert np.array_equal(np.concatenate(got_s), seq) and np.array_equal(np.concatenate(got_q), qual)
    assert lens == np.diff(idx.astype(np.int64)).tolist() and heads == hdr


def test_reader_blank_lines_and_empty_reads(tmp_path):
    path = str(tmp_path / "in.fq")
    open(path, "wb").write(b"\n@a\nACGT\n+a\nIIII\n\n\n@empty\n\n+\n\n@b\nNN\n+\n#!\n\n")
    r = hb.Reader(path)
    (s, qq, ix, ho, hl, ml), = list(r.batches(16, 64))
    r.close()
    assert bytes(s) == b"ACGTNN" and bytes(qq) == b"IIII#!" and ix.tolist() == [0, 4, 4, 6] and ml == 4


@pytest.mark.parametrize("text,what", [(b"@a\nACGT\n+\nIII\n", "quality has 3"), (b"a\nACGT\n+\nIIII\n", "'@'"),
                                       (b"@a\nACGT\n-\nIIII\n", "third line"), (b"@a\nACGT\n+\n", "truncated")])
def test_reader_rejects_malformed(tmp_path, text, what):
    path = str(tmp_path / "bad.fq")
    open(path, "wb").write(text)
    r = hb.Reader(path)
    with pytest.raises(ValueError, match=what):
        list(r.batches(16, 64))
    r.close()


@pytest.mark.parametrize("eol", [b"\n", b"\r\n"])
def test_text_chunks_end_on_record_boundaries(q, tmp_path, eol):
    """hq_reader_peek_text (device-parse path): every chunk is a run of whole records, also when quality lines start
    with '@' or '+' (the boundary test is '@' now and '+' two lines on)."""
    n = 3000
    cfg = q.synth_cfg((20, 90), seed=77)
    seq, qual, idx = q.synth_host(cfg, 0, n)
    qual = qual.copy()
    for i in range(n):  # hostile quality lines: '@' / '+' as first byte of every other read
        if idx[i + 1] > idx[i] and i % 2 == 0:
            qual[idx[i]] = ord("@") if i % 4 == 0 else ord("+")
    hdr = [b"@r%d +x @y" % i for i in range(n)]
    path = str(tmp_path / "c.fq")
    data = hb.write_fastq(path, hdr, seq, qual, idx, eol=eol)
    starts = set()
    pos = 0
    for i in range(n):
        starts.add(pos)
        pos += len(hdr[i]) + 2 * int(idx[i + 1] - idx[i]) + 1 + 4 * len(eol)
    assert pos == len(data)
    for budget in (700, 4096, 50000, len(data), 2 * len(data)):
        r = hb.Reader(path)
        chunks = r.text_chunks(budget)
        r.close()
        assert chunks[0][0] == 0 and sum(c[1] for c in chunks) == len(data)
        for off, ln in chunks:
            assert off in starts and 0 < ln <= max(budget, 1)
            assert off + ln == len(data) or (off + ln) in starts


@pytest.mark.parametrize("final_eol", [True, False])
def test_scan_is_stitched_across_thread_ranges(tmp_path, final_eol):
    """hq_reader_scan splits files above 1 MB over four threads: lines that straddle a range boundary (also one
    much longer than a range) must still be measured whole, '\\r' excluded."""
    rng = np.random.default_rng(3)
    lines = [b"x" * int(n) for n in rng.integers(0, 300, size=20000)]
    lines[7000] = b"y" * 900_000          # longer than a quarter of the file
    lines[12000] = b"z" * 1234 + b"\r"     # CRLF line: the '\\r' does not count
    lines[-1] = b"w" * 77
    data = b"\n".join(lines) + (b"\n" if final_eol else b"")
    assert len(data) > (1 << 20)
    path = str(tmp_path / "scan.txt")
    with open(path, "wb") as f:
        f.write(data)
    r = hb.Reader(path)
    mx, n = r.scan()
    r.close()
    assert mx == 900_000 and n == len(lines)
    # boundaries of the four ranges fall inside lines: move the long line around
    for where in (0, 3000, 19999):
        ls = [b"x" * int(k) for k in rng.integers(0, 300, size=20000)]
        ls[where] = b"q" * 700_001
        with open(path, "wb") as f:
            f.write(b"\r\n".join(ls) + b"\r\n")
        r = hb.Reader(path)
        mx, n = r.scan()
        r.close()
        assert mx == 700_001 and n == 20000


def test_writer_matches_record_layout(q, tmp_path):
    """header, sequence, '+', quality, each newline-terminated, in the order given by src."""
    seq, qual, idx = synth_reads(q, 300, seed=11)
    hdr = headers_for(300, b"w")
    path = str(tmp_path / "in.fq")
    hb.write_fastq(path, hdr, seq, qual, idx)
    r = hb.Reader(path)
    L = hb.lib()
    (s, qq, ix, ho, hl, ml), = list(r.batches(512, 1 << 16))
    order = np.array([5, 3, 299, 0, 17], dtype=np.uint32)
    # a "compacted" result: the selected reads packed back to back
    o_idx = np.zeros(len(order) + 1, np.uint32)
    o_seq, o_qual = [], []
    for k, i in enumerate(order):
        o_seq.append(s[ix[i]:ix[i + 1]])
        o_qual.append(qq[ix[i]:ix[i + 1]])
        o_idx[k + 1] = o_idx[k] + (ix[i + 1] - ix[i])
    o_seq, o_qual = np.concatenate(o_seq), np.concatenate(o_qual)
    b = hb.HostBatch(len(ix) - 1, ml, s.ctypes.data, qq.ctypes.data, ix.ctypes.data, ho.ctypes.data, hl.ctypes.data,
                     len(ix) - 1, len(s))
    w = C.c_void_p()
    out = str(tmp_path / "out.fq")
    assert L.hq_writer_open(out.encode(), C.byref(w)) == 0
    assert L.hq_writer_put(w, L.hq_reader_data(r.h), C.byref(b), o_seq.ctypes.data, o_qual.ctypes.data,
                           o_idx.ctypes.data, order.ctypes.data, 0, len(order)) == 0
    assert L.hq_writer_close(w) == 0
    r.close()
    assert open(out, "rb").read() == ro.fastq_text(hdr, seq, qual, idx, order=order.tolist())


# ------------------------------------------------------------------ report writer
def run_report(q, ob, tmp_path, args, seq, qual, idx, Lcap, name="sample.fq"):
    fq = tmp_path / name
    fq.write_bytes(b"@x\nA\n+\nI\n")  # only its name is used by the report
    outdir = tmp_path / "out"
    outdir.mkdir(exist_ok=True)
    o = hb.Options("stats", ["-f", str(fq), "-o", str(outdir)] + args)
    assert o.valid and o.runnable
    par = o.params(q.Params)
    if par.work & q.WORK_FILTER:
        hb.lib().hq_options_fill_unset(o.ptr)
        par = o.params(q.Params)
    kmers = bool(par.work & q.WORK_KMERS)
    lay = ob.layout(q.Layout, Lcap, kmers=int(kmers))
    blk, _, _ = ob.process_batch(par, lay, seq, qual, idx)
    assert hb.lib().hq_stats_report(C.byref(lay), blk.ctypes.data, o.ptr) == 0
    o.close()
    got = {p.name: p.read_text() for p in outdir.iterdir() if not p.name.endswith(".gnuplot")}
    return got, lay, blk, par, outdir


def test_report_micro_hand_computed(q, ob, tmp_path):
    """The 4-read micro FASTQ of tests/test_oracle.py: summary lines computed by hand from src/stats_report.c."""
    seq, qual, idx = pack(MICRO)
    got, lay, blk, par, _ = run_report(q, ob, tmp_path, [], seq, qual, idx, 8, name="micro.fq")
    rep = ro.Report(lay, blk, "micro.fq")
    assert got == rep.files()
    s = got["micro.fq.summary.txt"]
    n = len(MICRO)
    lens = [len(r[0]) for r in MICRO]
    assert f"Number of reads: {n}\n" in s and "Filter         : Disabled\n" in s
    mean = np.float32(1.0) * np.float32(sum(lens)) / np.float32(n)
    assert f"Read length (min., mean, max.): ({min(lens)}, {float(mean):0.2f}, {max(lens)})\n" in s
    allseq = b"".join(r[0] for r in MICRO).decode()
    tot = sum(allseq.count(c) for c in "ACGTN")
    for c in "ATGCN":
        v = np.float32(100.0) * np.float32(allseq.count(c)) / np.float32(tot)
        assert f"\t{c}: {float(v):0.2f} %\n" in s
    # length histogram: one line per length 1..max with the count (src/stats_report.c:178-180)
    lh = got["micro.fq.length.histogram.data"].splitlines()
    assert lh == [f"{L}\t{lens.count(L)}" for L in range(1, max(lens) + 1)]


@pytest.mark.parametrize("args,length", [([], 100), (["--kmers"], (30, 120)),
                                         (["--read-length-range", "50,", "--read-quality-range", "20,", "--max-N", "5",
                                           "--kmers"], (35, 150)),
                                         (["--quality-encoding", "phred64"], 64)])
def test_report_files_match_restated_reference(q, ob, tmp_path, args, length):
    cfg = q.synth_cfg(length, seed=99)
    seq, qual, idx = q.synth_host(cfg, 0, 4000)
    if "phred64" in args:
        qual = (qual + 31).astype(np.uint8)
    Lcap = 160
    got, lay, blk, par, outdir = run_report(q, ob, tmp_path, args, seq, qual, idx, Lcap)
    opts = {}
    for k in ("read-length-range", "read-quality-range"):
        if "--" + k in args:
            opts[k.replace("-", "_")] = args[args.index("--" + k) + 1]
    if "--max-N" in args:
        opts["max_N"] = int(args[args.index("--max-N") + 1])
    rep = ro.Report(lay, blk, "sample.fq", phred=64 if "phred64" in args else 33, kmers_on="--kmers" in args,
                    filter_on=bool(par.work & q.WORK_FILTER), opts=opts)
    exp = rep.files()
    assert sorted(got) == sorted(exp)
    for name in exp:
        assert got[name] == exp[name], f"{name} differs"
    # every data file has its gnuplot companion (src/stats_report.c:592-654); PNGs need gnuplot: unavailable here
    gp = sorted(p.name for p in outdir.iterdir() if p.name.endswith(".gnuplot"))
    assert "sample.fq.length.histogram.gnuplot" in gp and "sample.fq.nucleotides.gnuplot" in gp
    assert ("sample.fq.kmers.per.nt.gnuplot" in gp) == ("--kmers" in args)
    if "--kmers" in args:
        lines = got["sample.fq.kmers.txt"].splitlines()
        counts = [int(l.split("\t")[1]) for l in lines[1:]]
        assert len(counts) == 1024 and counts == sorted(counts, reverse=True)


# ------------------------------------------------------------------ the program without a GPU
def test_cli_fails_loudly_without_gpu(q, tmp_path):
    """No CPU fallback: on a box without a B200 the program prints the configuration and stops with the library's
    error; on a GPU box this test is skipped (the GPU tests run the program for real)."""
    n = C.c_int(0)
    if q.load_library().hpgq_device_count(C.byref(n)) == 0 and n.value > 0:
        pytest.skip("a GPU is visible")
    fq = tmp_path / "a.fq"
    fq.write_bytes(b"@r\nACGT\n+\nIIII\n")
    r = subprocess.run([hb.CLI, "stats", "-f", str(fq), "-o", str(tmp_path)], capture_output=True, text=True,
                       cwd=str(tmp_path))
    assert r.returncode == 1
    assert "PARAMETERS CONFIGURATION" in r.stdout and "FATAL" in r.stderr
    assert not (tmp_path / "a.fq.summary.txt").exists()
    r = subprocess.run([hb.CLI, "filter", "-f", str(fq)], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 255 and "Nothing to filter" in r.stdout
    r = subprocess.run([hb.CLI], capture_output=True, text=True)
    assert r.returncode == 255 and "Usage:" in r.stdout


def test_paired_end_options(tmp_path):
    """--fq1 / --fq2 (and the --fastq1 / --fastq2 synonyms, old/README:59-60): both mates or none, never with -f."""
    import host_binding as hb
    a, b = tmp_path / "m_1.fq", tmp_path / "m_2.fq"
    a.write_bytes(b"@r\nACGT\n+\nIIII\n")
    b.write_bytes(b"@r\nACGT\n+\nIIII\n")
    for k1, k2 in (("--fq1", "--fq2"), ("--fastq1", "--fastq2")):
        o = hb.Options("filter", [k1, str(a), k2, str(b), "--max-N", "1"])
        assert o.valid and o.runnable
        o.close()
    assert not hb.Options("filter", ["--fq1", str(a), "--max-N", "1"]).valid
    assert not hb.Options("filter", ["--fq1", str(a), "--fq2", str(b), "-f", str(a), "--max-N", "1"]).valid
    assert not hb.Options("stats", ["--fq1", str(a), "--fq2", str(tmp_path / "missing.fq")]).valid
    o = hb.Options("stats", ["--fq", str(a)])
    assert o.valid
    o.close()
