"""Device FASTQ text parser (hpgq_batch_submit_text, csrc/hpgq_parse.cu) against the oracle.

The parser replaces the record scan of fastq_fread_se() (reference call sites src/stats_fastq.c:183,
src/filter_fastq.c:114, src/edit_fastq.c:121): the same text must yield the same SoA batch, i.e. the same counters,
masks, trim records and compacted output as the host-packed batch, and the header spans the writer needs."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FILTER = dict(min_read_length=50, min_read_quality=20, max_N=5)
IGNORE = ("SLOW_READS", "BATCHES", "CG_FIXED_READS", "CG_MOVED_WORDS", "BAD_BATCH")


def make_text(seq, qual, idx, eol=b"\n", final_eol=True, hdr=lambda i: b"@read%d/%d len" % (i, i * 7919 % 1013)):
    parts, hoff, hlen, pos = [], [], [], 0
    n = len(idx) - 1
    sb, qb = seq.tobytes(), qual.tobytes()
    for i in range(n):
        h = hdr(i)
        rec = h + eol + sb[idx[i]:idx[i + 1]] + eol + b"+" + eol + qb[idx[i]:idx[i + 1]] + eol
        hoff.append(pos)
        hlen.append(len(h))
        parts.append(rec)
        pos += len(rec)
    text = b"".join(parts)
    if not final_eol:
        text = text[:-len(eol)]
    return text, np.array(hoff, dtype=np.uint32), np.array(hlen, dtype=np.uint32)


def diff_counters(q, got, exp):
    skip = {q.S[k] for k in IGNORE}
    return [int(i) for i in np.nonzero(got != exp)[0] if int(i) not in skip]


@pytest.mark.parametrize("length,eol,final_eol", [(100, b"\n", True), (150, b"\r\n", True), ((35, 260), b"\n", False),
                                                  ((1, 40), b"\r\n", False)])
def test_text_batch_equals_host_batch(q, ob, length, eol, final_eol):
    n = 30011
    cfg = q.synth_cfg(length)
    seq, qual, idx = q.synth_host(cfg, 5, n)
    text, hoff, hlen = make_text(seq, qual, idx, eol, final_eol)
    Lmax = length if isinstance(length, int) else length[1]
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=FILTER)
    with q.Ctx(0, max_reads=n + 5, max_bytes=int(idx[-1]) + 64, max_read_len=(Lmax + 63) // 64 * 64) as ctx:
        ctx.set_params(par)
        info = ctx.submit_text(text, 0)
        assert info.status == 0, f"status {info.status}, bad record {info.bad_record}"
        assert (info.n_reads, info.n_bases, info.n_lines) == (n, int(idx[-1]), 4 * n)
        assert info.max_len == int(np.diff(idx.astype(np.int64)).max())
        goff, glen = ctx.text_headers(n, 0)
        res = ctx.wait(0)
        got = ctx.counters()
        lay = ctx.layout
    assert np.array_equal(goff, hoff) and np.array_equal(glen, hlen)
    exp, emask, _ = ob.process_batch(par, lay, seq, qual, idx)
    assert np.array_equal(res["mask"], emask)
    assert not diff_counters(q, got, exp)


def test_text_batch_edit_filter_compact(q, ob):
    n = 20000
    cfg = q.synth_cfg(150, short_pct=3)
    seq, qual, idx = q.synth_host(cfg, 0, n)
    text, _, _ = make_text(seq, qual, idx)
    par = q.make_params(q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT, filter=FILTER,
                        edit=dict(left_length=10, min_left_quality=20, right_length=10, min_right_quality=20))
    with q.Ctx(0, max_reads=n, max_bytes=int(idx[-1]) + 64, max_read_len=192) as ctx:
        ctx.set_params(par)
        info = ctx.submit_text(text, 0)
        assert info.status == 0
        res = ctx.wait(0)
        lay = ctx.layout
    _, emask, etrim = ob.process_batch(par, lay, seq, qual, idx)
    exp = ob.compact(seq, qual, idx, emask, etrim)
    assert np.array_equal(res["mask"], emask)
    assert np.array_equal(res["trim"], etrim)
    assert res["n_passed"] == exp["n_passed"]
    for k in ("out_seq", "out_qual", "out_idx", "out_src"):
        assert np.array_equal(res[k], exp[k]), k


def test_text_batches_reuse_a_slot_and_accumulate(q, ob):
    cfg = q.synth_cfg(100)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=FILTER)
    n = 9000
    with q.Ctx(0, max_reads=n, max_bytes=n * 100, max_read_len=128) as ctx:
        ctx.set_params(par)
        lay = ctx.layout
        exp = None
        for b in range(3):
            seq, qual, idx = q.synth_host(cfg, b * n, n - 17 * b)
            text, _, _ = make_text(seq, qual, idx)
            info = ctx.submit_text(text, b % 2)
            assert info.status == 0 and info.n_reads == n - 17 * b
            ctx.wait(b % 2)
            exp, _, _ = ob.process_batch(par, lay, seq, qual, idx, counters=exp, finalize=False)
        got = ctx.counters()
    import ctypes as C
    ob.lib().oracle_finalize(C.byref(lay), exp.ctypes.data_as(C.c_void_p))
    assert not diff_counters(q, got, exp)


def test_malformed_chunks_are_reported_not_submitted(q):
    cfg = q.synth_cfg(60)
    n = 1000
    seq, qual, idx = q.synth_host(cfg, 0, n)
    text, _, _ = make_text(seq, qual, idx)
    par = q.make_params(q.WORK_STATS)
    with q.Ctx(0, max_reads=n, max_bytes=n * 60, max_read_len=64) as ctx:
        ctx.set_params(par)
        # a blank line between two records
        cut = text.index(b"@read500/")
        info = ctx.submit_text(text[:cut] + b"\n" + text[cut:], 0)
        assert info.status & q.TEXT_LINES
        # record 321 lost its '+'
        lines = text.split(b"\n")
        bad = list(lines)
        bad[4 * 321 + 2] = b"-"
        info = ctx.submit_text(b"\n".join(bad), 0)
        assert info.status == q.TEXT_RECORD and info.bad_record == 321
        # record 7: quality one byte short; record 900: header without '@'
        bad = list(lines)
        bad[4 * 7 + 3] = bad[4 * 7 + 3][:-1]
        bad[4 * 900] = b"read900"
        info = ctx.submit_text(b"\n".join(bad), 0)
        assert info.status == q.TEXT_RECORD and info.bad_record == 7
        # the slot is still free and the counters untouched: a good chunk goes through afterwards
        info = ctx.submit_text(text, 0)
        assert info.status == 0 and info.n_reads == n
        ctx.wait(0)
        got = ctx.counters()
        assert int(got[q.S["NUM_READS"]]) == n


def test_chunk_over_the_ctx_limits(q):
    cfg = q.synth_cfg(40)
    seq, qual, idx = q.synth_host(cfg, 0, 600)
    text, _, _ = make_text(seq, qual, idx)
    with q.Ctx(0, max_reads=500, max_bytes=600 * 40, max_read_len=64) as ctx:
        ctx.set_params(q.make_params(q.WORK_STATS))
        info = ctx.submit_text(text, 0)
        assert info.status & q.TEXT_TOO_MANY and info.n_reads == 600
    with q.Ctx(0, max_reads=1000, max_bytes=4096, max_read_len=64) as ctx:
        ctx.set_params(q.make_params(q.WORK_STATS))
        info = ctx.submit_text(text, 0)  # 24 000 bases into a ctx sized for 4 096
        assert info.status == q.TEXT_TOO_MANY and info.n_bases == 600 * 40


def test_bytes_that_look_like_line_breaks(q, ob):
    # '\v' (0x0B) and '\t' next to real line breaks must not be counted as breaks (exact per-byte compare)
    seq = np.frombuffer(b"ACGTNACGTN" * 4, dtype=np.uint8).copy()
    qual = np.frombuffer(b"IIII\x0bIIII\x0b" * 4, dtype=np.uint8).copy()
    idx = np.array([0, 10, 20, 30, 40], dtype=np.uint32)
    text, hoff, hlen = make_text(seq, qual, idx, hdr=lambda i: b"@h\x0b\x09%d" % i)
    par = q.make_params(q.WORK_STATS)
    with q.Ctx(0, max_reads=16, max_bytes=4096, max_read_len=64) as ctx:
        ctx.set_params(par)
        info = ctx.submit_text(text, 0)
        assert info.status == 0 and info.n_reads == 4 and info.n_lines == 16
        goff, glen = ctx.text_headers(4, 0)
        ctx.wait(0)
        got = ctx.counters()
        lay = ctx.layout
    assert np.array_equal(goff, hoff) and np.array_equal(glen, hlen)
    exp, _, _ = ob.process_batch(par, lay, seq, qual, idx)
    assert not diff_counters(q, got, exp)


@pytest.mark.parametrize("eol", [b"\n", b"\r\n"])
@pytest.mark.parametrize("work_filter", [True, False])
def test_output_records_composed_on_the_device(q, ob, eol, work_filter):
    """hpgq_batch_wait_text: passed.fq / failed.fq (edit.fq) bytes straight from the device = the writer's layout
    (header of the source read, sequence, '+', quality, LF-terminated) over the oracle's compacted batch."""
    n = 12345
    cfg = q.synth_cfg((20, 150), short_pct=2)
    seq, qual, idx = q.synth_host(cfg, 3, n)
    hdr = lambda i: b"@r%d %s" % (i, b"x" * (i % 23))
    text, _, _ = make_text(seq, qual, idx, eol=eol, hdr=hdr)
    work = q.WORK_EDIT | q.WORK_COMPACT | (q.WORK_FILTER if work_filter else 0)
    par = q.make_params(work, filter=FILTER if work_filter else None,
                        edit=dict(left_length=10, min_left_quality=20, right_length=10, min_right_quality=20))
    with q.Ctx(0, max_reads=n, max_bytes=int(idx[-1]) + 64, max_read_len=192) as ctx:
        ctx.set_params(par)
        info = ctx.submit_text(text, 0)
        assert info.status == 0
        res = ctx.wait_text(0)
        lay = ctx.layout
    _, emask, etrim = ob.process_batch(par, lay, seq, qual, idx)
    exp = ob.compact(seq, qual, idx, emask if work_filter else None, etrim)
    oseq, oqual, oidx, osrc = exp["out_seq"].tobytes(), exp["out_qual"].tobytes(), exp["out_idx"], exp["out_src"]

    def records(a, b):
        return b"".join(hdr(int(osrc[k])) + b"\n" + oseq[oidx[k]:oidx[k + 1]] + b"\n+\n" + oqual[oidx[k]:oidx[k + 1]] + b"\n"
                        for k in range(a, b))

    npass = exp["n_passed"] if work_filter else n
    assert res["n_passed"] == npass and res["n_reads"] == n
    assert res["passed_text"] == records(0, npass)
    assert res["failed_text"] == records(npass, n)
    assert res["n_edited"] == int((etrim[:, 1] != np.diff(idx.astype(np.int64))).sum())


@pytest.mark.parametrize("piece_mb,out_write", [(1, False), (1, True), (32, False)])
def test_text_from_a_file_to_files_through_the_pinned_rings(q, ob, tmp_path, monkeypatch, piece_mb, out_write):
    """hpgq_batch_submit_text_fd + hpgq_batch_text_launch + hpgq_batch_wait_text_fd: the chunk is read from the file in
    pieces of the pinned ring (1 MB pieces here: many pieces, the ring wraps several times), and the output records go
    to two descriptors, appended after what they already hold -- the same bytes as hpgq_batch_wait_text returns."""
    monkeypatch.setenv("HPGQ_PIECE_MB", str(piece_mb))
    monkeypatch.setenv("HPGQ_RING", "3")
    if out_write:
        monkeypatch.setenv("HPGQ_OUT_WRITE", "1")
    import os
    n = 40000
    cfg = q.synth_cfg((20, 150), short_pct=2)
    seq, qual, idx = q.synth_host(cfg, 11, n)
    text, _, _ = make_text(seq, qual, idx)
    src = tmp_path / "in.fq"
    lead = b"@skipped\nAC\n+\nII\n"
    src.write_bytes(lead + text)
    par = q.make_params(q.WORK_EDIT | q.WORK_COMPACT | q.WORK_FILTER, filter=FILTER,
                        edit=dict(left_length=10, min_left_quality=20, right_length=10, min_right_quality=20))
    with q.Ctx(0, max_reads=n, max_bytes=int(idx[-1]) + 64, max_read_len=192) as ctx:
        ctx.set_params(par)
        ctx.reserve_text(len(text), with_output=True)
        info = ctx.submit_text(text, 1)
        assert info.status == 0
        ref = ctx.wait_text(1)
        fd = os.open(src, os.O_RDONLY)
        fp = os.open(tmp_path / "passed.fq", os.O_RDWR | os.O_CREAT | os.O_TRUNC, 0o644)
        ff = os.open(tmp_path / "failed.fq", os.O_RDWR | os.O_CREAT | os.O_TRUNC, 0o644)
        os.write(fp, b"already here\n")
        try:
            for rep in range(2):  # twice: the second batch is appended behind the first
                ctx.submit_text_fd(fd, len(lead), len(text), 0)
                info = ctx.text_launch(0)
                assert info.status == 0 and info.n_reads == n
                res = ctx.wait_text_fd(fp, ff, 0)
                assert res["n_passed"] == ref["n_passed"] and res["n_edited"] == ref["n_edited"]
                assert res["passed_bytes"] == len(ref["passed_text"]) and res["failed_bytes"] == len(ref["failed_text"])
        finally:
            for d in (fd, fp, ff):
                os.close(d)
    assert len(ref["passed_text"]) > (2 << 20) and len(ref["failed_text"]) > 0
    assert (tmp_path / "passed.fq").read_bytes() == b"already here\n" + ref["passed_text"] * 2
    assert (tmp_path / "failed.fq").read_bytes() == ref["failed_text"] * 2
