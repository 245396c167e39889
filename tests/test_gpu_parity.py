"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.
Bit-exact for every integer output (counters, histograms, masks, trim coordinates, compacted bytes)."""
import numpy as np
import pytest

from test_oracle import (EDIT_OPTS, EDIT_READS, EDIT_TRIM, FILTER_MASK, FILTER_OPTS, FILTER_READS, MICRO, pack,
                         rand_batch)

pytestmark = pytest.mark.gpu

IGNORE = ("SLOW_READS", "BATCHES", "CG_LONG_RUNS", "CG_FIXED_READS", "CG_MOVED_WORDS", "BAD_BATCH")


def diff_report(q, lay, got, exp):
    bad = np.nonzero(got != exp)[0]
    ign = {q.S[n] for n in IGNORE}
    bad = [int(b) for b in bad if int(b) not in ign]
    if not bad:
        return ""
    out = []
    for b in bad[:12]:
        sec = max(i for i in range(q.SEC_COUNT) if int(lay.offset[i]) <= b)
        out.append(f"{q.SEC_NAMES[sec]}[{b - int(lay.offset[sec])}] got {int(got[b])} exp {int(exp[b])}")
    return f"{len(bad)} words differ: " + "; ".join(out)


def run_gpu(q, par, seq, qual, idx, Lcap, kmers=False, k2=0, cg_k=0):
    n = len(idx) - 1
    with q.Ctx(0, max_reads=max(n, 1), max_bytes=max(int(idx[-1]), 16), max_read_len=Lcap, kmers=kmers, k2=k2,
               cg_k=cg_k) as ctx:
        ctx.set_params(par)
        ctx.submit(seq, qual, idx, 0)
        res = ctx.wait(0)
        return ctx.counters(), res, ctx.layout, ctx.launch_count()


def check(q, ob, par, seq, qual, idx, Lcap, **kw):
    got, res, lay, launches = run_gpu(q, par, seq, qual, idx, Lcap, **kw)
    exp, emask, etrim = ob.process_batch(par, lay, seq, qual, idx)
    assert launches > 0 or len(idx) == 1  # an empty batch launches nothing
    assert np.array_equal(res["mask"], emask), f"mask differs at {np.nonzero(res['mask'] != emask)[0][:10]}"
    if par.work & q.WORK_EDIT:
        assert np.array_equal(res["trim"], etrim)
        assert res["n_edited"] == int(exp[q.S["NUM_EDITED"]])
    msg = diff_report(q, lay, got, exp)
    assert not msg, msg
    return got, res, exp, emask, etrim


@pytest.mark.parametrize("variant", [0, 1])
def test_micro_vectors(q, ob, variant):
    seq, qual, idx = pack(MICRO)
    check(q, ob, q.make_params(q.WORK_STATS, kernel_variant=variant), seq, qual, idx, 8)
    seq, qual, idx = pack(FILTER_READS)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=FILTER_OPTS, kernel_variant=variant)
    _, res, *_ = check(q, ob, par, seq, qual, idx, 8)
    assert list(res["mask"]) == FILTER_MASK


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("L", [100, 150, 250, 36, 7])
def test_stats_filter_fixed_length(q, ob, L, variant):
    cfg = q.synth_cfg(L)
    n = 60000 if variant == 0 else 20000
    seq, qual, idx = q.synth_host(cfg, 0, n)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, kernel_variant=variant,
                        filter=dict(min_read_length=min(50, L), min_read_quality=20, max_N=5))
    got, res, exp, *_ = check(q, ob, par, seq, qual, idx, 256)
    assert 0 < res["n_passed"] < n
    check(q, ob, q.make_params(q.WORK_STATS, kernel_variant=variant), seq, qual, idx, 256)


def test_stats_filter_full_options(q, ob):
    cfg = q.synth_cfg(150, short_pct=5)
    seq, qual, idx = q.synth_host(cfg, 7, 50000)
    flt = dict(min_read_length=60, max_read_length=150, min_read_quality=22, max_read_quality=39,
               max_out_of_quality=60, left_length=10, min_left_quality=30, max_left_quality=41, right_length=12,
               min_right_quality=15, max_right_quality=40, max_N=3)
    for variant in (0, 1):
        par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=flt, kernel_variant=variant)
        _, res, *_ = check(q, ob, par, seq, qual, idx, 160)
        assert 0 < res["n_passed"] < 50000
    par = q.make_params(q.WORK_FILTER, filter=flt)
    check(q, ob, par, seq, qual, idx, 160)


def test_ragged_and_edge_lengths(q, ob):
    rng = np.random.default_rng(5)
    seq, qual, idx = rand_batch(rng, 30000, 0, 130, p_n=0.03)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_length=20, min_read_quality=18, max_N=4,
                                                                  left_length=200, min_left_quality=19))
    check(q, ob, par, seq, qual, idx, 130)
    # empty batch, single read, all-empty reads
    for reads in ([], [(b"ACGT", b"IIII")], [(b"", b"")] * 5):
        s, ql, ix = pack(reads)
        check(q, ob, q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(max_N=1)), s, ql, ix, 16)
    # maximum length of the tiled kernel (640) and just beyond it (falls to the warp-per-read kernel)
    for L in (640, 641, 3000):
        seq, qual, idx = q.synth_host(q.synth_cfg(L), 0, 700)
        check(q, ob, q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_quality=20)), seq, qual, idx, L)


def test_lowercase_and_foreign_bytes_take_exact_path(q, ob):
    cfg = q.synth_cfg(100, lowercase_ppm=2000)
    seq, qual, idx = q.synth_host(cfg, 0, 40000)
    seq = seq.copy()
    seq[12345] = ord("R")
    seq[777] = 0xC3
    qual = qual.copy()
    qual[4321] = 200
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_quality=20, max_read_quality=40,
                                                                  max_out_of_quality=30, max_N=5))
    got, *_ = check(q, ob, par, seq, qual, idx, 128)
    assert got[q.S["SLOW_READS"]] > 0


@pytest.mark.parametrize("grid_cap", [None, 20, 7])
@pytest.mark.parametrize("lens", [(25, 114), (0, 60), (130, 250), (100, 100), (151, 151)])
def test_foreign_bytes_in_ragged_reads_every_grid(q, ob, lens, grid_cap, monkeypatch):
    """Ragged reads with 1 % lower-case bytes: most warps of the general tiled kernel take the exact path in most
    iterations.  Found by tools/stress_parity.py: with a per-byte if/else chain in that path the warp's cadence
    counter (a uniform register) ran ahead and per-warp records were lost or counted twice; which reads depended
    on the tile -> CTA assignment, hence the grid sizes."""
    if grid_cap:
        monkeypatch.setenv("HPGQ_GRID_CAP", str(grid_cap))
    rng = np.random.default_rng(31)
    seq, qual, idx = rand_batch(rng, 30340, lens[0], lens[1], p_n=0.05, lower=0.01)
    flt = dict(min_read_length=min(lens[0] + 11, lens[1]), min_read_quality=0, max_N=max(lens[1] // 20, 1))
    for work in (q.WORK_FILTER, q.WORK_STATS | q.WORK_FILTER, q.WORK_STATS):
        par = q.make_params(work, filter=flt if work & q.WORK_FILTER else None)
        got, res, exp, emask, _ = check(q, ob, par, seq, qual, idx, (lens[1] + 63) // 64 * 64)
        if work & q.WORK_FILTER:
            assert res["n_passed"] + res["n_failed"] == 30340 and 0 < res["n_passed"] < 30340
        if lens[0] == lens[1]:  # the aligned-word loop (device-resident batch stated as uniform) has the same path
            _device_check(q, ob, par, seq, qual, idx, (lens[1] + 63) // 64 * 64, uniform=True)
    assert got[q.S["SLOW_READS"]] > 0


def test_phred64(q, ob):
    cfg = q.synth_cfg(100)
    seq, qual, idx = q.synth_host(cfg, 0, 20000)
    qual = (qual + 31).astype(np.uint8)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, phred=64, filter=dict(min_read_quality=20, max_N=5))
    check(q, ob, par, seq, qual, idx, 128)


def test_multi_batch_accumulation_and_slots(q, ob):
    cfg = q.synth_cfg(150)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_length=50, min_read_quality=20, max_N=5))
    n, nb = 30000, 5
    with q.Ctx(0, max_reads=n, max_bytes=n * 150, max_read_len=160, n_slots=2) as ctx:
        ctx.set_params(par)
        lay = ctx.layout
        exp = np.zeros(lay.words, dtype=np.uint64)
        masks = {}
        for b in range(nb):
            slot = b % 2
            if b >= 2:
                masks[b - 2] = ctx.wait(slot)["mask"]
            seq, qual, idx = q.synth_host(cfg, b * n, n)
            ctx.submit(seq, qual, idx, slot)
            exp, emask, _ = ob.process_batch(par, lay, seq, qual, idx, counters=exp, finalize=False)
            masks[("e", b)] = emask
        for b in range(nb - 2, nb):
            masks[b] = ctx.wait(b % 2)["mask"]
        got = ctx.counters()
    ob.lib().oracle_finalize(__import__("ctypes").byref(lay), exp.ctypes.data_as(__import__("ctypes").c_void_p))
    for b in range(nb):
        assert np.array_equal(masks[b], masks[("e", b)])
    msg = diff_report(q, lay, got, exp)
    assert not msg, msg


def test_device_resident_generator_matches_host(q, ob):
    cfg = q.synth_cfg(150, short_pct=2)
    n = 50000
    seq, qual, idx = q.synth_host(cfg, 123, n)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_length=50, min_read_quality=20, max_N=5))
    with q.Ctx(0, max_reads=n, max_bytes=n * 150, max_read_len=160, host_staging=False) as ctx:
        ctx.set_params(par)
        b = ctx.synth_device(cfg, 123, n)
        assert b.n_bytes == int(idx[-1])
        assert np.array_equal(ctx.dev_download(b.seq, b.n_bytes), seq)
        assert np.array_equal(ctx.dev_download(b.qual, b.n_bytes), qual)
        for hint in (None, 0):  # with the caller's max-length hint and with the on-device measurement
            ctx.counters_reset()
            ctx.submit_device(b, 0, max_len_hint=hint)
            ctx.sync(0)
            got = ctx.counters()
            assert ctx.last_kernel_ms(0) > 0
            exp, *_ = ob.process_batch(par, ctx.layout, seq, qual, idx)
            msg = diff_report(q, ctx.layout, got, exp)
            assert not msg, msg
        ctx.free_batch(b)


@pytest.mark.parametrize("no_fuse", [None, 2, 1], ids=["one_pass", "one_pass_register_decisions", "two_pass"])
def test_edit_trim_filter_compact(q, ob, no_fuse, monkeypatch):
    """Both compaction paths: the one-pass kernel (decisions + look-back + output from the staged tile) and the
    two-pass path (decision kernel, scan, gather; HPGQ_NO_FUSE)."""
    if no_fuse != 1:
        monkeypatch.setenv("HPGQ_FUSE", "1")
    if no_fuse == 2:
        monkeypatch.setenv("HPGQ_FZ_SLOW", "1")
    seq, qual, idx = pack(EDIT_READS)
    par = q.make_params(q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT, edit=EDIT_OPTS, filter=dict(min_read_length=3))
    _, res, _, emask, etrim = check(q, ob, par, seq, qual, idx, 8)
    assert [tuple(int(x) for x in t) for t in res["trim"]] == EDIT_TRIM
    assert bytes(res["out_seq"]) == b"GTACACGTAC"
    cfg = q.synth_cfg(150, short_pct=3)
    seq, qual, idx = q.synth_host(cfg, 0, 80000)
    for work, flt in ((q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT, dict(min_read_length=50, min_read_quality=20, max_N=5)),
                      (q.WORK_EDIT | q.WORK_COMPACT, None), (q.WORK_FILTER | q.WORK_COMPACT, dict(min_read_quality=25))):
        par = q.make_params(work, edit=dict(left_length=10, min_left_quality=34, right_length=10, min_right_quality=20),
                            filter=flt)
        _, res, _, emask, etrim = check(q, ob, par, seq, qual, idx, 160)
        cp = ob.compact(seq, qual, idx, emask if work & q.WORK_FILTER else None, etrim if work & q.WORK_EDIT else None)
        assert res["n_passed"] == cp["n_passed"] and res["passed_bytes"] == cp["passed_bytes"]
        assert res["failed_bytes"] == cp["failed_bytes"]
        for k in ("out_idx", "out_src", "out_seq", "out_qual"):
            assert np.array_equal(res[k], cp[k]), k


def test_kmers_and_chaos_game(q, ob):
    seq, qual, idx = pack([(b"ACGTANGTACGT", b"I" * 12)])
    check(q, ob, q.make_params(q.WORK_KMERS, k2=3), seq, qual, idx, 16, kmers=True, k2=3)
    cfg = q.synth_cfg(250, cap_axis_runs=1)
    seq, qual, idx = q.synth_host(cfg, 0, 20000)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER | q.WORK_KMERS | q.WORK_CHAOS, k2=11, cg_k=10,
                        filter=dict(min_read_quality=20, max_N=5))
    got, *_ = check(q, ob, par, seq, qual, idx, 256, kmers=True, k2=11, cg_k=10)
    assert got[q.S["CG_LONG_RUNS"]] == 0
    # lower-case bytes: literal positional quality window of old/chaos_game.c:259
    rng = np.random.default_rng(11)
    seq, qual, idx = rand_batch(rng, 3000, 0, 90, p_n=0.02, lower=0.01)
    par = q.make_params(q.WORK_KMERS | q.WORK_CHAOS, k2=6, cg_k=5)
    check(q, ob, par, seq, qual, idx, 96, kmers=True, k2=6, cg_k=5)
    # a long axis run is reported (float and integer chaos-game indices may differ there)
    seq, qual, idx = pack([(b"T" * 70, b"I" * 70), (b"CCCCCCCCCCCCACGT", b"I" * 16)])
    got, _, lay, _ = run_gpu(q, q.make_params(q.WORK_CHAOS, cg_k=10), seq, qual, idx, 96, cg_k=10)
    assert got[q.S["CG_LONG_RUNS"]] > 0


def test_checksum_of_checksums_full_shape(q, ob):
    """Size-independent property at a bench-sized batch: column sums of the per-position tables equal the
    length histogram's suffix sums, base totals add up to the passed bases."""
    cfg = q.synth_cfg(150)
    n = 2_000_000
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_length=50, min_read_quality=20, max_N=5))
    with q.Ctx(0, max_reads=n, max_bytes=n * 150, max_read_len=160, host_staging=False) as ctx:
        ctx.set_params(par)
        b = ctx.synth_device(cfg, 0, n)
        ctx.submit_device(b, 0)
        ctx.sync(0)
        c = ctx.counters()
        lay = ctx.layout
        ctx.free_batch(b)
    S = q.S
    assert c[S["NUM_PASSED"]] + c[S["NUM_FAILED"]] == n and c[S["NUM_READS"]] == c[S["NUM_PASSED"]]
    lh = lay.section(c, "LEN_HIST").astype(np.int64)
    count_per_pos = lh[::-1].cumsum()[::-1][1:]  # reads with length > p
    tot = sum(lay.section(c, k).astype(np.int64) for k in ("POS_A", "POS_C", "POS_G", "POS_T", "POS_N"))
    assert np.array_equal(tot, count_per_pos[:len(tot)])
    assert sum(int(c[S[k]]) for k in ("NUM_A", "NUM_C", "NUM_G", "NUM_T", "NUM_N")) == int(c[S["ACC_LENGTH"]])
    assert int(lay.section(c, "SUMQ_BY_LEN").sum()) == int(lay.section(c, "POS_ACCQ").sum())
    assert int(lay.section(c, "QUAL_HIST").sum()) == int(c[S["NUM_READS"]]) == int(lay.section(c, "GC_HIST").sum())


def test_long_reads_group_and_position_kernels(q, ob):
    """Config C5 shape (variable 1-20 kb reads): the vectorised group kernel (decisions, per-read histograms) and the
    column-owner position kernel (per-position tables) against the oracle, with window filters, foreign bytes and a
    quality byte >= 128; plus a mixed batch where a few long reads sit among short ones."""
    cfg = q.synth_cfg((1000, 20000), seed=5)
    seq, qual, idx = q.synth_host(cfg, 0, 1500)
    seq, qual = seq.copy(), qual.copy()
    seq[1000:1010] = np.frombuffer(b"acgtnRYKM.", dtype=np.uint8)
    qual[50000] = 200
    for flt in (dict(min_read_length=2000, min_read_quality=30, max_N=80),
                dict(min_read_quality=20, max_read_quality=38, max_out_of_quality=4000, left_length=300,
                     min_left_quality=30, right_length=500, min_right_quality=22, max_right_quality=40)):
        par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=flt)
        _, res, _, emask, _ = check(q, ob, par, seq, qual, idx, 20032)
        assert 0 < int(emask.sum()) < len(emask)
    check(q, ob, q.make_params(q.WORK_STATS), seq, qual, idx, 20032)
    # a batch holding a read longer than hpgq_limits_t.max_read_len is refused loudly when statistics are on
    with pytest.raises(q.HpgqError, match="max_read_len"):
        run_gpu(q, q.make_params(q.WORK_STATS), seq, qual, idx, 9984)
    rng = np.random.default_rng(3)
    s1, q1, i1 = rand_batch(rng, 4000, 0, 300, p_n=0.02)
    s2, q2, i2 = rand_batch(rng, 30, 5000, 9000, p_n=0.01)
    seq = np.concatenate([s1, s2])
    qual = np.concatenate([q1, q2])
    idx = np.concatenate([i1, i2[1:] + i1[-1]]).astype(np.uint32)
    check(q, ob, q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_length=100, max_N=3)), seq, qual, idx, 9024)


@pytest.mark.parametrize("no_fuse", [None, 2, 1], ids=["one_pass", "one_pass_register_decisions", "two_pass"])
@pytest.mark.parametrize("lmax", [20, 60, 110, 300, 700, 2500])
def test_edit_group_kernel_all_lane_widths(q, ob, lmax, no_fuse, monkeypatch):
    """EDIT + FILTER (+ filter windows) on ragged reads: every lanes-per-read instantiation of the group kernel and
    of the one-pass kernel (reads of 2500 bp take the two-pass path either way)."""
    if no_fuse != 1:
        monkeypatch.setenv("HPGQ_FUSE", "1")
    if no_fuse == 2:
        monkeypatch.setenv("HPGQ_FZ_SLOW", "1")
    rng = np.random.default_rng(lmax)
    seq, qual, idx = rand_batch(rng, 6000, 0, lmax, p_n=0.03)
    qual = rng.integers(35, 74, size=len(qual)).astype(np.uint8)
    par = q.make_params(q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT,
                        edit=dict(left_length=7, min_left_quality=22, max_left_quality=25, right_length=9,
                                  min_right_quality=19, max_right_quality=24),
                        filter=dict(min_read_length=5, min_read_quality=20, max_read_quality=22, max_N=4,
                                    max_out_of_quality=lmax // 2 + 3, left_length=4, min_left_quality=15,
                                    max_left_quality=26, right_length=5, min_right_quality=14, max_right_quality=27))
    _, res, _, emask, etrim = check(q, ob, par, seq, qual, idx, lmax + 8)
    assert 0 < int(emask.sum()) < len(emask)
    cp = ob.compact(seq, qual, idx, emask, etrim)
    for k in ("out_idx", "out_src", "out_seq", "out_qual"):
        assert np.array_equal(res[k], cp[k]), k


@pytest.mark.parametrize("variant", [0, 1])
def test_committed_golden_fixture(q, variant):
    """The CUDA path against tests/golden/oracle_golden.json directly (no oracle in the loop): the section sums and
    position-weighted checksums committed by tests/golden/make_golden.py."""
    import json
    import os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "oracle_golden.json")) as f:
        g = json.load(f)
    seq, qual, idx = q.synth_host(q.synth_cfg(100), 0, g["n"])
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER | q.WORK_KMERS | q.WORK_CHAOS, k2=7, cg_k=6,
                        kernel_variant=variant, filter=dict(min_read_length=50, min_read_quality=20, max_N=5))
    got, res, lay, _ = run_gpu(q, par, seq, qual, idx, 128, kmers=True, k2=7, cg_k=6)
    assert int(res["mask"].sum()) == g["n_passed"]
    for name, (s, wsum) in g["sections"].items():
        sec = lay.section(got, name).astype(object)
        assert int(sum(sec)) == s, name
        assert int(sum((i + 1) * int(v) for i, v in enumerate(sec)) % (1 << 61)) == wsum, name


@pytest.mark.parametrize("ooq", [False, True])
def test_stats_on_mid_length_reads_register_kernel(q, ob, ooq):
    """Reads of 641..1009 bp with statistics: too long for the tiled kernel, short enough for the register-resident
    decision kernel (its GC-counting variants, with and without the out-of-quality count) + the position kernel."""
    rng = np.random.default_rng(77 + ooq)
    seq, qual, idx = rand_batch(rng, 3000, 0, 900, p_n=0.02)
    flt = dict(min_read_length=30, min_read_quality=18, max_N=12)
    if ooq:
        flt.update(max_out_of_quality=200, left_length=25, min_left_quality=16)
    _, _, _, emask, _ = check(q, ob, q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=flt), seq, qual, idx, 960)
    assert 0 < int(emask.sum()) < len(emask)
    check(q, ob, q.make_params(q.WORK_STATS), seq, qual, idx, 960)


# ---------------------------------------------------------------------------------------------------------------
# Round 2: the code paths the headline numbers come from.  The in-loop ladder flush (8-bit lanes -> shared u32
# tables) is reached by a warp only after ~250 of its iterations, i.e. by launches of millions of reads -- or by a
# small batch when HPGQ_GRID_CAP leaves it two CTAs.  Both kernels (uniform-length aligned-word loop and the general
# funnel-shift loop, HPGQ_NO_UNI=1) are compared with the oracle word for word.
class _env:
    def __init__(self, **kw):
        self.kw = kw

    def __enter__(self):
        import os
        self.old = {k: os.environ.get(k) for k in self.kw}
        for k, v in self.kw.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = str(v)

    def __exit__(self, *a):
        import os
        for k, v in self.old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def _device_check(q, ob, par, seq, qual, idx, Lcap, *, uniform=False, hint=None, lead=0):
    """Uploads the batch (optionally `lead` junk bytes in front, so that idx[0] != 0), runs it device-resident and
    compares counters + mask with the oracle."""
    n = len(idx) - 1
    if lead:
        junk = np.full(lead, ord("#"), dtype=np.uint8)
        seq_d, qual_d = np.concatenate([junk, seq]), np.concatenate([junk, qual])
        idx_d = (idx.astype(np.uint64) + lead).astype(np.uint32)
    else:
        seq_d, qual_d, idx_d = seq, qual, idx
    with q.Ctx(0, max_reads=max(n, 1), max_bytes=max(int(idx_d[-1]), 16), max_read_len=Lcap, host_staging=False) as ctx:
        ctx.set_params(par)
        b = ctx.upload_batch(seq_d, qual_d, idx_d, max_len=int(np.diff(idx).max()) if hint is None else hint)
        ctx.submit_device(b, 0, uniform=uniform)
        ctx.sync(0)
        got = ctx.counters()
        views = q.Result()
        q._ck(ctx.lib.hpgq_batch_device_views(ctx.h, 0, __import__("ctypes").byref(views)), "device_views")
        mask = ctx.dev_download(views.mask, n)
        lay = ctx.layout
        ctx.free_batch(b)
    exp, emask, _ = ob.process_batch(par, lay, seq, qual, idx)
    assert np.array_equal(mask, emask), f"mask differs at {np.nonzero(mask != emask)[0][:10]}"
    msg = diff_report(q, lay, got, exp)
    assert not msg, msg
    return got


C2_FILTER = dict(min_read_length=50, min_read_quality=20, max_N=5)


@pytest.mark.parametrize("no_uni", [None, 1])
def test_bench_shaped_launch_matches_oracle(q, ob, no_uni):
    """One launch of bench.py's shape: 2^23 x 150 bp, stats + filter, K = 5 / G = 8, all 148 CTAs -- every consumer
    warp runs ~470 iterations, past the in-loop flush at 248.  Whole counter block and mask against the oracle."""
    n = 1 << 23
    cfg = q.synth_cfg(150)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=C2_FILTER)
    with _env(HPGQ_NO_UNI=no_uni), q.Ctx(0, max_reads=n, max_bytes=n * 150, max_read_len=160, host_staging=False) as ctx:
        ctx.set_params(par)
        b = ctx.synth_device(cfg, 0, n)
        ctx.submit_device(b, 0, uniform=True)
        ctx.sync(0)
        got = ctx.counters()
        views = q.Result()
        q._ck(ctx.lib.hpgq_batch_device_views(ctx.h, 0, __import__("ctypes").byref(views)), "device_views")
        mask = ctx.dev_download(views.mask, n)
        seq = ctx.dev_download(b.seq, b.n_bytes)
        qual = ctx.dev_download(b.qual, b.n_bytes)
        idx = ctx.dev_download(b.idx, (n + 1) * 4).view(np.uint32)
        lay = ctx.layout
        ctx.free_batch(b)
    exp, emask, _ = ob.process_batch(par, lay, seq, qual, idx)
    assert np.array_equal(mask, emask)
    msg = diff_report(q, lay, got, exp)
    assert not msg, msg
    assert int(got[q.S["NUM_PASSED"]]) + int(got[q.S["NUM_FAILED"]]) == n


@pytest.mark.parametrize("no_uni", [None, 1])
@pytest.mark.parametrize("L,n", [(76, 70000), (150, 36000), (300, 18000), (600, 9500), (36, 140000), (100, 60000)])
def test_in_loop_flush_every_geometry(q, ob, L, n, no_uni):
    """Two CTAs (HPGQ_GRID_CAP): each consumer warp passes iteration 255 of every ladder cadence
    (G = 4, 8, 16, 32: MERGED and !MERGED), uniform reads and a ragged variant."""
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_length=min(50, L), min_read_quality=20, max_N=5))
    with _env(HPGQ_NO_UNI=no_uni, HPGQ_GRID_CAP=2):
        seq, qual, idx = q.synth_host(q.synth_cfg(L), 0, n)
        _device_check(q, ob, par, seq, qual, idx, 640, uniform=True)
        if no_uni:
            seq, qual, idx = q.synth_host(q.synth_cfg(L, short_pct=3, short_min=min(35, L // 2)), 0, n)
            _device_check(q, ob, par, seq, qual, idx, 640)


@pytest.mark.parametrize("no_uni", [None, 1])
def test_quality_sum_cell_carries_past_32_bits(q, ob, no_uni):
    """One CTA (HPGQ_GRID_CAP=1) sees 600 000 reads of 150 bp: its sum of quality bytes of that length passes 2^32, so
    the block-private cell's low word wraps and the carry of the two-word shared-memory add is exercised (a launch of
    more than ~110 M reads on all 148 SMs gets there too).  Whole counter block against the oracle."""
    n = 600_000
    seq, qual, idx = q.synth_host(q.synth_cfg(150), 0, n)
    assert int(qual.astype(np.uint64).sum()) > (1 << 32)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_length=50, min_read_quality=2, max_N=150))
    with _env(HPGQ_NO_UNI=no_uni, HPGQ_GRID_CAP=1):
        got = _device_check(q, ob, par, seq, qual, idx, 192, uniform=True)
    assert int(got[q.S["NUM_PASSED"]]) == n  # every read is in the sum


@pytest.mark.parametrize("L", [7, 36, 75, 96, 97, 100, 101, 106, 111, 112, 113, 149, 150, 151, 250, 251, 253, 300, 301, 302, 600, 637, 640])
def test_uniform_kernel_lengths_and_phases(q, ob, L):
    """Uniform-length kernel over the geometries' length range, odd lengths (where the group count does not keep the
    byte phase the kernel must not be chosen) and every start phase idx[0] mod 4 / mod 16."""
    n = 9000
    seq, qual, idx = q.synth_host(q.synth_cfg(L, lowercase_ppm=300), 0, n)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_length=5, min_read_quality=20, max_N=5))
    for lead in (0, 1, 2, 3, 13):
        _device_check(q, ob, par, seq, qual, idx, 640, uniform=True, lead=lead)
    full = dict(min_read_length=5, min_read_quality=22, max_read_quality=39, max_out_of_quality=L // 3 + 2,
                left_length=10, min_left_quality=30, max_left_quality=41, right_length=12, min_right_quality=15,
                max_right_quality=40, max_N=3)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=full)
    for lead in (0, 7):
        _device_check(q, ob, par, seq, qual, idx, 640, uniform=True, lead=lead)
    _device_check(q, ob, q.make_params(q.WORK_STATS), seq, qual, idx, 640, uniform=True, lead=2)


def test_false_length_statements_fail_loudly(q, ob):
    """hpgq_batch_t.reserved: a uniform-length claim or a max-length bound that the batch breaks is detected on the
    device and reported by hpgq_batch_sync / hpgq_counters_download, never silently wrong."""
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=C2_FILTER)
    seq, qual, idx = q.synth_host(q.synth_cfg(150, short_pct=2), 0, 50000)
    assert int(np.diff(idx).min()) < 150
    with pytest.raises(q.HpgqError, match="length statement"):
        _device_check(q, ob, par, seq, qual, idx, 160, uniform=True, hint=150)
    seq, qual, idx = q.synth_host(q.synth_cfg((100, 600)), 0, 50000)
    with pytest.raises(q.HpgqError, match="length statement"):
        _device_check(q, ob, par, seq, qual, idx, 640, hint=120)
    # the same batches with true statements
    _device_check(q, ob, par, seq, qual, idx, 640, hint=600)
    _device_check(q, ob, par, seq, qual, idx, 640, hint=0)


def test_compaction_checksum_matches_host_twin(q, ob):
    """hpgq_batch_checksum (device) == hpgq_checksum_host over the oracle's compacted arrays; order-sensitive."""
    seq, qual, idx = q.synth_host(q.synth_cfg(150, short_pct=3), 0, 40000)
    par = q.make_params(q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT, filter=C2_FILTER,
                        edit=dict(left_length=10, min_left_quality=20, right_length=10, min_right_quality=20))
    with q.Ctx(0, max_reads=40000, max_bytes=int(idx[-1]), max_read_len=160) as ctx:
        ctx.set_params(par)
        ctx.submit(seq, qual, idx, 0)
        res = ctx.wait(0)
        ck = ctx.checksum(0)
        lay = ctx.layout
    exp, emask, etrim = ob.process_batch(par, lay, seq, qual, idx)
    cp = ob.compact(seq, qual, idx, emask, etrim)
    assert ck["passed_bytes"] == cp["passed_bytes"] and ck["failed_bytes"] == cp["failed_bytes"]
    assert ck["n_passed"] == cp["n_passed"]
    assert ck["checksum"] == q.checksum_host(cp["out_seq"], cp["out_qual"], cp["out_idx"], cp["out_src"])
    swapped = cp["out_seq"].copy()
    swapped[[0, 1]] = swapped[[1, 0]]
    if swapped[0] != swapped[1]:
        assert ck["checksum"] != q.checksum_host(swapped, cp["out_qual"], cp["out_idx"], cp["out_src"])


@pytest.mark.parametrize("k", [2, 5, 10, 12])
def test_chaos_game_exact_on_long_axis_runs(q, ob, k):
    """Uncapped inputs (poly-T 70 + C-leading read, poly-G tails, runs with an N / a lower-case byte inside, runs
    across read boundaries, single-axis short reads in a row): the CUDA tables equal the oracle's literal double
    restatement AND the reference's own compiled chaos_game_fill_tables (oracle/_ref), cell for cell."""
    from test_oracle import axis_run_batches
    moved = 0
    for seq, qual, idx in axis_run_batches(3) + axis_run_batches(11):
        par = q.make_params(q.WORK_CHAOS, cg_k=k)
        got, _, lay, _ = run_gpu(q, par, seq, qual, idx, 512, cg_k=k)
        exp, _, _ = ob.process_batch(par, lay, seq, qual, idx)
        for name in ("CG_SEQ", "CG_Q"):
            assert np.array_equal(lay.section(got, name), lay.section(exp, name)), name
        assert got[q.S["CG_WORDS"]] == exp[q.S["CG_WORDS"]]
        moved += int(got[q.S["CG_MOVED_WORDS"]])
        ref = ob.chaos_reference(seq, qual, idx, k)
        if ref is not None:
            assert np.array_equal(lay.section(got, "CG_SEQ").astype(np.uint32), ref[0])
            assert np.array_equal((lay.section(got, "CG_Q") & 0xFFFFFFFF).astype(np.uint32), ref[1])
    if k >= 5:
        assert moved > 0, "the inputs must exercise the double / integer difference"


def test_chaos_game_exact_inside_a_large_batch(q, ob):
    """Long runs scattered through 60 000 ordinary reads (the warm-up of a flagged read starts from unknown state and
    must meet the true one), several batches accumulated, k = 10 with k-mers alongside."""
    rng = np.random.default_rng(21)
    cfg = q.synth_cfg(150)
    seq, qual, idx = q.synth_host(cfg, 0, 60000)
    seq = seq.copy()
    for r in rng.choice(60000, 400, replace=False):
        o, L = int(idx[r]), int(idx[r + 1] - idx[r])
        a = int(rng.integers(0, L - 50))
        n = int(rng.integers(40, L - a))
        seq[o + a:o + a + n] = rng.choice(np.frombuffer([b"AT", b"GT", b"T", b"A"][int(rng.integers(0, 4))], dtype=np.uint8), n)
    par = q.make_params(q.WORK_KMERS | q.WORK_CHAOS, k2=11, cg_k=10)
    got, *_ = check(q, ob, par, seq, qual, idx, 160, kmers=True, k2=11, cg_k=10)
    assert got[q.S["CG_FIXED_READS"]] >= 400


@pytest.mark.parametrize("L,n", [(150, 300000), (36, 500000), (1000, 40000)])
def test_one_pass_compaction_many_tiles(q, ob, L, n, monkeypatch):
    """Thousands of tiles through the decoupled look-back: filter-only, edit-only and edit + filter, most reads failing
    (a large failed stream through the side buffer), every read failing, every read passing."""
    monkeypatch.setenv("HPGQ_FUSE", "1")
    seq, qual, idx = q.synth_host(q.synth_cfg(L, short_pct=4, short_min=max(L // 4, 1)), 5, n)
    edit = dict(left_length=10, min_left_quality=30, right_length=12, min_right_quality=22)
    cases = [(q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT, dict(min_read_length=L // 2, min_read_quality=20, max_N=5)),
             (q.WORK_FILTER | q.WORK_COMPACT, dict(min_read_quality=36)),          # most reads fail
             (q.WORK_FILTER | q.WORK_COMPACT, dict(min_read_quality=60)),          # all fail
             (q.WORK_EDIT | q.WORK_COMPACT, None),                                 # all pass, trimmed
             (q.WORK_FILTER | q.WORK_COMPACT, dict(min_read_quality=25, max_out_of_quality=L // 3))]
    for work, flt in cases:
        par = q.make_params(work, edit=edit, filter=flt)
        _, res, _, emask, etrim = check(q, ob, par, seq, qual, idx, 1024)
        cp = ob.compact(seq, qual, idx, emask if work & q.WORK_FILTER else None, etrim if work & q.WORK_EDIT else None)
        assert res["n_passed"] == cp["n_passed"] and res["passed_bytes"] == cp["passed_bytes"]
        assert res["failed_bytes"] == cp["failed_bytes"]
        for k in ("out_idx", "out_src", "out_seq", "out_qual"):
            assert np.array_equal(res[k], cp[k]), k


@pytest.mark.parametrize("lmax", [30, 75, 150, 310, 630])
def test_one_pass_word_frame_decisions_on_ragged_reads(q, ob, lmax, monkeypatch):
    """The word-frame decision of the one-pass kernel (no filter windows, no out-of-quality limit) in each of its lane
    widths, on reads of every length from 0 up: reads shorter than the two trim windows fall back to the register
    code inside the same warp iteration."""
    monkeypatch.setenv("HPGQ_FUSE", "1")
    rng = np.random.default_rng(1000 + lmax)
    seq, qual, idx = rand_batch(rng, 20000, 0, lmax, p_n=0.04)
    qual = rng.integers(35, 74, size=len(qual)).astype(np.uint8)
    for edit, flt in ((dict(left_length=9, min_left_quality=22, max_left_quality=30, right_length=13, min_right_quality=19,
                            max_right_quality=28), dict(min_read_length=4, min_read_quality=18, max_read_quality=24, max_N=3)),
                      (dict(left_length=255, min_left_quality=21, right_length=3, min_right_quality=23), dict(max_N=1))):
        par = q.make_params(q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT, edit=edit, filter=flt)
        _, res, _, emask, etrim = check(q, ob, par, seq, qual, idx, lmax + 8)
        assert 0 < int(emask.sum()) < len(emask)
        cp = ob.compact(seq, qual, idx, emask, etrim)
        for k in ("out_idx", "out_src", "out_seq", "out_qual"):
            assert np.array_equal(res[k], cp[k]), k


@pytest.mark.parametrize("L", [363, 630, 150, 77])
def test_one_pass_word_frame_on_full_tiles(q, ob, L):
    """Tiles of the one-pass kernel that fill their staged span (uniform reads, 22 000 / L per tile): the word-frame
    decision loads a whole frame per read, for the tile's last reads beyond the staged bytes -- past the CTA's shared
    memory before the launcher left a frame of slack (illegal address at 60 x 363 bp, tools/stress_parity.py)."""
    rng = np.random.default_rng(L)
    seq, qual, idx = rand_batch(rng, 18000, L, L, p_n=0.03)
    edit = dict(left_length=4, min_left_quality=19, right_length=6, min_right_quality=20)
    for work, flt in ((q.WORK_EDIT | q.WORK_COMPACT, None),
                      (q.WORK_EDIT | q.WORK_FILTER | q.WORK_COMPACT, dict(min_read_length=L - 5, min_read_quality=22, max_N=L // 25))):
        par = q.make_params(work, edit=edit, filter=flt, kernel_variant=2)
        _, res, _, emask, etrim = check(q, ob, par, seq, qual, idx, (L + 63) // 64 * 64)
        cp = ob.compact(seq, qual, idx, emask if flt else None, etrim)
        for k in ("out_idx", "out_src", "out_seq", "out_qual"):
            assert np.array_equal(res[k], cp[k]), k


@pytest.mark.parametrize("k2,ck", [(11, 10), (7, 7), (9, 3), (3, 5), (13, 12)])
def test_kmers_and_chaos_game_single_reduction_path(q, ob, k2, ck, monkeypatch):
    """k-mers + chaos game without a filter: one packed reduction per base, split into K2 / CG_SEQ / CG_Q afterwards
    (k2 >= cg_k; (3, 5) keeps the two-table path).  Ns, long axis runs, low qualities below the base quality (values
    that do not fit the packed field), two batches accumulated, and the same through HPGQ_NO_KCPACK."""
    rng = np.random.default_rng(k2 * 16 + ck)
    seq, qual, idx = rand_batch(rng, 20000, 0, 180, p_n=0.02)
    qual = rng.integers(33, 74, size=len(qual)).astype(np.uint8)
    qual[rng.integers(0, len(qual), 500)] = 20  # below the base quality: the word's sum can wrap
    # stretches of very high quality bytes: window sums of 1024 and more do not fit the packed field either
    for start in rng.integers(0, len(qual) - 40, 60):
        qual[start:start + 40] = 200
    seq = seq.copy()
    o = int(idx[100])
    seq[o:o + 60] = ord("T")
    par = q.make_params(q.WORK_KMERS | q.WORK_CHAOS, k2=k2, cg_k=ck)
    for env in (None, "1"):
        if env:
            monkeypatch.setenv("HPGQ_NO_KCPACK", env)
        with q.Ctx(0, max_reads=20000, max_bytes=int(idx[-1]), max_read_len=192, kmers=True, k2=k2, cg_k=ck) as ctx:
            ctx.set_params(par)
            for _ in range(2):
                ctx.submit(seq, qual, idx, 0)
                ctx.wait(0)
            got = ctx.counters()
            lay = ctx.layout
        exp, _, _ = ob.process_batch(par, lay, seq, qual, idx, finalize=False)
        exp, _, _ = ob.process_batch(par, lay, seq, qual, idx, counters=exp)
        msg = diff_report(q, lay, got, exp)
        assert not msg, msg


def test_device_batch_larger_than_the_ctx_limit_is_refused(q):
    """hpgq_batch_submit_device with COMPACT: a batch whose bases exceed max_bytes_per_batch (the size of the
    compaction buffers) is refused, not written past the buffers."""
    seq, qual, idx = q.synth_host(q.synth_cfg(150), 0, 4000)
    par = q.make_params(q.WORK_FILTER | q.WORK_COMPACT, filter=C2_FILTER)
    with q.Ctx(0, max_reads=4000, max_bytes=4000 * 150 // 2, max_read_len=160, host_staging=False) as ctx:
        ctx.set_params(par)
        b = ctx.upload_batch(seq, qual, idx, max_len=150)
        with pytest.raises(q.HpgqError, match="max_bytes_per_batch"):
            ctx.submit_device(b, 0)
        ctx.free_batch(b)


def test_out_of_quality_count_ignores_the_neighbouring_read(q, ob):
    """A non-ASCII quality byte at the end of the PRECEDING read shares a 32-bit word with this read's first bytes; the
    long-read group kernel must not let it carry into this read's out-of-quality count."""
    rng = np.random.default_rng(9)
    seq, qual, idx = rand_batch(rng, 400, 1100, 1500, p_n=0.01)
    qual = rng.integers(40, 70, size=len(qual)).astype(np.uint8)
    for r in range(1, 400, 3):
        qual[int(idx[r]) - 1] = 250
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_quality=10, max_read_quality=36, max_out_of_quality=128))
    _, _, _, emask, _ = check(q, ob, par, seq, qual, idx, 1536)
    assert 0 < int(emask.sum()) < len(emask)


@pytest.mark.parametrize("two_pass", [False, True], ids=["one_pass", "two_pass"])
def test_long_reads_statistics_in_one_pass(q, ob, two_pass):
    """Statistics + a filter on length / mean quality / N count of reads too long for the tiled kernel: one pass over the
    batch (position tables of every read + per-read sums, decisions from the sums, failed reads subtracted again) against
    the oracle, and the two-pass path (HPGQ_NO_POSFUSE=1: decision kernel, masked position pass) on the same batches.
    Filters that fail most reads, none, and all of them; three batches on two slots, so that the per-read sums left
    behind by one batch would show in the next; a max-length statement the batch breaks."""
    env = dict(HPGQ_NO_POSFUSE="1") if two_pass else {}
    with _env(**env):
        rng = np.random.default_rng(19)
        seq, qual, idx = rand_batch(rng, 2500, 700, 9000, p_n=0.02)
        for flt in (dict(min_read_length=3000, min_read_quality=37, max_N=90),   # most reads fail
                    dict(min_read_length=1),                                       # none fails
                    dict(min_read_length=100000),                                  # all fail
                    dict(max_read_length=5000, max_read_quality=36, max_N=60)):
            par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=flt)
            _, _, _, emask, _ = check(q, ob, par, seq, qual, idx, 9024)
        assert int(emask.sum()) not in (0, len(emask))
        check(q, ob, q.make_params(q.WORK_STATS), seq, qual, idx, 9024)
        # batches of different sizes through the same slots: counters accumulate, the per-read sums start from zero
        par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=dict(min_read_length=2000, min_read_quality=36, max_N=70))
        batches = [rand_batch(rng, n, 650, hi, p_n=0.02) for n, hi in ((1200, 6000), (300, 9000), (900, 3000))]
        with q.Ctx(0, max_reads=1200, max_bytes=max(int(b[2][-1]) for b in batches) + 64, max_read_len=9024) as ctx:
            ctx.set_params(par)
            lay = ctx.layout
            exp = np.zeros(lay.words, dtype=np.uint64)
            got_masks, exp_masks = {}, {}
            for k, (s_, q_, i_) in enumerate(batches):
                if k >= 2:
                    got_masks[k - 2] = ctx.wait(k & 1)["mask"]
                ctx.submit(s_, q_, i_, k & 1)
                exp, exp_masks[k], _ = ob.process_batch(par, lay, s_, q_, i_, counters=exp, finalize=False)
            got_masks[1] = ctx.wait(1)["mask"]
            got_masks[2] = ctx.wait(0)["mask"]
            got = ctx.counters()
        ob.lib().oracle_finalize(__import__("ctypes").byref(lay), exp.ctypes.data_as(__import__("ctypes").c_void_p))
        for k in range(3):
            assert np.array_equal(got_masks[k], exp_masks[k]), k
        msg = diff_report(q, lay, got, exp)
        assert not msg, msg
        # a read longer than the stated maximum: refused, not silently short of its tail
        seq, qual, idx = rand_batch(rng, 400, 700, 8000, p_n=0.01)
        if not two_pass:
            with pytest.raises(q.HpgqError, match="length statement"):
                _device_check(q, ob, par, seq, qual, idx, 9024, hint=4096)
        _device_check(q, ob, par, seq, qual, idx, 9024, hint=8000)
