"""Pins the CPU oracle: hand-computed known answers for the restated (unpinned) per-read semantics, and the
REFERENCE's own chaos_game.c (compiled into oracle/_ref from /root/reference/old) for the chaos-game row.
SURVEY.md section 8c lists these vectors; the reference itself ships no tests or fixtures."""
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def pack(reads):
    seq = b"".join(r[0] for r in reads)
    qual = b"".join(r[1] for r in reads)
    idx = np.zeros(len(reads) + 1, dtype=np.uint32)
    idx[1:] = np.cumsum([len(r[0]) for r in reads])
    return np.frombuffer(seq, dtype=np.uint8).copy(), np.frombuffer(qual, dtype=np.uint8).copy(), idx


MICRO = [(b"ACGTN", b"IIIII"), (b"acGT", b"!!!!"), (b"GGCC", b"5555"), (b"A", b"#")]


def test_micro_stats_hand_computed(q, ob):
    seq, qual, idx = pack(MICRO)
    lay = ob.layout(q.Layout, 8)
    c, mask, _ = ob.process_batch(q.make_params(q.WORK_STATS), lay, seq, qual, idx)
    sec = lambda n: lay.section(c, n)
    assert c[q.S["NUM_READS"]] == 4 and c[q.S["ACC_LENGTH"]] == 14
    assert c[q.S["MIN_LENGTH"]] == 1 and c[q.S["MAX_LENGTH"]] == 5
    assert [c[q.S[k]] for k in ("NUM_A", "NUM_C", "NUM_G", "NUM_T", "NUM_N")] == [2, 3, 4, 2, 1]
    assert list(sec("LEN_HIST")[:9]) == [0, 1, 0, 0, 2, 1, 0, 0, 0]
    assert list(sec("SUMQ_BY_LEN")[:6]) == [0, 35, 0, 0, 132 + 212, 365]
    qh = sec("QUAL_HIST")
    assert {int(i): int(qh[i]) for i in np.nonzero(qh)[0]} == {33: 1, 35: 1, 53: 1, 73: 1}
    gc = sec("GC_HIST")
    assert {int(i): int(gc[i]) for i in np.nonzero(gc)[0]} == {0: 1, 25: 1, 40: 1, 100: 1}
    assert list(sec("POS_ACCQ")[:5]) == [194, 159, 159, 159, 73]
    assert list(sec("POS_A")[:5]) == [2, 0, 0, 0, 0]
    assert list(sec("POS_C")[:5]) == [0, 1, 1, 1, 0]
    assert list(sec("POS_G")[:5]) == [1, 1, 2, 0, 0]
    assert list(sec("POS_T")[:5]) == [0, 0, 0, 2, 0]
    assert list(sec("POS_N")[:5]) == [0, 0, 0, 0, 1]
    assert list(mask) == [1, 1, 1, 1]


FILTER_READS = [(b"ACGTN", b"IIIII"), (b"ACGT", b"5555"), (b"ACGT", b"4555"), (b"ACG", b"III"),
                (b"ACGTAC", b"IIIIII"), (b"ACGT", b"5J5I"), (b"ACGT", b"4J5J")]
FILTER_OPTS = dict(min_read_length=4, max_read_length=5, min_read_quality=20, max_read_quality=40, max_N=0,
                   max_out_of_quality=1)
FILTER_MASK = [0, 1, 0, 0, 0, 1, 0]


def test_filter_threshold_edges(q, ob):
    seq, qual, idx = pack(FILTER_READS)
    lay = ob.layout(q.Layout, 8)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER, filter=FILTER_OPTS)
    c, mask, _ = ob.process_batch(par, lay, seq, qual, idx)
    assert list(mask) == FILTER_MASK
    assert c[q.S["NUM_PASSED"]] == 2 and c[q.S["NUM_FAILED"]] == 5
    assert c[q.S["NUM_READS"]] == 2  # statistics cover passed reads only (src/stats_fastq.c:226-232)
    assert list(lay.section(c, "POS_ACCQ")[:4]) == [106, 53 + 74, 106, 53 + 73]


def test_filter_windows(q, ob):
    reads = [(b"ACGT", b"?@II"), (b"ACGT", b"??II"), (b"ACGT", b">?II"), (b"ACGT", b"II?>"), (b"A", b">")]
    seq, qual, idx = pack(reads)
    lay = ob.layout(q.Layout, 8)
    par = q.make_params(q.WORK_FILTER, filter=dict(left_length=2, min_left_quality=30, right_length=2,
                                                   min_right_quality=30))
    _, mask, _ = ob.process_batch(par, lay, seq, qual, idx)
    assert list(mask) == [1, 1, 0, 0, 0]


EDIT_READS = [(b"ACGTAC", b">?IIII"), (b"ACGT", b"II>>"), (b"AC", b">>"), (b"ACG", b">>>"), (b"ACGT", b"IIII"),
              (b"", b"")]
EDIT_OPTS = dict(left_length=2, min_left_quality=30, right_length=2, min_right_quality=30)
EDIT_TRIM = [(2, 4), (0, 2), (2, 0), (2, 0), (0, 4), (0, 0)]


def test_edit_trim_cases(q, ob):
    seq, qual, idx = pack(EDIT_READS)
    lay = ob.layout(q.Layout, 8)
    par = q.make_params(q.WORK_EDIT, edit=EDIT_OPTS)
    c, mask, trim = ob.process_batch(par, lay, seq, qual, idx)
    assert [tuple(int(x) for x in t) for t in trim] == EDIT_TRIM
    assert c[q.S["NUM_EDITED"]] == 4
    # edit then filter: the filter sees the TRIMMED read (src/edit_fastq.c:154-167)
    par = q.make_params(q.WORK_EDIT | q.WORK_FILTER, edit=EDIT_OPTS, filter=dict(min_read_length=3))
    _, mask, _ = ob.process_batch(par, lay, seq, qual, idx)
    assert list(mask) == [1, 0, 0, 0, 1, 0]
    cp = ob.compact(seq, qual, idx, mask, trim)
    assert bytes(cp["out_seq"]) == b"GTAC" + b"ACGT" + b"AC" and list(cp["out_src"]) == [0, 4, 1, 2, 3, 5]
    assert list(cp["out_idx"]) == [0, 4, 8, 10, 10, 10, 10] and cp["n_passed"] == 2


def test_kmers_hand_computed(q, ob):
    seq, qual, idx = pack([(b"ACGTANGTACGT", b"I" * 12)])
    lay = ob.layout(q.Layout, 16, kmers=1, k2=3)
    c, _, _ = ob.process_batch(q.make_params(q.WORK_KMERS, k2=3), lay, seq, qual, idx)
    k5 = lay.section(c, "K5")
    assert {int(i): int(k5[i]) for i in np.nonzero(k5)[0]} == {108: 1, 710: 1, 795: 1}
    k5p = lay.section(c, "K5_POS").reshape(1024, 16)
    assert k5p[108, 0] == 1 and k5p[710, 6] == 1 and k5p[795, 7] == 1 and k5p.sum() == 3
    k3 = lay.section(c, "K2")
    # 3-mers: ACG CGT GTA | GTA TAC ACG CGT
    exp = {0b000110: 2, 0b011011: 2, 0b101100: 2, 0b110001: 1}
    assert {int(i): int(k3[i]) for i in np.nonzero(k3)[0]} == exp


def rand_batch(rng, n, lmin, lmax, p_n=0.01, lower=0.0, cap_runs=True):
    lens = rng.integers(lmin, lmax + 1, size=n)
    idx = np.zeros(n + 1, dtype=np.uint32)
    idx[1:] = np.cumsum(lens)
    tot = int(idx[-1])
    seq = rng.choice(np.frombuffer(b"ACGTN", dtype=np.uint8), size=tot,
                     p=[(1 - p_n) / 4] * 4 + [p_n]).astype(np.uint8)
    if lower > 0:
        m = rng.random(tot) < lower
        seq[m] |= 0x20
    qual = rng.integers(35, 74, size=tot).astype(np.uint8)
    return seq, qual, idx


@pytest.mark.parametrize("k", [2, 5, 7, 10])
def test_chaos_restatement_equals_reference_code(q, ob, k):
    """oracle_chaos_fill vs the reference's chaos_game_fill_tables (oracle/_ref) -- the pinned row."""
    if ob.ref_lib() is None:
        pytest.skip("oracle/_ref not built (reference tree absent on this machine)")
    rng = np.random.default_rng(100 + k)
    for lower in (0.0, 0.02):
        seq, qual, idx = rand_batch(rng, 300, 0, 120, p_n=0.02, lower=lower)
        ts, tq, words = ob.chaos_literal(seq, qual, idx, k)
        rs, rq, rwords = ob.chaos_reference(seq, qual, idx, k)
        assert words == rwords
        assert np.array_equal(ts.astype(np.uint32), rs)
        assert np.array_equal((tq & 0xFFFFFFFF).astype(np.uint32), rq)  # reference cells are 32-bit


def test_chaos_hand_computed_k2(q, ob):
    # k=2 on "ACGT": words AC, CG, GT.  index = sum bit(n_i) 2^(i-1), latest base = MSB
    #   AC: x: A=1 then C=0 -> (0<<1 | 1) = 1 ; y: 0,0 -> 0          => cell (1,0)
    #   CG: x: 0,0 -> 0 ; y: C=0 then G=1 -> 2                        => cell (0,2)
    #   GT: x: G=0,T=1 -> 2 ; y: G=1,T=1 -> 3                         => cell (2,3)
    seq, qual, idx = pack([(b"ACGT", b"5678")])
    ts, tq, words = ob.chaos_literal(seq, qual, idx, 2, base_quality=33)
    assert words == 3
    exp = np.zeros((4, 4), dtype=np.uint64)
    exp[1, 0] = exp[0, 2] = exp[2, 3] = 1
    assert np.array_equal(ts.reshape(4, 4), exp)
    eq = np.zeros((4, 4), dtype=np.uint64)
    eq[1, 0] = 53 + 54 - 66
    eq[0, 2] = 54 + 55 - 66
    eq[2, 3] = 55 + 56 - 66
    assert np.array_equal(tq.reshape(4, 4), eq)


def test_chaos_integer_form_and_its_divergence(q, ob):
    rng = np.random.default_rng(7)
    cfg = q.synth_cfg(150, cap_axis_runs=1)
    seq, qual, idx = q.synth_host(cfg, 0, 2000)
    for k in (5, 10, 12):
        a = ob.chaos_literal(seq, qual, idx, k)
        b = ob.chaos_int(seq, qual, idx, k)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2] == b[2] and b[3] == 0
    # a 70-base poly-T followed by a C-leading read: the double state rounds up, the integer index does not
    seq, qual, idx = pack([(b"T" * 70, b"I" * 70), (b"C" * 12 + bytes(int(x) for x in rng.choice(list(b"ACGT"), 40)), b"I" * 52)])
    a = ob.chaos_literal(seq, qual, idx, 10)
    b = ob.chaos_int(seq, qual, idx, 10)
    assert b[3] > 0, "long axis run must be flagged"
    assert a[2] == b[2]


def test_generator_golden_fixture(q):
    """The committed fixture pins the generator (tests/golden/make_golden.py wrote it)."""
    with open(os.path.join(HERE, "golden", "synth_golden.json")) as f:
        g = json.load(f)
    for case in g["cases"]:
        cfg = q.synth_cfg(tuple(case["length"]) if isinstance(case["length"], list) else case["length"],
                          seed=case["seed"], cap_axis_runs=case["cap"], short_pct=case["short_pct"],
                          qual_shift=case["qual_shift"])
        seq, qual, idx = q.synth_host(cfg, case["first"], case["n"])
        assert bytes(seq[:case["head_len"]]).decode() == case["seq_head"]
        assert bytes(qual[:case["head_len"]]).decode() == case["qual_head"]
        assert int(idx[-1]) == case["total"]
        assert int(seq.astype(np.uint64).sum()) == case["seq_sum"] and int(qual.astype(np.uint64).sum()) == case["qual_sum"]


def test_oracle_golden_counters(q, ob):
    """Oracle counters on seeded synthetic reads against the committed fixture (regression pin)."""
    with open(os.path.join(HERE, "golden", "oracle_golden.json")) as f:
        g = json.load(f)
    cfg = q.synth_cfg(100)
    seq, qual, idx = q.synth_host(cfg, 0, g["n"])
    lay = ob.layout(q.Layout, 128, kmers=1, k2=7, cg_k=6)
    par = q.make_params(q.WORK_STATS | q.WORK_FILTER | q.WORK_KMERS | q.WORK_CHAOS, k2=7, cg_k=6,
                        filter=dict(min_read_length=50, min_read_quality=20, max_N=5))
    c, mask, _ = ob.process_batch(par, lay, seq, qual, idx)
    assert int(mask.sum()) == g["n_passed"]
    for name, (s, wsum) in g["sections"].items():
        sec = lay.section(c, name).astype(object)
        assert int(sum(sec)) == s, name
        assert int(sum((i + 1) * int(v) for i, v in enumerate(sec)) % (1 << 61)) == wsum, name


def axis_run_batches(seed=0):
    """Inputs on which the reference's double arithmetic leaves the integer index (old/chaos_game.c:200-251): axis
    runs of 40-120 bases (A/T, G/T, single letters) ending in and starting reads, crossing read boundaries, with an N
    or a lower-case byte inside, short single-axis reads in a row, and random reads in between."""
    rng = np.random.default_rng(seed)

    def rnd(n, letters=b"ACGT"):
        return bytes(int(x) for x in rng.choice(list(letters), n))

    def ql(n):
        return bytes(int(x) for x in rng.integers(35, 74, n))

    out = []
    # the survey's case: poly-T >= 60, then a C-leading read
    reads = [b"T" * 70, b"C" * 12 + rnd(40), rnd(80), b"ACGT" + b"G" * 66, b"CA" + rnd(60)]
    out.append(reads)
    # runs ending a read and going on in the next, tails of every length, N and lower case inside
    reads = []
    for i in range(60):
        run_letters = [b"AT", b"GT", b"T", b"A", b"G"][i % 5]
        n1, n2 = int(rng.integers(15, 70)), int(rng.integers(0, 60))
        reads.append(rnd(int(rng.integers(5, 60))) + rnd(n1, run_letters))
        reads.append(rnd(n2, run_letters) + rnd(int(rng.integers(1, 40)), b"CG" if run_letters in (b"AT", b"T", b"A") else b"CA")
                     + rnd(int(rng.integers(0, 50))))
        if i % 7 == 0:
            r = bytearray(rnd(90, run_letters))
            r[30] = ord("N")
            reads.append(bytes(r) + b"C" + rnd(30))
        if i % 11 == 0:
            r = bytearray(rnd(80, run_letters))
            r[41] = ord("a")
            reads.append(bytes(r) + rnd(20))
        reads.append(rnd(int(rng.integers(20, 100))))
    out.append(reads)
    # short reads that lie on one axis throughout, in a row (no read holds a run of 21), then a lower move
    reads = [rnd(int(rng.integers(8, 20)), b"AT") for _ in range(12)] + [b"CCG" + rnd(30)]
    reads += [b"G" * 15 for _ in range(8)] + [b"A" + rnd(25), b"", b"N" * 5, b"T" * 300, b"C"]
    out.append(reads)
    return [pack([(r, ql(len(r))) for r in reads]) for reads in out]


@pytest.mark.parametrize("k", [2, 5, 10, 12])
def test_chaos_restatement_equals_reference_code_on_long_axis_runs(q, ob, k):
    """The literal restatement against the reference's own compiled code where the double state matters; the integer
    form must differ somewhere on these inputs (otherwise they would not exercise anything)."""
    if ob.ref_lib() is None:
        pytest.skip("oracle/_ref not built (reference tree absent on this machine)")
    differs = False
    for seq, qual, idx in axis_run_batches(3):
        ts, tq, words = ob.chaos_literal(seq, qual, idx, k)
        rs, rq, rwords = ob.chaos_reference(seq, qual, idx, k)
        assert words == rwords
        assert np.array_equal(ts.astype(np.uint32), rs)
        assert np.array_equal((tq & 0xFFFFFFFF).astype(np.uint32), rq)
        try:
            its, itq, iw, _ = ob.chaos_int(seq, qual, idx, k)
            differs |= not np.array_equal(its, ts)
        except AssertionError:
            pass  # the integer twin declines reads with foreign bytes
    if k >= 5:
        assert differs
