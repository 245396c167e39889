"""Python restatement of the reference's report writer (src/stats_report.c) -- TEST INFRASTRUCTURE ONLY.

Written from the reference source, independently of hpg-fastq_b200/host/report.c, and used by the tests as the
checker of that file's output.  Input is the flat u64 counter block (oracle or GPU) plus the option strings.
Every float expression of the reference is float32 (`100.0f * a / b`, `1.0f * a / b`); numpy float32 scalars
reproduce them operation by operation.  Function by function:
  summary()            report_summary       src/stats_report.c:60-156
  length_histogram()   report_length        :160-203
  quality_histogram()  report_quality       :396-424 (its quality.per.nt.data is overwritten by report_nt_content)
  nt_content()         report_nt_content    :207-392
  kmers()              report_kmers         :489-535
Formats with a stray '%' ("%0.2f %\n", "%0.2f %)") print a literal '%' on glibc (SURVEY.md Appendix C probe 2).
"""
from __future__ import annotations

import math

import numpy as np

f32 = np.float32

SEC_NAMES = ["SCALARS", "LEN_HIST", "SUMQ_BY_LEN", "QUAL_HIST", "GC_HIST", "POS_ACCQ", "POS_A", "POS_C", "POS_G",
             "POS_T", "POS_N", "K5", "K5_POS", "K2", "CG_SEQ", "CG_Q"]
S = {n: i for i, n in enumerate(["NUM_READS", "NUM_PASSED", "NUM_FAILED", "NUM_EDITED", "ACC_LENGTH", "NUM_A", "NUM_C",
                                 "NUM_G", "NUM_T", "NUM_N", "MIN_LENGTH", "MAX_LENGTH"])}
MIN_VALUE, MAX_VALUE = 0, 100000


def _round(x) -> int:
    """C round(): half away from zero, on the double value of a float32."""
    x = float(x)
    return int(math.floor(abs(x) + 0.5) * (1 if x >= 0 else -1))


def _fmt2(x) -> str:
    """printf("%0.2f") of a float32 promoted to double."""
    return "%0.2f" % float(x)


def _pct(a, b):
    """100.0f * a / b in float32 (size_t operands converted to float first)."""
    with np.errstate(divide="ignore", invalid="ignore"):
        return f32(100.0) * f32(a) / f32(b)


def _as_i(v) -> int:
    """printf("%i") of a size_t: low 32 bits, signed."""
    v = int(v) & 0xFFFFFFFF
    return v - (1 << 32) if v >= (1 << 31) else v


def kmer_string(i: int, k: int = 5) -> str:
    return "".join("ACGT"[(i >> (2 * (k - 1 - j))) & 3] for j in range(k))


class Report:
    def __init__(self, lay, block, name, phred=33, kmers_on=False, filter_on=False, opts=None):
        self.lay, self.blk, self.name, self.phred = lay, np.asarray(block, dtype=np.uint64), name, phred
        self.kmers_on, self.filter_on = kmers_on, filter_on
        self.opts = opts or {}
        self.max_length = int(self.blk[S["MAX_LENGTH"]])
        lh = self.sec("LEN_HIST")
        # kh_count_quality_per_nt[j] = number of merged reads longer than j (src/stats_fastq.c:342-348)
        self.count = np.zeros(self.max_length + 1, dtype=np.uint64)
        run = 0
        for j in range(self.max_length - 1, -1, -1):
            run += int(lh[j + 1])
            self.count[j] = run

    def sec(self, name):
        i = SEC_NAMES.index(name)
        return self.blk[int(self.lay.offset[i]):int(self.lay.offset[i + 1])]

    def _norm_q(self, acc, count) -> int:
        with np.errstate(divide="ignore", invalid="ignore"):
            return _round(f32(1.0) * f32(acc) / f32(count) - f32(self.phred))

    def sorted_kmers(self):
        k5 = self.sec("K5")
        ids = sorted(range(1024), key=lambda i: (-int(k5[i]), i)) if len(k5) else list(range(1024))
        return [(i, kmer_string(i), int(k5[i]) if len(k5) else 0) for i in ids]

    # ------------------------------------------------------------------ summary.txt
    def summary(self) -> str:
        b, o = self.blk, self.opts
        nr, npass, nfail = int(b[S["NUM_READS"]]), int(b[S["NUM_PASSED"]]), int(b[S["NUM_FAILED"]])
        nA, nC, nG, nT, nN = (int(b[S[k]]) for k in ("NUM_A", "NUM_C", "NUM_G", "NUM_T", "NUM_N"))
        nn = nA + nC + nG + nT + nN
        out = ["-----------------------------------\n", "      FastQ quality report\n",
               "-----------------------------------\n", f"FastQ filename: {self.name}\n", "\n"]
        if self.filter_on:
            out.append("Filter options:\n")
            if o.get("read_length_range"):
                out.append(f"\tRead length range   : {o['read_length_range']}\n")
            if o.get("read_quality_range"):
                out.append(f"\tRead quality range  : {o['read_quality_range']}\n")
            if o.get("left_length", MIN_VALUE) != MIN_VALUE and o.get("left_quality_range"):
                out.append(f"\tLeft length         : {o['left_length']} nucleotides\n")
                out.append(f"\tLeft quality range  : {o['left_quality_range']}\n")
            if o.get("right_length", MIN_VALUE) != MIN_VALUE and o.get("right_quality_range"):
                out.append(f"\tRight length        : {o['right_length']} nucleotides\n")
                out.append(f"\tRight quality range : {o['right_quality_range']}\n")
            if o.get("max_N", MAX_VALUE) != MAX_VALUE:
                out.append(f"\tMax. number of Ns   : {o['max_N']}\n")
            if o.get("max_out_of_quality", MAX_VALUE) != MAX_VALUE and o.get("read_quality_range"):
                out.append(f"\tMax. out of quality : {o['max_out_of_quality']} nucletotides\n")
            out.append("\n")
            out.append(f"Number of reads in file  : {npass + nfail}\n")
            out.append(f"Number of processed reads: {nr} ({_fmt2(_pct(nr, npass + nfail))} %)\n")
        else:
            out.append("Filter         : Disabled\n")
            out.append(f"Number of reads: {nr}\n")
        out.append("\n")
        with np.errstate(divide="ignore", invalid="ignore"):
            mean_len = f32(1.0) * f32(int(b[S["ACC_LENGTH"]])) / f32(nr)
        out.append(f"Read length (min., mean, max.): ({int(b[S['MIN_LENGTH']])}, {_fmt2(mean_len)}, {self.max_length})\n")
        out.append("\n")
        # acc_quality = sum over reads of the float mean raw quality (src/stats_fastq.c:297); exact form: sum_L SUMQ[L]/L
        lh, sq = self.sec("LEN_HIST"), self.sec("SUMQ_BY_LEN")
        accq = 0.0
        for L in range(1, self.max_length + 1):
            if int(lh[L]):
                accq += float(int(sq[L])) / float(L)
        qual = self._norm_q(f32(accq), nr)
        out.append(f"Mean quality = {qual} [{chr(qual + self.phred)}]\n")
        out.append("\n")
        out.append("Nucleotide content (A, C, G, T, N)\n")
        for tag, v in (("A", nA), ("T", nT), ("G", nG), ("C", nC), ("N", nN)):
            out.append(f"\t{tag}: {_fmt2(_pct(v, nn))} %\n")
        out.append("GC content\n")
        out.append(f"\tCG: {_fmt2(_pct(nG + nC, nn))} %\n")
        out.append("\n")
        out.append("Mean quality per nucleotide position\n")
        acc = self.sec("POS_ACCQ")
        for k in range(self.max_length):
            q = self._norm_q(int(acc[k]), int(self.count[k]))
            out.append(f"\tpos. {k + 1}: {q} [{chr(q + self.phred)}]\t")
            if (k + 1) % 5 == 0:
                out.append("\n")
        out.append("\n")
        if self.kmers_on:
            out.append("K-mers (top 20)\n")
            out.append("\tSequence\tCount\n")
            for _, s, c in self.sorted_kmers()[:21]:  # the reference's loop bound is 21 (src/stats_report.c:147)
                out.append(f"\t{s}\t\t{c}\n")
        return "".join(out)

    # ------------------------------------------------------------------ data files
    def length_histogram(self) -> str:
        lh = self.sec("LEN_HIST")
        return "".join(f"{i}\t{_as_i(lh[i])}\n" for i in range(1, self.max_length + 1))

    def quality_histogram(self) -> str:
        qh = self.sec("QUAL_HIST")
        keys = [k for k in range(256) if int(qh[k])]
        if not keys:
            return ""
        return "".join(f"{i - self.phred}\t{_as_i(qh[i])}\n" for i in range(min(keys), max(keys) + 1))

    def gc_histogram(self) -> str:
        gc = self.sec("GC_HIST")
        return "".join(f"{i}\t{_as_i(gc[i])}\n" for i in range(1, 100) if int(gc[i]))

    def _nt(self, k):
        return tuple(int(self.sec(n)[k]) for n in ("POS_A", "POS_T", "POS_G", "POS_C", "POS_N"))

    def gc_per_nt(self) -> str:
        out = []
        for k in range(self.max_length):
            a, t, g, c, n = self._nt(k)
            v = _pct(g + c, a + t + g + c + n)
            if v > f32(1.0):
                out.append(f"{k + 1}\t{_fmt2(v)}\n")
        return "".join(out)

    def quality_per_nt(self) -> str:
        acc = self.sec("POS_ACCQ")
        return "".join(f"{k}\t{self._norm_q(int(acc[k]), int(self.count[k]))}\n" for k in range(self.max_length))

    def nucleotides(self) -> str:
        out = []
        for k in range(self.max_length):
            a, t, g, c, n = self._nt(k)
            tot = a + t + g + c + n
            out.append(f"{k + 1}\t" + "\t".join(_fmt2(_pct(v, tot)) for v in (a, t, g, c, n)) + "\n")
        return "".join(out)

    def kmers_txt(self) -> str:
        return "# Sequence\tCount\n" + "".join(f"{s}\t{c}\n" for _, s, c in self.sorted_kmers())

    def kmers_per_nt(self) -> str:
        top = self.sorted_kmers()[:5]
        kp = self.sec("K5_POS")
        Lcap = int(self.lay.max_read_len)
        # counter_by_pos has one entry per read position; a k-mer never seen has size 0 (src/stats_fastq.c:396-405)
        num_cols = self.max_length if any(c for _, _, c in top) else 0
        out = []
        for i in range(num_cols):
            out.append(f"{i + 1}\t" + "\t".join(str(int(kp[kid * Lcap + i])) if c else "0" for kid, _, c in top) + "\n")
        return "".join(out)

    def files(self) -> dict:
        d = {"summary.txt": self.summary(), "length.histogram.data": self.length_histogram(),
             "read.quality.histogram.data": self.quality_histogram(), "GC.histogram.data": self.gc_histogram(),
             "GC.per.nt.data": self.gc_per_nt(), "quality.per.nt.data": self.quality_per_nt(),
             "nucleotides.data": self.nucleotides()}
        if self.kmers_on:
            d["kmers.txt"] = self.kmers_txt()
            d["kmers.per.nt.data"] = self.kmers_per_nt()
        return {f"{self.name}.{k}": v for k, v in d.items()}


def fastq_text(headers, seq, qual, idx, order=None, trim=None) -> bytes:
    """FASTQ records (header, sequence, '+', quality) of reads `order` (default all)."""
    out = []
    n = len(idx) - 1
    for i in (range(n) if order is None else order):
        o, e = int(idx[i]), int(idx[i + 1])
        if trim is not None:
            o, e = o + int(trim[i][0]), o + int(trim[i][0]) + int(trim[i][1])
        out.append(headers[i] + b"\n" + bytes(seq[o:e]) + b"\n+\n" + bytes(qual[o:e]) + b"\n")
    return b"".join(out)
