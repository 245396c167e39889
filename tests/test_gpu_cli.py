"""GPU tests of the `hpg-fastq` program (C host over the C ABI): the report files, passed/failed/edit FASTQ files
and RESULTS blocks of `stats`, `filter` and `edit` against the CPU oracle + the Python restatement of
src/stats_report.c, byte for byte, on a synthetic FASTQ file."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import host_binding as hb
import report_oracle as ro

pytestmark = pytest.mark.gpu

N_READS = 30000


@pytest.fixture(scope="module")
def fastq(q, tmp_path_factory):
    d = tmp_path_factory.mktemp("cli")
    cfg = q.synth_cfg((35, 150), seed=2024)
    seq, qual, idx = q.synth_host(cfg, 0, N_READS)
    hdr = [b"@SYN." + str(i).encode() + b" " + str(i).encode() + b"/1" for i in range(N_READS)]
    path = str(d / "reads.fq")
    hb.write_fastq(path, hdr, seq, qual, idx)
    return dict(dir=d, path=path, seq=seq, qual=qual, idx=idx, hdr=hdr)


def run_cli(fastq, command, args, sub):
    out = fastq["dir"] / sub
    out.mkdir(exist_ok=True)
    r = subprocess.run([hb.CLI, command, "-f", fastq["path"], "-o", str(out)] + args, capture_output=True, text=True,
                       cwd=str(out), timeout=600)
    assert r.returncode == 0, f"{command} {args}: rc={r.returncode}\n{r.stdout[-2000:]}\n{r.stderr[-2000:]}"
    assert (out / "hpg-fastq.log").exists()
    return out, r.stdout


def oracle_for(q, ob, fastq, command, args, Lcap=192):
    o = hb.Options(command, ["-f", fastq["path"]] + args)
    assert o.valid and o.runnable
    par = o.params(q.Params)
    if command == "stats" and (par.work & q.WORK_FILTER):
        hb.lib().hq_options_fill_unset(o.ptr)
        par = o.params(q.Params)
    o.close()
    kmers = bool(par.work & q.WORK_KMERS)
    lay = ob.layout(q.Layout, Lcap, kmers=int(kmers), k2=par.k2 if kmers else 0, cg_k=par.cg_k)
    blk, mask, trim = ob.process_batch(par, lay, fastq["seq"], fastq["qual"], fastq["idx"])
    return par, lay, blk, mask, trim


def arg_opts(args):
    opts = {}
    for k in ("read-length-range", "read-quality-range", "left-quality-range", "right-quality-range"):
        if "--" + k in args:
            opts[k.replace("-", "_")] = args[args.index("--" + k) + 1]
    for k in ("left-length", "right-length", "max-N", "max-out-of-quality"):
        if "--" + k in args:
            opts[k.replace("-", "_")] = int(args[args.index("--" + k) + 1])
    return opts


STATS_CASES = [
    ("plain", []),
    ("kmers_small_batches", ["--kmers", "--batch-size", "1000"]),
    ("filter", ["--read-length-range", "50,", "--read-quality-range", "20,", "--max-N", "5"]),
    ("filter_windows", ["--read-length-range", "40,140", "--read-quality-range", "18,40", "--max-out-of-quality", "30",
                        "--left-length", "10", "--left-quality-range", "20,", "--right-length", "12",
                        "--right-quality-range", "15,45", "--kmers", "--batch-size", "7777"]),
]


@pytest.mark.parametrize("name,args", STATS_CASES, ids=[c[0] for c in STATS_CASES])
def test_stats_report_files(q, ob, fastq, name, args):
    out, stdout = run_cli(fastq, "stats", args, "stats_" + name)
    par, lay, blk, mask, _ = oracle_for(q, ob, fastq, "stats", args)
    filter_on = bool(par.work & q.WORK_FILTER)
    rep = ro.Report(lay, blk, "reads.fq", kmers_on="--kmers" in args, filter_on=filter_on, opts=arg_opts(args))
    exp = rep.files()
    got = {p.name: p.read_text() for p in out.iterdir() if p.name.startswith("reads.fq.") and not p.name.endswith(".gnuplot")}
    assert sorted(got) == sorted(exp)
    for k in exp:
        assert got[k] == exp[k], f"{k} differs from the restated reference report"
    if filter_on:
        npass = int(mask.sum())
        assert f"\tSo, statistics were computed for {npass} of {N_READS} reads.\n" in stdout
        assert 0 < npass < N_READS
    else:
        assert "Filtering: disabled" in stdout


def test_stats_extensions_k11_and_chaos(q, ob, fastq):
    """--kmers2 11 / --chaos-game 7: tables in the extension files equal the oracle's sections."""
    args = ["--kmers", "--kmers2", "11", "--chaos-game", "7"]
    out, _ = run_cli(fastq, "stats", args, "stats_ext")
    par, lay, blk, _, _ = oracle_for(q, ob, fastq, "stats", args)
    k2 = lay.section(blk, "K2")
    lines = (out / "reads.fq.kmers.k11.txt").read_text().splitlines()[1:]
    got = {l.split("\t")[0]: int(l.split("\t")[1]) for l in lines}
    nz = np.nonzero(k2)[0]
    assert len(got) == len(nz)
    for i in nz[:: max(1, len(nz) // 500)]:
        assert got[ro.kmer_string(int(i), 11)] == int(k2[i])
    assert sum(got.values()) == int(k2.sum())
    freq = np.loadtxt(out / "reads.fq.chaos.k7.freq.data", dtype=np.uint64)
    qual = np.loadtxt(out / "reads.fq.chaos.k7.quality.data", dtype=np.uint64)
    assert np.array_equal(freq.reshape(-1), lay.section(blk, "CG_SEQ")[:128 * 128])
    assert np.array_equal(qual.reshape(-1), lay.section(blk, "CG_Q")[:128 * 128])


@pytest.mark.parametrize("k", [6, 8])
def test_chaos_game_images_equal_the_reference_code(q, ob, fastq, k):
    """--chaos-game K --chaos-game-signature F: the three PGM files are the ones the reference's own
    chaos_game_write_table_images (old/chaos_game.c:407-465, oracle/_ref) writes from the same tables."""
    import shutil
    import struct
    import tempfile
    ref = ob.ref_lib()
    if ref is None:
        pytest.skip("oracle/_ref was not built")
    dim = 1 << k
    work = tempfile.mkdtemp(prefix="cg", dir="/tmp")  # short paths: the reference formats names into 100-byte buffers
    try:
        rng = np.random.default_rng(k)
        gs = rng.integers(0, 40, size=(dim, dim), dtype=np.uint32)
        gs_path = os.path.join(work, "ref.gs")
        with open(gs_path, "wb") as f:
            f.write(b"ref.gs".ljust(180, b"\0") + struct.pack("<4I", k, dim, dim, int(gs.sum())))
            f.write(gs.astype("<u4").tobytes())
        args = ["--chaos-game", str(k), "--chaos-game-signature", gs_path]
        out, _ = run_cli(fastq, "stats", args, f"stats_cgimg{k}")
        par, lay, blk, _, _ = oracle_for(q, ob, fastq, "stats", args)
        seq = lay.section(blk, "CG_SEQ")[:dim * dim].astype(np.uint32)
        tq = lay.section(blk, "CG_Q")[:dim * dim].astype(np.uint32)
        rdif = np.zeros(dim * dim, dtype=np.int32)
        hi, lo, rw = C.c_int(0), C.c_int(0), C.c_uint(0)
        refdir = os.path.join(work, "r")
        os.mkdir(refdir)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        assert ref.ref_chaos_post(C.c_int(k), p(seq), p(tq), C.c_uint(int(blk[q.S["CG_WORDS"]])), gs_path.encode(),
                                  fastq["path"].encode(), refdir.encode(), p(rdif), C.byref(hi), C.byref(lo), C.byref(rw)) == 0
        names = sorted(os.listdir(refdir))
        assert names == [f"reads.fq_k={k}_FG.pgm", f"reads.fq_k={k}_FG_dif.pgm", f"reads.fq_k={k}_QQ.pgm"]
        for n in names:
            assert open(os.path.join(refdir, n), "rb").read() == (out / n).read_bytes(), n
        txt = (out / f"reads.fq.chaos.k{k}.diff.txt").read_text()
        assert f"Interval of variation of diff matrix values = [{hi.value}, {lo.value}]" in txt
    finally:
        shutil.rmtree(work, ignore_errors=True)


FILTER_ARGS = ["--read-length-range", "50,", "--read-quality-range", "20,", "--max-N", "5"]


@pytest.mark.parametrize("extra", [[], ["--batch-size", "999"]], ids=["one_batch", "many_batches"])
def test_filter_outputs(q, ob, fastq, extra):
    out, stdout = run_cli(fastq, "filter", FILTER_ARGS + extra, "filter_" + str(len(extra)))
    _, _, _, mask, _ = oracle_for(q, ob, fastq, "filter", FILTER_ARGS)
    passed = np.nonzero(mask)[0].tolist()
    failed = np.nonzero(mask == 0)[0].tolist()
    assert passed and failed
    f = fastq
    assert (out / "passed.fq").read_bytes() == ro.fastq_text(f["hdr"], f["seq"], f["qual"], f["idx"], order=passed)
    assert (out / "failed.fq").read_bytes() == ro.fastq_text(f["hdr"], f["seq"], f["qual"], f["idx"], order=failed)
    assert f"Num. passed reads: {len(passed)} ({out}/passed.fq)\n" in stdout
    assert f"Num. failed reads: {len(failed)} ({out}/failed.fq)\n" in stdout


EDIT_ARGS = ["--left-length", "10", "--left-quality-range", "34,", "--right-length", "10", "--right-quality-range", "22,"]


@pytest.mark.parametrize("filt", [[], FILTER_ARGS + ["--batch-size", "4096"]], ids=["edit_only", "edit_filter"])
def test_edit_outputs(q, ob, fastq, filt):
    out, stdout = run_cli(fastq, "edit", EDIT_ARGS + filt, "edit_" + str(len(filt)))
    par, _, blk, mask, trim = oracle_for(q, ob, fastq, "edit", EDIT_ARGS + filt)
    f = fastq
    lens = np.diff(f["idx"].astype(np.int64))
    n_edited = int((trim[:, 1] != lens).sum())
    assert 0 < n_edited < N_READS
    assert f"Num. edited reads : {n_edited}\n" in stdout
    if filt:
        passed = np.nonzero(mask)[0].tolist()
        failed = np.nonzero(mask == 0)[0].tolist()
        assert passed and failed
        assert (out / "edit.fq").read_bytes() == ro.fastq_text(f["hdr"], f["seq"], f["qual"], f["idx"], order=passed, trim=trim)
        assert (out / "failed.fq").read_bytes() == ro.fastq_text(f["hdr"], f["seq"], f["qual"], f["idx"], order=failed, trim=trim)
        assert f"\tNum. passed reads : {len(passed)} ({out}/edit.fq)\n" in stdout
    else:
        assert (out / "edit.fq").read_bytes() == ro.fastq_text(f["hdr"], f["seq"], f["qual"], f["idx"], trim=trim)
        assert not (out / "failed.fq").exists()


def test_multi_gpu_round_robin_same_files(q, ob, fastq):
    """--gpus 2: batches dealt round-robin, counters merged by one NCCL allreduce, outputs in input order."""
    n = C.c_int(0)
    q.load_library().hpgq_device_count(C.byref(n))
    if n.value < 2:
        pytest.skip("needs 2 GPUs")
    args = ["--kmers", "--batch-size", "1500", "--read-quality-range", "20,"]
    one, _ = run_cli(fastq, "stats", args, "mg_stats1")
    two, _ = run_cli(fastq, "stats", args + ["--gpus", "2"], "mg_stats2")
    for p in one.iterdir():
        if p.name.startswith("reads.fq.") and not p.name.endswith(".gnuplot"):
            assert (two / p.name).read_bytes() == p.read_bytes(), p.name
    f1, _ = run_cli(fastq, "filter", FILTER_ARGS + ["--batch-size", "1500"], "mg_f1")
    f2, _ = run_cli(fastq, "filter", FILTER_ARGS + ["--batch-size", "1500", "--gpus", "2"], "mg_f2")
    assert (f1 / "passed.fq").read_bytes() == (f2 / "passed.fq").read_bytes()
    assert (f1 / "failed.fq").read_bytes() == (f2 / "failed.fq").read_bytes()


def _files(d):
    # the .gnuplot scripts embed the output directory name
    return {f: open(os.path.join(d, f), "rb").read() for f in sorted(os.listdir(d))
            if f != "hpg-fastq.log" and not f.endswith(".gnuplot")}


@pytest.mark.parametrize("command,args", [("stats", ["--kmers", "--batch-size", "4000"]),
                                          ("filter", ["--read-length-range", "50,", "--max-N", "3", "--batch-size", "5000"]),
                                          ("edit", ["--left-length", "8", "--left-quality-range", "25,", "--batch-size", "3000"])])
def test_device_parser_and_host_reader_give_the_same_files(q, fastq, command, args):
    """Default: records are cut on the GPU (hpgq_batch_submit_text); --host-parse: by hq_reader_next.  Same files."""
    dev, _ = run_cli(fastq, command, args, f"devparse_{command}")
    host, _ = run_cli(fastq, command, args + ["--host-parse"], f"hostparse_{command}")
    log = open(dev / "hpg-fastq.log").read()
    assert " 0 device-parsed" not in log and "device-parsed + 0 host-parsed" in log, log
    assert "0 device-parsed" in open(host / "hpg-fastq.log").read()
    a, b = _files(dev), _files(host)
    assert a.keys() == b.keys() and len(a) > 0
    for k in a:
        assert a[k] == b[k], k


def test_device_parser_hands_irregular_input_back(q, fastq, tmp_path):
    """CRLF line ends are cut on the device; a blank line in the middle of the file sends the rest to the host reader."""
    n = 6000
    idx = fastq["idx"][:n + 1]
    path = str(tmp_path / "crlf.fq")
    data = hb.write_fastq(path, fastq["hdr"][:n], fastq["seq"], fastq["qual"], idx, eol=b"\r\n")
    cut = data.index(b"@SYN.4100 ")
    with open(path, "wb") as f:
        f.write(data[:cut] + b"\r\n" + data[cut:] + b"\r\n\r\n")
    outs = {}
    for mode, extra in (("dev", []), ("host", ["--host-parse"])):
        out = tmp_path / mode
        out.mkdir()
        r = subprocess.run([hb.CLI, "filter", "-f", path, "-o", str(out), "--read-quality-range", "20,", "--batch-size", "1500"]
                           + extra, capture_output=True, text=True, cwd=str(out), timeout=600)
        assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
        outs[mode] = _files(out)
    log = open(tmp_path / "dev" / "hpg-fastq.log").read()
    assert "handed the input back" in log and " 0 device-parsed" not in log
    assert outs["dev"] == outs["host"]
    assert outs["dev"]["passed.fq"].count(b"\n+\n") + outs["dev"]["failed.fq"].count(b"\n+\n") == n


def test_paired_end_invocation_equals_two_single_runs(q, fastq, tmp_path):
    """SURVEY.md 8f rank 4, as far as the reference pins it: `--fq1 A --fq2 B` runs both mate files in one invocation
    and "the outputs are calculated seperately for both files" (old/README:181): per-mate .valid / .invalid files
    (old/README:78,115-118) equal passed.fq / failed.fq of single-file runs, per-mate report files equal the
    single-file reports; the legacy action flags --filter / --qc select the command."""
    f = fastq
    n2 = 9000
    m2 = str(tmp_path / "mate_2.fq")
    hb.write_fastq(m2, [h.replace(b"/1", b"/2") for h in f["hdr"][:n2]], f["seq"], np.clip(f["qual"].astype(np.int32) - 2, 33, 126).astype(np.uint8),
                   f["idx"][:n2 + 1])
    singles = {}
    for tag, path in (("a", f["path"]), ("b", m2)):
        out = tmp_path / ("single_" + tag)
        out.mkdir()
        r = subprocess.run([hb.CLI, "filter", "-f", path, "-o", str(out)] + FILTER_ARGS, capture_output=True, text=True, cwd=str(out))
        assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
        singles[tag] = out
    out = tmp_path / "paired"
    out.mkdir()
    r = subprocess.run([hb.CLI, "--filter", "--fq1", f["path"], "--fastq2", m2, "-o", str(out)] + FILTER_ARGS, capture_output=True,
                       text=True, cwd=str(out))
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    for tag, path in (("a", f["path"]), ("b", m2)):
        base = os.path.basename(path)
        assert (out / (base + ".valid")).read_bytes() == (singles[tag] / "passed.fq").read_bytes()
        assert (out / (base + ".invalid")).read_bytes() == (singles[tag] / "failed.fq").read_bytes()
    assert r.stdout.count("Num. passed reads:") == 2
    # stats (--qc): per-mate report files, identical to the single-file runs
    reps = {}
    for tag, args in (("one_a", ["stats", "-f", f["path"]]), ("one_b", ["stats", "-f", m2]), ("two", ["--qc", "--fq1", f["path"], "--fq2", m2])):
        d = tmp_path / ("stats_" + tag)
        d.mkdir()
        r = subprocess.run([hb.CLI] + args + ["-o", str(d), "--kmers"], capture_output=True, text=True, cwd=str(d))
        assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
        reps[tag] = _files(d)
    both = dict(reps["one_a"])
    both.update(reps["one_b"])
    assert reps["two"] == both and len(both) == len(reps["one_a"]) + len(reps["one_b"])
