"""The C host program (host/uvaia_b200_main.c -> uvaia_b200/bin/uvaia_b200): the `uvaia` command line over the C ABI.

CPU: option handling and the loud failure without a GPU.  GPU: the program is run on the golden cases exactly like the
reference binary was when the goldens were made (same options, same xz inputs) and its two output files are compared
with the reference's: CSV byte for byte, alignment dump by record names and sequences."""
import lzma
import os

import subprocess

import numpy as np
import pytest

from helpers import ROOT, Case, case_names

BIN = os.path.join(ROOT, "uvaia_b200", "bin", "uvaia_b200")
NON_H1 = [n for n in case_names() if not Case(n).h1]


def run(args, cwd=None, env=None):
    return subprocess.run([BIN] + args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, cwd=cwd, timeout=600, env=env)


def test_binary_is_built():
    assert os.path.exists(BIN), "build it with `make -C uvaia_b200/csrc` (or __graft_entry__.build())"


def test_help_version_and_usage_errors():
    p = run(["--version"])
    assert p.returncode == 0 and b"uvaia_b200" in p.stdout
    p = run(["--help"])
    assert p.returncode == 0 and b"--nbest" in p.stdout and b"--ref_ambiguity" in p.stdout
    assert run([]).returncode == 1                                   # no reference, no query
    assert run(["-r", "x.fa"]).returncode == 1                       # no query file
    assert run(["--no-such-option", "-r", "x.fa", "q.fa"]).returncode == 1
    p = run(["-r", "/nonexistent/r.fa", "/nonexistent/q.fa"])
    assert p.returncode == 1 and b"cannot open file" in p.stderr
    assert b"--keep_resolved" in run(["--help"]).stdout


def test_no_gpu_means_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("this box has a GPU")
    c = Case("syn_default")
    p = run(c.args + ["-r", c.ref_files[0], c.query_file])
    assert p.returncode == 1
    assert b"Query database composed of" in p.stderr                 # the queries were read and filtered on the host ...
    assert b"no CPU fallback" in p.stderr                            # ... and the search refuses to run without the device


def read_fasta_text(raw):
    recs, name = [], None
    for ln in raw.decode().split("\n"):
        if ln.startswith(">"):
            name = ln[1:]
            recs.append([name, ""])
        elif ln:
            recs[-1][1] += ln
    return recs


@pytest.mark.gpu
@pytest.mark.parametrize("name", case_names())
def test_host_program_reproduces_the_reference_outputs(name, tmp_path):
    c = Case(name)
    args = list(c.args)
    for rf in c.ref_files:
        args += ["-r", rf]
    prefix = str(tmp_path / "out")
    p = run(args + ["-o", prefix, "--batch", "257", c.query_file])
    assert p.returncode == 0, p.stderr.decode(errors="replace")[-2000:]
    csv = lzma.open(prefix + ".csv.xz", "rb").read().decode()
    assert csv == c.expected_csv
    aln = read_fasta_text(lzma.open(prefix + ".aln.xz", "rb").read())
    assert [r[0].encode() for r in aln] == c.expected_aln_names
    part = c.meta.get("reference_stderr_partition_line")
    if part:
        assert part[-1].encode() in p.stderr


@pytest.mark.gpu
def test_host_program_streams_with_drops(tmp_path):
    """a resident budget smaller than the database: push / search / drop cycles must not move the table"""
    c = Case("syn_n100")
    args = list(c.args)
    for rf in c.ref_files:
        args += ["-r", rf]
    prefix = str(tmp_path / "out")
    p = run(args + ["-o", prefix, "--batch", "100", "--max_resident", "300", c.query_file])
    assert p.returncode == 0, p.stderr.decode(errors="replace")[-2000:]
    assert lzma.open(prefix + ".csv.xz", "rb").read().decode() == c.expected_csv


@pytest.mark.gpu
@pytest.mark.parametrize("name,extra", [("syn_default", ["--batch", "700", "--max_resident", "1500"]), ("syn_ref_filter", []), ("syn_two_files", ["--batch", "333"]),
                                       ("amb_default", ["--batch", "512"]), ("syn_acgt", []), ("keep_default", ["--batch", "96"])])
def test_host_program_cache_gives_the_same_outputs(name, extra, tmp_path):
    """--cache: the first run writes the packed references, the second one reads them instead of the FASTA files; both
    must give the reference's table and alignment names.  A changed -A threshold makes the cache stale."""
    c = Case(name)
    args = list(c.args)
    for rf in c.ref_files:
        args += ["-r", rf]
    cache = str(tmp_path / "refs.uvg")
    for run_no in range(2):
        prefix = str(tmp_path / f"out{run_no}")
        p = run(args + extra + ["-o", prefix, "--cache", cache, c.query_file])
        assert p.returncode == 0, p.stderr.decode(errors="replace")[-2000:]
        assert (b"will be saved to the cache" if run_no == 0 else b"are read from the cache") in p.stderr
        assert lzma.open(prefix + ".csv.xz", "rb").read().decode() == c.expected_csv
        aln = read_fasta_text(lzma.open(prefix + ".aln.xz", "rb").read())
        assert [r[0].encode() for r in aln] == c.expected_aln_names
        assert os.path.exists(cache) and not os.path.exists(cache + ".tmp")
    if "-A" not in args:
        p = run(args + ["-A", "0.37", "-o", str(tmp_path / "out2"), "--cache", cache, c.query_file])
        assert p.returncode == 0 and b"will be saved to the cache" in p.stderr          # stale: written again
    p = run(args + ["-x", "-o", str(tmp_path / "out3"), "--cache", cache, c.query_file])
    assert p.returncode == 0 and b"--cache is ignored with -x" in p.stderr


def _messy_fasta(rng, n_rec, L):
    """records with the hazards the reader has to get right: CRLF, wrapped lines, blank and 1-byte lines, comments, text before
    the first header, blanks inside sequence lines, '>' inside a header line, lower case, empty records"""
    out = [b"ACGT stray text before any header\n", b"# a comment with > inside\n"]
    alphabet = np.frombuffer(b"ACGTNacgtn-RYKM", np.uint8)
    for i in range(n_rec):
        eol = b"\r\n" if i % 7 == 3 else (b"\r" if i % 11 == 5 else b"\n")
        hdr = b">rec%d|x" % i if i % 5 else b"  junk>rec%d after junk" % i
        out.append(hdr + eol)
        if i % 13 == 6:
            continue                                                     # header without sequence
        seq = alphabet[rng.integers(0, len(alphabet), size=L)].tobytes()
        width = int(rng.integers(7, 120)) if i % 3 else L
        for a in range(0, L, width):
            line = seq[a:a + width]
            if i % 9 == 4 and len(line) > 4:
                line = line[:2] + b" \t" + line[2:]
            out.append(line + eol)
            if i % 4 == 1 and a == 0:
                out.append(b"; comment inside a record" + eol + eol + b"A" + eol)    # comment, empty line, 1-byte line (dropped)
    return b"".join(out)


def test_parallel_parser_delivers_what_the_serial_one_does(tmp_path):
    """raw FASTA files are parsed by several threads over slices that begin at header lines; --parse_only prints a digest of
    the delivered records (no GPU needed): it must not depend on the number of threads or on where the slices are cut"""
    rng = np.random.default_rng(5)
    path = str(tmp_path / "messy.fa")
    open(path, "wb").write(_messy_fasta(rng, 900, 333))
    def digest(env):
        e = dict(os.environ); e.update(env)
        p = subprocess.run([BIN, "--parse_only", "-r", path], stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=e)
        assert p.returncode == 0, p.stderr
        f = p.stdout.decode().split()
        return f[1], f[2], f[3]
    serial = digest({"UVG_PARSE_THREADS": "1"})
    assert serial[2] == "parser=serial" and serial[0] == "records=900"
    for threads, window in (("2", "100000000"), ("5", "20000"), ("16", "3000"), ("3", "700"), ("32", "1")):
        got = digest({"UVG_PARSE_THREADS": threads, "UVG_PARSE_MIN_BYTES": "1", "UVG_PARSE_WINDOW": window})
        assert got[2] == "parser=parallel"
        assert got[:2] == serial[:2], (threads, window)
    # compressed input falls back to the serial reader
    import gzip as gz
    with gz.open(path + ".gz", "wb") as fh:
        fh.write(open(path, "rb").read())
    p = subprocess.run([BIN, "--parse_only", "-r", path + ".gz"], stdout=subprocess.PIPE, env=dict(os.environ, UVG_PARSE_MIN_BYTES="1"))
    f = p.stdout.decode().split()
    assert f[3] == "parser=serial" and (f[1], f[2]) == serial[:2]


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["hygiene", "syn_default", "syn_two_files", "syn_exclude_self", "syn_ref_filter", "real_c1_small"])
def test_host_program_parallel_parser_on_raw_files(name, tmp_path):
    """raw reference files go through the multi-threaded parser (forced here on small files, with windows of a few records):
    same table and alignment dump as from the compressed files"""
    c = Case(name)
    args = list(c.args)
    for i, rf in enumerate(c.ref_files):
        raw = str(tmp_path / f"r{i}.fa")
        open(raw, "wb").write(lzma.open(rf, "rb").read())
        args += ["-r", raw]
    prefix = str(tmp_path / "out")
    env = dict(os.environ, UVG_PARSE_THREADS="5", UVG_PARSE_MIN_BYTES="1", UVG_PARSE_WINDOW="50000", UVG_PARSE_TIMING="1")
    p = subprocess.run([BIN] + args + ["-o", prefix, "--batch", "257", c.query_file], stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env)
    assert p.returncode == 0, p.stderr.decode(errors="replace")[-2000:]
    assert b"parallel parse of" in p.stderr
    assert lzma.open(prefix + ".csv.xz", "rb").read().decode() == c.expected_csv
    aln = read_fasta_text(lzma.open(prefix + ".aln.xz", "rb").read())
    assert [r[0].encode() for r in aln] == c.expected_aln_names


# ---- `uvaia_b200 ball`: the radius search program (src/ball.c) ----
from test_oracle_ball import BallCase, ball_case_names  # noqa: E402


def test_ball_mode_usage():
    p = run(["ball", "--help"])
    assert p.returncode == 0 and b"--distance" in p.stdout
    assert run(["ball"]).returncode == 1
    assert run(["ball", "-r", "/nonexistent/r.fa", "/nonexistent/q.fa"]).returncode == 1


@pytest.mark.gpu
@pytest.mark.parametrize("name", ball_case_names())
def test_host_program_ball_mode_reproduces_uvaiaball(name, tmp_path):
    c = BallCase(name)
    args = ["ball"] + list(c.args)
    for rf in c.ref_files:
        args += ["-r", rf]
    prefix = str(tmp_path / "ball")
    p = run(args + ["-o", prefix, "--batch", "301", c.query_file])
    assert p.returncode == 0, p.stderr.decode(errors="replace")[-2000:]
    aln = read_fasta_text(lzma.open(prefix + ".aln.xz", "rb").read())
    assert [r[0].encode() for r in aln] == c.expected
    for ln in c.meta.get("reference_stderr", []):
        if "Query database now composed" in ln:
            assert ln.strip().encode() in p.stderr             # same query set after the pruning


# ---- --devices: the same program with the references row-sharded over several GPUs inside one process ----
@pytest.mark.gpu
@pytest.mark.parametrize("name", [n for n in NON_H1 if n not in ("real_c1_full_reversed",)])
def test_host_program_on_two_gpus(name, tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    c = Case(name)
    args = list(c.args)
    for rf in c.ref_files:
        args += ["-r", rf]
    prefix = str(tmp_path / "out")
    p = run(args + ["-o", prefix, "--batch", "257", "--devices", "0,1", c.query_file], env=dict(os.environ, UVG_MG_ROUND="64"))   # small rounds: both GPUs get rows
    assert p.returncode == 0, p.stderr.decode(errors="replace")[-2000:]
    assert b"row-sharded over 2 GPUs" in p.stderr
    csv = lzma.open(prefix + ".csv.xz", "rb").read().decode()
    assert csv == c.expected_csv
    aln = read_fasta_text(lzma.open(prefix + ".aln.xz", "rb").read())
    assert [r[0].encode() for r in aln] == c.expected_aln_names
