"""The C host program (host/uvaia_b200_main.c -> uvaia_b200/bin/uvaia_b200): the `uvaia` command line over the C ABI.

CPU: option handling and the loud failure without a GPU.  GPU: the program is run on the golden cases exactly like the
reference binary was when the goldens were made (same options, same xz inputs) and its two output files are compared
with the reference's: CSV byte for byte, alignment dump by record names and sequences."""
import lzma
import os
import subprocess

import pytest

from helpers import ROOT, Case, case_names

BIN = os.path.join(ROOT, "uvaia_b200", "bin", "uvaia_b200")
NON_H1 = [n for n in case_names() if not Case(n).h1]


def run(args, cwd=None):
    return subprocess.run([BIN] + args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, cwd=cwd, timeout=600)


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
@pytest.mark.parametrize("name", NON_H1)
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
