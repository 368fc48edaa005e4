"""Command line contract of `leon` (boundary (1) of SURVEY.md 8b) where scripts depend on it: messages and exit codes of the
reference for the argument errors (C/Leon.cpp:177-180, :427-431, src/main.cpp:43-49), the summary keys (:726-767), and
`-d -test-file` (:2035-2110).  Host only: decompression needs no GPU."""
import os
import subprocess

import numpy as np
import pytest

from container_util import result_from_golden, ROOT
from golden_util import load_case

CLI = os.path.join(ROOT, "leon_b200", "bin", "leon")


def _run(args, cwd=None):
    import __graft_entry__ as g
    g.build()
    return subprocess.run([CLI] + args, cwd=cwd, capture_output=True, text=True, timeout=300)


def test_installed_under_the_reference_name():
    _run(["-h"])
    assert os.path.exists(CLI) and os.path.exists(os.path.join(ROOT, "leon_b200", "bin", "leon_b200"))


def test_neither_or_both_modes_is_a_message_and_success(tmp_path):
    for extra in ([], ["-c", "-d"]):
        r = _run(["-file", str(tmp_path / "x.fasta")] + extra)
        assert r.returncode == 0
        assert r.stdout == "Choose one option among -c (compress) or -d (decompress)\n\n"


def test_unknown_extension_is_a_message_and_success(tmp_path):
    (tmp_path / "x.txt").write_text(">r\nACGT\n")
    r = _run(["-file", str(tmp_path / "x.txt"), "-c"])
    assert r.returncode == 0
    assert r.stdout.startswith("Input format: unknown. Input extension must be one among fasta (.fa, .fasta) or fastq (.fq, .fastq)")
    assert not (tmp_path / "x.txt.leon").exists()


def test_failures_are_exceptions_with_exit_failure(tmp_path):
    r = _run(["-file", str(tmp_path / "missing.leon"), "-d"])
    assert r.returncode == 1 and r.stderr.startswith("EXCEPTION: ")
    r = _run(["-file", str(tmp_path / "missing.fastq"), "-c"])
    assert r.returncode == 1 and r.stderr.startswith("EXCEPTION: ")


def test_decompress_summary_and_test_file(tmp_path):
    """`leon -file x.fasta.leon -d -test-file`: the reference's output name, its summary keys, and the check of the decoded
    file against <dir>/x.fasta: OK for the original, the reference's message for a different one."""
    import leon_b200
    c = load_case("fa_k31_blocks")
    offs = c["offsets"].astype(np.int64)
    n = len(offs) - 1
    names = ["read_%d/1" % i for i in range(n)]
    text = np.frombuffer("".join(names).encode(), np.uint8)
    hoffs = np.concatenate([[0], np.cumsum([len(x) for x in names])]).astype(np.uint64)
    h = leon_b200.encode_headers(text, hoffs, c["rpb"], 0)
    lf = str(tmp_path / "x.fasta.leon")
    leon_b200.write_container(lf, result_from_golden(c), c["k"], True, header=h)
    with open(tmp_path / "x.fasta", "wb") as f:
        for i in range(n):
            f.write(b">" + names[i].encode() + b"\n" + c["bases"][offs[i]:offs[i + 1]].tobytes() + b"\n")
    r = _run(["-file", lf, "-d", "-test-file", "-nb-cores", "2"])
    assert r.returncode == 0, r.stderr
    assert (tmp_path / "x.fasta.d").exists()
    for key in ("Input format", "# blocks", "# sequences", "# nucleotides", "Output filename", "Time", "Speed", "Checking decompressed file",
                "Original file", "New file"):
        assert key in r.stdout, key
    assert r.stdout.rstrip().endswith("OK")
    assert (tmp_path / "x.fasta.d").read_bytes() == (tmp_path / "x.fasta").read_bytes()
    # a different original: reported, still exit 0 (the reference only prints)
    data = (tmp_path / "x.fasta").read_bytes().replace(b"read_7/1", b"read_X/1")
    (tmp_path / "x.fasta").write_bytes(data)
    r = _run(["-file", lf, "-d", "-test-file"])
    assert r.returncode == 0 and "Sequence with a different header" in r.stdout and "OK" not in r.stdout.split("Checking")[1]
