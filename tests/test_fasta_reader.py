"""The library's threaded FASTA reader (csrc/host/fasta_reader.cpp, what the executable uses)
against the oracle's restatement of read_input (fbg.cpp:103-201): rows, header list and the
WARNING / NOTICE lines, on the reference's fixtures and on awkward files."""
import os

import numpy as np
import pytest

import founderblockgraphs_b200 as pkg
from founderblockgraphs_b200 import api
from oracle import oracle as O
from util import FIXTURES

MODES = [(1, True), (1, False), (0, False), (3, False), (2, False)]


def check(path, threads=(0, 1, 3, 16)):
    for gl, el in MODES:
        for paths in (False, True):
            rows, ids, msgs = O.read_fasta(path, gap_limit=gl, elastic=el, output_paths=paths)
            for t in threads:
                msa, ids2, msgs2 = api.read_fasta(path, gl, el, paths, threads=t)
                if rows and len(rows[0]):
                    assert [bytes(r) for r in msa] == rows, (path, gl, el, t)
                else:
                    assert msa.size == 0
                assert ids2 == ids and msgs2 == msgs, (path, gl, el, t)


def test_reference_fixtures():
    for fx in sorted(os.listdir(FIXTURES)):
        check(os.path.join(FIXTURES, fx))


def test_awkward_files(tmp_path):
    cases = {
        "crlf": b">a\r\nAC-T\r\n>b\r\nACGT\r\n",
        "no_final_newline": b">a\nACGT\n>b\nAC--",
        "wrapped": b">a\nAC\nGT\nAA\n>b\nACG\nTAA\n>c\nA\nC\nG\nT\nA\nA\n",
        "gap_run_across_lines": b">a\nAC-\n-\n-T\n>b\nACGGGT\n>c\nA----T\n",
        "empty_lines": b">a\n\nAC\n\nGT\n\n>b\nACGT\n\n\n",
        "first_line_not_header": b"xyz\nACGT\n>b\nACGT\n",
        "header_only_last": b">a\nACGT\n>b\n",
        "mismatched": b">a\nACGT\n>b\nACG\n>c\nACGTA\n>d\nTTTT\n",
        "gt_inside_line": b">a\nAC>T\n>b\nA>GT\n",
        "all_filtered": b">a\nA--T\n>b\nA--T\n",
        "single": b">only\nACGTACGT",
        "header_then_eof": b">a",
        "blank_first": b"\nACGT\n>b\nACGT\n",
        "odd_names": b">s q\t|1| >x\nAC\n>\nA\n>>\nAC\n",
    }
    for name, data in cases.items():
        p = str(tmp_path / name)
        with open(p, "wb") as fh:
            fh.write(data)
        check(p)


@pytest.mark.parametrize("seed", range(6))
def test_random_files_all_thread_counts(tmp_path, seed):
    rng = np.random.default_rng(70 + seed)
    m, n = int(rng.integers(1, 60)), int(rng.integers(1, 400))
    width = int(rng.choice([0, 1, 7, 60]))
    rows = []
    for i in range(m):
        k = n if rng.random() > 0.1 else int(rng.integers(0, n + 3))
        r = np.frombuffer(b"ACGT-", np.uint8)[rng.choice(5, size=k, p=[0.22, 0.22, 0.22, 0.22, 0.12])]
        rows.append(r.tobytes())
    p = str(tmp_path / "r.fasta")
    O.write_fasta(p, rows, names=[b"row %d" % i for i in range(m)], width=width)
    check(p, threads=(1, 2, 5, 64))


def test_large_file_default_threads(tmp_path):
    """Above 4 MB the reader goes parallel on its own; chunk borders fall inside lines and headers."""
    rng = np.random.default_rng(3)
    m, n = 300, 20_000
    base = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, size=n)]
    p = str(tmp_path / "big.fasta")
    with open(p, "wb") as fh:
        for i in range(m):
            r = base.copy()
            r[rng.integers(0, n, size=20)] = 45
            fh.write(b">h%d some description\n" % i)
            b = r.tobytes()
            if i % 3 == 0:
                fh.write(b"\n".join(b[o:o + 70] for o in range(0, n, 70)) + b"\n")
            else:
                fh.write(b + b"\n")
    assert os.path.getsize(p) > (4 << 20)
    for gl, el in ((1, True), (2, False)):
        rows, ids, msgs = O.read_fasta(p, gap_limit=gl, elastic=el, output_paths=True)
        msa, ids2, msgs2 = api.read_fasta(p, gl, el, True)
        assert [bytes(r) for r in msa] == rows and ids2 == ids and msgs2 == msgs
        msa1, _ = pkg.read_input(p, gl, el, True)
        assert np.array_equal(msa, msa1)


def test_unreadable_file(tmp_path):
    with pytest.raises(api.FbgError):
        api.read_fasta(str(tmp_path / "does_not_exist.fasta"), 1, True)
