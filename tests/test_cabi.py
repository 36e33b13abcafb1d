"""CPU: the C-ABI library loads and exports every symbol include/fbg_cuda.h declares;
host-side (no GPU) behaviour of the drop-in CLI: flag validation, messages, exit codes."""
import os
import re
import subprocess

import pytest

import founderblockgraphs_b200 as pkg
from util import FIXTURES, ROOT, golden_cases


def header_symbols():
    text = open(os.path.join(ROOT, "include", "fbg_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fbg_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = pkg.load_library()
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), s
    assert sorted(pkg.api.EXPORTS) == syms


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.FbgError):
        pkg.Context(0)


def run_cli(args):
    pr = pkg.main(args, capture_output=True)
    return pr.returncode, pr.stderr.decode("utf8", "replace"), pr.stdout.decode()


def test_cli_flag_validation_matches_reference(tmp_path):
    inp = os.path.join(FIXTURES, "msa.fasta")
    out = str(tmp_path / "o.xgfa")
    # messages and exit codes captured from the reference binary (golden cases) + fbg.cpp:3308-3334
    for c in golden_cases():
        if c["fixture"] != "msa.fasta" or c["flags"] not in (["--gfa"], ["--elastic"]):
            continue
        code, err, _ = run_cli(["--input", inp, "--output", out] + c["flags"])
        assert code == c["exit"] == 1 and err.strip() == c["stderr"][0]
    code, err, _ = run_cli(["--input", inp, "--output", out, "--gap-limit=-2"])
    assert code == 1 and err == "Gap limit needs to be non-negative.\n"
    code, err, _ = run_cli(["--input", inp, "--output", out, "-p"])
    assert code == 1 and err == "Output of original sequences as paths without option --elastic is not implemented!\n"
    code, err, _ = run_cli(["--input", inp, "--output", out, "--elastic", "--gfa", "--heuristic-subset=0"])
    assert code == 1 and err == "wrong value for --heuristic-subset!\n"
    code, err, _ = run_cli(["--output", out])
    assert code == 1 and "'--input' option required" in err
    code, err, _ = run_cli(["--input", inp])
    assert code == 1 and "'--output' option required" in err
    code, err, _ = run_cli(["--input", inp, "--output", out, "--gap-limit=x"])
    assert code == 1 and "invalid numeric value: x" in err
    code, err, _ = run_cli(["--input", inp, "--outp", out])
    assert code == 1 and "ambiguous" in err
    code, err, so = run_cli(["--version"])
    assert code == 0 and so == "founderblockgraphs 0.5\n"
    code, err, so = run_cli(["--help"])
    assert code == 0 and so.startswith("Usage: founderblockgraph --input=MSA.fasta") and "--heuristic-subset" not in so
    code, err, so = run_cli(["--full-help"])
    assert code == 0 and "--heuristic-subset" in so


def test_cli_row_filter_messages(tmp_path):
    # gap-limit=1 on test/msa.fasta: every row has a gap -> 4 NOTICE lines + "Unable to read" 
    inp = os.path.join(FIXTURES, "msa.fasta")
    case = [c for c in golden_cases() if c["fixture"] == "msa.fasta" and c["flags"] == ["--gap-limit=1"]][0]
    code, err, _ = run_cli(["--input", inp, "--output", str(tmp_path / "x"), "--gap-limit=1"])
    assert code == 1
    assert [l for l in err.splitlines() if l.startswith(("NOTICE", "Unable"))] == [l for l in case["stderr"] if l.startswith(("NOTICE", "Unable"))]


def test_python_read_input_mirror():
    from oracle import oracle as O
    for fx in sorted(os.listdir(FIXTURES)):
        for gl, el in ((1, True), (1, False), (0, False), (3, False)):
            rows, ids, _ = O.read_fasta(os.path.join(FIXTURES, fx), gap_limit=gl, elastic=el, output_paths=True)
            msa, ids2 = pkg.read_input(os.path.join(FIXTURES, fx), gl, el, True)
            assert [bytes(r) for r in msa] == rows and ids2 == ids
