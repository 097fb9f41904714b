"""CPU tests of the `metabuli classify` command line (metabuli_work_b200/host/metabuli_main.cpp): the positional rules and
file checks of the reference's classify.cpp:45-139 are applied before any device work, and without a CUDA device the command
fails loudly instead of falling back to the CPU."""
import os
import subprocess

import pytest

import helpers

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "metabuli_work_b200", "metabuli")


def _run(*args):
    p = subprocess.run([CLI, *args], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=120)
    return p.returncode, p.stdout, p.stderr


@pytest.fixture(scope="module", autouse=True)
def _built():
    import metabuli_work_b200.build as b
    b.build()
    assert os.path.exists(CLI)


def test_usage_and_option_errors(golden, tmp_path):
    rc, _, err = _run()
    assert rc == 1 and "usage: metabuli classify" in err
    rc, _, err = _run("build")                                   # only `classify` is in scope
    assert rc == 1 and "usage" in err
    db, r1 = golden.dbdir("plain_paired"), os.path.join(helpers.DATA, "reads_1.fq")
    rc, _, err = _run("classify", r1, db, str(tmp_path), "job", "--no-such-flag", "1")
    assert rc == 1 and "unknown option --no-such-flag" in err
    rc, _, err = _run("classify", r1, db, str(tmp_path), "job", "--min-score")
    assert rc == 1 and "missing value for --min-score" in err


def test_positional_rules(golden, tmp_path):
    db = golden.dbdir("plain_paired")
    r1, r2 = os.path.join(helpers.DATA, "reads_1.fq"), os.path.join(helpers.DATA, "reads_2.fq")
    # --seq-mode 2 (the default) wants two read files, modes 1 and 3 one
    rc, _, err = _run("classify", r1, db, str(tmp_path), "job")
    assert rc == 1 and "expected 5 positional arguments for --seq-mode 2, got 4" in err
    rc, _, err = _run("classify", "--seq-mode", "1", r1, r2, db, str(tmp_path), "job")
    assert rc == 1 and "expected 4 positional arguments for --seq-mode 1, got 5" in err


def test_file_checks(golden, tmp_path):
    db = golden.dbdir("plain_paired")
    r1, r2 = os.path.join(helpers.DATA, "reads_1.fq"), os.path.join(helpers.DATA, "reads_2.fq")
    rc, _, err = _run("classify", r1, str(tmp_path / "nope.fq"), db, str(tmp_path), "job")
    assert rc == 1 and "nope.fq does not exist" in err
    empty = tmp_path / "emptydb"
    empty.mkdir()
    rc, _, err = _run("classify", r1, r2, str(empty), str(tmp_path), "job")
    assert rc == 1 and "diffIdx file is missing" in err
    # a DB directory without its taxonomy, and --taxonomy-path pointing somewhere without one
    part = tmp_path / "partdb"
    part.mkdir()
    for f in ("diffIdx", "info", "split"):
        os.symlink(os.path.join(db, f), part / f)
    rc, _, err = _run("classify", r1, r2, str(part), str(tmp_path), "job")
    assert rc == 1 and "taxonomyDB file is missing in " + str(part) in err
    rc, _, err = _run("classify", r1, r2, db, str(tmp_path), "job", "--taxonomy-path", str(empty))
    assert rc == 1 and "taxonomyDB file is missing in " + str(empty) in err


def test_no_cpu_fallback(golden, tmp_path):
    """Valid arguments on a machine without a CUDA device: an error, a non-zero exit code, no result files."""
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a CUDA device is present")
    except ImportError:
        pass
    db = golden.dbdir("plain_paired")
    r1, r2 = os.path.join(helpers.DATA, "reads_1.fq"), os.path.join(helpers.DATA, "reads_2.fq")
    out = tmp_path / "out"
    rc, _, err = _run("classify", r1, r2, db, str(out), "job")
    assert rc != 0 and "Error:" in err
    assert not os.path.exists(out / "job_classifications.tsv") and not os.path.exists(out / "job_report.tsv")
