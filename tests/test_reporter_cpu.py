"""F6 / A19 pinned to the REFERENCE's own Reporter (src/commons/Reporter.cpp:38-83,146-253, compiled into
oracle/_ref/ref_harness by oracle/Makefile): tests/golden/data/reports/ holds the files it wrote for every golden case
(tests/golden/make_golden2.py).  The product's host writers (host/mb_host_io.cpp) get the reference's per-read results
(<case>.results.tsv) and must reproduce those files byte for byte: classifications with and without --lineage 1, the
Kraken-style report and the Krona chart (the part after the static prelude, which is not copied into this repository)."""
import os

import numpy as np
import pytest

import helpers

REPORTS = os.path.join(helpers.DATA, "reports")


@pytest.fixture(scope="module")
def mb():
    import __graft_entry__ as ge
    ge.build()
    import metabuli_work_b200.binding as b
    return b


def results_from_golden(mb, rows):
    res = np.zeros(len(rows), dtype=mb.RESULT_DT)
    ids, cnt = [], []
    for i, (is_c, name, tax, org, qlen, bits, tc) in enumerate(rows):
        res[i]["classification"] = tax
        res[i]["score"] = np.uint32(bits).view(np.float32)
        res[i]["is_classified"] = is_c
        res[i]["query_length"] = qlen
        res[i]["taxcnt_begin"] = len(ids)
        for item in tc.split():
            t, c = item.split(":")
            ids.append(int(t))
            cnt.append(int(c))
        res[i]["taxcnt_end"] = len(ids)
    return res, np.array(ids + [0], dtype=np.int32), np.array(cnt + [0], dtype=np.int32)


def test_every_case_has_reference_reports(golden):
    for case in golden.names():
        for suffix in ("_classifications.tsv", "_report.tsv", "_lineage_classifications.tsv", "_krona_tail.html"):
            assert os.path.exists(os.path.join(REPORTS, case + suffix)), case + suffix


@pytest.mark.parametrize("case", ["plain_paired", "sync_paired", "legacy_paired", "plain_long", "sync_long", "plain_single",
                                  "euk_paired", "euk_long", "redund_paired", "sync4_paired", "sync6_paired", "sync4_long", "legacy_long",
                                  "mask_paired", "mask_paired_p05", "mask_long"])
def test_host_writers_equal_reference_reporter(golden, mb, tmp_path, case):
    rows = golden.results(case)
    res, ids, cnt = results_from_golden(mb, rows)
    names = [r[1] for r in rows]
    out = str(tmp_path)
    prelude = os.path.join(out, "prelude.html")
    with open(prelude, "wb") as f:
        f.write(b"<html><!-- stand-in for krona_prelude.html -->\n")
    mb.write_outputs(golden.dbdir(case), out, "plain", names, res, ids, cnt, lineage=False, krona_prelude=prelude)
    mb.write_outputs(golden.dbdir(case), out, "lineage", names, res, ids, cnt, lineage=True)

    def rd(p):
        with open(p, "rb") as f:
            return f.read()
    assert rd(os.path.join(out, "plain_classifications.tsv")) == rd(os.path.join(REPORTS, case + "_classifications.tsv"))
    assert rd(os.path.join(out, "plain_report.tsv")) == rd(os.path.join(REPORTS, case + "_report.tsv"))
    assert rd(os.path.join(out, "lineage_classifications.tsv")) == rd(os.path.join(REPORTS, case + "_lineage_classifications.tsv"))
    assert rd(os.path.join(out, "lineage_report.tsv")) == rd(os.path.join(REPORTS, case + "_report.tsv"))
    assert rd(os.path.join(out, "plain_krona.html")) == rd(prelude) + rd(os.path.join(REPORTS, case + "_krona_tail.html"))
    # without a prelude the chart data is still there
    assert rd(os.path.join(out, "lineage_krona.html")).endswith(rd(os.path.join(REPORTS, case + "_krona_tail.html")))
