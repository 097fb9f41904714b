"""Shared helpers of the test-suite: golden fixture access and result comparison."""
import hashlib
import json
import os

import numpy as np

import orc  # oracle/orc.py (tests may use the oracle; the product may not)

HERE = os.path.dirname(os.path.abspath(__file__))
DATA = os.path.join(HERE, "golden", "data")


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


class Golden:
    def __init__(self):
        with open(os.path.join(DATA, "digests.json")) as f:
            self.cases = json.load(f)

    def names(self):
        return sorted(self.cases)

    def dbdir(self, case):
        return os.path.join(DATA, self.cases[case]["db"])

    def reads(self, case):
        """-> (names, mates1, mates2 or None)"""
        files = self.cases[case]["reads"]
        names, s1 = orc.read_fastx(os.path.join(DATA, files[0]))
        s2 = orc.read_fastx(os.path.join(DATA, files[1]))[1] if len(files) > 1 else None
        return names, s1, s2

    def results(self, case):
        """Reference per-read results: list of (is_classified, name, taxid_internal, taxid_org, qlen, score_bits, taxcnt_str)."""
        rows = []
        with open(os.path.join(DATA, case + ".results.tsv")) as f:
            for line in f:
                c = line.rstrip("\n").split("\t")
                rows.append((int(c[0]), c[1], int(c[2]), int(c[3]), int(c[4]), int(c[5], 16), c[6]))
        return rows


def taxcnt_string(ids, cnt, b, e):
    s = " ".join(f"{a}:{c}" for a, c in zip(ids[b:e], cnt[b:e]))
    return s + " " if s else s


def check_results_against_golden(rows, res, ids, cnt):
    """res: array with RESULT_DT-like fields. Returns list of mismatching read indices."""
    bad = []
    assert len(rows) == len(res)
    for i, row in enumerate(rows):
        r = res[i]
        bits = int(np.float32(r["score"]).view(np.uint32))
        tc = taxcnt_string(ids, cnt, int(r["taxcnt_begin"]), int(r["taxcnt_end"]))
        if not (row[0] == int(r["is_classified"]) and row[2] == int(r["classification"]) and row[4] == int(r["query_length"])
                and row[5] == bits and row[6] == tc):
            bad.append(i)
    return bad
