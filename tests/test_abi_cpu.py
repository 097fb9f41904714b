"""CPU-side checks of the boundary: the library loads, exports every symbol include/metabuli_b200.h
declares, record layouts match the reference's, and the product refuses to run without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def mb():
    import __graft_entry__ as ge
    ge.build()
    import metabuli_work_b200.binding as b
    return b


def test_exports_every_declared_symbol(mb):
    header = open(os.path.join(ROOT, "include", "metabuli_b200.h")).read()
    declared = set(re.findall(r"\b(mb_[a-z_0-9]+)\s*\(", header))
    declared -= {"mb_ctx"}
    assert declared, "no declarations found"
    lib = C.CDLL(mb.LIB_CUDA)
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert missing == []
    assert set(mb.ABI_SYMBOLS) == declared


def test_record_layouts(mb):
    # sizeof(Kmer)==16, Match payload 24 B, QueryKmerInfo bit layout (Kmer.h:11-16)
    assert mb.KMER_DT.itemsize == 16 and mb.MATCH_DT.itemsize == 24
    assert C.sizeof(mb.Params) == 64
    p = mb.Params()
    C.CDLL(mb.LIB_CUDA).mb_default_params(C.byref(p))
    assert (p.kmer_format, p.seq_mode, p.min_cons_cnt, p.min_cons_cnt_euk, p.smer_len) == (1, 2, 4, 9, 5)
    assert abs(p.tie_ratio - 0.95) < 1e-7


def test_no_cpu_fallback(mb, golden):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(mb.MetabuliError) as e:
        mb.Classifier(golden.dbdir("plain_paired"))
    assert "no CUDA device" in str(e.value)


def test_product_does_not_touch_oracle():
    """Nothing under metabuli_work_b200/ or include/ may reference oracle/ (the oracle is test infrastructure)."""
    bad = []
    for base in ("metabuli_work_b200", "include"):
        for d, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".so", ".o", ".pyc")) or f == "metabuli":
                    continue
                txt = open(os.path.join(d, f), errors="ignore").read()
                if re.search(r"oracle/|import orc|liborc|mb_oracle", txt):
                    bad.append(os.path.join(d, f))
    assert bad == []


def test_pack_reads_layout(mb):
    b, o, l = mb.pack_reads([b"ACGT", b"", b"GG"], [b"TT", b"CCC", b""])
    assert bytes(b) == b"ACGTTTCCCGG"
    assert list(l) == [4, 2, 0, 3, 2, 0]
    assert list(o) == [0, 4, 6, 6, 9, 11]
