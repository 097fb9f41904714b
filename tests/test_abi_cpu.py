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


def test_rebalanced_cuts():
    """Shard boundaries that equalise measured per-shard times (host/mb_host.cpp rebalancedCuts): equal times keep equal
    shards, a slow shard shrinks, the new boundaries cut the piecewise-constant work density into equal parts."""
    import numpy as np
    import metabuli_work_b200.binding as mb
    assert np.allclose(mb.rebalanced_cuts([3.0, 3.0, 3.0, 3.0]), [0.25, 0.5, 0.75])
    c = mb.rebalanced_cuts([2.0, 1.0])
    assert np.allclose(c, [0.375])                       # half of the work (1.5) lies at 3/4 of the first shard
    # second round from moved boundaries: density 2/0.375 and 1/0.625 ... the cut keeps equalising
    c2 = mb.rebalanced_cuts([1.5, 1.5], c)
    assert np.allclose(c2, c)
    t = np.array([40.0, 39, 21, 28, 31, 29, 35, 24])
    cuts = mb.rebalanced_cuts(t)
    assert len(cuts) == 7 and np.all(np.diff(cuts) > 0) and 0 < cuts[0] and cuts[-1] < 1
    # work in each new shard under the model = total / 8
    edges = np.concatenate([[0], cuts, [1]])
    old = np.linspace(0, 1, 9)
    dens = t / np.diff(old)
    work = [sum(dens[g] * max(0.0, min(b, old[g + 1]) - max(a, old[g])) for g in range(8)) for a, b in zip(edges[:-1], edges[1:])]
    assert np.allclose(work, t.sum() / 8)
