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


def kmer_sort_key(values, sort_bits, kmer_format):
    """The key mb_sort_kmers orders by (SortT, csrc/mb_common.cuh): leading amino-acid bits, or for format 2 by default
    (kmer_sort_bits 21) the five leading residues as a base-21 number plus two bits of the sixth."""
    values = np.asarray(values, dtype=np.uint64)
    if not sort_bits:
        sort_bits = 21 if kmer_format == 2 else 32
    if sort_bits == 31:                  # seven residues base 21, one bit of the eighth (sort_t_packed7)
        aa = values >> np.uint64(24)
        t = (aa >> np.uint64(35)) & np.uint64(31)
        for sh in (30, 25, 20, 15, 10, 5):
            t = t * np.uint64(21) + ((aa >> np.uint64(sh)) & np.uint64(31))
        return (t << np.uint64(1)) | ((aa & np.uint64(31)) >= np.uint64(11)).astype(np.uint64)
    if sort_bits != 21:
        return values >> np.uint64(24 + 40 - sort_bits)
    hi = (values >> np.uint64(32)).astype(np.uint64)
    t = hi >> np.uint64(27)
    for sh in (22, 17, 12, 7):
        t = t * np.uint64(21) + ((hi >> np.uint64(sh)) & np.uint64(31))
    return (t << np.uint64(2)) | ((((hi >> np.uint64(2)) & np.uint64(31)) * np.uint64(49)) >> np.uint64(8))


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


# ---- host emulation of the device logic (tests/emu) ----------------------------------------------------
import ctypes as C
import subprocess

EMU_DIR = os.path.join(HERE, "emu")
ROOT = os.path.dirname(HERE)


def emu_lib():
    so = os.path.join(EMU_DIR, "libemu.so")
    srcs = [os.path.join(EMU_DIR, "emu.cpp")] + [os.path.join(ROOT, "metabuli_work_b200", "csrc", f)
                                                 for f in ("mb_common.cuh", "mb_score.cuh", "mb_tables.cu")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["/usr/bin/g++", "-O2", "-std=gnu++17", "-fPIC", "-shared", "-Wno-unknown-pragmas", "-o", so,
                               srcs[0], "-x", "c++", srcs[3]])
    L = C.CDLL(so)
    vp, sz, i32, u32 = C.c_void_p, C.c_size_t, C.c_int32, C.c_uint32
    L.emu_extract.restype = sz
    L.emu_extract.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, u32, vp, sz]
    L.emu_hamming.argtypes = [vp, vp, vp, sz, vp, vp]
    L.emu_hamming_lut.argtypes = [vp, vp, vp, sz, vp, vp]
    L.emu_lca.restype = i32
    L.emu_lca.argtypes = [vp, vp, i32, i32, i32, i32]
    L.emu_score.restype = C.c_int64
    L.emu_score.argtypes = [vp, sz, vp, u32, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float,
                            C.c_float, vp, vp, vp, vp, i32, i32, i32, vp, vp, vp, sz]
    L.emu_set_fast.argtypes = [C.c_int]
    L.emu_is_syncmer.argtypes = [vp, sz, C.c_int, vp]
    L.emu_sort_t.argtypes = [C.c_int, C.c_int, vp, sz, vp]
    L.emu_fast_fallbacks.restype = C.c_long
    return L


def dense_taxonomy(db):
    """Dense arrays of mb_load_taxonomy derived from the ORACLE's taxonomy (test side)."""
    L = orc.lib()
    mx = L.orc_tax_max_taxid(db.h)
    parent = np.full(mx + 1, -1, dtype=np.int32)
    species_of = np.zeros(mx + 1, dtype=np.int32)
    species_rank_of = np.zeros(mx + 1, dtype=np.int32)
    for t in range(mx + 1):
        if t == 0:
            continue
        if L.orc_tax_rank(db.h, t) == b"" and L.orc_tax_name(db.h, t) == b"":
            continue
        parent[t] = L.orc_tax_parent(db.h, t)
        species_of[t] = L.orc_tax_species_of(db.h, t)
        species_rank_of[t] = L.orc_tax_at_rank(db.h, t, b"species")
    depth = np.full(mx + 1, -1, dtype=np.int32)
    for t in range(mx + 1):
        if parent[t] < 0:
            continue
        d, x = 0, t
        while parent[x] != x and d < 1000:
            x = parent[x]
            d += 1
        depth[t] = d
    node0 = 1
    return parent, depth, species_of, species_rank_of, mx, node0, L.orc_tax_eukaryota(db.h)


def build_edge_records(taxid, species):
    """Target records that stress the DB writer: values far apart (five-word deltas), and one amino-acid group so large that
    several split thresholds fall inside it (the reference's writer has one pending flag: they share a split).  taxid /
    species: the committed build_input columns, sampled for plausible taxa.  Deterministic."""
    rng = np.random.default_rng(5)
    big = rng.integers(0, 1 << 63, 5000, dtype=np.uint64) << np.uint64(1)
    grp = (np.uint64(0x123456789A) << np.uint64(24)) | rng.integers(0, 1 << 24, 20000, dtype=np.uint64)
    vv = np.concatenate([big, grp])
    pick = rng.integers(0, len(taxid), len(vv))
    return vv, np.ascontiguousarray(taxid[pick]), np.ascontiguousarray(species[pick])
