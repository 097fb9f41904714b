"""ctypes binding of oracle/_build/liborc.so — TEST INFRASTRUCTURE ONLY.

May be imported from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs, never from metabuli_work_b200/.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "liborc.so")
REF_HARNESS = os.path.join(HERE, "_ref", "ref_harness")

KMER_DT = np.dtype([("value", "<u8"), ("qinfo", "<u8")])
MATCH_DT = np.dtype([("qinfo", "<u8"), ("targetId", "<i4"), ("speciesId", "<i4"), ("dnaEncoding", "<u4"),
                     ("rightEndHamming", "<u2"), ("hamming", "u1"), ("pad", "u1")])
RESULT_DT = np.dtype([("classification", "<i4"), ("score", "<f4"), ("is_classified", "<i4"),
                      ("query_length", "<i4"), ("taxcnt_begin", "<i4"), ("taxcnt_end", "<i4")])
assert KMER_DT.itemsize == 16 and MATCH_DT.itemsize == 24 and RESULT_DT.itemsize == 24


class Params(C.Structure):
    _fields_ = [("kmer_format", C.c_int), ("syncmer", C.c_int), ("smer_len", C.c_int), ("seq_mode", C.c_int),
                ("min_cons_cnt", C.c_int), ("min_cons_cnt_euk", C.c_int), ("min_score", C.c_float),
                ("min_sp_score", C.c_float), ("tie_ratio", C.c_float), ("skip_redundancy", C.c_int),
                ("accession_level", C.c_int), ("threads", C.c_int)]


def build(force=False):
    if force or not os.path.exists(LIB_PATH):
        subprocess.check_call(["make", "-C", HERE, "-s"])
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        vp, sz, i32, i64, u32 = C.c_void_p, C.c_size_t, C.c_int32, C.c_int64, C.c_uint32
        L.orc_db_open.restype = vp
        L.orc_db_open.argtypes = [C.c_char_p]
        L.orc_db_close.argtypes = [vp]
        L.orc_last_error.restype = C.c_char_p
        L.orc_default_params.argtypes = [C.POINTER(Params)]
        L.orc_db_apply_parameters.argtypes = [vp, C.POINTER(Params)]
        L.orc_db_num_kmers.restype = sz
        L.orc_db_num_kmers.argtypes = [vp]
        L.orc_db_num_words.restype = sz
        L.orc_db_num_words.argtypes = [vp]
        L.orc_db_decode.restype = sz
        L.orc_db_decode.argtypes = [vp, vp, vp, sz]
        for name in ("orc_tax_max_taxid", "orc_tax_eukaryota"):
            getattr(L, name).restype = i32
            getattr(L, name).argtypes = [vp]
        for name in ("orc_tax_parent", "orc_tax_species_of", "orc_tax_original"):
            getattr(L, name).restype = i32
            getattr(L, name).argtypes = [vp, i32]
        for name in ("orc_tax_lca", "orc_tax_is_ancestor"):
            getattr(L, name).restype = i32
            getattr(L, name).argtypes = [vp, i32, i32]
        L.orc_tax_at_rank.restype = i32
        L.orc_tax_at_rank.argtypes = [vp, i32, C.c_char_p]
        for name in ("orc_tax_rank", "orc_tax_name"):
            getattr(L, name).restype = C.c_char_p
            getattr(L, name).argtypes = [vp, i32]
        L.orc_query_kmer_number.restype = i64
        L.orc_query_kmer_number.argtypes = [i64]
        L.orc_max_covered_length.restype = i64
        L.orc_max_covered_length.argtypes = [i64]
        L.orc_extract.restype = sz
        L.orc_extract.argtypes = [C.POINTER(Params), vp, vp, vp, u32, vp, sz]
        L.orc_sort_kmers.argtypes = [vp, sz, C.c_int]
        L.orc_match_kmers.restype = sz
        L.orc_match_kmers.argtypes = [vp, C.POINTER(Params), vp, sz, vp, sz]
        L.orc_sort_matches.argtypes = [vp, sz, C.c_int]
        L.orc_score.restype = i64
        L.orc_score.argtypes = [vp, C.POINTER(Params), vp, sz, vp, u32, vp, vp, vp, sz]
        L.orc_write_classifications_lineage.restype = C.c_int
        L.orc_write_classifications_lineage.argtypes = [vp, C.c_char_p, C.POINTER(C.c_char_p), vp, u32, vp, vp]
        L.orc_write_classifications.restype = C.c_int
        L.orc_write_classifications.argtypes = [vp, C.c_char_p, C.POINTER(C.c_char_p), vp, u32, vp, vp, C.c_int]
        L.orc_write_report.restype = C.c_int
        L.orc_write_report.argtypes = [vp, C.c_char_p, vp, u32]
        L.orc_write_db_files.restype = C.c_int
        L.orc_write_db_files.argtypes = [C.c_char_p, vp, vp, sz, C.c_int, C.c_int]
        L.orc_build_index.restype = i64
        L.orc_build_index.argtypes = [vp, vp, vp, vp, sz, C.c_int, C.c_int, C.c_char_p]
        L.orc_mask_sequences.argtypes = [vp, vp, vp, u32, C.c_float, vp]
        L.orc_mask_lr.restype = C.c_double
        L.orc_mask_lr.argtypes = [C.c_int]
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def default_params(**kw):
    p = Params()
    lib().orc_default_params(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p


class Db:
    def __init__(self, dbdir):
        self.h = lib().orc_db_open(dbdir.encode())
        if not self.h:
            raise RuntimeError(lib().orc_last_error().decode())
        self.dir = dbdir

    def close(self):
        if self.h:
            lib().orc_db_close(self.h)
            self.h = None

    def params(self, **kw):
        p = default_params(**kw)
        lib().orc_db_apply_parameters(self.h, C.byref(p))
        return p

    def num_kmers(self):
        return lib().orc_db_num_kmers(self.h)

    def decode(self):
        n = self.num_kmers()
        v = np.zeros(n, dtype=np.uint64)
        t = np.zeros(n, dtype=np.uint32)
        got = lib().orc_db_decode(self.h, _ptr(v), _ptr(t), n)
        assert got == n, (got, n)
        return v, t


class Batch:
    """Packed reads: concatenated bases + per-(read, mate) offsets/lengths."""

    def __init__(self, mates1, mates2=None):
        n = len(mates1)
        self.n = n
        off = np.zeros(2 * n, dtype=np.uint64)
        ln = np.zeros(2 * n, dtype=np.uint32)
        chunks = []
        pos = 0
        for i in range(n):
            for m, lst in enumerate((mates1, mates2)):
                if lst is None:
                    continue
                s = lst[i]
                if isinstance(s, str):
                    s = s.encode()
                off[2 * i + m] = pos
                ln[2 * i + m] = len(s)
                chunks.append(s)
                pos += len(s)
        self.seqs = np.frombuffer(b"".join(chunks) + b"\0" * 8, dtype=np.uint8).copy()
        self.off, self.len = off, ln


def build_index(db, values, taxids, species, dbdir, split_num=4096, kmer_format=2):
    """DB-build inner loop: writes dbdir/{diffIdx,info,split}; -> number of DB k-mers."""
    v = np.ascontiguousarray(values, dtype=np.uint64)
    t = np.ascontiguousarray(taxids, dtype=np.int32)
    s = np.ascontiguousarray(species, dtype=np.int32)
    os.makedirs(dbdir, exist_ok=True)
    n = lib().orc_build_index(db.h, _ptr(v), _ptr(t), _ptr(s), len(v), split_num, kmer_format, dbdir.encode())
    assert n >= 0
    return int(n)


def mask(batch, mask_prob=0.9, want_probs=False):
    """--mask 1: the batch with every mate masked (tantan), optionally the per-base repeat probabilities."""
    import copy
    out = copy.copy(batch)
    out.seqs = batch.seqs.copy()
    probs = np.zeros(len(out.seqs), dtype=np.float32) if want_probs else None
    lib().orc_mask_sequences(_ptr(out.seqs), _ptr(out.off), _ptr(out.len), 2 * batch.n, mask_prob, _ptr(probs) if want_probs else None)
    return (out, probs) if want_probs else out


def extract(p, batch):
    cap = int(sum(max(0, lib().orc_query_kmer_number(int(x))) for x in batch.len)) + 16
    out = np.zeros(cap, dtype=KMER_DT)
    n = lib().orc_extract(C.byref(p), _ptr(batch.seqs), _ptr(batch.off), _ptr(batch.len), batch.n, _ptr(out), cap)
    assert n <= cap
    return out[:n].copy()


def sort_kmers(k, threads=1):
    k = np.ascontiguousarray(k)
    lib().orc_sort_kmers(_ptr(k), len(k), threads)
    return k


def match(db, p, sorted_kmers, cap=None):
    cap = cap or max(1 << 16, 8 * len(sorted_kmers))
    while True:
        out = np.zeros(cap, dtype=MATCH_DT)
        n = lib().orc_match_kmers(db.h, C.byref(p), _ptr(sorted_kmers), len(sorted_kmers), _ptr(out), cap)
        if n <= cap:
            return out[:n].copy()
        cap = n


def sort_matches(m, threads=1):
    m = np.ascontiguousarray(m)
    lib().orc_sort_matches(_ptr(m), len(m), threads)
    return m


def score(db, p, sorted_matches, batch):
    res = np.zeros(batch.n, dtype=RESULT_DT)
    cap = max(1024, len(sorted_matches))
    ids = np.zeros(cap, dtype=np.int32)
    cnt = np.zeros(cap, dtype=np.int32)
    used = lib().orc_score(db.h, C.byref(p), _ptr(sorted_matches), len(sorted_matches), _ptr(batch.len), batch.n,
                           _ptr(res), _ptr(ids), _ptr(cnt), cap)
    assert used >= 0
    return res, ids[:used], cnt[:used]


def classify(db, p, batch, threads=1):
    """Whole path on the CPU; returns (kmers_sorted, matches_sorted, results, taxcnt_ids, taxcnt_counts)."""
    k = sort_kmers(extract(p, batch), threads)
    m = sort_matches(match(db, p, k), threads)
    res, ids, cnt = score(db, p, m, batch)
    return k, m, res, ids, cnt


def write_classifications(db, path, names, res, ids, cnt, append=False):
    arr = (C.c_char_p * len(names))(*[n.encode() if isinstance(n, str) else n for n in names])
    rc = lib().orc_write_classifications(db.h, path.encode(), arr, _ptr(res), len(names), _ptr(ids), _ptr(cnt), int(append))
    assert rc == 0


def write_classifications_lineage(db, path, names, res, ids, cnt):
    arr = (C.c_char_p * len(names))(*[n.encode() if isinstance(n, str) else n for n in names])
    assert lib().orc_write_classifications_lineage(db.h, path.encode(), arr, _ptr(res), len(names), _ptr(ids), _ptr(cnt)) == 0


def write_report(db, path, res):
    assert lib().orc_write_report(db.h, path.encode(), _ptr(res), len(res)) == 0


def read_fastx(path, limit=None):
    """Minimal FASTA/FASTQ reader (names up to the first whitespace, like kseq)."""
    names, seqs = [], []
    with open(path, "rb") as f:
        first = f.read(1)
        f.seek(0)
        if first == b"@":
            while True:
                h = f.readline()
                if not h:
                    break
                s = f.readline().rstrip(b"\r\n")
                f.readline()
                f.readline()
                names.append(h[1:].split()[0].decode())
                seqs.append(s)
                if limit and len(names) >= limit:
                    break
        else:
            name, buf = None, []
            for line in f:
                if line.startswith(b">"):
                    if name is not None:
                        names.append(name)
                        seqs.append(b"".join(buf))
                        if limit and len(names) >= limit:
                            name = None
                            break
                    name, buf = line[1:].split()[0].decode(), []
                else:
                    buf.append(line.strip())
            if name is not None:
                names.append(name)
                seqs.append(b"".join(buf))
    return names, seqs
