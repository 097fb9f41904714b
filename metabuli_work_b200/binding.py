"""ctypes binding of the C ABI (include/metabuli_b200.h) and of the host library (host/mb_host.h).

No compute happens here; every call goes to the native libraries.  There is no fallback: a missing
library raises at import time.
"""
import ctypes as C
import os

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_CUDA = os.path.join(PKG, "libmetabuli_b200.so")
LIB_HOST = os.path.join(PKG, "libmbhost.so")
CLI = os.path.join(PKG, "metabuli")

for _p in (LIB_CUDA, LIB_HOST):
    if not os.path.exists(_p):
        raise ImportError(f"{_p} is missing: run `python -m metabuli_work_b200.build` (there is no CPU fallback)")

KMER_DT = np.dtype([("value", "<u8"), ("qinfo", "<u8")])
MATCH_DT = np.dtype([("qinfo", "<u8"), ("targetId", "<i4"), ("speciesId", "<i4"), ("dnaEncoding", "<u4"),
                     ("rightEndHamming", "<u2"), ("hamming", "u1"), ("pad", "u1")])
RESULT_DT = np.dtype([("classification", "<i4"), ("score", "<f4"), ("is_classified", "<i4"), ("query_length", "<i4"),
                      ("hamming_dist", "<i4"), ("taxcnt_begin", "<i4"), ("taxcnt_end", "<i4"), ("n_matches", "<i4")])
assert KMER_DT.itemsize == 16 and MATCH_DT.itemsize == 24 and RESULT_DT.itemsize == 32

MB_OK, MB_E_CUDA, MB_E_ARG, MB_E_CAPACITY, MB_E_DATA = 0, -1, -2, -3, -4
MB_NOALLOC_MAX_KMERS = 1 << 22          # include/metabuli_b200.h


class Params(C.Structure):
    _fields_ = [("device", C.c_int32), ("kmer_format", C.c_int32), ("syncmer", C.c_int32), ("smer_len", C.c_int32),
                ("seq_mode", C.c_int32), ("min_cons_cnt", C.c_int32), ("min_cons_cnt_euk", C.c_int32),
                ("min_score", C.c_float), ("min_sp_score", C.c_float), ("tie_ratio", C.c_float),
                ("skip_redundancy", C.c_int32), ("accession_level", C.c_int32), ("last_shard", C.c_int32),
                ("kmer_sort_bits", C.c_int32), ("mask_mode", C.c_int32), ("mask_prob", C.c_float)]


class PhaseTimes(C.Structure):
    _fields_ = [("h2d_ms", C.c_float), ("extract_ms", C.c_float), ("sort_kmers_ms", C.c_float), ("match_ms", C.c_float),
                ("sort_matches_ms", C.c_float), ("score_ms", C.c_float), ("d2h_ms", C.c_float), ("total_ms", C.c_float),
                ("n_bases", C.c_uint64), ("n_kmers", C.c_uint64), ("n_matches", C.c_uint64), ("n_reads", C.c_uint64),
                ("db_words", C.c_uint64), ("db_kmers", C.c_uint64), ("kernel_launches", C.c_uint32),
                ("match_retries", C.c_uint32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class KernelTime(C.Structure):
    _fields_ = [("name", C.c_char * 32), ("ms", C.c_float), ("launches", C.c_uint32)]


class PeerMem(C.Structure):
    """mb_peer_mem: one rank's receive side (device addresses valid in the process that holds the struct)."""
    _fields_ = [("kmers", C.c_void_p), ("matches", C.c_void_p), ("counters", C.c_void_p), ("kmer_cap", C.c_uint64),
                ("match_cap", C.c_uint64)]


# every symbol include/metabuli_b200.h declares
ABI_SYMBOLS = ["mb_default_params", "mb_create", "mb_destroy", "mb_last_error", "mb_global_error", "mb_set_params",
               "mb_load_db", "mb_decode_db", "mb_load_taxonomy", "mb_classify_batch", "mb_stage_reads",
               "mb_extract_kmers", "mb_sort_kmers", "mb_match_kmers", "mb_sort_matches", "mb_score_reads",
               "mb_fetch_results", "mb_classify_staged", "mb_copy_kmers", "mb_copy_matches", "mb_set_kmers",
               "mb_set_matches", "mb_kmers_device", "mb_matches_device", "mb_set_kmers_device", "mb_set_matches_device",
               "mb_partition_kmers", "mb_partition_matches", "mb_get_timings", "mb_stream",
               "mb_export_kmers_device", "mb_export_matches_device", "mb_set_read_id_base", "mb_set_profile",
               "mb_get_kernel_times", "mb_peer_alloc", "mb_ipc_export", "mb_ipc_open", "mb_ipc_close", "mb_peer_attach",
               "mb_prefetch_reads", "mb_stage_fastq", "mb_prefetch_fastq", "mb_alloc_host", "mb_free_host", "mb_fetch_fastq_names", "mb_fetch_staged_reads", "mb_extract_kmers_routed", "mb_take_routed_kmers", "mb_match_kmers_routed", "mb_take_routed_matches",
               "mb_sort_kmers_noalloc", "mb_mem_info", "mb_enable_peer_access", "mb_mask_reads", "mb_build_index", "mb_fetch_results_async", "mb_fetch_wait"]

_cuda = C.CDLL(LIB_CUDA, mode=C.RTLD_GLOBAL)
_host = C.CDLL(LIB_HOST)
vp, sz, i32, u32, u64 = C.c_void_p, C.c_size_t, C.c_int32, C.c_uint32, C.c_uint64
szp = C.POINTER(C.c_size_t)


def _sig(lib, name, res, args):
    f = getattr(lib, name)
    f.restype = res
    f.argtypes = args
    return f


_sig(_cuda, "mb_default_params", None, [C.POINTER(Params)])
_sig(_cuda, "mb_create", vp, [C.POINTER(Params)])
_sig(_cuda, "mb_destroy", None, [vp])
_sig(_cuda, "mb_last_error", C.c_char_p, [vp])
_sig(_cuda, "mb_global_error", C.c_char_p, [])
_sig(_cuda, "mb_set_params", C.c_int, [vp, C.POINTER(Params)])
_sig(_cuda, "mb_load_db", C.c_int, [vp, vp, sz, vp, sz])
_sig(_cuda, "mb_decode_db", C.c_int, [vp, vp, sz, szp])
_sig(_cuda, "mb_load_taxonomy", C.c_int, [vp, vp, vp, vp, i32, i32, i32])
_sig(_cuda, "mb_classify_batch", C.c_int, [vp, vp, sz, vp, vp, u32, vp, vp, vp, sz, szp])
_sig(_cuda, "mb_stage_reads", C.c_int, [vp, vp, sz, vp, vp, u32])
_sig(_cuda, "mb_prefetch_reads", C.c_int, [vp, vp, sz, vp, vp, u32])
_sig(_cuda, "mb_extract_kmers", C.c_int, [vp, szp])
_sig(_cuda, "mb_sort_kmers", C.c_int, [vp])
_sig(_cuda, "mb_match_kmers", C.c_int, [vp, szp])
_sig(_cuda, "mb_sort_matches", C.c_int, [vp])
_sig(_cuda, "mb_score_reads", C.c_int, [vp, szp])
_sig(_cuda, "mb_fetch_results", C.c_int, [vp, vp, vp, vp, sz])
_sig(_cuda, "mb_classify_staged", C.c_int, [vp, szp])
_sig(_cuda, "mb_copy_kmers", C.c_int, [vp, vp, sz, szp])
_sig(_cuda, "mb_copy_matches", C.c_int, [vp, vp, sz, szp])
_sig(_cuda, "mb_set_kmers", C.c_int, [vp, vp, sz])
_sig(_cuda, "mb_set_matches", C.c_int, [vp, vp, sz])
_sig(_cuda, "mb_kmers_device", C.c_int, [vp, C.POINTER(vp), szp])
_sig(_cuda, "mb_matches_device", C.c_int, [vp, C.POINTER(vp), szp])
_sig(_cuda, "mb_set_kmers_device", C.c_int, [vp, vp, sz])
_sig(_cuda, "mb_set_matches_device", C.c_int, [vp, vp, sz])
_sig(_cuda, "mb_export_kmers_device", C.c_int, [vp, vp, sz, szp])
_sig(_cuda, "mb_export_matches_device", C.c_int, [vp, vp, sz, szp])
_sig(_cuda, "mb_set_read_id_base", C.c_int, [vp, u32])
_sig(_cuda, "mb_partition_kmers", C.c_int, [vp, vp, C.c_int, vp])
_sig(_cuda, "mb_partition_matches", C.c_int, [vp, u32, C.c_int, vp])
_sig(_cuda, "mb_get_timings", C.c_int, [vp, C.POINTER(PhaseTimes)])
_sig(_cuda, "mb_mem_info", C.c_int, [vp, szp, szp])
_sig(_cuda, "mb_enable_peer_access", C.c_int, [vp, C.c_int])
_sig(_cuda, "mb_sort_kmers_noalloc", C.c_int, [vp, vp, vp, sz, vp])
_sig(_cuda, "mb_stream", vp, [vp])
_sig(_cuda, "mb_set_profile", C.c_int, [vp, C.c_int])
_sig(_cuda, "mb_get_kernel_times", C.c_int, [vp, C.POINTER(KernelTime), sz, szp])
_sig(_cuda, "mb_peer_alloc", C.c_int, [vp, sz, sz, C.POINTER(PeerMem)])
_sig(_cuda, "mb_ipc_export", C.c_int, [vp, vp, vp])
_sig(_cuda, "mb_ipc_open", C.c_int, [vp, vp, C.POINTER(vp)])
_sig(_cuda, "mb_ipc_close", C.c_int, [vp, vp])
_sig(_cuda, "mb_peer_attach", C.c_int, [vp, u32, u32, C.POINTER(PeerMem), vp, u32])
class FastqIndex(C.Structure):
    _fields_ = [("n_reads", C.c_uint32), ("reserved", C.c_uint32), ("consumed1", C.c_uint64), ("consumed2", C.c_uint64),
                ("n_bases", C.c_uint64)]


_sig(_cuda, "mb_stage_fastq", C.c_int, [vp, vp, sz, vp, sz, u32, C.c_int, C.POINTER(FastqIndex)])
_sig(_cuda, "mb_prefetch_fastq", C.c_int, [vp, vp, sz, vp, sz])
_sig(_cuda, "mb_alloc_host", vp, [sz])
_sig(_cuda, "mb_free_host", None, [vp])
_sig(_cuda, "mb_fetch_fastq_names", C.c_int, [vp, vp, vp])
_sig(_cuda, "mb_fetch_staged_reads", C.c_int, [vp, vp, vp, vp])
_sig(_cuda, "mb_mask_reads", C.c_int, [vp])
_sig(_cuda, "mb_fetch_results_async", C.c_int, [vp, vp, vp, vp, C.c_size_t])
_sig(_cuda, "mb_fetch_wait", C.c_int, [vp])
_sig(_cuda, "mb_build_index", C.c_int, [vp, vp, vp, vp, C.c_size_t, C.c_int, vp, C.c_size_t, C.POINTER(C.c_size_t), vp, C.c_size_t,
                                          C.POINTER(C.c_size_t), vp])
_sig(_cuda, "mb_extract_kmers_routed", C.c_int, [vp, szp])
_sig(_cuda, "mb_take_routed_kmers", C.c_int, [vp, szp])
_sig(_cuda, "mb_match_kmers_routed", C.c_int, [vp, szp])
_sig(_cuda, "mb_take_routed_matches", C.c_int, [vp, szp])

_sig(_host, "mbh_last_error", C.c_char_p, [])
_sig(_host, "mbh_open", vp, [C.c_char_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int])
_sig(_host, "mbh_open_cuts", vp, [C.c_char_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, vp])
_sig(_host, "mbh_rebalanced_cuts", None, [vp, C.c_int, vp, vp])
_sig(_host, "mbh_close", None, [vp])
_sig(_host, "mbh_ctx", vp, [vp])
_sig(_host, "mbh_params", None, [vp, C.POINTER(Params)])
_sig(_host, "mbh_num_bounds", C.c_int, [vp])
_sig(_host, "mbh_bounds", None, [vp, vp])
_sig(_host, "mbh_original_taxid", i32, [vp, i32])
_sig(_host, "mbh_rank", C.c_char_p, [vp, i32])
_sig(_host, "mbh_db_kmers", u64, [vp])
_sig(_host, "mbh_db_words", u64, [vp])
_sig(_host, "mbh_write_outputs", C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, u32, C.POINTER(C.c_char_p), vp, vp, vp])
_sig(_host, "mbh_read_fastx", C.c_int, [C.c_char_p, C.c_char_p, u64, u64, vp, u64, vp, vp, u64, vp, u64, C.POINTER(u32), C.POINTER(u64),
                                        C.POINTER(u64)])
_sig(_host, "mbh_classify_files", C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_int,
                                            C.c_float, C.c_float, u64])


class MetabuliError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"[{code}] {msg}")
        self.code = code


def _ptr(a):
    if a is None:
        return None
    return a.ctypes.data_as(C.c_void_p)


def rebalanced_cuts(times, current=None):
    """Shard boundaries (fractions of the split table) that equalise the measured per-shard times."""
    t = np.ascontiguousarray(times, dtype=np.float64)
    cur = None if current is None else np.ascontiguousarray(current, dtype=np.float64)
    out = np.zeros(len(t) - 1, dtype=np.float64)
    _host.mbh_rebalanced_cuts(_ptr(t), len(t), _ptr(cur) if cur is not None else None, _ptr(out))
    return out


def host_read_fastx(path1, path2=None, max_reads=1 << 62, max_bases=1 << 62):
    """The C++ host reader on its own (no GPU): first batch of the file(s) -> (names, bases uint8[], offsets uint64[2n],
    lens uint32[2n]) in the layout of mb_stage_reads (the second mate of a single-end read has length 0)."""
    n_reads, n_bases, names_len = u32(0), u64(0), u64(0)
    cap_b, cap_r, cap_n = 1 << 16, 1 << 10, 1 << 14
    while True:
        bases = np.zeros(cap_b, dtype=np.uint8)
        offsets = np.zeros(2 * cap_r, dtype=np.uint64)
        lens = np.zeros(2 * cap_r, dtype=np.uint32)
        names = np.zeros(cap_n, dtype=np.uint8)
        rc = _host.mbh_read_fastx(os.fsencode(path1), os.fsencode(path2) if path2 else None, int(max_reads), int(max_bases),
                                  _ptr(bases), cap_b, _ptr(offsets), _ptr(lens), cap_r, _ptr(names), cap_n,
                                  C.byref(n_reads), C.byref(n_bases), C.byref(names_len))
        if rc == 2:
            cap_b, cap_r, cap_n = max(cap_b, n_bases.value), max(cap_r, n_reads.value), max(cap_n, names_len.value)
            continue
        if rc != 0:
            raise MetabuliError(rc, _host.mbh_last_error().decode())
        n = n_reads.value
        nm = names[:names_len.value].tobytes().decode().split("\n")[:n]
        return nm, bases[:n_bases.value].copy(), offsets[:2 * n].copy(), lens[:2 * n].copy()


def pack_reads(mates1, mates2=None):
    """-> (bases uint8[], offsets uint64[2n], lens uint32[2n]) in the layout of mb_classify_batch."""
    n = len(mates1)
    lens = np.zeros(2 * n, dtype=np.uint32)
    l1 = np.fromiter((len(s) for s in mates1), dtype=np.uint32, count=n)
    lens[0::2] = l1
    if mates2 is not None:
        lens[1::2] = np.fromiter((len(s) for s in mates2), dtype=np.uint32, count=n)
    offsets = np.zeros(2 * n, dtype=np.uint64)
    if n:
        offsets[1:] = np.cumsum(lens[:-1], dtype=np.uint64)
    chunks = []
    for i in range(n):
        s = mates1[i]
        chunks.append(s.encode() if isinstance(s, str) else s)
        if mates2 is not None:
            s = mates2[i]
            chunks.append(s.encode() if isinstance(s, str) else s)
    bases = np.frombuffer(b"".join(chunks), dtype=np.uint8).copy() if chunks else np.zeros(0, dtype=np.uint8)
    return bases, offsets, lens


class Context:
    """Methods shared by every owner of an mb_ctx: the per-batch calls that Classifier::startClassify makes
    (src/commons/Classifier.cpp:105-119), one Python method per C-ABI entry point."""

    ctx = None
    n_reads = 0

    def _ck(self, rc):
        if rc != MB_OK:
            raise MetabuliError(rc, _cuda.mb_last_error(self.ctx).decode())

    def set_params(self, **kw):
        p = self.params
        self._overrides = dict(getattr(self, "_overrides", {}), **kw)       # earlier overrides stay in force
        for k, v in self._overrides.items():
            setattr(p, k, v)
        self._ck(_cuda.mb_set_params(self.ctx, C.byref(p)))

    def decode_db(self):
        n = self.db_kmers()
        out = np.zeros(n, dtype=np.uint64)
        got = C.c_size_t(0)
        self._ck(_cuda.mb_decode_db(self.ctx, _ptr(out), n, C.byref(got)))
        return out[:got.value]

    # ---- stages ----
    def stage(self, bases, offsets, lens, bases_ptr=None):
        self.n_reads = len(lens) // 2
        self._keep = (bases, offsets, lens)
        bp = bases_ptr if bases_ptr is not None else _ptr(bases)
        self._ck(_cuda.mb_stage_reads(self.ctx, bp, len(bases), _ptr(offsets), _ptr(lens), self.n_reads))

    def prefetch(self, bases, offsets, lens, bases_ptr=None):
        """Start the host->device copy of the NEXT batch on the copy stream; a later stage() with the same arrays adopts it."""
        self._keep_next = (bases, offsets, lens)
        bp = bases_ptr if bases_ptr is not None else _ptr(bases)
        self._ck(_cuda.mb_prefetch_reads(self.ctx, bp, len(bases), _ptr(offsets), _ptr(lens), len(lens) // 2))

    def stage_fastq(self, text1, text2=None, max_reads=0xFFFFFFFF, final=True):
        """Raw FASTQ text (bytes / uint8 arrays) -> staged batch, parsed on the device (mb_stage_fastq).
        Returns (n_reads, consumed1, consumed2, n_bases)."""
        t1 = np.frombuffer(text1, dtype=np.uint8) if isinstance(text1, (bytes, bytearray)) else text1
        t2 = None if text2 is None else (np.frombuffer(text2, dtype=np.uint8) if isinstance(text2, (bytes, bytearray)) else text2)
        self._keep = (t1, t2)
        idx = FastqIndex()
        self._ck(_cuda.mb_stage_fastq(self.ctx, _ptr(t1) if len(t1) else None, len(t1),
                                      None if t2 is None else (_ptr(t2) if len(t2) else _ptr(np.zeros(1, dtype=np.uint8))),
                                      0 if t2 is None else len(t2), max_reads, 1 if final else 0, C.byref(idx)))
        self.n_reads = idx.n_reads
        return idx.n_reads, idx.consumed1, idx.consumed2, idx.n_bases

    def prefetch_fastq(self, text1, text2=None):
        """Start the host->device copy of the NEXT chunk's text on the copy stream; a later stage_fastq() with the same
        arrays adopts it."""
        self._keep_next = (text1, text2)
        self._ck(_cuda.mb_prefetch_fastq(self.ctx, _ptr(text1), len(text1), None if text2 is None else _ptr(text2),
                                         0 if text2 is None else len(text2)))

    def fastq_names(self, text1):
        """Names of the reads staged by stage_fastq, cut out of text1 through the device-computed spans."""
        off = np.zeros(self.n_reads, dtype=np.uint64)
        ln = np.zeros(self.n_reads, dtype=np.uint32)
        self._ck(_cuda.mb_fetch_fastq_names(self.ctx, _ptr(off), _ptr(ln)))
        t = bytes(text1)
        return [t[int(o):int(o) + int(l)] for o, l in zip(off, ln)]

    def staged_reads(self, n_bases):
        bases = np.zeros(n_bases, dtype=np.uint8)
        offsets = np.zeros(2 * self.n_reads, dtype=np.uint64)
        lens = np.zeros(2 * self.n_reads, dtype=np.uint32)
        self._ck(_cuda.mb_fetch_staged_reads(self.ctx, _ptr(bases) if n_bases else None, _ptr(offsets) if self.n_reads else None,
                                             _ptr(lens) if self.n_reads else None))
        return bases, offsets, lens

    def build_index(self, values, taxids, species, split_num=4096):
        """DB-build inner loop -> (diffIdx words u16, info u32, split table as a (split_num, 3) u64 array)."""
        v = np.ascontiguousarray(values, dtype=np.uint64)
        t = np.ascontiguousarray(taxids, dtype=np.int32)
        s = np.ascontiguousarray(species, dtype=np.int32)
        n = len(v)
        words = np.zeros(5 * n + 8, dtype=np.uint16)
        info = np.zeros(n + 8, dtype=np.uint32)
        splits = np.zeros((split_num, 3), dtype=np.uint64)
        nw, nk = C.c_size_t(0), C.c_size_t(0)
        self._ck(_cuda.mb_build_index(self.ctx, _ptr(v), _ptr(t), _ptr(s), n, split_num, _ptr(words), len(words), C.byref(nw), _ptr(info),
                                      len(info), C.byref(nk), _ptr(splits)))
        return words[:nw.value].copy(), info[:nk.value].copy(), splits

    def mask_reads(self):
        """--mask 1 on the staged batch (tap; extract() applies it itself when mask_mode is set)."""
        self._ck(_cuda.mb_mask_reads(self.ctx))

    def extract(self):
        n = C.c_size_t(0)
        self._ck(_cuda.mb_extract_kmers(self.ctx, C.byref(n)))
        return n.value

    def sort_kmers(self):
        self._ck(_cuda.mb_sort_kmers(self.ctx))

    def match(self):
        n = C.c_size_t(0)
        self._ck(_cuda.mb_match_kmers(self.ctx, C.byref(n)))
        return n.value

    def sort_matches(self):
        self._ck(_cuda.mb_sort_matches(self.ctx))

    def score(self):
        n = C.c_size_t(0)
        self._ck(_cuda.mb_score_reads(self.ctx, C.byref(n)))
        return n.value

    def classify_staged(self):
        n = C.c_size_t(0)
        self._ck(_cuda.mb_classify_staged(self.ctx, C.byref(n)))
        return n.value

    def fetch(self, used):
        res = np.zeros(self.n_reads, dtype=RESULT_DT)
        ids = np.zeros(used + 1, dtype=np.int32)
        cnt = np.zeros(used + 1, dtype=np.int32)
        self._ck(_cuda.mb_fetch_results(self.ctx, _ptr(res), _ptr(ids), _ptr(cnt), used + 1))
        return res, ids[:used], cnt[:used]

    def fetch_into(self, res, ids, cnt):
        """mb_fetch_results into caller-owned (e.g. pinned) arrays."""
        self._ck(_cuda.mb_fetch_results(self.ctx, _ptr(res), _ptr(ids), _ptr(cnt), len(ids)))

    def fetch_async(self, res, ids, cnt):
        """mb_fetch_results_async into caller-owned (page-locked) arrays; fetch_wait() before they are read."""
        self._ck(_cuda.mb_fetch_results_async(self.ctx, _ptr(res), _ptr(ids), _ptr(cnt), len(ids)))

    def fetch_wait(self):
        self._ck(_cuda.mb_fetch_wait(self.ctx))

    def kmers(self):
        n = C.c_size_t(0)
        rc = _cuda.mb_copy_kmers(self.ctx, None, 0, C.byref(n))
        if rc not in (MB_OK, MB_E_CAPACITY):
            self._ck(rc)
        out = np.zeros(n.value, dtype=KMER_DT)
        self._ck(_cuda.mb_copy_kmers(self.ctx, _ptr(out), n.value, C.byref(n)))
        return out

    def matches(self):
        n = C.c_size_t(0)
        rc = _cuda.mb_copy_matches(self.ctx, None, 0, C.byref(n))
        if rc not in (MB_OK, MB_E_CAPACITY):
            self._ck(rc)
        out = np.zeros(n.value, dtype=MATCH_DT)
        self._ck(_cuda.mb_copy_matches(self.ctx, _ptr(out), n.value, C.byref(n)))
        return out

    def sort_kmers_noalloc(self, d_kmers, d_tmp, n, stream=None):
        """mb_sort_kmers_noalloc on caller-owned device buffers (device addresses as ints)."""
        self._ck(_cuda.mb_sort_kmers_noalloc(self.ctx, C.c_void_p(d_kmers), C.c_void_p(d_tmp), n, C.c_void_p(stream) if stream else None))

    def mem_info(self):
        f, t = C.c_size_t(0), C.c_size_t(0)
        self._ck(_cuda.mb_mem_info(self.ctx, C.byref(f), C.byref(t)))
        return f.value, t.value

    def set_kmers(self, k):
        k = np.ascontiguousarray(k, dtype=KMER_DT)
        self._ck(_cuda.mb_set_kmers(self.ctx, _ptr(k), len(k)))

    def set_matches(self, m):
        m = np.ascontiguousarray(m, dtype=MATCH_DT)
        self._ck(_cuda.mb_set_matches(self.ctx, _ptr(m), len(m)))

    def kmers_device(self):
        p, n = C.c_void_p(0), C.c_size_t(0)
        self._ck(_cuda.mb_kmers_device(self.ctx, C.byref(p), C.byref(n)))
        return p.value or 0, n.value

    def matches_device(self):
        p, n = C.c_void_p(0), C.c_size_t(0)
        self._ck(_cuda.mb_matches_device(self.ctx, C.byref(p), C.byref(n)))
        return p.value or 0, n.value

    def set_kmers_device(self, dptr, n):
        self._ck(_cuda.mb_set_kmers_device(self.ctx, C.c_void_p(dptr), n))

    def set_matches_device(self, dptr, n):
        self._ck(_cuda.mb_set_matches_device(self.ctx, C.c_void_p(dptr), n))

    def export_kmers_device(self, dptr, cap):
        n = C.c_size_t(0)
        self._ck(_cuda.mb_export_kmers_device(self.ctx, C.c_void_p(dptr), cap, C.byref(n)))
        return n.value

    def export_matches_device(self, dptr, cap):
        n = C.c_size_t(0)
        self._ck(_cuda.mb_export_matches_device(self.ctx, C.c_void_p(dptr), cap, C.byref(n)))
        return n.value

    def set_read_id_base(self, base):
        self._ck(_cuda.mb_set_read_id_base(self.ctx, base))

    def partition_kmers(self, bounds):
        bounds = np.ascontiguousarray(bounds, dtype=np.uint64)
        counts = np.zeros(len(bounds) + 1, dtype=np.uint64)
        self._ck(_cuda.mb_partition_kmers(self.ctx, _ptr(bounds), len(bounds) + 1, _ptr(counts)))
        return counts

    def partition_matches(self, reads_per_rank, n_ranks):
        counts = np.zeros(n_ranks, dtype=np.uint64)
        self._ck(_cuda.mb_partition_matches(self.ctx, reads_per_rank, n_ranks, _ptr(counts)))
        return counts

    # ---- multi-GPU, fused form: producers write into the destination rank's receive buffer over peer memory ----
    def peer_alloc(self, kmer_cap, match_cap):
        m = PeerMem()
        self._ck(_cuda.mb_peer_alloc(self.ctx, int(kmer_cap), int(match_cap), C.byref(m)))
        return m

    def ipc_export(self, dptr):
        h = (C.c_ubyte * 64)()
        self._ck(_cuda.mb_ipc_export(self.ctx, C.c_void_p(dptr), h))
        return bytes(h)

    def ipc_open(self, handle):
        h = (C.c_ubyte * 64).from_buffer_copy(handle)
        p = C.c_void_p(0)
        self._ck(_cuda.mb_ipc_open(self.ctx, h, C.byref(p)))
        return p.value

    def ipc_close(self, dptr):
        self._ck(_cuda.mb_ipc_close(self.ctx, C.c_void_p(dptr)))

    def peer_attach(self, rank, peers, bounds, reads_per_rank):
        """peers: list of PeerMem, one per rank, addresses valid in this process (entry `rank` = own peer_alloc)."""
        arr = (PeerMem * len(peers))(*peers)
        bounds = np.ascontiguousarray(bounds, dtype=np.uint64)
        self._ck(_cuda.mb_peer_attach(self.ctx, rank, len(peers), arr, _ptr(bounds) if len(bounds) else None, int(reads_per_rank)))

    def extract_routed(self):
        n = C.c_size_t(0)
        self._ck(_cuda.mb_extract_kmers_routed(self.ctx, C.byref(n)))
        return n.value

    def take_routed_kmers(self):
        """-> (k-mers that arrived, overflowed?)"""
        n = C.c_size_t(0)
        rc = _cuda.mb_take_routed_kmers(self.ctx, C.byref(n))
        if rc == MB_E_CAPACITY:
            return n.value, True
        self._ck(rc)
        return n.value, False

    def match_routed(self):
        n = C.c_size_t(0)
        self._ck(_cuda.mb_match_kmers_routed(self.ctx, C.byref(n)))
        return n.value

    def take_routed_matches(self):
        n = C.c_size_t(0)
        rc = _cuda.mb_take_routed_matches(self.ctx, C.byref(n))
        if rc == MB_E_CAPACITY:
            return n.value, True
        self._ck(rc)
        return n.value, False

    def timings(self):
        t = PhaseTimes()
        self._ck(_cuda.mb_get_timings(self.ctx, C.byref(t)))
        return t.as_dict()

    def stream(self):
        return _cuda.mb_stream(self.ctx)

    def set_profile(self, on):
        self._ck(_cuda.mb_set_profile(self.ctx, int(on)))

    def kernel_times(self):
        """-> {kernel name: (ms summed over the launches of the last batch, launches)}"""
        arr = (KernelTime * 64)()
        n = C.c_size_t(0)
        self._ck(_cuda.mb_get_kernel_times(self.ctx, arr, 64, C.byref(n)))
        return {arr[i].name.decode(): (arr[i].ms, arr[i].launches) for i in range(n.value) if arr[i].launches}

    # ---- whole path, host buffers in / host buffers out (mb_classify_batch) ----
    def classify(self, bases, offsets, lens, bases_ptr=None):
        self.n_reads = len(lens) // 2
        res = np.zeros(self.n_reads, dtype=RESULT_DT)
        cap = max(1024, 8 * self.n_reads)
        while True:
            ids = np.zeros(cap, dtype=np.int32)
            cnt = np.zeros(cap, dtype=np.int32)
            used = C.c_size_t(0)
            bp = bases_ptr if bases_ptr is not None else _ptr(bases)
            rc = _cuda.mb_classify_batch(self.ctx, bp, len(bases), _ptr(offsets), _ptr(lens), self.n_reads, _ptr(res),
                                         _ptr(ids), _ptr(cnt), cap, C.byref(used))
            if rc == MB_E_CAPACITY and used.value > cap:
                cap = used.value
                continue
            self._ck(rc)
            return res, ids[:used.value], cnt[:used.value]


    def classify_into(self, bases_ptr, n_bases, offsets, lens, res, ids, cnt):
        """mb_classify_batch with caller-owned (e.g. pinned) buffers; returns the number of taxCnt entries.
        Raises MetabuliError(MB_E_CAPACITY) when ids/cnt are too small."""
        self.n_reads = len(lens) // 2
        used = C.c_size_t(0)
        rc = _cuda.mb_classify_batch(self.ctx, bases_ptr, n_bases, _ptr(offsets), _ptr(lens), self.n_reads, _ptr(res),
                                     _ptr(ids), _ptr(cnt), len(ids), C.byref(used))
        self._ck(rc)
        return used.value


class Classifier(Context):
    """A database directory opened through the C++ host library (host/mb_host.h): parses taxonomyDB,
    taxID_list, db.parameters, cuts the shard and makes everything resident on the GPU."""

    def __init__(self, dbdir, device=0, seq_mode=2, min_score=0.0, min_sp_score=0.0, shard=0, n_shards=1, cut_fractions=None):
        if cut_fractions is not None:
            self._cuts = np.ascontiguousarray(cut_fractions, dtype=np.float64)
            assert len(self._cuts) == n_shards - 1
            self.h = _host.mbh_open_cuts(os.fsencode(dbdir), device, seq_mode, min_score, min_sp_score, shard, n_shards, _ptr(self._cuts))
        else:
            self.h = _host.mbh_open(os.fsencode(dbdir), device, seq_mode, min_score, min_sp_score, shard, n_shards)
        if not self.h:
            raise MetabuliError(-1, _host.mbh_last_error().decode())
        self.ctx = _host.mbh_ctx(self.h)
        self.n_reads = 0

    def close(self):
        if self.h:
            _host.mbh_close(self.h)
            self.h = None

    @property
    def params(self):
        p = Params()
        _host.mbh_params(self.h, C.byref(p))
        return p

    def shard_bounds(self):
        n = _host.mbh_num_bounds(self.h)
        out = np.zeros(max(n, 1), dtype=np.uint64)
        _host.mbh_bounds(self.h, _ptr(out))
        return out[:n]

    def db_kmers(self):
        return _host.mbh_db_kmers(self.h)

    def db_words(self):
        return _host.mbh_db_words(self.h)

    def original_taxid(self, t):
        return _host.mbh_original_taxid(self.h, int(t))


class RawContext(Context):
    """An mb_ctx created straight from arrays (mb_create / mb_load_taxonomy / mb_load_db): what a host that
    already holds the database in memory would do.  bench.py uses it for the synthetic workload."""

    def __init__(self, **kw):
        p = Params()
        _cuda.mb_default_params(C.byref(p))
        for k, v in kw.items():
            setattr(p, k, v)
        self._params = p
        self.ctx = _cuda.mb_create(C.byref(p))
        if not self.ctx:
            raise MetabuliError(-1, _cuda.mb_global_error().decode())
        self._n_db = 0
        self._n_words = 0

    def close(self):
        if self.ctx:
            _cuda.mb_destroy(self.ctx)
            self.ctx = None

    @property
    def params(self):
        return self._params

    def load_taxonomy(self, parent, species_of, species_rank_of, node0_taxid=1, eukaryota_taxid=0):
        parent = np.ascontiguousarray(parent, dtype=np.int32)
        species_of = np.ascontiguousarray(species_of, dtype=np.int32)
        species_rank_of = np.ascontiguousarray(species_rank_of, dtype=np.int32)
        self._ck(_cuda.mb_load_taxonomy(self.ctx, _ptr(parent), _ptr(species_of), _ptr(species_rank_of), len(parent) - 1,
                                        node0_taxid, eukaryota_taxid))

    def load_db(self, words, info):
        words = np.ascontiguousarray(words, dtype=np.uint16)
        info = np.ascontiguousarray(info, dtype=np.uint32)
        self._ck(_cuda.mb_load_db(self.ctx, _ptr(words), len(words), _ptr(info), len(info)))
        self._n_db, self._n_words = len(info), len(words)

    def db_kmers(self):
        return self._n_db

    def db_words(self):
        return self._n_words


def classify_files(reads1, reads2, dbdir, outdir, jobid, seq_mode=2, device=0, min_score=0.0, min_sp_score=0.0, batch_reads=0):
    rc = _host.mbh_classify_files(os.fsencode(reads1), os.fsencode(reads2 or ""), os.fsencode(dbdir), os.fsencode(outdir),
                                  jobid.encode(), seq_mode, device, min_score, min_sp_score, batch_reads)
    if rc != 0:
        raise MetabuliError(rc, _host.mbh_last_error().decode())


def write_outputs(dbdir, outdir, jobid, names, res, ids, cnt, lineage=False, krona_prelude=None):
    """The host Reporter alone (no GPU needed): classifications / report / krona files from per-read results."""
    res = np.ascontiguousarray(res, dtype=RESULT_DT)
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    cnt = np.ascontiguousarray(cnt, dtype=np.int32)
    arr = (C.c_char_p * len(names))(*[n.encode() if isinstance(n, str) else n for n in names])
    rc = _host.mbh_write_outputs(os.fsencode(dbdir), os.fsencode(outdir), jobid.encode(), int(lineage),
                                 os.fsencode(krona_prelude) if krona_prelude else None, len(names), arr, _ptr(res), _ptr(ids), _ptr(cnt))
    if rc != 0:
        raise MetabuliError(rc, _host.mbh_last_error().decode())
