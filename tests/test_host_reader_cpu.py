"""CPU tests of the C++ host reader (FastxReader + loadBatch, metabuli_work_b200/host/mb_host_io.cpp) through the
mbh_read_fastx tap: the record model of kseq (lib/mmseqs/src/commons/KSeqWrapper.h:30 — name up to the first whitespace,
sequence lines joined, gzip transparently) and the batch bounds of runClassify's host path.  The names are also compared
with the ones the reference's own reader produced (the results.tsv fixtures written by oracle/_ref/ref_harness)."""
import gzip
import os

import numpy as np
import pytest

import helpers
import orc


@pytest.fixture(scope="module")
def mb():
    import metabuli_work_b200.binding as b
    return b


def _records(names, bases, offsets, lens, paired):
    out = []
    for i, n in enumerate(names):
        a = bases[int(offsets[2 * i]):int(offsets[2 * i]) + int(lens[2 * i])].tobytes()
        b = bases[int(offsets[2 * i + 1]):int(offsets[2 * i + 1]) + int(lens[2 * i + 1])].tobytes()
        out.append((n, a, b) if paired else (n, a))
    return out


@pytest.mark.parametrize("case", ["plain_paired", "plain_single", "plain_long", "mask_paired", "mask_long"])
def test_golden_inputs(golden, mb, case):
    files = [os.path.join(helpers.DATA, f) for f in golden.cases[case]["reads"]]
    paired = len(files) > 1
    names, s1, s2 = golden.reads(case)
    got = _records(*mb.host_read_fastx(files[0], files[1] if paired else None), paired)
    want = [(n, a, b) for n, a, b in zip(names, s1, s2)] if paired else list(zip(names, s1))
    assert got == want
    # the names the reference's reader saw (kseq through the harness), in file order
    assert [g[0] for g in got] == [r[1] for r in golden.results(case)]
    # a single-end read's second mate is empty and sits where the next read starts
    if not paired:
        _, bases, offsets, lens = mb.host_read_fastx(files[0])
        assert not lens[1::2].any() and np.array_equal(offsets[1::2], offsets[0::2] + lens[0::2])


RECORDS = [("r1", b"ACGTACGTTGCA" * 9), ("second/1", b"TTGACC"), ("empty", b""), ("n4", b"NNNNACGTNNNN" * 21), ("last", b"GATTACA" * 40)]


def _fasta(path, width, opener=open, header_extra=" some description\twith a tab", trailing_newline=True, crlf=False):
    nl = b"\r\n" if crlf else b"\n"
    with opener(path, "wb") as f:
        for i, (n, s) in enumerate(RECORDS):
            f.write(b">" + n.encode() + header_extra.encode() + nl)
            lines = [s[k:k + width] for k in range(0, len(s), width)] if width else [s]
            body = nl.join(lines)
            last = i + 1 == len(RECORDS)
            f.write(body + (nl if (not last or trailing_newline) else b""))


def _fastq(path, width, opener=open, crlf=False):
    nl = b"\r\n" if crlf else b"\n"
    with opener(path, "wb") as f:
        for n, s in RECORDS:
            f.write(b"@" + n.encode() + b" 1:N:0" + nl)
            lines = [s[k:k + width] for k in range(0, len(s), width)] if width else [s]
            f.write(nl.join(lines) + nl + b"+" + nl)
            # qualities that start with '@' and '+': only their LENGTH ends a record
            q = (b"@+" * len(s))[:len(s)]
            qlines = [q[k:k + width] for k in range(0, len(q), width)] if width else [q]
            f.write(nl.join(qlines) + nl)


@pytest.mark.parametrize("kind,width,gz,crlf", [("fa", 0, False, False), ("fa", 60, False, False), ("fa", 7, True, False), ("fa", 60, False, True),
                                                 ("fq", 0, False, False), ("fq", 50, False, False), ("fq", 0, True, False), ("fq", 11, True, True)])
def test_record_model(mb, tmp_path, kind, width, gz, crlf):
    """Wrapped FASTA and FASTQ, gzip, CR LF, descriptions behind the name, an empty record, qualities that look like headers."""
    path = str(tmp_path / ("x." + kind + (".gz" if gz else "")))
    (_fasta if kind == "fa" else _fastq)(path, width, opener=gzip.open if gz else open, crlf=crlf)
    got = _records(*mb.host_read_fastx(path), False)
    assert got == [(n, s) for n, s in RECORDS]


def test_fasta_without_final_newline(mb, tmp_path):
    path = str(tmp_path / "x.fa")
    _fasta(path, 60, trailing_newline=False, header_extra="")
    assert _records(*mb.host_read_fastx(path), False) == [(n, s) for n, s in RECORDS]


def test_batch_bounds_and_ragged_mates(mb, tmp_path):
    """loadBatch stops at --batch-reads reads, or before the first read that starts at or beyond the base bound; a second file
    with fewer records leaves empty second mates (KmerExtractor pairs by position)."""
    p1, p2 = str(tmp_path / "a.fq"), str(tmp_path / "b.fq")
    _fastq(p1, 0)
    with open(p2, "wb") as f:
        for n, s in RECORDS[:3]:
            f.write(b"@" + n.encode() + b"\n" + s[::-1] + b"\n+\n" + b"I" * len(s) + b"\n")
    got = _records(*mb.host_read_fastx(p1, p2), True)
    assert got == [(n, s, RECORDS[i][1][::-1] if i < 3 else b"") for i, (n, s) in enumerate(RECORDS)]
    assert len(mb.host_read_fastx(p1, max_reads=2)[0]) == 2
    first = len(RECORDS[0][1])
    assert len(mb.host_read_fastx(p1, max_bases=first)[0]) == 1            # the bound is checked before a read is taken
    assert len(mb.host_read_fastx(p1, max_bases=first + 1)[0]) == 2
    with pytest.raises(mb.MetabuliError):
        mb.host_read_fastx(str(tmp_path / "missing.fq"))
