"""GPU test of mb_sort_kmers_noalloc (the allocation-free entry point added): its output must be
bit-identical (the tolerance of every integer test here) to mb_sort_kmers' on the same k-mers, on two calls in a row that
share the static scratch array.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mb():
    import metabuli_work_b200.binding as b
    return b


def _random_kmers(mb, rng, n, fmt):
    k = np.zeros(n, dtype=mb.KMER_DT)
    if fmt == 2:
        # format-2 metamers: eight 5-bit residue codes in 0..20, then 24 DNA bits; few distinct leading residues so that
        # equal keys are common and stability is exercised
        aa = np.zeros(n, dtype=np.uint64)
        for j in range(8):
            hi = 21 if j >= 2 else 3
            aa = (aa << np.uint64(5)) | rng.integers(0, hi, n, dtype=np.uint64)
        k["value"] = (aa << np.uint64(24)) | rng.integers(0, 1 << 24, n, dtype=np.uint64)
    else:
        k["value"] = (rng.integers(0, 21 ** 8 // 1000, n, dtype=np.uint64) << np.uint64(24)) | rng.integers(0, 1 << 24, n, dtype=np.uint64)
    k["qinfo"] = np.arange(n, dtype=np.uint64)          # original position: makes every record unique, shows stability
    return k


def _sort_existing(ctx, k):
    ctx.set_kmers(k)
    ctx.sort_kmers()
    return ctx.kmers()


def _sort_noalloc(mb, ctx, k, stream=None):
    import torch
    n = len(k)
    d_k = torch.from_numpy(k.view(np.uint8).reshape(-1).copy()).cuda()
    d_t = torch.empty_like(d_k)
    torch.cuda.synchronize()
    ctx.sort_kmers_noalloc(d_k.data_ptr(), d_t.data_ptr(), n, stream)
    torch.cuda.synchronize()
    out = d_k.cpu().numpy().view(mb.KMER_DT).copy()
    del d_k, d_t
    return out


@pytest.mark.parametrize("fmt,bits", [(2, 0), (2, 40), (1, 0)])
def test_noalloc_sort_matches_sort_kmers(mb, fmt, bits):
    rng = np.random.default_rng(77 + fmt + bits)
    ctx = mb.RawContext(kmer_format=fmt, kmer_sort_bits=bits)
    try:
        # two calls in a row on the shared scratch: different sizes and contents, the second smaller than the first (what
        # the first call left in the scratch beyond the second call's tiles must not matter), then the first again
        sizes = [300_001, 70_003, 300_001]
        inputs = [_random_kmers(mb, rng, n, fmt) for n in sizes]
        expect = [_sort_existing(ctx, k) for k in inputs]
        for k, e in zip(inputs, expect):
            got = _sort_noalloc(mb, ctx, k)
            assert got.tobytes() == e.tobytes()
        # stream-ordered form on the context's own stream handle
        got = _sort_noalloc(mb, ctx, inputs[1], stream=ctx.stream())
        assert got.tobytes() == expect[1].tobytes()
    finally:
        ctx.close()


@pytest.mark.parametrize("fmt,bits", [(2, 0), (2, 31), (2, 40), (1, 0)])
def test_sort_is_the_stable_sort_by_key_at_tile_edges(mb, fmt, bits):
    """mb_sort_kmers and mb_sort_kmers_noalloc against numpy's stable sort by the same key (tests/helpers.kmer_sort_key), at
    sizes around the radix pass' 4096-record tile: one record, a partial tile, exactly one tile (the bulk-copy path), one
    more, several tiles and a partial one.  Equal keys are common, so a rank that is not stable shows."""
    import helpers
    rng = np.random.default_rng(1000 + 10 * fmt + bits)
    ctx = mb.RawContext(kmer_format=fmt, kmer_sort_bits=bits)
    try:
        for n in (1, 31, 4095, 4096, 4097, 8192, 3 * 4096 + 5, 40 * 4096 + 777):
            k = _random_kmers(mb, rng, n, fmt)
            key = helpers.kmer_sort_key(k["value"], bits, fmt)
            expect = k[np.argsort(key, kind="stable")]
            got = _sort_existing(ctx, k)
            assert got.tobytes() == expect.tobytes(), (n, "mb_sort_kmers")
            got = _sort_noalloc(mb, ctx, k)
            assert got.tobytes() == expect.tobytes(), (n, "mb_sort_kmers_noalloc")
    finally:
        ctx.close()


def test_noalloc_limits(mb):
    import torch
    ctx = mb.RawContext(kmer_format=2)
    try:
        buf = torch.zeros(64, dtype=torch.uint8, device="cuda")
        with pytest.raises(mb.MetabuliError) as e:
            ctx.sort_kmers_noalloc(buf.data_ptr(), buf.data_ptr() + 32, mb.MB_NOALLOC_MAX_KMERS + 1)
        assert e.value.code == mb.MB_E_CAPACITY and "MB_NOALLOC_MAX_KMERS" in str(e.value)
        with pytest.raises(mb.MetabuliError):
            ctx.sort_kmers_noalloc(buf.data_ptr(), buf.data_ptr(), 2)          # same buffer twice
        ctx.sort_kmers_noalloc(0, 0, 0)                                        # empty input: nothing to do
        # the largest admitted size runs (scratch sized for it)
        n = mb.MB_NOALLOC_MAX_KMERS
        rng = np.random.default_rng(3)
        k = _random_kmers(mb, rng, n, 2)
        got = _sort_noalloc(mb, ctx, k)
        assert got.tobytes() == _sort_existing(ctx, k).tobytes()
    finally:
        ctx.close()
