"""Synthetic workload of BASELINE.json configs[1] for bench.py (bench infrastructure, not product):
1M synthetic 2x150 paired reads vs a 10k-genome synthetic SYNCMER database.

Everything is generated on the GPU box at run time from a seed (there is no network and nothing this
large can travel): a 7-rank taxonomy, 10 000 genomes of 200-500 kb that diverge along it, the database
(pseudo-genes -> syncmer metamers, computed here with torch and independently of the library -> (value, species)
de-duplication with LCA -> 15-bit delta stream + info, the on-disk format of IndexCreator.cpp:811-886),
and Illumina-like read pairs (1 % substitutions, 2 % random pairs).  torch is used as plumbing for the
random numbers, sorting and gathers of this SET-UP; the timed region only runs the library.
"""
import ctypes as C
import time

import numpy as np
import torch

CODE2ASCII = np.frombuffer(b"ACTG", dtype=np.uint8)   # library base codes: A0 C1 T2 G3


class Taxonomy:
    """root(1) > domain(2) > phylum(3) > class(4) > order(5) > F families > G genera > S species > T strains.
    Internal tax IDs are dense; dense arrays are what mb_load_taxonomy takes."""

    def __init__(self, F, G, S, T):
        self.F, self.G, self.S, self.T = F, G, S, T
        ids = [(0, -1, ""), (1, 1, "no rank"), (2, 1, "superkingdom"), (3, 2, "phylum"), (4, 3, "class"), (5, 4, "order")]
        names = {1: "root", 2: "Bacteria", 3: "Synthphyla", 4: "Synthclassia", 5: "Synthorderales"}
        nxt = 6
        self.family, self.genus, self.species, self.strain = [], [], [], []
        for f in range(F):
            fam = nxt; nxt += 1
            ids.append((fam, 5, "family")); names[fam] = f"Fam{f}"
            self.family.append(fam)
            for g in range(G):
                gen = nxt; nxt += 1
                ids.append((gen, fam, "genus")); names[gen] = f"Gen{f}_{g}"
                self.genus.append(gen)
                for s in range(S):
                    sp = nxt; nxt += 1
                    ids.append((sp, gen, "species")); names[sp] = f"Gen{f}_{g} sp{s}"
                    self.species.append(sp)
                    for t in range(T):
                        st = nxt; nxt += 1
                        ids.append((st, sp, "no rank")); names[st] = f"Gen{f}_{g} sp{s} str{t}"
                        self.strain.append(st)
        self.nodes = ids
        self.names = names
        n = nxt
        self.parent = np.full(n, -1, dtype=np.int32)
        self.rank = [""] * n
        for t, p, r in ids[1:]:
            self.parent[t] = p
            self.rank[t] = r
        # getTaxIdAtRank(t, "species") (TaxonomyWrapper.cpp:479-498): walk up while the rank is below species
        self.species_rank_of = np.zeros(n, dtype=np.int32)
        order = {"no rank": -1, "species": 4, "genus": 8, "family": 12, "order": 17, "class": 21, "phylum": 24, "superkingdom": 28}
        for t in range(2, n):
            cur, steps = t, 0
            while steps < 30 and order[self.rank[cur]] < 4:
                cur = self.parent[cur]
                steps += 1
            self.species_rank_of[t] = t if steps == 30 else cur
        # taxId2speciesId for the DB's taxa (strains) and their ancestors up to species (KmerMatcher.cpp:56-120)
        self.species_of = np.zeros(n, dtype=np.int32)
        for st in self.strain:
            sp = self.parent[st]
            self.species_of[st] = sp
            self.species_of[sp] = sp
        self.genome_tax = np.array(self.strain, dtype=np.int64)
        self.genome_species = self.parent[self.genome_tax].astype(np.int64)
        self.n = n

    def use_internal_ids(self, ext2int):
        """The DB stores the INTERNAL tax IDs the reference's TaxonomyWrapper assigned (tax_map.tsv written by
        oracle/_ref/ref_mkdb); order and `info` must use them."""
        m = np.zeros(self.n, dtype=np.int64)
        for e, i in ext2int.items():
            if e < self.n:
                m[e] = i
        self.genome_tax = m[self.genome_tax]
        self.genome_species = m[self.genome_species]
        self.int_of = m


def _mutate(seq, rate, gen):
    """Substitute a fraction `rate` of the codes (uint8 0..3) with a different base, in place, chunked by rows."""
    rows = max(1, (1 << 28) // max(1, seq.shape[-1]))
    flat = seq.view(-1, seq.shape[-1])
    thr = int(rate * 65536)
    for r0 in range(0, flat.shape[0], rows):
        blk = flat[r0:r0 + rows]
        r = torch.randint(0, 65536 * 4, blk.shape, generator=gen, device=blk.device, dtype=torch.int32)
        hit = (r >> 2) < thr
        shift = (r & 3).clamp_(max=2).to(torch.uint8) + 1
        blk.copy_(torch.where(hit, (blk + shift) & 3, blk))
    return seq


def make_genomes(tax, seed, len_min, len_max, device):
    """-> (codes uint8[total] packed genomes on `device`, starts int64[n+1] on the host)."""
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    L = len_max
    root = torch.randint(0, 4, (1, L), generator=gen, device=device, dtype=torch.uint8)
    fam = _mutate(root.repeat(tax.F, 1), 0.25, gen)
    genus = _mutate(fam.repeat_interleave(tax.G, dim=0), 0.10, gen)
    del fam
    species = _mutate(genus.repeat_interleave(tax.S, dim=0), 0.06, gen)
    del genus
    n = len(tax.strain)
    lens = torch.randint(len_min, len_max + 1, (n,), generator=gen, device=device, dtype=torch.int64)
    starts = torch.zeros(n + 1, dtype=torch.int64, device=device)
    starts[1:] = torch.cumsum(lens, 0)
    total = int(starts[-1].item())
    codes = torch.empty(total, dtype=torch.uint8, device=device)
    rows = max(1, (1 << 29) // L)
    pos = torch.arange(L, device=device)
    for g0 in range(0, n, rows):
        g1 = min(n, g0 + rows)
        sp_idx = torch.arange(g0, g1, device=device) // tax.T
        blk = _mutate(species[sp_idx].clone(), 0.01, gen)
        keep = pos.unsqueeze(0) < lens[g0:g1].unsqueeze(1)
        codes[int(starts[g0].item()):int(starts[g1].item())] = blk[keep]
    return codes, starts.cpu().numpy()


def _encode_deltas(values_sorted_u):
    """values as int64 holding the uint64 bit pattern, ascending in UNSIGNED order -> uint16 numpy delta stream
    (IndexCreator::getDiffIdx, IndexCreator.cpp:868-886): 15-bit groups, most significant first, last flagged."""
    v = values_sorted_u
    d = torch.empty_like(v)
    d[0] = v[0]
    d[1:] = v[1:] - v[:-1]
    def lsr(x, k):                                      # logical shift right on int64 bit patterns
        return (x >> k) & ((1 << (64 - k)) - 1) if k else x

    nw = torch.ones_like(d)
    for k in (15, 30, 45, 60):
        nw += (lsr(d, k) != 0).to(torch.int64)
    end = torch.cumsum(nw, 0)
    total = int(end[-1].item())
    words = torch.zeros(total, dtype=torch.int32, device=v.device)
    for j in range(5):
        m = nw > j
        idx = end[m] - 1 - j
        w = (lsr(d[m], 15 * j) & 0x7FFF).to(torch.int32)
        if j == 0:
            w |= 0x8000
        words[idx] = w
    return words.to(torch.int16).cpu().numpy().view(np.uint16), end


FLIP = -(1 << 63)


def genetic_code_tables():
    """codonF / codonR of the standard genetic code over ARNDCQEGHILKMFPSTWYVX with 3-bit codon ids, indexed by
    b0 | b1 << 2 | b2 << 4 over base codes A0 C1 T2 G3: cell = aa | cid << 5 (GeneticCode.h:34-193; the reverse table is the
    codon comp(b2) comp(b1) comp(b0), KmerScanner.h:94-98).  Built here from the TCAG table, independently of the library."""
    alphabet = "ARNDCQEGHILKMFPSTWYVX"
    tcag = "FFLLSSSSYYXXCCXWLLLLPPPPHHQQRRRRIIIMTTTTNNKKSSRRVVVVAAAADDEEGGGG"
    t2c = [2, 1, 0, 3]
    aa = np.zeros((4, 4, 4), dtype=np.int64)
    cid = np.zeros((4, 4, 4), dtype=np.int64)
    for a in range(4):
        for b in range(4):
            for c in range(4):
                x, y, z = t2c[a], t2c[b], t2c[c]
                aa[x, y, z] = alphabet.index(tcag[a * 16 + b * 4 + c])
                cid[x, y, z] = z
    cid[0, 3, 3], cid[0, 3, 0] = 4, 5          # Arg AGG, AGA
    cid[2, 2, 3], cid[2, 2, 0] = 4, 5          # Leu TTG, TTA
    cid[0, 3, 2], cid[0, 3, 1] = 6, 7          # Ser AGT, AGC
    cid[2, 3, 0] = 5                           # stop TGA
    f = np.zeros(64, dtype=np.int64)
    r = np.zeros(64, dtype=np.int64)
    for i in range(64):
        b0, b1, b2 = i & 3, (i >> 2) & 3, (i >> 4) & 3
        f[i] = aa[b0, b1, b2] | (cid[b0, b1, b2] << 5)
        r0, r1, r2 = b2 ^ 2, b1 ^ 2, b0 ^ 2
        r[i] = aa[r0, r1, r2] | (cid[r0, r1, r2] << 5)
    return f, r


def syncmer_metamers(codes, starts_rel, device, smer_len=5, syncmer=True, kmer_format=2):
    """Format-2 metamers (MetamerScanner::next, KmerScanner.h:82-117) of every 24-base window of the packed genomes `codes`
    (uint8 codes on `device`) on both strands, filtered by the closed-syncmer rule of SyncmerScanner::next
    (SyncmerScanner.h:36-102).  Pure torch (set-up plumbing); the library under test is not involved.
    -> (value int64 [uint64 bit pattern], window start inside its genome int64, genome index int64, strand*3 + start % 3 int64)."""
    n = codes.numel()
    f_np, r_np = genetic_code_tables()
    cf = torch.from_numpy(f_np).to(device)
    cr = torch.from_numpy(r_np).to(device)
    c = codes.to(torch.int64)
    idx3 = c[:-2] | (c[1:-1] << 2) | (c[2:] << 4)              # codon starting at every base
    del c
    m = n - 23                                                  # window starts
    st = torch.from_numpy(np.asarray(starts_rel, dtype=np.int64)).to(device)
    p = torch.arange(m, device=device, dtype=torch.int64)
    g = torch.searchsorted(st, p, right=True) - 1
    # frame range of the reference (KmerExtractor.cpp:312-386, LocalUtil.h:45-59): a window of phase b = start % 3 must end
    # inside [b, b + cov - 1], cov = L - 2 / L - 4 / L - 3 for L % 3 = 2 / 1 / 0
    glen = st[g + 1] - st[g]
    rem = glen % 3
    cov = torch.where(rem == 2, glen - 2, torch.where(rem == 1, glen - 4, glen - 3))
    rel = p - st[g]
    inside = rel + 23 <= rel % 3 + cov - 1
    p = rel
    out = []
    s = smer_len
    last = 8 - s
    mask = (1 << (5 * s)) - 1
    for strand in (0, 1):
        tab = (cr if strand else cf)[idx3]
        aa = torch.zeros(m, dtype=torch.int64, device=device)
        dna = torch.zeros(m, dtype=torch.int64, device=device)
        if kmer_format == 2:
            for j in range(8):
                cj = tab[3 * j:3 * j + m]
                sh = j if strand else 7 - j                     # reverse strand: translation runs from the high end down
                aa |= (cj & 31) << (5 * sh)
                dna |= (cj >> 5) << (3 * sh)
        else:
            # legacy format 1 (OldMetamerScanner::next, KmerScanner.h:137-181): codons consumed from the far end, amino-acid
            # part a base-21 number; forward strand: highest stored codon most significant, reverse strand: lowest
            order = range(8) if strand else range(7, -1, -1)
            for j in order:
                cj = tab[3 * j:3 * j + m]
                aa = aa * 21 + (cj & 31)
                dna = (dna << 3) | (cj >> 5)
        del tab
        if syncmer:
            v0 = (aa >> (5 * last)) & mask
            vl = aa & mask
            mid = torch.full_like(aa, 1 << 40)
            for i in range(1, last):
                mid = torch.minimum(mid, (aa >> (5 * (last - i))) & mask)
            keep = inside & ((v0 <= torch.minimum(mid, vl)) | (vl < torch.minimum(mid, v0)))
            del v0, vl, mid
        else:
            keep = inside
        value = (aa << 24) | dna
        out.append((value[keep], p[keep], g[keep], p[keep] % 3 + 3 * strand))
        del aa, dna, value, keep
    return tuple(torch.cat([o[k] for o in out]) for k in range(4))


def build_db(tax, codes, starts, device, genomes_per_chunk=200, block=1500, log=None, syncmer=True, kmer_format=2):
    """-> dict(values int64 tensor (uint64 bit patterns, unsigned-ascending), info int64 tensor).  K-mers are the syncmer
    metamers of the genomes (syncmer_metamers above, torch only); one (strand, frame) is kept per 1.5-kb block, the density of
    a gene-finder driven build."""
    n = len(starts) - 1
    vals, taxs = [[] for _ in range(8)], [[] for _ in range(8)]
    t0 = time.time()
    for g0 in range(0, n, genomes_per_chunk):
        g1 = min(n, g0 + genomes_per_chunk)
        a, b = int(starts[g0]), int(starts[g1])
        value, pos, g_rel, frame = syncmer_metamers(codes[a:b], starts[g0:g1 + 1] - a, device, syncmer=syncmer, kmer_format=kmer_format)
        genome = g_rel + g0
        chosen = ((genome * 7919 + (pos // block) * 104729) >> 3) % 6
        keep = frame == chosen
        vk, gk = value[keep], genome[keep]
        tb = (vk >> 61) & 7                              # value range (top three bits): sorted separately below
        for bucket in range(8):
            mb_ = tb == bucket
            vals[bucket].append(vk[mb_].clone())
            taxs[bucket].append(gk[mb_].to(torch.int32))
        del value, pos, g_rel, frame, genome, chosen, keep, vk, gk, tb
    gtax_of = torch.from_numpy(tax.genome_tax).to(device)
    gsp_of = torch.from_numpy(tax.genome_species).to(device)
    # sorted and de-duplicated in eight value ranges (top three bits), one after another: torch sorts (and masks) at most
    # 2^31 - 1 elements at a time, and a GTDB-scale build has more k-mers than that before the de-duplication
    outs_v, outs_t = [], []
    for bucket in range(8):
        if not vals[bucket] or sum(len(x) for x in vals[bucket]) == 0:
            continue
        v = torch.cat(vals[bucket])
        g = torch.cat(taxs[bucket]).long()
        vals[bucket], taxs[bucket] = None, None
        gtax = gtax_of[g]
        gsp = gsp_of[g]
        del g
        # order (value unsigned, species, taxid): Kmer.h:77-87
        o1 = torch.argsort(gsp * (1 << 24) + gtax, stable=True)
        v, gtax, gsp = v[o1], gtax[o1], gsp[o1]
        o2 = torch.argsort(v ^ FLIP, stable=True)
        v, gtax, gsp = v[o2], gtax[o2], gsp[o2]
        del o1, o2
        # one entry per (value, species); tax id = LCA of the group's tax ids = the strain if unique, else the species
        first = torch.ones(len(v), dtype=torch.bool, device=device)
        first[1:] = (v[1:] != v[:-1]) | (gsp[1:] != gsp[:-1])
        gid = torch.cumsum(first.to(torch.int64), 0) - 1
        ng = int(gid[-1].item()) + 1
        tmin = torch.full((ng,), 1 << 40, dtype=torch.int64, device=device).scatter_reduce_(0, gid, gtax, "amin")
        tmax = torch.zeros(ng, dtype=torch.int64, device=device).scatter_reduce_(0, gid, gtax, "amax")
        o_sp = gsp[first]
        outs_v.append(v[first])
        outs_t.append(torch.where(tmin == tmax, tmin, o_sp))
        del v, gtax, gsp, first, gid, tmin, tmax, o_sp
    out_v = torch.cat(outs_v)
    out_t = torch.cat(outs_t)
    del outs_v, outs_t
    if log:
        log(f"db: {len(out_v)} k-mers from {n} genomes in {time.time() - t0:.1f}s")
    return {"values": out_v, "info": out_t}


def shard_db(db, shard, n_shards):
    """Cut at amino-acid group boundaries nearest to equal k-mer counts (what the `split` table encodes,
    IndexCreator.cpp:811-866) -> (words uint16 numpy, info uint32 numpy, bounds uint64 numpy[n_shards-1], n_kmers)."""
    v, t = db["values"], db["info"]
    n = len(v)
    cuts = [0]
    aa = v >> 24
    for g in range(1, n_shards):
        i = g * n // n_shards
        # first index >= i that starts a new AA group
        win = aa[i - 1:i + (1 << 16)]
        ch = torch.nonzero(win[1:] != win[:-1])
        i = i + int(ch[0].item()) if len(ch) else n
        cuts.append(i)
    cuts.append(n)
    bounds = np.array([int(v[c].item()) & 0xFFFFFFFFFF000000 for c in cuts[1:-1]], dtype=np.uint64)   # amino-acid part only
    a, b = cuts[shard], cuts[shard + 1]
    words, _ = _encode_deltas(v[a:b])
    info = t[a:b].to(torch.int32).cpu().numpy().astype(np.uint32)
    return words, info, bounds, b - a


def make_reads(codes, starts, n_pairs, seed, device, read_len=150, sub_rate=0.01, random_frac=0.02):
    """-> (bases uint8 numpy [n_pairs*2*read_len] ASCII, offsets uint64[2n], lens uint32[2n], genome index int64[n])."""
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    n = len(starts) - 1
    st = torch.from_numpy(starts).to(device)
    g = torch.randint(0, n, (n_pairs,), generator=gen, device=device)
    glen = st[g + 1] - st[g]
    insert = torch.randint(300, 501, (n_pairs,), generator=gen, device=device)
    off = (torch.rand(n_pairs, generator=gen, device=device, dtype=torch.float64) * (glen - insert).double()).long()
    base = st[g] + off
    ar = torch.arange(read_len, device=device)
    m1 = codes[(base.unsqueeze(1) + ar.unsqueeze(0))]
    m2 = codes[(base + insert - 1).unsqueeze(1) - ar.unsqueeze(0)] ^ 2          # reverse complement of the far end
    flip = torch.rand(n_pairs, generator=gen, device=device) < 0.5
    r1 = torch.where(flip.unsqueeze(1), m2, m1)
    r2 = torch.where(flip.unsqueeze(1), m1, m2)
    pair = torch.cat([r1, r2], dim=1)
    _mutate(pair, sub_rate, gen)
    rnd = torch.rand(n_pairs, generator=gen, device=device) < random_frac
    noise = torch.randint(0, 4, pair.shape, generator=gen, device=device, dtype=torch.uint8)
    pair = torch.where(rnd.unsqueeze(1), noise, pair)
    ascii_lut = torch.from_numpy(CODE2ASCII.copy()).to(device)
    bases = ascii_lut[pair.long()].reshape(-1).cpu().numpy()
    offsets = (np.arange(2 * n_pairs, dtype=np.uint64) * np.uint64(read_len))
    lens = np.full(2 * n_pairs, read_len, dtype=np.uint32)
    return bases, offsets, lens, g.cpu().numpy()


def write_fastq_pair(path1, path2, bases, read_len, n_pairs):
    b = bases[:n_pairs * 2 * read_len].reshape(n_pairs, 2, read_len)
    qual = b"I" * read_len
    with open(path1, "wb") as f1, open(path2, "wb") as f2:
        for i in range(n_pairs):
            f1.write(b"@r%d/1\n" % i + b[i, 0].tobytes() + b"\n+\n" + qual + b"\n")
            f2.write(b"@r%d/2\n" % i + b[i, 1].tobytes() + b"\n+\n" + qual + b"\n")


def write_taxdump(tax, outdir):
    import os
    os.makedirs(outdir, exist_ok=True)
    with open(f"{outdir}/nodes.dmp", "w") as fn, open(f"{outdir}/names.dmp", "w") as fm:
        for t, p, r in tax.nodes[1:]:
            fn.write(f"{t}\t|\t{p}\t|\t{r}\t|\t\t|\t0\t|\n")
            fm.write(f"{t}\t|\t{tax.names[t]}\t|\t\t|\tscientific name\t|\n")
    open(f"{outdir}/merged.dmp", "w").close()


def write_db_dir(dbdir, db, tax, split_num=4096, syncmer=1, smer_len=5, kmer_format=2):
    """Writes diffIdx / info / split / taxID_list / db.parameters of a Metabuli database directory (taxonomyDB comes
    from ref_mkdb).  Format: IndexCreator::writeTargetFilesAndSplits + getDiffIdx (IndexCreator.cpp:811-886) and
    writeDbParameters (:1245-1266)."""
    import os
    v, t = db["values"], db["info"]
    n = len(v)
    words, end = _encode_deltas(v)
    words.tofile(os.path.join(dbdir, "diffIdx"))
    t.to(torch.int32).cpu().numpy().astype(np.uint32).tofile(os.path.join(dbdir, "info"))
    # split table: after the k-th equal-count boundary, the first k-mer whose amino-acid part differs
    aa = v >> 24
    change = torch.nonzero(aa[1:] != aa[:-1]).flatten() + 1          # indices that start a new AA group
    size = n // (split_num - 1)
    splits = np.zeros((split_num, 3), dtype=np.uint64)
    si = 1
    bidx = (torch.arange(1, split_num, device=v.device, dtype=torch.int64) * size - 1).clamp_(max=n - 1)
    nxt = torch.searchsorted(change, bidx, right=True)                # first group start > boundary index
    has = nxt < len(change)
    j = torch.where(has, change[nxt.clamp(max=max(len(change) - 1, 0))] if len(change) else bidx, torch.full_like(bidx, -1))
    bidx_h, j_h = bidx.cpu().numpy(), j.cpu().numpy()
    jv = v[j.clamp(min=0)].cpu().numpy().astype(np.uint64)
    je = end[j.clamp(min=0)].cpu().numpy().astype(np.uint64)
    for k in range(split_num - 1):
        if size == 0:
            break
        jk = int(j_h[k])
        if jk < 0:
            break
        # a later boundary reached before this one resolved overwrites it (single `pending` state)
        if k + 1 < split_num - 1 and jk > int(bidx_h[k + 1]):
            continue
        if si < split_num:
            splits[si] = (jv[k], je[k], jk + 1)
            si += 1
    splits.tofile(os.path.join(dbdir, "split"))
    with open(os.path.join(dbdir, "taxID_list"), "w") as f:
        for x in sorted(set(int(x) for x in tax.genome_tax)):
            f.write(f"{x}\n")
    with open(os.path.join(dbdir, "db.parameters"), "w") as f:
        f.write("DB_name\tsynthetic\nCreation_date\t2026-01-01\n")
        f.write("Metabuli commit used to create the DB\tsynthetic\n")
        f.write("Reduced_alphabet\t0\nAccession_level\t0\nMask_mode\t0\nMask_prob\t0.900000\nSkip_redundancy\t1\n")
        f.write(f"Syncmer\t{syncmer}\n")
        if syncmer:
            f.write(f"Syncmer_len\t{smer_len}\n")
        if kmer_format != 1:             # legacy databases carry no Kmer_format line (classify.cpp:13, updateDB.cpp:16)
            f.write(f"Kmer_format\t{kmer_format}\n")
    return len(words)


def make_long_reads(codes, starts, n_reads, seed, device, len_min, len_max, sub=0.02, ins=0.01, dele=0.02, chunk=2000):
    """ONT-like reads (BASELINE.json configs[3]: sub / ins / del in the ratio 2:1:2) or contigs (sub only): fragments of
    random genomes, errors applied per base, half of them reverse-complemented.
    -> (bases uint8 numpy ASCII, offsets uint64[2n] with an empty second mate, lens uint32[2n])."""
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    n = len(starts) - 1
    st = torch.from_numpy(np.asarray(starts, dtype=np.int64)).to(device)
    ascii_lut = torch.from_numpy(CODE2ASCII.copy()).to(device)
    out_bases, out_lens = [], []
    for c0 in range(0, n_reads, chunk):
        c = min(chunk, n_reads - c0)
        g = torch.randint(0, n, (c,), generator=gen, device=device)
        glen = st[g + 1] - st[g]
        L = torch.randint(len_min, len_max + 1, (c,), generator=gen, device=device)
        L = torch.minimum(L, glen - 1)
        off = (torch.rand(c, generator=gen, device=device, dtype=torch.float64) * (glen - L).double()).long()
        rstart = torch.zeros(c + 1, dtype=torch.int64, device=device)
        rstart[1:] = torch.cumsum(L, 0)
        tot = int(rstart[-1].item())
        rid = torch.repeat_interleave(torch.arange(c, device=device), L)
        pos = torch.arange(tot, device=device) - rstart[rid]
        src = codes[st[g][rid] + off[rid] + pos]
        r = torch.rand(tot, generator=gen, device=device)
        is_sub = r < sub
        is_ins = (r >= sub) & (r < sub + ins)
        is_del = (r >= sub + ins) & (r < sub + ins + dele)
        shift = torch.randint(1, 4, (tot,), generator=gen, device=device, dtype=torch.uint8)
        src = torch.where(is_sub, (src + shift) & 3, src)
        cnt = (~is_del).to(torch.int64) + is_ins.to(torch.int64)
        opos = torch.cumsum(cnt, 0) - cnt
        ototal = int((opos[-1] + cnt[-1]).item()) if tot else 0
        out = torch.zeros(ototal, dtype=torch.uint8, device=device)
        keep = ~is_del
        out[opos[keep]] = src[keep]
        extra = torch.randint(0, 4, (int(is_ins.sum().item()),), generator=gen, device=device, dtype=torch.uint8)
        out[opos[is_ins] + 1] = extra
        # new read boundaries
        csum = torch.zeros(tot + 1, dtype=torch.int64, device=device)
        csum[1:] = torch.cumsum(cnt, 0)
        nstart = csum[rstart]
        nlen = nstart[1:] - nstart[:-1]
        # reverse complement of half of the reads
        flip = torch.rand(c, generator=gen, device=device) < 0.5
        orid = torch.repeat_interleave(torch.arange(c, device=device), nlen)
        p = torch.arange(ototal, device=device)
        rel = p - nstart[orid]
        srcp = torch.where(flip[orid], nstart[orid] + nlen[orid] - 1 - rel, p)
        out = out[srcp] ^ (flip[orid].to(torch.uint8) * 2)
        out_bases.append(ascii_lut[out.long()].cpu().numpy())
        out_lens.append(nlen.cpu().numpy())
    bases = np.concatenate(out_bases) if out_bases else np.zeros(0, dtype=np.uint8)
    rl = np.concatenate(out_lens).astype(np.uint32)
    lens = np.zeros(2 * n_reads, dtype=np.uint32)
    lens[0::2] = rl
    offsets = np.zeros(2 * n_reads, dtype=np.uint64)
    ends = np.cumsum(rl.astype(np.uint64))
    offsets[0::2][1:] = ends[:-1]
    offsets[1::2] = ends
    return bases, offsets, lens


def write_fasta(path, bases, offsets, lens, n_reads):
    with open(path, "wb") as f:
        for i in range(n_reads):
            o, l = int(offsets[2 * i]), int(lens[2 * i])
            f.write(b">r%09d\n" % i)
            f.write(bases[o:o + l].tobytes())
            f.write(b"\n")
