"""Seeded synthetic inputs for the five BASELINE.json configs.

Everything is flat SoA numpy so that the CPU checkers and the CUDA engine read identical bytes:

  RegionBatch
    read_bases/base_q/ins_q/del_q/gcp : uint8[sum R]   five per-read byte arrays, concatenated
    read_off int64[n_reads], read_len int32[n_reads]
    hap_bases uint8[sum H], hap_off int64[n_haps], hap_len int32[n_haps]
    reg_read_start int64[n_regions+1], reg_hap_start int64[n_regions+1]
  Pairs are implied: every read of a region x every haplotype of that region, results read-major
  per region (out[r * nHaps + h], VectorLoglessPairHMM.cpp:123-125), regions concatenated.

Quality values are what the kernel sees, i.e. *after* PairHMMLikelihoodCalculationEngine's capping
(base q <18 -> 6, ins/del floor 6, gcp == 10; PairHMMLikelihoodCalculationEngine.cpp:256-292).
"""
from dataclasses import dataclass
import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)


@dataclass
class RegionBatch:
    read_bases: np.ndarray
    base_q: np.ndarray
    ins_q: np.ndarray
    del_q: np.ndarray
    gcp: np.ndarray
    read_off: np.ndarray
    read_len: np.ndarray
    hap_bases: np.ndarray
    hap_off: np.ndarray
    hap_len: np.ndarray
    reg_read_start: np.ndarray
    reg_hap_start: np.ndarray
    name: str = ""

    @property
    def n_regions(self):
        return int(self.reg_read_start.shape[0] - 1)

    @property
    def n_reads(self):
        return int(self.read_len.shape[0])

    @property
    def n_haps(self):
        return int(self.hap_len.shape[0])

    def region_out_offsets(self):
        nr = np.diff(self.reg_read_start)
        nh = np.diff(self.reg_hap_start)
        off = np.zeros(self.n_regions + 1, np.int64)
        np.cumsum(nr * nh, out=off[1:])
        return off

    @property
    def n_pairs(self):
        return int(self.region_out_offsets()[-1])

    def cells(self):
        """sum over pairs of R*H (the GCUPS numerator)."""
        tot = 0
        for g in range(self.n_regions):
            r0, r1 = self.reg_read_start[g], self.reg_read_start[g + 1]
            h0, h1 = self.reg_hap_start[g], self.reg_hap_start[g + 1]
            tot += int(self.read_len[r0:r1].astype(np.int64).sum()) * int(self.hap_len[h0:h1].astype(np.int64).sum())
        return tot

    def pair_lists(self):
        """explicit (pair_read, pair_hap) uint32 arrays in output order."""
        pr, ph = [], []
        for g in range(self.n_regions):
            r0, r1 = int(self.reg_read_start[g]), int(self.reg_read_start[g + 1])
            h0, h1 = int(self.reg_hap_start[g]), int(self.reg_hap_start[g + 1])
            pr.append(np.repeat(np.arange(r0, r1, dtype=np.uint32), h1 - h0))
            ph.append(np.tile(np.arange(h0, h1, dtype=np.uint32), r1 - r0))
        if not pr:
            return np.zeros(0, np.uint32), np.zeros(0, np.uint32)
        return np.concatenate(pr), np.concatenate(ph)

    def as_flat(self):
        """view with explicit pair lists, the shape the CPU checkers take (oracle_py._Lib.compute)."""
        pr, ph = self.pair_lists()
        return FlatPairs(self.read_bases, self.base_q, self.ins_q, self.del_q, self.gcp, self.read_off,
                         self.read_len, self.hap_bases, self.hap_off, self.hap_len, pr, ph)

    def subset(self, regions):
        """a new batch holding only the given region indices (used for bounded CPU samples)."""
        rb, bq, iq, dq, gc, rl, hb, hl = [], [], [], [], [], [], [], []
        rrs, rhs = [0], [0]
        for g in regions:
            r0, r1 = int(self.reg_read_start[g]), int(self.reg_read_start[g + 1])
            h0, h1 = int(self.reg_hap_start[g]), int(self.reg_hap_start[g + 1])
            a, b = int(self.read_off[r0]), int(self.read_off[r1 - 1] + self.read_len[r1 - 1])
            for src, dst in ((self.read_bases, rb), (self.base_q, bq), (self.ins_q, iq), (self.del_q, dq), (self.gcp, gc)):
                dst.append(src[a:b])
            rl.append(self.read_len[r0:r1])
            a, b = int(self.hap_off[h0]), int(self.hap_off[h1 - 1] + self.hap_len[h1 - 1])
            hb.append(self.hap_bases[a:b])
            hl.append(self.hap_len[h0:h1])
            rrs.append(rrs[-1] + (r1 - r0))
            rhs.append(rhs[-1] + (h1 - h0))
        rl = np.concatenate(rl).astype(np.int32)
        hl = np.concatenate(hl).astype(np.int32)
        return RegionBatch(np.concatenate(rb), np.concatenate(bq), np.concatenate(iq), np.concatenate(dq),
                           np.concatenate(gc), _offsets(rl), rl, np.concatenate(hb), _offsets(hl), hl,
                           np.asarray(rrs, np.int64), np.asarray(rhs, np.int64), self.name + "/subset")


@dataclass
class FlatPairs:
    read_bases: np.ndarray
    base_q: np.ndarray
    ins_q: np.ndarray
    del_q: np.ndarray
    gcp: np.ndarray
    read_off: np.ndarray
    read_len: np.ndarray
    hap_bases: np.ndarray
    hap_off: np.ndarray
    hap_len: np.ndarray
    pair_read: np.ndarray
    pair_hap: np.ndarray


def _offsets(lens):
    off = np.zeros(len(lens), np.int64)
    if len(lens) > 1:
        np.cumsum(lens[:-1], out=off[1:])
    return off


# --------------------------------------------------------------------------------------------
# generators
# --------------------------------------------------------------------------------------------
def _rand_bases(rng, n, n_rate=0.001):
    b = ACGT[rng.integers(0, 4, size=n)]
    if n_rate > 0 and n > 0:
        b = b.copy()
        b[rng.random(n) < n_rate] = ord("N")
    return b


def _variant_hap(rng, hap0, target_len):
    """hap0 with 1-3 SNVs and one 1-5 bp indel, padded/trimmed to target_len."""
    h = hap0.copy()
    for _ in range(int(rng.integers(1, 4))):
        p = int(rng.integers(0, len(h)))
        h[p] = ACGT[(int(np.searchsorted(ACGT, h[p]) if h[p] in ACGT else 0) + int(rng.integers(1, 4))) % 4]
    p = int(rng.integers(1, max(2, len(h) - 6)))
    n = int(rng.integers(1, 6))
    if rng.random() < 0.5:
        h = np.concatenate([h[:p], _rand_bases(rng, n, 0.0), h[p:]])
    else:
        h = np.concatenate([h[:p], h[p + n:]])
    if len(h) < target_len:
        h = np.concatenate([h, _rand_bases(rng, target_len - len(h), 0.0)])
    return np.ascontiguousarray(h[:target_len])


def _quals(rng, shape, profile):
    n = int(np.prod(shape))
    if profile == "clean":
        bq = np.clip(np.rint(rng.normal(32.0, 6.0, size=n)), 2, 41).astype(np.uint8)
        bq[bq < 18] = 6
        iq = np.full(n, 45, np.uint8)
        dq = np.full(n, 45, np.uint8)
        m = rng.random(n) < 0.10
        iq[m] = rng.integers(20, 41, size=int(m.sum()))
        m = rng.random(n) < 0.10
        dq[m] = rng.integers(20, 41, size=int(m.sum()))
    elif profile == "lowq":
        bq = rng.integers(18, 26, size=n).astype(np.uint8)
        bq[rng.random(n) < 0.70] = 6
        iq = rng.integers(10, 31, size=n).astype(np.uint8)
        dq = rng.integers(10, 31, size=n).astype(np.uint8)
    else:
        raise ValueError(profile)
    gc = np.full(n, 10, np.uint8)
    return bq.reshape(shape), iq.reshape(shape), dq.reshape(shape), gc.reshape(shape)


def _region(rng, n_reads, hap_lens, R, profile, sub_rate=None, foreign_frac=0.0):
    """one region: haps (list of arrays) + [n_reads, R] read arrays. `foreign_frac` of the reads are
    random sequence unrelated to any haplotype (mis-mapped reads): these are the pairs whose float
    likelihood underflows 1e-28 and take the double-precision rescue."""
    H0 = int(hap_lens[0])
    hap0 = _rand_bases(rng, H0)
    haps = [hap0] + [_variant_hap(rng, hap0, int(L)) for L in hap_lens[1:]]
    nh = len(haps)
    Hmax = max(len(h) for h in haps)
    hmat = np.empty((nh, max(Hmax, R)), np.uint8)
    hmat[:] = ACGT[rng.integers(0, 4, size=hmat.shape)]
    for i, h in enumerate(haps):
        hmat[i, :len(h)] = h
    src = rng.integers(0, nh, size=n_reads)
    lens = np.array([len(h) for h in haps])[src]
    start = (rng.random(n_reads) * np.maximum(lens - R, 0).astype(np.float64)).astype(np.int64)
    idx = start[:, None] + np.arange(R)[None, :]
    bases = hmat[src[:, None], idx]
    bq, iq, dq, gc = _quals(rng, (n_reads, R), profile)
    # sequencing errors at 10^(-q/10) (clean) and/or an explicit per-read substitution rate
    perr = np.power(10.0, -bq.astype(np.float64) / 10.0)
    if sub_rate is not None:
        perr = np.maximum(perr, rng.uniform(sub_rate[0], sub_rate[1], size=(n_reads, 1)))
    err = rng.random((n_reads, R)) < perr
    bases = np.where(err, ACGT[rng.integers(0, 4, size=(n_reads, R))], bases)
    if foreign_frac > 0:
        foreign = rng.random(n_reads) < foreign_frac
        bases = np.where(foreign[:, None], ACGT[rng.integers(0, 4, size=(n_reads, R))], bases)
    bases = np.where(rng.random((n_reads, R)) < 0.001, np.uint8(ord("N")), bases).astype(np.uint8)
    return haps, bases, bq, iq, dq, gc


def make_batch(seed, n_reads_per_region, hap_lens_per_region, R, profile="clean", sub_rate=None, name="",
               foreign_frac=0.0, region_ids=None):
    """Every region is generated from its own RNG stream (seed, region id), so any subset of a job can be
    generated by any rank without generating the rest (used by the LPT sharding of bench.py)."""
    rb, bq, iq, dq, gc, hb, hl, rl = [], [], [], [], [], [], [], []
    rrs, rhs = [0], [0]
    if region_ids is None:
        region_ids = range(len(n_reads_per_region))
    for gid, nr, hls in zip(region_ids, n_reads_per_region, hap_lens_per_region):
        rng = np.random.default_rng([int(seed), int(gid)])
        haps, b, q, i, d, c = _region(rng, int(nr), hls, R, profile, sub_rate, foreign_frac)
        rb.append(b.reshape(-1)); bq.append(q.reshape(-1)); iq.append(i.reshape(-1))
        dq.append(d.reshape(-1)); gc.append(c.reshape(-1))
        rl.append(np.full(int(nr), R, np.int32))
        for h in haps:
            hb.append(h)
            hl.append(len(h))
        rrs.append(rrs[-1] + int(nr))
        rhs.append(rhs[-1] + len(haps))
    rl = np.concatenate(rl) if rl else np.zeros(0, np.int32)
    hl = np.asarray(hl, np.int32)
    cat = lambda xs: np.concatenate(xs) if xs else np.zeros(0, np.uint8)
    return RegionBatch(cat(rb), cat(bq), cat(iq), cat(dq), cat(gc), _offsets(rl), rl, cat(hb), _offsets(hl), hl,
                       np.asarray(rrs, np.int64), np.asarray(rhs, np.int64), name)


SEED0 = 0xB200


def config1(profile="clean"):
    """C1: one region, 64 reads x 8 haps, 150 bp x 300 bp (BASELINE.json configs[0])."""
    return make_batch(SEED0 + 1, [64], [np.full(8, 300)], 150, profile, name="C1-" + profile)


def config2_shapes(n_regions, seed_extra=0):
    """region shapes of C2 without generating any bases: (reads per region, list of haplotype lengths, cost)
    with cost = sum(read lengths) * sum(haplotype lengths) = DP cells of the region."""
    rng = np.random.default_rng(SEED0 + 2 + 7919 * seed_extra)
    nr = np.clip(rng.poisson(200, size=n_regions), 50, 400)
    nh = rng.integers(4, 17, size=n_regions)
    base = rng.integers(250, 501, size=n_regions)
    hls = [np.clip(b + rng.integers(-10, 11, size=k), 250, 500) for b, k in zip(base, nh)]
    cost = np.array([int(n) * 150 * int(h.sum()) for n, h in zip(nr, hls)], np.int64)
    return nr, hls, cost


def config2(n_regions=10000, profile="clean", seed_extra=0, select=None):
    """C2: WES-like, ~200 reads x 4-16 haps per region, R=150, H in [250,500] (configs[1]). `select` = indices
    of the regions to actually generate (default all)."""
    nr, hls, _ = config2_shapes(n_regions, seed_extra)
    ids = np.arange(n_regions) if select is None else np.asarray(select, np.int64)
    return make_batch(SEED0 + 2 + 104729 * seed_extra, nr[ids], [hls[i] for i in ids], 150, profile,
                      name="C2-%s-%d" % (profile, len(ids)), foreign_frac=0.02, region_ids=ids)


def lpt_shard(costs, n_shards):
    """Longest-processing-time-first split of work items by cost: returns a list of index arrays, one per
    shard (cell-count-balanced split of regions over GPUs, no collective)."""
    import heapq
    costs = np.asarray(costs, np.int64)
    order = np.argsort(-costs, kind="stable")
    heap = [(0, s) for s in range(n_shards)]
    heapq.heapify(heap)
    out = [[] for _ in range(n_shards)]
    for i in order:
        load, s = heapq.heappop(heap)
        out[s].append(int(i))
        heapq.heappush(heap, (load + int(costs[i]), s))
    return [np.sort(np.asarray(o, np.int64)) for o in out]


def config3(n_reads=2000, n_haps=32, profile="lowq"):
    """C3: high-depth stress, 2000 reads x 32 haps, R=150, H~600, rescue heavy (configs[2])."""
    rng = np.random.default_rng(SEED0 + 3)
    return make_batch(SEED0 + 3, [n_reads], [rng.integers(590, 611, size=n_haps)], 150, profile,
                      sub_rate=(0.05, 0.25), name="C3-" + profile, foreign_frac=0.30)


def config4(n_regions=64, n_reads=128, n_haps=8, profile="clean"):
    """C4: R=250, H log-uniform in [1000,2000], wide variance (configs[3])."""
    rng = np.random.default_rng(SEED0 + 4)
    hls = [np.exp(rng.uniform(np.log(1000), np.log(2000), size=n_haps)).astype(np.int64) for _ in range(n_regions)]
    return make_batch(SEED0 + 4, [n_reads] * n_regions, hls, 250, profile, name="C4-" + profile,
                      foreign_frac=0.02)


def config5_shapes(n_regions=25000, shard=0):
    """region shapes of C5 without generating any bases: (list of haplotype lengths, cost = DP cells per region)"""
    rng = np.random.default_rng(SEED0 + 5 + 7919 * shard)
    hls = [rng.integers(300, 501, size=10) for _ in range(n_regions)]
    cost = np.array([200 * 150 * int(h.sum()) for h in hls], np.int64)
    return hls, cost


def config5(n_regions=25000, profile="clean", shard=0, select=None):
    """C5: chr20-scale, regions x 200 reads x 10 haps, R=150, H in [300,500] (configs[4]); `shard`
    reseeds so every rank / shard gets distinct regions; `select` = indices of the regions to actually generate
    (one rank's part of an LPT split of the whole job)."""
    hls, _ = config5_shapes(n_regions, shard)
    ids = np.arange(n_regions) if select is None else np.asarray(select, np.int64)
    return make_batch(SEED0 + 5 + 104729 * shard, [200] * len(ids), [hls[i] for i in ids], 150, profile,
                      name="C5-%s-%d-s%d" % (profile, len(ids), shard), foreign_frac=0.02, region_ids=ids)


def snv_family(n_regions=64, n_reads=64, n_haps=12, R=150, H=300, profile="clean", seed_extra=0, max_snvs=3,
               dup_frac=0.15, indel_frac=0.0):
    """The workload the reference's trie path is built for (VectorLoglessPairHMM.cpp:35: nHaps / nDistinctLens > 3):
    a region's haplotypes are the reference haplotype plus variants that differ from it (or from another variant)
    by a few SNVs only, so most have the same length and share long prefixes. `dup_frac` of the haplotypes repeat an
    earlier one exactly, `indel_frac` carry a 1-5 bp indel on top (another length: never shared). H may be an int or
    a per-region list. Reads as in the other configs."""
    Hs = [int(H)] * n_regions if np.isscalar(H) else [int(x) for x in H]
    b = make_batch(SEED0 + 7 + seed_extra, [n_reads] * n_regions, [np.full(n_haps, h) for h in Hs], R, profile,
                   name="snv-family-" + profile, foreign_frac=0.1)
    hb, hl = [], []
    for g in range(n_regions):
        rng = np.random.default_rng([SEED0 + 77 + seed_extra, g])
        h0 = int(b.reg_hap_start[g])
        fam = [b.hap_bases[int(b.hap_off[h0]):int(b.hap_off[h0]) + int(b.hap_len[h0])].copy()]
        for _ in range(1, n_haps):
            par = fam[int(rng.integers(0, len(fam)))].copy()
            if rng.random() >= dup_frac and len(par):
                for _ in range(int(rng.integers(1, max_snvs + 1))):
                    p = int(rng.integers(0, len(par)))
                    par[p] = ACGT[(int(np.searchsorted(ACGT, par[p]) if par[p] in ACGT else 0) + int(rng.integers(1, 4))) % 4]
            if rng.random() < indel_frac and len(par) > 8:
                p, n = int(rng.integers(1, len(par) - 6)), int(rng.integers(1, 6))
                par = np.concatenate([par[:p], _rand_bases(rng, n, 0.0), par[p:]]) if rng.random() < 0.5 else np.concatenate([par[:p], par[p + n:]])
            fam.append(np.ascontiguousarray(par))
        perm = rng.permutation(n_haps)                      # the caller's order is arbitrary
        for k in perm:
            hb.append(fam[int(k)]); hl.append(len(fam[int(k)]))
    b.hap_len = np.asarray(hl, np.int32)
    b.hap_bases = np.concatenate(hb) if hb else np.zeros(0, np.uint8)
    b.hap_off = _offsets(b.hap_len)
    return b


def long_reads(seed=4321, uniform_gcp=True, profile="clean"):
    """reads longer than one warp's register block (384 rows), mixed with ordinary ones in the same regions: lengths
    around the stripe boundaries of the long-read kernel (256-row stripes) and well beyond, haplotypes shorter and
    longer than the reads, N bases. uniform_gcp=False perturbs the gap-continuation bytes (general kernels)."""
    rng = np.random.default_rng(seed)
    regions = [([385, 150, 400, 7], [300, 600, 601]), ([512, 513, 384, 511], [700, 90]), ([768, 769, 1025, 33], [1200, 1, 40]),
               ([2049, 151, 640], [2500, 2100]), ([1000] * 3 + [150] * 6, [1500] * 5)]
    rb, bq, iq, dq, gc, rl, hb, hl, rrs, rhs = [], [], [], [], [], [], [], [], [0], [0]
    for Rs, Hs in regions:
        base = _rand_bases(rng, max(Hs) + max(Rs), 0.002)
        haps = []
        for k, h in enumerate(Hs):
            hh = base[:h].copy()
            if k and h > 8:
                hh[int(rng.integers(0, h))] = ACGT[int(rng.integers(0, 4))]
            haps.append(hh); hb.append(hh); hl.append(h)
        for r, R in enumerate(Rs):
            st = int(rng.integers(0, max(1, max(Hs) - R + 1)))
            b = base[st:st + R].copy()
            if r % 4 == 3:
                b = _rand_bases(rng, R)                    # unrelated read: underflows in float -> double pass
            mut = rng.random(R) < 0.01
            b[mut] = ACGT[rng.integers(0, 4, size=int(mut.sum()))]
            q, i_, d_, c_ = _quals(rng, (R,), profile)
            if not uniform_gcp:
                c_ = c_.copy(); c_[rng.random(R) < 0.3] = 7
            rb.append(b); bq.append(q); iq.append(i_); dq.append(d_); gc.append(c_); rl.append(R)
        rrs.append(rrs[-1] + len(Rs)); rhs.append(rhs[-1] + len(Hs))
    rl = np.asarray(rl, np.int32); hl = np.asarray(hl, np.int32)
    return RegionBatch(np.concatenate(rb), np.concatenate(bq), np.concatenate(iq), np.concatenate(dq),
                       np.concatenate(gc), _offsets(rl), rl, np.concatenate(hb), _offsets(hl), hl,
                       np.asarray(rrs, np.int64), np.asarray(rhs, np.int64), "long-reads-" + profile + ("" if uniform_gcp else "-gcp"))


def edge_cases(seed=1234):
    """ragged edge-case batch: R in {1,2,15,...,255}, H in {1..700} incl. R>H, N bases, lower-case and
    junk bytes (behave as 'A'), quals with the high bit set, per-base gcp != 10, quals 0/6/93/127."""
    rng = np.random.default_rng(seed)
    Rs = [1, 2, 15, 16, 17, 31, 32, 33, 48, 63, 64, 65, 95, 96, 97, 127, 128, 129, 150, 151, 159, 160, 161, 191, 192,
          193, 223, 224, 225, 250, 255, 256]
    rb, bq, iq, dq, gc, rl, hb, hl, rrs, rhs = [], [], [], [], [], [], [], [], [0], [0]
    for gi, R in enumerate(Rs):
        nh = 4
        Hs = [1, max(1, R // 2), R + 3, int(rng.integers(R, 700))][:nh]
        base = _rand_bases(rng, max(Hs) + R, 0.01)
        nr = 5
        for h in Hs:
            hh = base[:h].copy()
            if h > 4 and gi % 3 == 0:
                hh[int(rng.integers(0, h))] = ord("a")      # lower case -> behaves as 'A'
                hh[int(rng.integers(0, h))] = ord("x")      # junk -> behaves as 'A'
            hb.append(hh); hl.append(h)
        for r in range(nr):
            st = int(rng.integers(0, max(1, max(Hs) - R + 1)))
            b = base[st:st + R].copy()
            if len(b) < R:
                b = np.concatenate([b, _rand_bases(rng, R - len(b))])
            mut = rng.random(R) < (0.02 if r < 3 else 0.3)
            b[mut] = ACGT[rng.integers(0, 4, size=int(mut.sum()))]
            q = rng.choice(np.array([0, 6, 18, 20, 30, 41, 60, 93, 127], np.uint8), size=R)
            i_ = rng.choice(np.array([0, 6, 10, 30, 45, 93, 127], np.uint8), size=R)
            d_ = rng.choice(np.array([0, 6, 10, 30, 45, 93, 127], np.uint8), size=R)
            c_ = rng.choice(np.array([1, 6, 10, 10, 10, 40], np.uint8), size=R)
            if r == 4:  # high bit set: must be masked with &127 (ReadForPairHMM.cpp:35)
                q = q | 0x80; i_ = i_ | 0x80; d_ = d_ | 0x80; c_ = c_ | 0x80
            rb.append(b); bq.append(q.astype(np.uint8)); iq.append(i_.astype(np.uint8))
            dq.append(d_.astype(np.uint8)); gc.append(c_.astype(np.uint8)); rl.append(R)
        rrs.append(rrs[-1] + nr); rhs.append(rhs[-1] + nh)
    rl = np.asarray(rl, np.int32); hl = np.asarray(hl, np.int32)
    return RegionBatch(np.concatenate(rb), np.concatenate(bq), np.concatenate(iq), np.concatenate(dq),
                       np.concatenate(gc), _offsets(rl), rl, np.concatenate(hb), _offsets(hl), hl,
                       np.asarray(rrs, np.int64), np.asarray(rhs, np.int64), "edge")


# --------------------------------------------------------------------------------------------
# generating a large job on several host processes (every region has its own RNG stream, so any split gives the same bytes)
# --------------------------------------------------------------------------------------------
def concat(batches, name=""):
    """batches that hold consecutive region ranges of one job -> one RegionBatch"""
    batches = list(batches)
    cat = lambda f, dt: (np.concatenate([getattr(b, f) for b in batches]) if batches else np.zeros(0, dt))
    rl = cat("read_len", np.int32).astype(np.int32)
    hl = cat("hap_len", np.int32).astype(np.int32)
    rrs, rhs = [np.zeros(1, np.int64)], [np.zeros(1, np.int64)]
    r0 = h0 = 0
    for b in batches:
        rrs.append(b.reg_read_start[1:] + r0); rhs.append(b.reg_hap_start[1:] + h0)
        r0 += b.n_reads; h0 += b.n_haps
    return RegionBatch(cat("read_bases", np.uint8), cat("base_q", np.uint8), cat("ins_q", np.uint8), cat("del_q", np.uint8),
                       cat("gcp", np.uint8), _offsets(rl), rl, cat("hap_bases", np.uint8), _offsets(hl), hl,
                       np.concatenate(rrs), np.concatenate(rhs), name or (batches[0].name if batches else ""))


def _gen_part(args):
    which, n_regions, ids, kw = args
    return {"c2": config2, "c5": config5}[which](n_regions, select=ids, **kw)


def generate_parallel(which, n_regions, procs=0, select=None, **kw):
    """config2 / config5 generated on `procs` host processes (0 = as many as there are cores, at most 16): the same bytes
    as the serial generator, in a fraction of the time (the bench's workloads are 2-5 million reads)"""
    import multiprocessing as mp
    import os
    ids = np.arange(n_regions) if select is None else np.asarray(select, np.int64)
    procs = procs or max(1, min(16, (os.cpu_count() or 1)))
    if procs <= 1 or len(ids) < 4 * procs:
        return _gen_part((which, n_regions, ids if select is not None else None, kw))
    parts = [p for p in np.array_split(ids, procs * 4) if len(p)]
    with mp.get_context("fork").Pool(procs) as pool:
        out = pool.map(_gen_part, [(which, n_regions, p, kw) for p in parts])
    b = concat(out)
    b.name = "%s-%d" % (which.upper(), len(ids))
    return b
