"""Prefix sharing between same-length haplotypes (the GPU analogue of the reference's trie path,
M2/trie/buildTreeUtils.cpp:155-258 + avx-pairhmm-template.h:475-721).

CPU part: the host-side plan (pairhmm_cuda_prefix_plan) is replayed by a small model of the checkpoint slots -- every
sweep must resume from a checkpoint written under exactly its own prefix. GPU part: the prefix-sharing kernel gives the
same bits as the per-pair sweep (as the reference's trie path does against its per-pair path, SURVEY.md 8c probe 2) and
stays within tolerance of the oracle."""
import numpy as np
import pytest

import fgdp_b200
from fgdp_b200 import synth
from util import assert_parity

TILE = 4          # kTile: columns per tile of the haplotype word layout
SLOTS = 4         # kPfxSlots


def _haps_of(b, g):
    h0, h1 = int(b.reg_hap_start[g]), int(b.reg_hap_start[g + 1])
    return [b.hap_bases[int(b.hap_off[h]):int(b.hap_off[h]) + int(b.hap_len[h])] for h in range(h0, h1)]


def _replay(haps, unit_haps):
    """checks a plan; returns (columns swept, columns of a sweep without sharing)"""
    p = fgdp_b200.prefix_plan(haps, unit_haps)
    n = len(haps)
    assert sorted(p["order"].tolist()) == list(range(n))
    lens = [len(haps[i]) for i in p["order"]]
    assert lens == sorted(lens), "sweep order is by length first"
    hmax = max([len(h) for h in haps] + [0])
    if unit_haps > 0:                             # as plan_region: runs of (nearly) equal length, balanced_unit_haps()
        g0 = min(unit_haps, 32)
        units = (max(1, n) + g0 - 1) // g0
        group = (max(1, n) + units - 1) // units
    else:                                         # default_unit_haps (engine.cu): runs of equal length
        g0 = min(max(1, n), max(1, 16384 // (hmax + 1)), 32)
        units = (max(1, n) + g0 - 1) // g0
        group = (max(1, n) + units - 1) // units
    slots = {}
    swept = shared = 0
    for q in range(n):
        hp = bytes(haps[int(p["order"][q])])
        H = len(hp)
        code_tiles = (H + TILE - 1) // TILE
        ts, rs, tv, ss = int(p["t_start"][q]), int(p["restore_slot"][q]), int(p["t_save"][q]), int(p["save_slot"][q])
        if q % group == 0:
            slots = {}                    # a new work unit starts with no checkpoints
            assert ts == 0 and rs == -1
        if ts == 0:
            assert rs == -1
        else:
            assert 0 <= rs < SLOTS and rs in slots, "resumes from a slot nobody wrote"
            ck_hap, ck_tiles = slots[rs]
            assert ck_tiles == ts and len(ck_hap) == H, "initial Y row is 2^s/H: only equal lengths share (template.h:115)"
            assert ts <= code_tiles
            ncol = min(H, ts * TILE)
            assert ck_hap[:ncol] == hp[:ncol], "checkpoint was written under another prefix"
            if ts * TILE > H:             # the checkpoint covers the padded last tile: the whole haplotype is equal
                assert ck_hap == hp
        if tv:
            assert ts < tv <= code_tiles and 0 <= ss < SLOTS
            assert ss != rs, "a sweep never overwrites the checkpoint it came from (a later haplotype may need it)"
            slots[ss] = (hp, tv)
        swept += H - min(H, ts * TILE)
        shared += min(H, ts * TILE)
    assert shared == p["shared_cols"]
    return swept, sum(len(h) for h in haps)


def test_plan_replay_on_snv_families():
    tot_swept = tot_full = 0
    for H, n_haps, unit in ((300, 12, 0), (301, 20, 0), (37, 9, 0), (5, 6, 0), (3, 6, 0), (1, 4, 0), (300, 40, 0),
                            (300, 12, 5), (150, 16, 3), (300, 12, 1)):
        b = synth.snv_family(6, 2, n_haps, 20, H, indel_frac=0.2, seed_extra=H + n_haps)
        for g in range(b.n_regions):
            s, f = _replay(_haps_of(b, g), unit)
            tot_swept += s; tot_full += f
    assert tot_swept < 0.85 * tot_full, "SNV families share a good part of their columns"


def test_plan_uses_the_packers_balanced_groups():
    """10 haplotypes at 8 per unit are swept as two runs of 5 (balanced_unit_haps), not 8 + 2: the plan the test hook
    returns is the one the packer runs -- sharing must not cross position 5"""
    b = synth.snv_family(1, 2, 10, 20, 120, seed_extra=77)
    haps = _haps_of(b, 0)
    p = fgdp_b200.prefix_plan(haps, 8)
    assert int(p["t_start"][5]) == 0 and int(p["restore_slot"][5]) == -1
    _replay(haps, 8)


def test_plan_special_cases():
    rng = np.random.default_rng(3)
    a = synth.ACGT[rng.integers(0, 4, 64)]
    # no haplotypes / one / empty haplotypes / identical haplotypes / all lengths different
    assert fgdp_b200.prefix_plan([])["shared_cols"] == 0
    _replay([a], 0)
    _replay([a[:0], a[:0], a], 0)
    s, f = _replay([a, a.copy(), a.copy()], 0)
    assert s == 64 and f == 192                                  # duplicates are swept once
    s, f = _replay([a[:10], a[:11], a[:12], a[:13]], 0)
    assert s == f                                                # different lengths never share
    # a deep chain: hap k differs from the first at position 8*k -> the stack of checkpoints is limited to SLOTS
    chain = []
    for k in range(1, 8):
        c = a.copy(); c[8 * k] = ord("N")
        chain.append(c)
    s, f = _replay(chain + [a], 0)
    assert s < f
    # lower-case and unknown bytes are compared as bytes: equal bytes => equal base codes (pairhmm_common.cc:26)
    lo = a.copy(); lo[40] = ord("a")
    _replay([a, lo, lo.copy()], 0)


# ---- GPU ----------------------------------------------------------------------------------------------------------
def _run(b, **kw):
    e = fgdp_b200.Engine(**kw)
    try:
        e.resident_load(b)
        e.resident_run()
        raw = e.resident_fetch_raw()
        out, resc = e.resident_fetch()
        st = e.stats()
    finally:
        e.close()
    return raw, out, resc, st


def _batches():
    yield synth.config1("clean")                                            # 8 haplotypes of 300 bases
    yield synth.config1("lowq")
    yield synth.snv_family(24, 40, 12, 150, 300)
    yield synth.snv_family(8, 30, 40, 150, 203, indel_frac=0.2)             # > 32 haplotypes: several units per read
    yield synth.snv_family(12, 33, 9, 101, [1, 2, 3, 4, 5, 7, 8, 9, 31, 64, 127, 129], profile="lowq", indel_frac=0.1)
    yield synth.snv_family(6, 21, 16, 250, 1001, max_snvs=1, dup_frac=0.3)
    yield synth.snv_family(3, 2, 6, 37, 300)                                # few reads: haplotype runs are split
    yield synth.edge_cases()
    yield synth.config3(96, 6)
    yield synth.config2(12)


@pytest.mark.gpu
def test_prefix_sharing_is_bit_identical_to_the_per_pair_sweep():
    shared_somewhere = 0
    for b in _batches():
        raw0, out0, resc0, st0 = _run(b, dedup_reads=False)
        # 2 = share whatever the plan finds (1 would skip batches that share little); unit_haps: these batches are small, the
        # packer would otherwise cut the haplotype runs into single-haplotype units to fill the GPU
        raw1, out1, resc1, st1 = _run(b, dedup_reads=False, prefix_sharing=2, unit_haps=7 if b.name.endswith("lowq") else 32)
        bad = np.nonzero(raw0.view(np.uint32) != raw1.view(np.uint32))[0]
        assert bad.size == 0, "%s: %d of %d FP32 sums differ in bits, first at %s" % (b.name, bad.size, raw0.size, bad[:8])
        assert np.array_equal(out0.view(np.uint64), out1.view(np.uint64)) and np.array_equal(resc0, resc1), b.name
        assert st0["cells_swept"] == st0["cells"] and st1["cells"] == st0["cells"]
        assert st1["cells_swept"] <= st0["cells_swept"]
        shared_somewhere += st1["cells_swept"] < st0["cells_swept"]
    assert shared_somewhere >= 6


@pytest.mark.gpu
@pytest.mark.parametrize("lane_group", [8, 16, 32])
def test_prefix_sharing_every_lane_group_width(lane_group):
    for R, H in ((1, 9), (17, 40), (64, 64), (150, 300), (251, 90), (384, 500)):
        if (R + lane_group - 1) // lane_group > 12:
            continue
        b = synth.snv_family(5, 11, 10, R, H, seed_extra=R)
        raw0, out0, resc0, _ = _run(b, lane_group=lane_group)
        raw1, out1, resc1, st1 = _run(b, lane_group=lane_group, prefix_sharing=2, unit_haps=32)
        assert np.array_equal(raw0.view(np.uint32), raw1.view(np.uint32)), (lane_group, R, H)
        assert np.array_equal(out0.view(np.uint64), out1.view(np.uint64)) and np.array_equal(resc0, resc1)
        assert st1["cells_swept"] < st1["cells"]


@pytest.mark.gpu
def test_prefix_sharing_against_the_oracle_through_the_region_call(oracle):
    """pairhmm_cuda_compute / _compute_batch (host buffers, chunking, result scatter back to the caller's haplotype
    order) with sharing on"""
    e = fgdp_b200.Engine(prefix_sharing=True, unit_haps=32)
    try:
        for b in (synth.snv_family(16, 50, 12, 150, 300), synth.snv_family(4, 64, 33, 150, 120, profile="lowq", indel_frac=0.3),
                  synth.config1("clean")):
            want, wresc, _ = oracle.compute(b.as_flat(), threads=4)
            got, gresc = e.compute_batch(b)
            assert_parity(got, gresc, want, wresc, what="prefix " + b.name)
        b = synth.snv_family(1, 64, 8, 150, 300)
        h0, h1 = int(b.reg_hap_start[0]), int(b.reg_hap_start[1])
        reads = []
        for r in range(b.n_reads):
            o, n = int(b.read_off[r]), int(b.read_len[r])
            reads.append(tuple(a[o:o + n] for a in (b.read_bases, b.base_q, b.ins_q, b.del_q, b.gcp)))
        want, wresc, _ = oracle.compute(b.as_flat(), threads=4)
        e.initialize(150, 300)
        got, gresc = e.compute(reads, _haps_of(b, 0))
        assert_parity(got, gresc, want, wresc, what="prefix region call")
        assert 0 < e.stats()["cells_swept"] < e.stats()["cells"]
    finally:
        e.close()


@pytest.mark.gpu
def test_prefix_sharing_across_many_chunks_and_mixed_batches(oracle):
    """a batch cut into many launch groups (chunk_cells), SNV families and indel-only regions interleaved: the choice is made
    per launch group, the results do not depend on it"""
    a, c = synth.snv_family(10, 40, 12, 150, 300), synth.config2(10)
    parts = []
    for g in range(10):
        parts += [a.subset([g]), c.subset([g])]
    cat = lambda f: np.concatenate([getattr(p, f) for p in parts])
    b = synth.RegionBatch(cat("read_bases"), cat("base_q"), cat("ins_q"), cat("del_q"), cat("gcp"), None, cat("read_len"),
                          cat("hap_bases"), None, cat("hap_len"),
                          np.concatenate([[0], np.cumsum([p.n_reads for p in parts])]).astype(np.int64),
                          np.concatenate([[0], np.cumsum([p.n_haps for p in parts])]).astype(np.int64), "mixed")
    b.read_off = synth._offsets(b.read_len); b.hap_off = synth._offsets(b.hap_len)
    want, wresc, _ = oracle.compute(b.as_flat(), threads=4)
    outs = []
    for kw in ({}, {"prefix_sharing": True, "unit_haps": 32, "chunk_cells": 150_000_000}, {"prefix_sharing": 2, "unit_haps": 32, "chunk_cells": 400_000_000}):
        e = fgdp_b200.Engine(**kw)
        try:
            got, gresc = e.compute_batch(b)
            st = e.stats()
        finally:
            e.close()
        assert_parity(got, gresc, want, wresc, what="mixed %s" % kw)
        outs.append((got, gresc, st))
    for got, gresc, _ in outs[1:]:
        assert np.array_equal(got.view(np.uint64), outs[0][0].view(np.uint64)) and np.array_equal(gresc, outs[0][1])
    assert outs[1][2]["cells_swept"] < outs[0][2]["cells_swept"] and outs[2][2]["cells_swept"] <= outs[1][2]["cells_swept"]


@pytest.mark.gpu
def test_prefix_sharing_is_taken_only_where_it_pays():
    """prefix_sharing=True (config.reserved[4] == 1): a batch whose haplotypes all differ in length is swept by the plain
    packed kernel, an SNV family by the sharing one"""
    for b, shares in ((synth.config2(12), False), (synth.snv_family(24, 40, 12, 150, 300), True)):
        _, out0, resc0, st0 = _run(b, dedup_reads=False, unit_haps=32)
        _, out1, resc1, st1 = _run(b, dedup_reads=False, unit_haps=32, prefix_sharing=True)
        assert np.array_equal(out0.view(np.uint64), out1.view(np.uint64)) and np.array_equal(resc0, resc1)
        assert (st1["cells_swept"] < st0["cells_swept"]) == shares, b.name
