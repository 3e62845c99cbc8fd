"""Smith-Waterman with backtrace: the C restatement (oracle/sw_oracle.c) against the reference's own
aligner (golden vectors + live oracle/_ref/libref_sw.so), and -- marked gpu -- the CUDA aligner against the oracle."""
import os

import numpy as np
import pytest

import oracle_py

HERE = os.path.dirname(os.path.abspath(__file__))


def _golden():
    z = np.load(os.path.join(HERE, "golden", "sw.npz"))
    refs = [bytes(z["ref_bases"][int(o):int(o) + int(n)]) for o, n in zip(z["ref_off"], z["ref_len"])]
    alts = [bytes(z["alt_bases"][int(o):int(o) + int(n)]) for o, n in zip(z["alt_off"], z["alt_len"])]
    cases = [(tuple(int(x) for x in z["par%d" % k][:4]), int(z["par%d" % k][4]), [str(c) for c in z["cig%d" % k]], z["off%d" % k])
             for k in range(int(z["n_cases"]))]
    return refs, alts, cases


def test_sw_oracle_matches_reference_golden():
    """every CIGAR string and offset equal to the committed outputs of the reference's SmithWaterman_align"""
    refs, alts, cases = _golden()
    p = oracle_py.SWPairs(refs, alts)
    assert len(cases) == 12
    for params, strat, cig, off in cases:
        c, o = oracle_py.oracle_sw(p, params, strat)
        assert c == cig and np.array_equal(o, off), (params, strat)


def test_sw_oracle_vs_live_reference():
    if not oracle_py.reference_sw_available():
        pytest.skip("oracle/_ref/libref_sw.so not built (no /root/reference on this machine)")
    import sys
    sys.path.insert(0, os.path.join(HERE, "golden"))
    from make_golden import sw_pairs
    refs, alts = sw_pairs(99, 300)
    p = oracle_py.SWPairs(refs, alts)
    for params in ((10, -15, -30, -5), (200, -150, -260, -11)):
        for strat in (9, 10, 11, 12):
            c, o = oracle_py.oracle_sw(p, params, strat)
            rc, ro, _ = oracle_py.reference_sw(p, params, strat, threads=4)
            assert c == rc and np.array_equal(o, ro), (params, strat)


def test_sw_known_alignments():
    """hand-checkable cases"""
    p = oracle_py.SWPairs([b"ACGTACGTAC", b"AAAACCCCGGGG", b"ACGT"], [b"ACGTACGTAC", b"AAAAGGGG", b"TTACGTTT"])
    c, o = oracle_py.oracle_sw(p, (10, -15, -30, -5), oracle_py.SW_SOFTCLIP)
    assert c[0] == "10M" and o[0] == 0
    assert c[1] == "4M4D4M" and o[1] == 0
    assert c[2] == "2S4M2S" and o[2] == 0


@pytest.mark.gpu
@pytest.mark.parametrize("strategy", [9, 10, 11, 12])
def test_cuda_sw_equals_oracle(strategy):
    """smithwaterman_cuda_align_batch: CIGAR text and offset identical to the oracle (itself identical to the reference)"""
    from fgdp_b200 import smithwaterman as sw
    refs, alts, cases = _golden()
    for params, strat, cig, off in cases:
        if strat != strategy:
            continue
        c, o, ms = sw.align_batch(refs, alts, params, strat)
        bad = [k for k in range(len(refs)) if c[k] != cig[k] or o[k] != off[k]]
        assert not bad, (params, strat, bad[:3], [(c[k], cig[k], int(o[k]), int(off[k])) for k in bad[:3]])


@pytest.mark.gpu
def test_cuda_sw_batch_of_read_realignments():
    """the call Mutect2Cpp makes after the PairHMM (AlignmentUtils.cpp:498): every read against a haplotype, parameters
    (10, -15, -30, -5), SOFTCLIP -- a few thousand pairs through one launch, plus the single-pair drop-in entry point"""
    import sys
    sys.path.insert(0, os.path.join(HERE, "golden"))
    from make_golden import sw_pairs
    from fgdp_b200 import smithwaterman as sw
    refs, alts = sw_pairs(7, 3000)
    p = oracle_py.SWPairs(refs, alts)
    want_c, want_o = oracle_py.oracle_sw(p)
    c, o, ms = sw.align_batch(refs, alts)
    assert c == want_c and np.array_equal(o, want_o)
    assert ms > 0
    import ctypes as C
    lib = sw._lib()
    for k in (0, 5, 17):
        r = np.frombuffer(refs[k], np.uint8).copy(); a = np.frombuffer(alts[k], np.uint8).copy()
        n = 2 * max(len(r), len(a))
        buf = np.zeros(n, np.uint8)
        off = lib.smithwaterman_cuda_align(r.ctypes.data_as(C.c_void_p), len(r), a.ctypes.data_as(C.c_void_p), len(a),
                                           buf.ctypes.data_as(C.c_void_p), n, 10, -15, -30, -5, 9)
        assert off == want_o[k] and bytes(buf).split(b"\0")[0].decode() == want_c[k]
    with pytest.raises(Exception):
        sw.align_batch([b""], [b"A"])


@pytest.mark.gpu
@pytest.mark.parametrize("min_group", ["1", None])
def test_cuda_sw_every_kernel_class(min_group, monkeypatch):
    """alt lengths 1..300 cover the eight register-blocked kernels (K = ceil(alt/32) columns per lane) and the
    anti-diagonal kernel; SW_CUDA_MIN_GROUP=1 launches every class on its own instead of merging the small ones"""
    from fgdp_b200 import smithwaterman as sw
    if min_group:
        monkeypatch.setenv("SW_CUDA_MIN_GROUP", min_group)
    rng = np.random.default_rng(41)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    refs, alts = [], []
    for k in range(2400):
        H, R = int(rng.integers(1, 420)), 1 + (k * 7) % 300
        hap = rng.choice(acgt, H)
        if k % 5 == 0 or H <= R:
            read = rng.choice(acgt, R)
        else:
            st = int(rng.integers(0, H - R))
            read = hap[st:st + R].copy()
            m = rng.random(R) < 0.04
            read[m] = rng.choice(acgt, int(m.sum()))
            if R > 20 and k % 3 == 0:
                read = np.concatenate([read[:R // 2], read[R // 2 + 3:], rng.choice(acgt, 3)])
        refs.append(bytes(hap)); alts.append(bytes(read))
    p = oracle_py.SWPairs(refs, alts)
    for params, strat in (((10, -15, -30, -5), 9), ((25, -50, -110, -6), 10), ((3, -1, -4, -3), 11), ((200, -150, -260, -11), 12)):
        want_c, want_o = oracle_py.oracle_sw(p, params, strat)
        c, o, ms = sw.align_batch(refs, alts, params, strat)
        bad = [k for k in range(len(refs)) if c[k] != want_c[k] or o[k] != want_o[k]]
        assert not bad, (params, strat, len(bad), [(len(refs[k]), len(alts[k]), c[k], want_c[k]) for k in bad[:3]])


@pytest.mark.gpu
def test_cuda_sw_longest_sequences():
    """haplotype-against-reference sized pairs up to SW_CUDA_MAX_LEN = 4096 on both sides (the anti-diagonal kernel with
    one warp per CTA: its shared memory holds eight rows of max_alt ints), next to short ones in the same call; one base
    more is refused"""
    from fgdp_b200 import smithwaterman as sw
    rng = np.random.default_rng(77)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    refs, alts = [], []
    for H, R in ((4096, 4096), (3000, 1200), (700, 4096), (4096, 33), (520, 500), (64, 40)):
        hap = rng.choice(acgt, H)
        src = hap[:R] if R <= H else np.concatenate([hap, rng.choice(acgt, R - H)])
        read = src.copy()
        m = rng.random(R) < 0.02
        read[m] = rng.choice(acgt, int(m.sum()))
        if R > 300:
            read = np.concatenate([read[:100], read[110:R // 2], rng.choice(acgt, 10), read[R // 2:]])
        refs.append(bytes(hap)); alts.append(bytes(read[:4096]))
    p = oracle_py.SWPairs(refs, alts)
    for params, strat in (((10, -15, -30, -5), 9), ((200, -150, -260, -11), 10), ((25, -50, -110, -6), 12)):
        want_c, want_o = oracle_py.oracle_sw(p, params, strat)
        c, o, ms = sw.align_batch(refs, alts, params, strat)
        assert c == want_c and np.array_equal(o, want_o), (params, strat)
    with pytest.raises(Exception):
        sw.align_batch([b"A" * 4097], [b"A" * 10])


@pytest.mark.gpu
def test_repeated_calls_reuse_the_thread_context():
    """per-region use: many small calls of changing size on one host thread (the stream and the device buffers are kept
    between calls and grow), a release in the middle, and a second host thread with its own context"""
    import threading
    from fgdp_b200 import smithwaterman as sw
    acgt = np.frombuffer(b"ACGT", np.uint8)

    def one_round(seed, results):
        r = np.random.default_rng(seed)
        out = []
        for n, hl, rl in ((3, 60, 30), (200, 400, 150), (1, 17, 9), (64, 900, 300), (5, 40, 40)):
            refs = [bytes(r.choice(acgt, hl + int(r.integers(0, 9)))) for _ in range(n)]
            alts = [bytes(r.choice(acgt, rl + int(r.integers(0, 5)))) for _ in range(n)]
            out.append((refs, alts, sw.align_batch(refs, alts)[:2]))
        results.append(out)

    res = []
    one_round(1, res)
    sw.release()
    one_round(1, res)
    th = threading.Thread(target=one_round, args=(1, res)); th.start(); th.join()
    assert len(res) == 3
    for k in (1, 2):
        for (_, _, (c0, o0)), (_, _, (c1, o1)) in zip(res[0], res[k]):
            assert c0 == c1 and np.array_equal(o0, o1)
    import oracle_py
    for refs, alts, (c, o) in res[0]:
        wc, wo = oracle_py.oracle_sw(oracle_py.SWPairs(refs, alts))
        assert c == wc and np.array_equal(o, wo)
