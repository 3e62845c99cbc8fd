"""pairhmm_cuda_forward_pairs_noalloc (include/pairhmm_cuda.h): the allocation-free entry point added
-- its scratch is a static __device__ array, the caller passes device inputs and outputs. Checked through
the C ABI against the existing entry point (pairhmm_cuda_compute_flat) on two calls in a row, within the tolerance of the
other parity tests (1e-4 in log10; identical rescue flags). Collected last on purpose
(file name)."""
import ctypes as C

import numpy as np
import pytest

import fgdp_b200
from fgdp_b200 import synth
from util import TOL

pytestmark = pytest.mark.gpu

MAX_HAP = 2048      # PAIRHMM_CUDA_NOALLOC_MAX_HAP_LEN


def _batch(seed):
    """one region: reads of 1 .. 1000 bases (below and above the 256-row stripe and the 384-row register limit) x haplotypes
    of 1 .. 2048 bases; a mix of good and poor qualities so that both the FP32 and the FP64 route are taken"""
    rng = np.random.default_rng(seed)
    rlens = [1, 2, 31, 150, 150, 151, 255, 256, 257, 384, 385, 600, 1000]
    hlens = [1, 7, 150, 300, 301, 700, 2047, MAX_HAP]
    acgt = np.frombuffer(b"ACGT", np.uint8)
    hap0 = acgt[rng.integers(0, 4, MAX_HAP)]
    haps = []
    for n in hlens:
        h = hap0[:n].copy()
        if n > 20:
            h[rng.integers(0, n, 3)] = acgt[rng.integers(0, 4, 3)]
        haps.append(h)
    reads, bq, iq, dq, gc = [], [], [], [], []
    for k, n in enumerate(rlens):
        st = int(rng.integers(0, MAX_HAP - n + 1))
        r = hap0[st:st + n].copy()
        if k % 3 == 2:                                   # unrelated read: underflows in float on the long haplotypes
            r = acgt[rng.integers(0, 4, n)]
        if n > 10:
            r[int(rng.integers(0, n))] = ord("N")
        reads.append(r)
        q = rng.integers(25, 41, n) if k % 2 == 0 else rng.integers(6, 20, n)
        bq.append(q.astype(np.uint8)); iq.append(rng.integers(10, 46, n).astype(np.uint8)); dq.append(rng.integers(10, 46, n).astype(np.uint8))
        gc.append(np.full(n, 10, np.uint8) if k % 4 else rng.integers(8, 13, n).astype(np.uint8))
    rl = np.array(rlens, np.int32); hl = np.array(hlens, np.int32)
    off = lambda l: np.concatenate([[0], np.cumsum(l)[:-1]]).astype(np.int64)
    return synth.RegionBatch(np.concatenate(reads), np.concatenate(bq), np.concatenate(iq), np.concatenate(dq), np.concatenate(gc),
                             off(rl), rl, np.concatenate(haps), off(hl), hl, np.array([0, len(rl)], np.int64),
                             np.array([0, len(hl)], np.int64), "noalloc-%d" % seed)


def _call(lib, b, pair_read, pair_hap, max_hap=MAX_HAP, use_double=False):
    """device buffers are the CALLER's (torch here); the entry point itself allocates nothing"""
    import torch
    dev = torch.device("cuda", 0)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    plane = int(b.read_bases.shape[0])
    planes = up(np.concatenate([b.read_bases, b.base_q, b.ins_q, b.del_q, b.gcp]))
    roff = up(b.read_off.astype(np.uint32).view(np.int32)); rlen = up(b.read_len.astype(np.int32))
    hb = up(b.hap_bases); hoff = up(b.hap_off.astype(np.uint64).view(np.int64)); hlen = up(b.hap_len.astype(np.int32))
    pr = up(pair_read.astype(np.uint32).view(np.int32)); ph = up(pair_hap.astype(np.uint32).view(np.int32))
    n = int(pair_read.shape[0])
    out = torch.full((n,), 12345.0, dtype=torch.float64, device=dev); resc = torch.full((n,), 77, dtype=torch.uint8, device=dev)
    err = C.create_string_buffer(256)
    vp = C.c_void_p
    lib.pairhmm_cuda_forward_pairs_noalloc.argtypes = [C.c_int, vp, vp, C.c_uint64, vp, vp, vp, vp, vp, vp, vp, C.c_uint32, C.c_int32,
                                                       C.c_int32, vp, vp, C.c_char_p, C.c_size_t]
    torch.cuda.synchronize()
    st = lib.pairhmm_cuda_forward_pairs_noalloc(0, None, planes.data_ptr(), plane, roff.data_ptr(), rlen.data_ptr(), hb.data_ptr(),
                                                hoff.data_ptr(), hlen.data_ptr(), pr.data_ptr(), ph.data_ptr(), n, int(max_hap),
                                                int(use_double), out.data_ptr(), resc.data_ptr(), err, 256)
    torch.cuda.synchronize()
    return st, out.cpu().numpy(), resc.cpu().numpy(), err.value.decode()


def test_noalloc_entry_point_equals_the_existing_entry_point_twice_in_a_row(engine):
    lib = fgdp_b200.load_library()
    for call, seed in enumerate((11, 12)):               # two calls in a row share the static scratch: the second must not see the first
        b = _batch(seed)
        want, wresc = engine.compute_batch(b)            # existing entry point: pairhmm_cuda_compute_flat
        pr, ph = b.pair_lists()
        order = np.arange(len(pr)) if call == 0 else np.arange(len(pr))[::-1].copy()    # another hand-out order the second time
        st, got, resc, err = _call(lib, b, pr[order], ph[order])
        assert st == 0, err
        w, wr = want[order], wresc[order]
        fin = np.isfinite(w)
        assert np.array_equal(np.isfinite(got), fin), call
        assert np.array_equal(got[~fin], w[~fin])
        assert np.abs(got[fin] - w[fin]).max() <= TOL, (call, np.abs(got[fin] - w[fin]).max())
        assert np.array_equal(resc, wr), (call, np.nonzero(resc != wr)[0][:5])
        assert wr.any() and not wr.all(), "both routes exercised"
    # use_double: every pair through the FP64 pass, as a use_double handle
    e = fgdp_b200.Engine(use_double=True)
    try:
        want, wresc = e.compute_batch(b)
    finally:
        e.close()
    st, got, resc, err = _call(lib, b, pr, ph, use_double=True)
    fin = np.isfinite(want)
    assert st == 0 and resc.all() and np.array_equal(np.isfinite(got), fin) and np.abs(got[fin] - want[fin]).max() <= TOL


def test_noalloc_entry_point_refuses_what_its_static_scratch_cannot_hold(engine):
    lib = fgdp_b200.load_library()
    b = _batch(13)
    pr, ph = b.pair_lists()
    st, got, resc, err = _call(lib, b, pr, ph, max_hap=MAX_HAP + 1)
    assert st == -1 and "PAIRHMM_CUDA_NOALLOC_MAX_HAP_LEN" in err and (got == 12345.0).all(), err      # nothing was launched
    st, got, resc, err = _call(lib, b, pr[:0], ph[:0])
    assert st == 0
    # the engine's own entry points are unaffected by the static arrays' use
    want, wresc = engine.compute_batch(b)
    again, aresc = engine.compute_batch(b)
    assert np.array_equal(want, again) and np.array_equal(wresc, aresc)
