"""Multi-device handles (config.device_mask) and the shared batcher (config.batch_window_us): GPU tests.
The host-side plan is covered on the CPU by tests/test_host.py::test_multi_device_plan_balances_cells_and_keeps_regions_whole."""
import threading

import numpy as np
import pytest

import fgdp_b200
from fgdp_b200 import synth
from util import assert_parity, raw_quals

pytestmark = pytest.mark.gpu


def _region_lists(b, g):
    r0, r1 = int(b.reg_read_start[g]), int(b.reg_read_start[g + 1])
    h0, h1 = int(b.reg_hap_start[g]), int(b.reg_hap_start[g + 1])
    reads = [tuple(a[int(b.read_off[r]):int(b.read_off[r]) + int(b.read_len[r])] for a in (b.read_bases, b.base_q, b.ins_q, b.del_q, b.gcp))
             for r in range(r0, r1)]
    haps = [b.hap_bases[int(b.hap_off[h]):int(b.hap_off[h]) + int(b.hap_len[h])] for h in range(h0, h1)]
    return reads, haps


def _masks():
    n = fgdp_b200.device_count()
    return [1] + ([(1 << n) - 1] if n > 1 else [])


@pytest.mark.parametrize("mask", _masks() if fgdp_b200.device_count() else [1])
def test_multi_device_handle_equals_single_device(engine, mask):
    """one handle over the devices of `mask` (1 GPU: the same code path with one child): batch, region, submit / wait,
    normalized, raw and resident calls give the bits of a single-device handle, in the caller's ONE result array"""
    m = fgdp_b200.Engine(device_mask=mask)
    try:
        for b in (synth.config2(40), synth.config3(300, 12), synth.edge_cases(), synth.long_reads()):
            want, wresc = engine.compute_batch(b)
            got, gresc = m.compute_batch(b)
            assert np.array_equal(got, want) and np.array_equal(gresc, wresc), b.name
            m.resident_load(b)
            ms, msf = m.resident_run()
            assert ms > 0
            rg, rr = m.resident_fetch()
            assert np.array_equal(rg, want) and np.array_equal(rr, wresc), b.name
        st = m.stats()
        assert st["kernel_launches"] > 0 and st["cells"] > 0
        # a region worth several devices is cut into row ranges (BASELINE configs[2]); small ones go to the devices in turn
        c3 = synth.config3(1200, 32)
        reads, haps = _region_lists(c3, 0)
        want, wresc = engine.compute(reads, haps)
        got, gresc = m.compute(reads, haps)
        assert np.array_equal(got, want) and np.array_equal(gresc, wresc)
        b = synth.config2(6)
        tickets = []
        for g in range(b.n_regions):
            reads, haps = _region_lists(b, g)
            w, _ = engine.compute(reads, haps)
            t, out, resc = m.submit(reads, haps)
            tickets.append((t, out, w, resc))      # the result arrays must stay alive until the matching wait
        for t, out, w, _ in tickets:
            m.wait(t)
            assert np.array_equal(out, w)
        # fused epilogue and raw reads through the split
        b = synth.config2(12)
        wn, wr, ws = engine.compute_batch_normalized(b)
        gn, gr, gs = m.compute_batch_normalized(b)
        assert np.array_equal(gn, wn) and np.array_equal(gs["best_hap"], ws["best_hap"]) and np.array_equal(gs["poorly_modeled"], ws["poorly_modeled"])
        bq, mapq, _, _ = raw_quals(b, 3, False)
        wr0, wr1 = engine.compute_batch_raw(b, bq, mapq)
        gr0, gr1 = m.compute_batch_raw(b, bq, mapq)
        assert np.array_equal(gr0, wr0) and np.array_equal(gr1, wr1)
        with pytest.raises(fgdp_b200.PairHMMError):
            m.wait(12345)
    finally:
        m.close()


def test_multi_device_handle_uses_every_device():
    n = fgdp_b200.device_count()
    if n < 2:
        pytest.skip("one GPU visible")
    m = fgdp_b200.Engine(device_mask=(1 << n) - 1)
    try:
        b = synth.config2(60 * n)
        m.compute_batch(b)
        st = m.stats()
        assert st["cells"] == b.cells()
    finally:
        m.close()
    with pytest.raises(fgdp_b200.PairHMMError):
        fgdp_b200.Engine(device_mask=1 << 20)


def test_shared_batcher_merges_regions_of_concurrent_callers(engine, oracle):
    """config.batch_window_us: blocking region calls from T host threads (one handle each, as one PairHMM per Mutect2Cpp
    worker) go through one process-wide batcher; every caller gets exactly its own region's results (bits of a direct
    call), and regions of different callers did share launch groups (fewer chunks than regions)"""
    T, per_thread = 8, 12
    b = synth.config2(T * per_thread)
    regions = [_region_lists(b, g) for g in range(b.n_regions)]
    want = [engine.compute(r, h) for r, h in regions]
    engines = [fgdp_b200.Engine(batch_window_us=200) for _ in range(T)]
    got = [None] * b.n_regions
    errors = []

    def worker(t):
        try:
            for k in range(per_thread):
                g = t * per_thread + k
                got[g] = engines[t].compute(*regions[g])
        except Exception as ex:      # noqa: BLE001
            errors.append(ex)
    try:
        for rnd in range(2):
            th = [threading.Thread(target=worker, args=(t,)) for t in range(T)]
            [x.start() for x in th]; [x.join() for x in th]
            assert not errors, errors
            for g in range(b.n_regions):
                assert np.array_equal(got[g][0], want[g][0]) and np.array_equal(got[g][1], want[g][1]), g
        st = engines[0].stats()      # the shared backend's counters as of this handle's last call
        assert st["regions"] >= b.n_regions and st["kernel_launches"] > 0
        # bad input of one caller is refused for that caller only
        with pytest.raises(fgdp_b200.PairHMMError):
            ra, ha, keep = engines[1]._structs(*regions[0])
            engines[1]._check(engines[1].lib.pairhmm_cuda_compute(engines[1].h, ra, len(regions[0][0]), ha, len(regions[0][1]), None, None))
        out, resc = engines[2].compute(*regions[3])
        assert np.array_equal(out, want[3][0])
        # a lone caller still works (a group per region) and the oracle agrees
        flat = b.subset([5]).as_flat()
        w, wr, _ = oracle.compute(flat, threads=2)
        o, r = engines[3].compute(*regions[5])
        assert_parity(o, r, w, wr, what="batched call vs oracle")
    finally:
        for e in engines:
            e.close()
