"""The drop-in adapter EXECUTED: integration/CudaPairHMM.cpp compiled against the reference's REAL
headers and linked, in integration/replay/libm2_replay.so, with the reference's own PairHMMLikelihoodCalculationEngine.cpp,
VectorLoglessPairHMM.cpp, SAMRecord.cpp, Haplotype.cpp, AssemblyResultSet.cpp, AlleleLikelihoods.h ... (oracle/Makefile, target
`replay`; the reference sources are compiled unmodified where they lie). The harness (oracle/ref_m2_driver.cpp) builds real
SAMRecord / Haplotype objects, runs the reference's computeReadLikelihoods once with its own VectorLoglessPairHMM and once with
the engine's `pairHMM` member replaced by CudaPairHMM -- the one-line edit of INTEGRATION.md -- and hands back every cell of the
SampleMatrix in input order.

The .so is built in the container that has /root/reference and travels to the GPU box prebuilt."""
import numpy as np
import pytest

import oracle_py
from util import TOL
from test_oracle import _m2_region_cases


@pytest.fixture(scope="module")
def replay():
    if not oracle_py.replay_available():
        pytest.skip("integration/replay/libm2_replay.so not built (needs /root/reference at build time: make -C oracle replay)")
    return oracle_py.ReferenceM2(replay=True)


def test_replay_harness_links_the_adapter_and_the_product_library(replay):
    """CPU: the harness loads, exports its entry points, and depends on libpairhmm_cuda.so (not on anything under oracle/)"""
    import subprocess
    for sym in ("ref_m2_region", "ref_m2_prepass", "ref_m2_last_error"):
        assert hasattr(replay.lib, sym)
    ldd = subprocess.run(["ldd", replay.path], capture_output=True, text=True).stdout
    assert "libpairhmm_cuda.so" in ldd and "oracle" not in ldd.replace("/oracle/..", "")
    nm = subprocess.run(["nm", "-C", "--defined-only", replay.path], capture_output=True, text=True).stdout
    assert "CudaPairHMM::computeLog10Likelihoods" in nm and "VectorLoglessPairHMM::computeLog10Likelihoods" in nm


def test_replay_reference_arm_reproduces_goldens(replay):
    """CPU: the harness' reference arm (VectorLoglessPairHMM inside the reference's own engine) gives the committed goldens"""
    for name, d, out0, out1, removed1, order in list(_m2_region_cases())[:8]:
        o0, _, od = replay.region(stage=0, use_cuda=False, **d)
        assert np.allclose(o0, out0, rtol=0, atol=1e-9) and np.array_equal(od, order), name   # same code; AVX width may differ per host


def test_adapter_turns_a_failed_status_into_std_exception(replay):
    """CPU (no GPU here): CudaPairHMM's constructor throws std::runtime_error when pairhmm_cuda_create fails -- the exception
    Mutect2Cpp's worker catches to skip the region (M2/main.cpp:303-310); the harness reports it as an error string"""
    import fgdp_b200
    if fgdp_b200.device_count() > 0:
        pytest.skip("a GPU is visible: covered by the gpu tests")
    name, d, *_ = next(iter(_m2_region_cases()))
    with pytest.raises(RuntimeError, match="pairhmm_cuda_create failed"):
        replay.region(stage=0, use_cuda=True, **d)


@pytest.mark.gpu
@pytest.mark.parametrize("stage", [0, 1])
def test_cuda_adapter_fills_the_same_sample_matrix_as_the_reference(replay, stage):
    """GPU: every golden region (per-pair and trie regions, duplicate reads, 1-3 samples, BI/BD arrays on all / every other
    read, low mapping qualities, other PCR model, use_double, ragged edge regions; the matrix's allele order always differs
    from the haplotype list's) through the reference's engine with CudaPairHMM as its pairHMM: every SampleMatrix cell
    within 1e-4 of the reference's own VectorLoglessPairHMM run (live, same harness) and of the committed golden; after
    normalizeLikelihoods + filterPoorlyModeledEvidence the same reads survive."""
    n = 0
    for name, d, out0, out1, removed1, order in _m2_region_cases():
        want = out0 if stage == 0 else out1
        got, removed, od = replay.region(stage=stage, use_cuda=True, **d)
        ref, rremoved, rod = replay.region(stage=stage, use_cuda=False, **d)
        assert np.array_equal(od, order) and np.array_equal(rod, order), name
        for w, wrem, what in ((want, removed1, "golden"), (ref, rremoved, "live reference")):
            if stage == 1:
                # a read whose best likelihood is within 1e-4 of the filter threshold may fall on either side
                thr = np.minimum(2.0, np.ceil(d["read_len"] * 0.02)) * -4.0
                sure = np.abs(out0.max(axis=1) - thr) > TOL
                assert np.array_equal(removed[sure], wrem[sure]), (name, what)
                keep = (removed == 0) & (wrem == 0)
            else:
                assert not removed.any()
                keep = np.ones(len(removed), bool)
            assert np.isfinite(got[keep]).all() and np.isfinite(w[keep]).all(), (name, what)
            if keep.any():
                assert np.abs(got[keep] - w[keep]).max() <= TOL, (name, what, np.abs(got[keep] - w[keep]).max())
        n += 1
    assert n >= 20


@pytest.mark.gpu
def test_cuda_adapter_one_engine_many_regions(replay):
    """GPU: regions of different shapes back to back through fresh engines (initialize once per region, one compute per
    sample -- PHLCE.cpp:88-91,120), twice: results are reproducible and independent of what ran before"""
    cases = list(_m2_region_cases())
    first = {}
    for rnd in range(2):
        for name, d, out0, *_ in cases[:10]:
            got, _, _ = replay.region(stage=0, use_cuda=True, **d)
            if rnd == 0:
                first[name] = got
            else:
                assert np.array_equal(got, first[name]), name
