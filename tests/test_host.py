"""CPU tests: the C-ABI library loads and exports what include/pairhmm_cuda.h declares (no compute without a GPU),
the engine's uploaded tables equal the reference's, host-side sharding logic, the bench's JSON contract."""
import ctypes
import json
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import fgdp_b200
from fgdp_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "pairhmm_cuda.h")).read()
    declared = sorted(set(re.findall(r"\b(pairhmm_cuda_[a-z_0-9]+)\s*\(", hdr)))
    assert len(declared) >= 19
    lib = fgdp_b200.load_library()
    for name in declared:
        assert hasattr(lib, name), "libpairhmm_cuda.so does not export %s" % name
    assert sorted(fgdp_b200.EXPORTED_SYMBOLS) == declared
    assert lib.pairhmm_cuda_abi_version() == 2
    m = re.search(r"#define PAIRHMM_CUDA_ABI_VERSION (\d+)", hdr)
    assert m and int(m.group(1)) == 2
    # the ctypes mirror of pairhmm_cuda_config has the size the library assumes (default_config clears the whole struct)
    cfg = fgdp_b200.pairhmm.Config()
    ctypes.memset(ctypes.byref(cfg), 0xff, ctypes.sizeof(cfg))
    lib.pairhmm_cuda_default_config(ctypes.byref(cfg))
    assert cfg.device == -1 and cfg.dedup_reads == 1 and cfg.device_mask == 0 and cfg.batch_window_us == 0
    assert list(cfg.reserved) == [0] * 8 and list(cfg.reserved2) == [0] * 6 and ctypes.sizeof(cfg) == 96
    # include/smithwaterman_cuda.h (the Smith-Waterman entry points live in the same library)
    from fgdp_b200 import smithwaterman
    hdr = open(os.path.join(ROOT, "include", "smithwaterman_cuda.h")).read()
    declared = sorted(set(re.findall(r"\b(smithwaterman_cuda_[a-z_0-9]+)\s*\(", hdr)))
    assert declared == sorted(smithwaterman.SW_SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), "libpairhmm_cuda.so does not export %s" % name


def test_product_library_does_not_link_the_oracle():
    out = subprocess.run(["ldd", fgdp_b200.lib_path()], capture_output=True, text=True).stdout
    assert "oracle" not in out and "ref_pairhmm" not in out
    syms = subprocess.run(["nm", "-D", fgdp_b200.lib_path()], capture_output=True, text=True).stdout
    assert "oracle_" not in syms and "initNative" not in syms


def test_engine_tables_match_reference_golden():
    from fgdp_b200.pairhmm import host_tables
    z = np.load(os.path.join(ROOT, "tests", "golden", "tables.npz"))
    pf, mf, pd, md = host_tables()
    for got, key in ((pf, "ph2pr_f"), (mf, "mm_f"), (pd, "ph2pr_d"), (md, "mm_d")):
        assert np.array_equal(got.view(np.uint8), z[key].view(np.uint8)), key


def test_no_cpu_fallback_without_gpu():
    if fgdp_b200.device_count() > 0:
        pytest.skip("a GPU is visible")
    with pytest.raises(fgdp_b200.PairHMMError) as e:
        fgdp_b200.Engine()
    assert e.value.status == -3            # PAIRHMM_ERR_NO_DEVICE: loud failure, nothing routed to a CPU path
    with pytest.raises(fgdp_b200.PairHMMError):
        fgdp_b200.CudaPairHMM()


def test_default_config_and_null_handling():
    lib = fgdp_b200.load_library()
    from fgdp_b200.pairhmm import Config
    cfg = Config()
    lib.pairhmm_cuda_default_config(ctypes.byref(cfg))
    assert (cfg.device, cfg.use_double, cfg.exact_fp32, cfg.dedup_reads) == (-1, 0, 0, 1)
    assert lib.pairhmm_cuda_create(ctypes.byref(cfg), None) == -1
    assert lib.pairhmm_cuda_initialize(None, 10, 10) == -1
    assert lib.pairhmm_cuda_destroy(None) == 0


def test_multi_device_plan_balances_cells_and_keeps_regions_whole():
    """pairhmm_cuda_device_plan = what a multi-device handle does with a batch: LPT over whole regions, a
    region is cut into row ranges only when it alone exceeds 1 / (4 x devices) of the batch; every read lands on exactly
    one device; loads are balanced; one device = identity"""
    b = synth.config2(400)
    nr = np.diff(b.reg_read_start).astype(np.int32)
    rl = np.array([b.read_len[b.reg_read_start[g]:b.reg_read_start[g + 1]].sum() for g in range(b.n_regions)], np.uint64)
    hl = np.array([b.hap_len[b.reg_hap_start[g]:b.reg_hap_start[g + 1]].sum() for g in range(b.n_regions)], np.uint64)
    cost = rl.astype(np.float64) * hl.astype(np.float64)
    pr, pf, pn, pd = fgdp_b200.device_plan(rl, hl, nr, 1)
    assert np.array_equal(pr, np.arange(b.n_regions)) and not pf.any() and np.array_equal(pn, nr) and not pd.any()
    for nd in (2, 4, 8):
        pr, pf, pn, pd = fgdp_b200.device_plan(rl, hl, nr, nd)
        assert len(pr) == b.n_regions, "400 WES-like regions: none is heavy enough to be cut"
        assert np.array_equal(pr, np.arange(b.n_regions)) and np.array_equal(pn, nr)
        load = np.array([cost[pr[pd == d]].sum() for d in range(nd)])
        assert load.min() > 0 and load.max() / load.mean() < 1.02, load
    # BASELINE configs[2]: ONE region of 2000 reads x 32 haplotypes -> cut into row ranges that cover every read once
    c3 = synth.config3()
    nr3 = np.array([c3.n_reads], np.int32)
    for nd in (2, 8):
        pr, pf, pn, pd = fgdp_b200.device_plan(np.array([c3.read_len.sum()], np.uint64), np.array([c3.hap_len.sum()], np.uint64), nr3, nd)
        assert len(pr) >= 4 * nd and not pr.any()
        assert pf[0] == 0 and np.array_equal(pf[1:], np.cumsum(pn)[:-1]) and pn.sum() == c3.n_reads and pn.min() > 0
        per_dev = np.array([pn[pd == d].sum() for d in range(nd)])
        assert per_dev.max() - per_dev.min() <= c3.n_reads // (4 * nd) + 1, per_dev
    # a heavy region among light ones, empty regions, zero-length inputs
    rl2 = np.array([30000, 150, 0, 150, 150], np.uint64); hl2 = np.array([4000, 300, 300, 0, 300], np.uint64)
    nr2 = np.array([200, 1, 0, 1, 1], np.int32)
    pr, pf, pn, pd = fgdp_b200.device_plan(rl2, hl2, nr2, 4)
    assert (pr == 0).sum() >= 4 and pn[pr == 0].sum() == 200 and set(pr.tolist()) == {0, 1, 2, 3, 4}
    assert len(fgdp_b200.device_plan(np.zeros(0, np.uint64), np.zeros(0, np.uint64), np.zeros(0, np.int32), 8)[0]) == 0


def test_lpt_shard_balances_cells():
    _, _, cost = synth.config2_shapes(4000)
    for n in (2, 4, 8):
        shards = synth.lpt_shard(cost, n)
        assert sorted(np.concatenate(shards).tolist()) == list(range(4000))
        loads = np.array([cost[s].sum() for s in shards], np.float64)
        assert loads.max() / loads.mean() < 1.001


def test_select_generates_the_same_regions():
    full = synth.config2(12)
    part = synth.config2(12, select=[2, 5, 11])
    sub = full.subset([2, 5, 11])
    for f in ("read_bases", "base_q", "ins_q", "del_q", "gcp", "hap_bases", "read_len", "hap_len"):
        assert np.array_equal(getattr(part, f), getattr(sub, f)), f


def test_bench_reference_arm_json_contract():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0", "--cpu-sample-regions", "3"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "GCUPS" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["higher_is_better"] is True


def test_reference_side_adapter_compiles():
    """integration/CudaPairHMM.cpp (the class a Mutect2Cpp maintainer adds) compiles against the
    reference's real PairHMM.h; skipped where /root/reference does not exist (the GPU box)."""
    import subprocess
    import pytest
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([os.path.join(root, "integration", "check.sh")], capture_output=True, text=True)
    if r.returncode == 77:
        pytest.skip("reference tree absent")
    assert r.returncode == 0, r.stdout + r.stderr


def test_streaming_copy_equals_memcpy(tmp_path):
    """csrc/hostcopy.cpp (the packer's and the scatter's copies with non-temporal stores): every size around the thresholds,
    every source / destination alignment, bytes before and after the destination untouched. Host code: compiled and run here."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    csrc = os.path.join(root, "fast-genomic-data-processing_b200", "csrc")
    main = tmp_path / "t.cpp"
    main.write_text(r'''
#include <cstdio>
#include <cstring>
#include <vector>
#include "hostcopy.h"
int main() {
    std::vector<unsigned char> src(1 << 16), dst(1 << 16), want(1 << 16);
    for (size_t i = 0; i < src.size(); ++i) src[i] = (unsigned char)(i * 131 + 7);
    const size_t sizes[] = {0, 1, 31, 32, 33, 127, 128, 129, 511, 512, 513, 640, 1000, 4096, 4097, 30000, 60000};
    long checked = 0;
    for (size_t n : sizes)
        for (size_t da = 0; da < 67; da += (n > 4096 ? 13 : 1))
            for (size_t sa = 0; sa < 35; sa += (n > 4096 ? 11 : 3)) {
                std::memset(dst.data(), 0xEE, dst.size()); std::memset(want.data(), 0xEE, want.size());
                std::memcpy(want.data() + 64 + da, src.data() + sa, n);
                phmm::copy_stream(dst.data() + 64 + da, src.data() + sa, n);
                phmm::copy_stream_fence();
                if (std::memcmp(dst.data(), want.data(), dst.size())) { std::printf("FAIL n=%zu da=%zu sa=%zu\n", n, da, sa); return 1; }
                ++checked;
            }
    std::printf("ok %ld\n", checked);
    return 0;
}
''')
    exe = tmp_path / "t"
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", csrc, str(main), os.path.join(csrc, "hostcopy.cpp"), "-o", str(exe)], check=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.startswith("ok"), r.stdout + r.stderr
