"""world_size-2 CPU test (gloo) of the N > 1 path of bench.py: every rank derives the same LPT split of the
global job from the region shapes, generates only its own regions, and the union is the whole job with no overlap.
The per-rank compute is the oracle here (no GPU in this container); on the GPU box it is the CUDA engine."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_REGIONS = 6


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import fgdp_b200  # noqa: F401
    from fgdp_b200 import synth
    import oracle_py
    _, _, cost = synth.config2_shapes(N_REGIONS)
    mine = synth.lpt_shard(cost, world)[rank]
    b = synth.config2(N_REGIONS, select=mine)
    out, resc, _ = oracle_py.load_oracle().compute(b.as_flat(), threads=2)
    cells = torch.tensor([float(b.cells())], dtype=torch.float64)
    dist.all_reduce(cells, op=dist.ReduceOp.SUM)
    t = torch.tensor([0.1 * (rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)           # the bench's max-over-ranks of the step time
    dist.barrier()
    q.put((rank, mine.tolist(), out, float(cells.item()), float(t.item())))
    dist.destroy_process_group()


def test_two_ranks_cover_the_job_once():
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import fgdp_b200  # noqa: F401
    from fgdp_b200 import synth
    import oracle_py
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=300) for _ in procs])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    full = synth.config2(N_REGIONS)
    want, _, _ = oracle_py.load_oracle().compute(full.as_flat(), threads=4)
    off = full.region_out_offsets()
    seen = []
    for rank, mine, out, cells, tmax in res:
        assert cells == float(full.cells()) and abs(tmax - 0.2) < 1e-12
        pos = 0
        for g in mine:
            n = int(off[g + 1] - off[g])
            assert np.array_equal(out[pos:pos + n], want[off[g]:off[g + 1]])
            pos += n
        seen += mine
    assert sorted(seen) == list(range(N_REGIONS))
