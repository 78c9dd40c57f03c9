"""CPU, world_size 2 over gloo: the multi-GPU contract of bench.py / DESIGN.md §5 on the oracle — ranks trace disjoint
sample partitions (sample s -> rank s mod N), the (sum, count) accumulators are sum-reduced to rank 0, and the
result equals the single-rank render: counts exactly, sums to fp32 re-association."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out_path):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    sys.path.insert(0, os.path.join(ROOT, "path-tracing_b200"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle_py import Oracle
    import scenes
    from util import with_planes
    orc = Oracle()
    sc = with_planes(scenes.cornell(), orc)
    W, H, spp = 24, 16, 12
    sm, _, cnt, segs, _ = orc.render_accum(sc, orc.config(W, H, spp, 8, 1e-4, -1.0, 42, rank, world), threads=2)
    buf = torch.from_numpy(np.concatenate([sm.ravel(), cnt.ravel().astype(np.float32)]))   # the layout bench.py reduces
    seg = torch.tensor([float(segs)], dtype=torch.float64)
    dist.reduce(buf, 0, op=dist.ReduceOp.SUM)
    dist.reduce(seg, 0, op=dist.ReduceOp.SUM)
    if rank == 0:
        full_sm, _, full_cnt, full_segs, _ = orc.render_accum(sc, orc.config(W, H, spp, 8, 1e-4, -1.0, 42), threads=2)
        got_sm = buf[:W * H * 3].numpy().reshape(H, W, 3)
        got_cnt = buf[W * H * 3:].numpy().reshape(H, W)
        ok = (np.array_equal(got_cnt.astype(np.int32), full_cnt) and np.allclose(got_sm, full_sm, rtol=1e-5, atol=1e-6)
              and int(seg.item()) == full_segs and full_cnt.sum() > 0)
        open(out_path, "w").write("ok" if ok else "mismatch")
    dist.destroy_process_group()


def test_spp_partition_reduce_world2(tmp_path):
    out = str(tmp_path / "result.txt")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    assert open(out).read() == "ok"
