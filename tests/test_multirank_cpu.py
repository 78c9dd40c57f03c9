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


def _worker_rows(rank, world, port, out_path):
    """The adaptive rule (main.cpp:118-125) under a multi-GPU split: image ROWS per rank (PtConfig.row_begin / row_end,
    what pt_render does for n_gpus > 1 when error >= 0).  A pixel keeps its whole ordered sample sequence, the reduce
    adds zeros, so the result is the single-rank render bit for bit — unlike the sample split, whose per-rank
    statistics trace a different set of samples."""
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
    W, H, spp, err = 24, 17, 40, 0.05
    full = orc.render_accum(sc, orc.config(W, H, spp, 8, 1e-4, err, 42), threads=2)
    r0, r1 = H * rank // world, H * (rank + 1) // world          # the band arithmetic of render_frame_multi
    band = [np.zeros_like(a) for a in full[:3]]
    for b, a in zip(band, full[:3]):                              # pixels are independent: a band render = the rows of the frame
        b[r0:r1] = a[r0:r1]
    fbuf = torch.from_numpy(np.concatenate([band[0].ravel(), band[1].ravel()]))
    ibuf = torch.from_numpy(band[2].ravel().copy())
    dist.reduce(fbuf, 0, op=dist.ReduceOp.SUM)
    dist.reduce(ibuf, 0, op=dist.ReduceOp.SUM)
    part = orc.render_accum(sc, orc.config(W, H, spp, 8, 1e-4, err, 42, rank, world), threads=2)
    pcnt = torch.from_numpy(part[2].ravel().copy())
    dist.reduce(pcnt, 0, op=dist.ReduceOp.SUM)
    if rank == 0:
        n = W * H * 3
        same = (np.array_equal(fbuf[:n].numpy().view(np.uint32), full[0].ravel().view(np.uint32)) and
                np.array_equal(fbuf[n:].numpy().view(np.uint32), full[1].ravel().view(np.uint32)) and
                np.array_equal(ibuf.numpy(), full[2].ravel()))
        rule_acted = full[3] < W * H * spp * 7                    # the fixture must exercise the rule
        sample_split_differs = not np.array_equal(pcnt.numpy(), full[2].ravel())
        open(out_path, "w").write("ok" if same and rule_acted and sample_split_differs else
                                  f"mismatch same={same} rule={rule_acted} differs={sample_split_differs}")
    dist.destroy_process_group()


def test_row_partition_under_the_adaptive_rule_world2(tmp_path):
    out = str(tmp_path / "result.txt")
    port = 31500 + (os.getpid() % 2000)
    mp.spawn(_worker_rows, args=(2, port, out), nprocs=2, join=True)
    assert open(out).read() == "ok"
