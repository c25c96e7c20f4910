"""world_size-2 gloo test (CPU) of the multi-GPU host logic: sample/row sharding plans and the framebuffer reduce.  The per-rank
"renderer" here is the CPU oracle in per-(pixel,sample) stream mode (this is a test; the product's renderer is the CUDA kernel)
so the sharding invariance that the streams are designed to give can be checked without a GPU."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_plans(R):
    from ray_tracing_b200 import _dist
    for total in (1, 2, 7, 8, 100, 1024):
        for world in (1, 2, 3, 4, 8):
            got = [_dist.shard_samples(total, r, world) for r in range(world)]
            assert sum(n for _, n in got) == total
            pos = 0
            for s0, n in got:
                assert s0 == pos
                pos += n
            assert max(n for _, n in got) - min(n for _, n in got) <= 1
    rows = [_dist.shard_rows(40, r, 3) for r in range(3)]
    assert rows[0][0] == 0 and rows[-1][1] == 40 and all(rows[i][1] == rows[i + 1][0] for i in range(2))
    p = _dist.plan(2, 40, 3, 8)  # fewer samples than ranks -> row bands
    assert p["spp"] == 2 and p["region"][0] == 15 and p["region"][1] == 20


def _worker(rank, world, port, out_path, total_spp):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import rtb200_loader
    rtb200_loader.load()
    from ray_tracing_b200 import _dist
    import oracle as O
    W, H = 24, 16
    sc = O.OracleScene("cornell")
    pl = _dist.plan(total_spp, H, rank, world)
    region = None
    if pl["region"] is not None:
        region = (pl["region"][0], pl["region"][1], 0, W)
    part, _ = sc.render(W, H, pl["spp"], depth=20, prec=O.PREC_F32, mode=1, seed=5, s0=pl["s0"], region=region, threads=2)
    t = torch.from_numpy(part)
    _dist.reduce_framebuffer(t, 0)
    if rank == 0:
        np.save(out_path, t.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("total_spp", [6, 1])
def test_two_rank_reduce_equals_single_rank(R, O, tmp_path, total_spp):
    out = str(tmp_path / "fb.npy")
    port = 29500 + (os.getpid() % 2000) + total_spp
    mp.spawn(_worker, args=(2, port, out, total_spp), nprocs=2, join=True)
    got = np.load(out)
    full, _ = O.OracleScene("cornell").render(24, 16, total_spp, depth=20, prec=O.PREC_F32, mode=1, seed=5)
    assert np.allclose(got, full, rtol=1e-12, atol=1e-12)
    assert got.sum() > 0
