"""Multi-GPU host logic (pyvspg.multi) on CPU: world_size-2 gloo processes render their shards with the CPU oracle
standing in for the device render, combine the fp64 films with the same reduce bench.py issues over NCCL, and rank 0
compares with the single-process film."""
import os
import socket
import sys

import numpy as np
import pytest

import helpers as H
from pyvspg import multi


def test_shards_partition_the_work():
    for world in (1, 2, 4, 8):
        seen = []
        for r in range(world):
            b, e, s = multi.sample_pass_shard(r, world, 13)
            seen += list(range(b, e, s))
        assert sorted(seen) == list(range(13))
        rows = []
        for r in range(world):
            for y0, y1 in multi.band_shard(r, world, 1080, band=8):
                rows += list(range(y0, y1))
        assert sorted(rows) == list(range(1080))
        for mode in ("samples", "bands"):
            assert sum(multi.paths_of_shard(1920, 1080, 16, r, world, mode) for r in range(world)) == 1920 * 1080 * 16
    with pytest.raises(ValueError):
        multi.sample_pass_shard(2, 2, 4)


def _worker(rank, world, port, mode, out_path):
    import torch
    import torch.distributed as dist
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (os.path.join(root, "vspg-pbrt-v4_b200"), os.path.join(root, "oracle"), os.path.join(root, "tests")):
        sys.path.insert(0, p)
    import orc
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    cfg = H.small("C2", res=24, spp=4)
    film = np.zeros((cfg.height, cfg.width, 4), dtype=np.float64)

    def render(b, e, s, window, seed):
        orc.render(cfg.scene, cfg.guiding, cfg.ip, b, e, spp_stride=s, window=window, seed=seed, nthreads=2, film=film)

    multi.render_shard(render, cfg.width, cfg.height, cfg.spp, rank, world, seed=3, mode=mode, band=4)
    t = torch.from_numpy(film.reshape(-1))
    multi.combine_film(t, world, dst=0)
    if rank == 0:
        np.save(out_path, film)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["samples", "bands"])
def test_two_rank_gloo_render_equals_single_process(orc, tmp_path, mode):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "film.npy")
    mp.spawn(_worker, args=(2, port, mode, out), nprocs=2, join=True)
    got = np.load(out)
    cfg = H.small("C2", res=24, spp=4)
    want, cnt = orc.render(cfg.scene, cfg.guiding, cfg.ip, 0, cfg.spp, seed=3, nthreads=4)
    assert np.array_equal(got[..., 3], want[..., 3])                 # every pixel received exactly spp samples
    if mode == "bands":
        assert np.array_equal(got, want)                              # disjoint supports: the reduce is a gather
    else:
        assert np.allclose(got, want, rtol=1e-12, atol=0)             # fp64 sums in a different association
