"""N > 1 host logic on CPU: world_size-2 gloo. Each rank renders its block of frame indices with the oracle
(standing in for its GPU), the (sum, count) images are reduced exactly as bench.py does over NCCL, and the result
must equal a single-rank render of the union of the frames up to fp32 summation order."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from prosper_b200 import partition


def test_rank_frames_partition_is_exact_cover():
    spp, G = 16, 8
    seen = []
    for g in range(G):
        first, n = partition.rank_frames(g, G, spp)
        seen += list(range(first, first + n))
    assert seen == list(range(1, G * spp + 1))  # the frame set one GPU would render for G*spp samples
    with pytest.raises(ValueError):
        partition.rank_frames(0, 8, 512)  # 4096 frames: past the reference's period
    with pytest.raises(ValueError):
        partition.rank_frames(2, 2, 4)


def test_sum_mean_round_trip():
    rng = np.random.default_rng(0)
    img = rng.random((4, 5, 4)).astype(np.float32)
    img[..., 3] = 7
    back = partition.sum_to_mean(partition.mean_to_sum(img))
    assert np.allclose(back, img, rtol=1e-6)
    zero = np.zeros((2, 2, 4), np.float32)
    assert np.array_equal(partition.sum_to_mean(zero), zero)


def _worker(rank, world_size, port, spp, w, h, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    import oracle
    from prosper_b200 import scene_data as sd, scene_zoo
    scene = scene_zoo.test_scene(w, h)
    orc = oracle.Oracle(scene)
    first, n = partition.rank_frames(rank, world_size, spp)
    mean, _ = orc.render(sd.reference_pc(frame_index=first, max_bounces=4), w, h, n, threads=2)
    t = torch.from_numpy(partition.mean_to_sum(mean))
    partition.reduce_sum_image(t, dst=0)
    if rank == 0:
        np.save(out_path, partition.sum_to_mean(t.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_reduce_equals_single_rank(tmp_path):
    import oracle
    from prosper_b200 import scene_data as sd, scene_zoo
    w, h, spp = 48, 32, 3
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "reduced.npy")
    mp.spawn(_worker, args=(2, port, spp, w, h, out), nprocs=2, join=True)
    reduced = np.load(out)
    orc = oracle.Oracle(scene_zoo.test_scene(w, h))
    single, _ = orc.render(sd.reference_pc(frame_index=1, max_bounces=4), w, h, 2 * spp, threads=2)
    assert np.all(reduced[..., 3] == 2 * spp)
    assert np.allclose(reduced[..., :3], single[..., :3], rtol=2e-5, atol=1e-6)
