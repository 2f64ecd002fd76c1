"""Multi-GPU host logic on CPU: sample-range arithmetic and the accumulator reduce over a world-size-2 gloo group
(the data path itself needs no other collective)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cudapathtracer_b200 import sharding


@pytest.mark.parametrize("world,spp", [(1, 64), (2, 64), (4, 1024), (8, 1024), (8, 3), (3, 10)])
def test_sample_ranges_are_disjoint_and_cover(world, spp):
    seen = []
    for r in range(world):
        first, n = sharding.sample_range(r, world, spp)
        seen += list(range(first, first + n))
    assert seen == list(range(spp))
    sizes = [sharding.sample_range(r, world, spp)[1] for r in range(world)]
    assert max(sizes) - min(sizes) <= 1


def test_weak_ranges():
    assert [sharding.weak_sample_range(r, 64) for r in range(3)] == [(0, 64), (64, 64), (128, 64)]
    with pytest.raises(ValueError):
        sharding.sample_range(2, 2, 8)


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # each rank "renders" its sample range: a stand-in accumulator whose value is a known function of the sample index
        first, n = sharding.sample_range(rank, world, 10)
        acc = torch.zeros(6 * 3, dtype=torch.float32)
        for s in range(first, first + n):
            acc += torch.arange(18, dtype=torch.float32) * (s + 1)
        sharding.reduce_accumulators(acc, dst=0)
        if rank == 0:
            np.save(out, sharding.normalise(acc, 10).numpy())
    finally:
        dist.destroy_process_group()


def test_reduce_of_accumulators_gloo_world2(tmp_path):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = str(tmp_path / "img.npy")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    want = np.arange(18, dtype=np.float32) * sum(range(1, 11)) / 10.0
    assert np.allclose(got, want)


def test_group_needs_a_gpu(lib):
    """cupt_group is the native multi-GPU path (NCCL owned by the library): without a CUDA device it fails loudly, like
    cupt_create — there is no CPU stand-in"""
    import ctypes as C
    from cudapathtracer_b200 import api
    try:
        n = C.c_int(0)
        if C.CDLL("libcudart.so.12").cudaGetDeviceCount(C.byref(n)) == 0 and n.value > 0:
            pytest.skip("a GPU is present")
    except OSError:
        pass
    with pytest.raises(api.CuptError):
        api.Group(devices=[0])
