"""CPU tier, world size 2 (gloo): the host-side logic of the multi-GPU path.

bench.py splits samples over ranks (rank g of N renders absolute sample indices
(k*N + g)*spp + [1..spp] of step k) and combines the fp32 sums with one reduce.
Here the two ranks produce their partial sums with the host simulation of the
kernels (tests/hostsim, Philox policy) and reduce them with torch.distributed /
gloo; the result must equal one process rendering all samples, up to fp32
re-association."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_path):
    sys.path.insert(0, str(ROOT))
    import torch
    import torch.distributed as dist
    from tests import cases
    from tests.util import HostSim
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    hs = HostSim()
    sc = cases.frame_scenes()["mixed"]
    w, h, spp, steps = 32, 18, 3, 2
    total = torch.zeros(h, w, 3)
    for k in range(steps):
        first = (k * world + rank) * spp + 1                   # bench.py::step
        frames, _ = hs.render(sc, w, h, bsdf_mode=0, rng_mode=0, max_bounces=4, frame_first=first, n_frames=spp)
        acc = torch.from_numpy(frames.sum(axis=0, dtype=np.float32))
        dist.reduce(acc, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            total += acc
    if rank == 0:
        np.save(out_path, (total / float(spp * steps * world)).numpy())
    dist.destroy_process_group()


def test_spp_split_reduce_equals_single_process(tmp_path, hostsim):
    import torch.multiprocessing as mp
    from tests import cases
    out = tmp_path / "mean.npy"
    mp.spawn(_worker, args=(2, _free_port(), str(out)), nprocs=2, join=True)
    got = np.load(out)
    sc = cases.frame_scenes()["mixed"]
    frames, _ = hostsim.render(sc, 32, 18, bsdf_mode=0, rng_mode=0, max_bounces=4, frame_first=1, n_frames=12)
    want = frames.mean(axis=0, dtype=np.float64)
    assert np.isfinite(got).all() and got.mean() > 0
    assert np.allclose(got, want, rtol=1e-5, atol=1e-7)         # same 12 samples per pixel, different summation order
