"""CPU tier: the N>1 host logic -- contiguous batch partition, ragged final gather and the
max-over-ranks timing reduction -- on a world-size-2 gloo group.  The per-rank compute is a numpy stand-in (no GPU here):
what is tested is that sharding any per-unit function over G ranks and
gathering gives exactly the unsharded result, for any G, because units never interact."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from nntoolkitcore_b200 import sharding  # noqa: E402


@pytest.mark.parametrize("n,world", [(0, 1), (1, 1), (7, 2), (8, 2), (4096, 8), (65536, 8), (5, 8), (1000, 3)])
def test_unit_range_is_a_contiguous_exact_partition(n, world):
    cuts = [sharding.unit_range(n, r, world) for r in range(world)]
    assert cuts[0][0] == 0 and cuts[-1][1] == n
    for (a, b), (c, d) in zip(cuts, cuts[1:]):
        assert b == c and a <= b and c <= d
    sizes = [b - a for a, b in cuts]
    assert max(sizes) - min(sizes) <= 1 and sum(sizes) == n


def test_unit_range_rejects_nonsense():
    for bad in [(4, 2, 2), (4, -1, 2), (-1, 0, 1), (4, 0, 0)]:
        with pytest.raises(ValueError):
            sharding.unit_range(*bad)


def _per_unit(units: np.ndarray) -> np.ndarray:
    """stand-in stage: framed magnitude spectrum per clip (numpy); any function that treats units independently will do"""
    frames = np.lib.stride_tricks.sliding_window_view(units, 64, axis=1)[:, ::16]
    return np.abs(np.fft.rfft(frames * np.hanning(64).astype(np.float32), axis=-1)).astype(np.float32)


def _worker(rank, world, port, n_units, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(123)                      # same seed on every rank: the "global" batch
        units = rng.standard_normal((n_units, 400)).astype(np.float32)
        local = sharding.shard_apply(_per_unit, units, rank, world)
        first, last = sharding.unit_range(n_units, rank, world)
        assert local.shape[0] == last - first
        full = sharding.gather_to_root(torch.from_numpy(local), n_units, dist)
        slow = sharding.max_over_ranks(10.0 + rank, dist)     # rank-dependent "device time"
        assert slow == 10.0 + world - 1
        if rank == 0:
            np.save(os.path.join(out_dir, "gathered.npy"), full.numpy())
        else:
            assert full is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("n_units", [9, 16])      # ragged (5 + 4) and even split
def test_two_rank_gloo_shard_and_gather_equals_unsharded(tmp_path, n_units):
    mp.spawn(_worker, args=(2, _free_port(), n_units, str(tmp_path)), nprocs=2, join=True)
    got = np.load(tmp_path / "gathered.npy")
    rng = np.random.default_rng(123)
    want = _per_unit(rng.standard_normal((n_units, 400)).astype(np.float32))
    assert got.shape == want.shape
    assert np.array_equal(got, want), "sharded + gathered result must be bit-identical to the unsharded one"


def test_single_process_paths_are_identity():
    x = torch.arange(12.0).reshape(4, 3)
    assert sharding.gather_to_root(x, 4, None) is x
    assert sharding.max_over_ranks(3.5, None) == 3.5
