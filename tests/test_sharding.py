"""Host-side multi-GPU logic on CPU: byte-balanced shard bounds and the size/offset gather over a
world_size-2 gloo group (no data-path collective, one size gather)."""
import importlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

S = importlib.import_module("hipcomp-core_b200.sharding")


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("kind", ["equal", "ragged", "one_big", "few"])
def test_shard_bounds_cover_and_balance(world, kind):
    rng = np.random.default_rng(world)
    sizes = {"equal": np.full(1000, 65536), "ragged": rng.integers(1, 65536, 1000),
             "one_big": np.concatenate([[10_000_000], rng.integers(1, 100, 50)]), "few": np.array([5, 7, 9])}[kind]
    b = S.shard_bounds(sizes, world)
    assert len(b) == world + 1 and b[0] == 0 and b[-1] == len(sizes) and (np.diff(b) >= 0).all()
    if kind in ("equal", "ragged"):
        per = np.array([sizes[b[r]:b[r + 1]].sum() for r in range(world)])
        assert per.max() - per.min() <= 2 * sizes.max(), "ranges are balanced by bytes to within a chunk"


def test_shard_bounds_empty():
    assert list(S.shard_bounds([], 4)) == [0, 0, 0, 0, 0]


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(123)
        raw_sizes = rng.integers(1000, 65536, 37)  # identical on both ranks
        lo, hi = S.my_range(raw_sizes, rank, world)
        comp = raw_sizes[lo:hi] // (2 + rank)  # stand-in for this rank's compressed sizes
        sizes, offsets, first, base = S.gather_sizes(comp)
        q.put((rank, lo, hi, sizes.tolist(), offsets.tolist(), first, base))
    finally:
        dist.destroy_process_group()


def test_size_gather_world2_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    (r0, lo0, hi0, sz0, of0, f0, b0), (r1, lo1, hi1, sz1, of1, f1, b1) = res
    assert lo0 == 0 and hi0 == lo1 and hi1 == 37
    assert sz0 == sz1 and of0 == of1, "every rank sees the same global layout"
    rng = np.random.default_rng(123)
    raw = rng.integers(1000, 65536, 37)
    expect = np.concatenate([raw[lo0:hi0] // 2, raw[lo1:hi1] // 3])
    assert sz0 == expect.tolist()
    assert of0 == (np.cumsum(expect) - expect).tolist()
    assert (f0, b0) == (0, 0) and f1 == hi0 and b1 == int(expect[:hi0].sum())
