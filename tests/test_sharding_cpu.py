"""N > 1 host logic on CPU: shard plan and the final token gather over gloo, world_size 2."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from eventgpt_b200 import sharding


def test_shard_bounds_cover_and_balance():
    for n in (1, 7, 32, 256, 257):
        for world in (1, 2, 4, 8):
            blocks = [sharding.shard_bounds(n, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_bounds(8, 2, 2)


def test_shard_events_ragged():
    off = np.array([0, 10, 10, 35, 100, 101])
    (s, e), local, (lo, hi) = sharding.shard_events(off, 2, 0)
    assert (s, e) == (0, 35) and local.tolist() == [0, 10, 10, 35] and (lo, hi) == (0, 3)
    (s, e), local, (lo, hi) = sharding.shard_events(off, 2, 1)
    assert (s, e) == (35, 101) and local.tolist() == [0, 65, 66] and (lo, hi) == (3, 5)


def _worker(rank, world, port, n_total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = sharding.shard_bounds(n_total, world, rank)
    # "tokens" of window w are filled with w: the gather must restore global window order
    tok = torch.stack([torch.full((6, 4), float(w)) for w in range(lo, hi)]) if hi > lo else torch.zeros((0, 6, 4))
    out = sharding.gather_tokens(tok, n_total)
    ok = tuple(out.shape) == (n_total, 6, 4) and all(bool((out[w] == w).all()) for w in range(n_total))
    q.put((rank, ok))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [8, 7])
def test_gather_tokens_gloo_world2(n_total):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_total, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]
