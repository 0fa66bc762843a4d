"""world_size-2 gloo tests of the multi-GPU host logic (shard ranges + the 16-byte winner exchange)."""
import importlib
import os
import socket

import numpy as np
import pytest

PKG = "catenary-model-estimation-and-mpc-control-for-rov-tethered-systems_b200"


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, costs, q):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import catenary_b200  # noqa: F401
    d = importlib.import_module(PKG + ".distributed")
    lo, hi = d.shard_range(len(costs), rank, world)
    local = np.asarray(costs[lo:hi])
    li = int(np.argmin(np.where(np.isnan(local), np.inf, local)))   # stands in for the per-GPU kernel argmin (NaN last)
    best = d.exchange_winner(d.pack_winner(float(local[li]), lo + li))
    q.put((rank, best))
    dist.destroy_process_group()


@pytest.mark.parametrize("case", ["random", "tie_across_ranks", "nan"])
def test_exchange_winner_gloo_world2(built_lib, case):
    import torch.multiprocessing as mp
    rng = np.random.default_rng(11)
    costs = rng.uniform(1, 2, 101)
    if case == "tie_across_ranks":
        costs[7] = costs[80] = 0.5
    if case == "nan":
        costs[60] = np.nan
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, costs, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    want = int(np.argmin(np.where(np.isnan(costs), np.inf, costs)))   # MPC winner order: a NaN cost never wins
    for rank, (idx, cost) in res:
        assert idx == want, (rank, idx, want)
        assert cost == costs[want]


def test_shard_range_partition(built_lib):
    d = importlib.import_module(PKG + ".distributed")
    for total in (0, 1, 7, 4194304, 4097):
        for world in (1, 2, 3, 8):
            spans = [d.shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
