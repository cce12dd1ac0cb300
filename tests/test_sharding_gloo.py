"""world_size-2 gloo test of the object-sharding host logic (no GPU): partition, gather, order."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from qsp_slam_b200 import sharding


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_obj, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    costs = [(i % 5) + 1 for i in range(n_obj)]
    ranks = sharding.balanced_order(costs, world)
    mine = ranks[rank]
    # fake per-object records: row i carries its object id everywhere
    local = torch.tensor([[float(i)] * sharding.RECORD_LEN for i in mine], dtype=torch.float32).reshape(-1, sharding.RECORD_LEN)
    allr = sharding.gather_records(local, [len(r) for r in ranks])
    ordered = sharding.scatter_back(allr.numpy(), ranks)
    # the preallocated form gives the same rows, also when called repeatedly (buffers reused)
    g = sharding.RecordGatherer([len(r) for r in ranks], "cpu")
    for _ in range(2):
        g.gather(local)
        assert torch.equal(g.ordered(), allr)
    if rank == 0:
        q.put(ordered[:, 0].tolist())
    dist.barrier()
    dist.destroy_process_group()


def test_partition_covers_everything():
    for n in (0, 1, 7, 32, 1024):
        for w in (1, 2, 4, 8):
            p = sharding.partition(n, w)
            assert p[0][0] == 0 and p[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(p, p[1:]))
            sizes = [hi - lo for lo, hi in p]
            assert max(sizes) - min(sizes) <= 1


def test_balanced_order_is_a_permutation_and_balanced():
    costs = list(np.random.default_rng(0).integers(1, 100, 57))
    r = sharding.balanced_order(costs, 4)
    assert sorted(i for x in r for i in x) == list(range(57))
    loads = [sum(costs[i] for i in x) for x in r]
    assert max(loads) - min(loads) <= max(costs)


def test_balanced_equal_counts_and_cost_proxy():
    costs = list(np.random.default_rng(1).integers(1, 100, 64).astype(float))
    r = sharding.balanced_equal_counts(costs, 8)
    assert sorted(i for x in r for i in x) == list(range(64)) and all(len(x) == 8 for x in r)
    loads = [sum(costs[i] for i in x) for x in r]
    assert max(loads) - min(loads) <= max(costs)
    r3 = sharding.balanced_equal_counts(costs[:10], 3)
    assert sorted(len(x) for x in r3) == [3, 3, 4]
    # the cost proxy counts the depth samples inside the unit ball at the initial pose
    from qsp_slam_b200 import synth
    T = np.eye(4, dtype=np.float32)
    T[:3, :3] *= 2.0
    T[2, 3] = 10.0
    rays = np.array([[0.0, 0.0, 1.0], [1.0, 0.0, 1.0]], np.float32)      # through the centre / missing the ball
    c = sharding.inball_cost(dict(t_cam_obj=T, rays=rays), 101)
    assert abs(c - 100.0) < 1e-6                                          # the whole depth range of the central ray
    assert sharding.inball_cost(dict(t_cam_obj=T, rays=np.zeros((0, 3), np.float32)), 50) == 0.0


def test_gather_two_ranks_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    n_obj = 11
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_obj, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert got == [float(i) for i in range(n_obj)]


def test_refined_equal_count_dealing_narrows_the_spread():
    """refine_equal_counts: same number of objects per rank, every object dealt once, spread of the rank sums far below
    the snake's (which is bounded by one object's cost only)."""
    import numpy as np
    from qsp_slam_b200 import sharding
    rng = np.random.default_rng(3)
    for world, per in ((8, 32), (4, 32), (2, 5), (3, 1)):
        costs = (1.0 + 0.3 * rng.random(world * per)).tolist()
        snake = sharding.balanced_equal_counts(costs, world)
        fine = sharding.refine_equal_counts(costs, snake)
        assert sorted(i for r in fine for i in r) == list(range(world * per))
        assert all(len(r) == per for r in fine)
        spread = lambda sh: max(sum(costs[i] for i in r) for r in sh) - min(sum(costs[i] for i in r) for r in sh)
        assert spread(fine) <= spread(snake) + 1e-12
        if per >= 5:
            assert spread(fine) < 0.02 * max(costs)
        assert sharding.refine_equal_counts(costs, snake) == fine          # deterministic
