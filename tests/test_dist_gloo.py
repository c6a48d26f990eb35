"""world_size-2 `gloo` test (CPU) of the multi-GPU host logic: batch sharding + ONE all-gather of the packed
per-rank records + winner selection equals the single-process selection over the whole batch.
The packing / selection kernels themselves are CUDA (tested in test_gpu_parity.py); here they are stood in by
numpy restatements that live in this test file only."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NZ = 18 * 8 - 6


def _np_pack(costs, xu, index_offset):
    """numpy restatement of i7_pack_best: [costs | best_cost, best_index (global), XU_best]; NaN costs never win,
    ties keep the lowest index."""
    c = np.where(np.isnan(costs), np.inf, costs)
    bi = int(np.argmin(c))
    return np.concatenate([costs, [costs[bi], float(bi + index_offset)], xu[bi]])


def _np_reduce(gathered, world):
    """numpy restatement of i7_reduce_best: record with the lowest best_cost (first rank wins ties)."""
    g = np.asarray(gathered).reshape(world, -1)
    tails = g[:, -(2 + NZ):]
    return tails[int(np.argmin(tails[:, 0]))]


def _worker(rank, world, port, batch_total, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from indy7_mpc_b200 import dist as i7d
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(7)
        costs = rng.random(batch_total)
        costs[3] = np.nan
        xu = rng.standard_normal((batch_total, NZ))
        lo, hi = i7d.shard_range(batch_total, rank, world)
        packed = torch.from_numpy(_np_pack(costs[lo:hi], xu[lo:hi], lo))
        assert packed.numel() == i7d.packed_len(hi - lo, NZ)
        best, gathered = i7d.gather_best(packed, hi - lo, world,
                                         lambda g, w: torch.from_numpy(_np_reduce(g.numpy(), w)))
        allc = i7d.all_costs(gathered, world, hi - lo, NZ).numpy()
        q.put((rank, best.numpy().copy(), allc.copy()))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.timeout(180)
def test_sharded_gather_equals_single_process():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    world, batch_total = 2, 16
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, batch_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=150) for _ in procs]
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    rng = np.random.default_rng(7)
    costs = rng.random(batch_total)
    costs[3] = np.nan
    xu = rng.standard_normal((batch_total, NZ))
    want = _np_pack(costs, xu, 0)[batch_total:]          # single-process selection over the whole batch
    for rank, best, allc in res:
        assert np.array_equal(best, want), rank          # bit-identical on every rank
        assert np.array_equal(np.nan_to_num(allc, nan=-1.0), np.nan_to_num(costs, nan=-1.0))
        assert int(best[1]) == int(np.nanargmin(costs))


def test_shard_range_and_world1():
    sys.path.insert(0, ROOT)
    import torch
    from indy7_mpc_b200 import dist as i7d
    assert [i7d.shard_range(4096, r, 8) for r in (0, 7)] == [(0, 512), (3584, 4096)]
    with pytest.raises(ValueError):
        i7d.shard_range(10, 0, 4)
    with pytest.raises(ValueError):
        i7d.shard_range(8, 2, 2)
    packed = torch.arange(4 + 2 + NZ, dtype=torch.float64)
    best, g = i7d.gather_best(packed, 4, 1, None)        # world 1: no collective, tail of the record
    assert best.numel() == 2 + NZ and best[0] == 4.0 and g is packed


class _FakeEngine:
    """Stands in for Engine.p2p_create / p2p_connect (CUDA IPC) in the collective set-up logic."""
    def __init__(self, fail_create=False, fail_connect=False):
        self.fail_create, self.fail_connect, self.connected = fail_create, fail_connect, None

    def p2p_create(self, world, rank):
        if self.fail_create:
            raise RuntimeError("no peer access")
        return bytes([rank]) * 64

    def p2p_connect(self, handles):
        if self.fail_connect:
            raise RuntimeError("cannot map the peer buffer")
        self.connected = list(handles)


def _connect_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from indy7_mpc_b200 import dist as i7d
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        good = _FakeEngine()
        r_ok = i7d.connect_peers(good, rank, world)
        r_create = i7d.connect_peers(_FakeEngine(fail_create=(rank == 1)), rank, world)
        r_connect = i7d.connect_peers(_FakeEngine(fail_connect=(rank == 0)), rank, world)
        q.put((rank, r_ok, good.connected, r_create, r_connect))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_connect_peers_is_a_collective_decision():
    """dist.connect_peers: every rank gets every handle in rank order; a failure on ONE rank (creating or mapping a
    buffer) makes EVERY rank return False (the caller keeps the NCCL path) and strands nobody."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_connect_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=150) for _ in procs]
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    for rank, r_ok, handles, r_create, r_connect in res:
        assert r_ok is True and handles == [bytes([0]) * 64, bytes([1]) * 64]
        assert r_create is False and r_connect is False
