"""Multi-GPU plumbing of the batched path: the batch of independent SQP problems is sharded over the ranks of one
box (one process per GPU, torch.distributed), and ONE collective per control tick exchanges
[local costs | best_cost, best_index, XU_best]  (the reference analogue is the host loop
`find_best_trajectory_id`, src/gato_mpc_batch_sample.py:254-300, and `find_best_idx`, gato_controller.py:109-118).

The data path has no other collective.  The functions here only move tensors; packing and the final selection are
CUDA kernels of the library (i7_pack_best / i7_reduce_best) passed in by the caller.
"""
import os


def rank_info():
    """(rank, world_size, local_rank) from the torchrun environment; (0, 1, 0) when launched plainly."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def shard_range(batch_total, rank, world):
    """Contiguous block of problems owned by `rank`.  The batch must divide evenly (the drivers use powers of 2)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    if batch_total % world:
        raise ValueError(f"batch {batch_total} does not divide across {world} ranks")
    per = batch_total // world
    return rank * per, (rank + 1) * per


def packed_len(batch_local, nz):
    """Length of the per-rank record produced by i7_pack_best: costs (B) + best_cost + best_index + XU_best (nz)."""
    return batch_local + 2 + nz


def gather_best(packed, batch_local, world, reduce_fn, gathered=None, group=None):
    """One all-gather of the packed per-rank records, then the winner selection `reduce_fn(gathered, world)`
    (Engine.reduce_best on the GPU).  Returns (best_record [best_cost, global_index, XU_best], gathered).
    With world == 1 no collective is issued and the tail of `packed` is the answer."""
    import torch
    import torch.distributed as dist
    n = packed.numel()
    if world == 1:
        return packed[batch_local:], packed
    if gathered is None:
        gathered = torch.empty(world * n, dtype=packed.dtype, device=packed.device)
    if gathered.numel() != world * n:
        raise ValueError("gathered buffer has the wrong size")
    dist.all_gather_into_tensor(gathered, packed.contiguous(), group=group)
    return reduce_fn(gathered, world), gathered


def all_costs(gathered, world, batch_local, nz):
    """Per-problem costs of the whole batch, in global problem order, from the gathered records."""
    stride = packed_len(batch_local, nz)
    return gathered.view(world, stride)[:, :batch_local].reshape(-1)


def connect_peers(engine, rank, world, group=None):
    """Sets up the peer-memory exchange of `engine` (i7_p2p_*): every rank allocates its exchange buffer, the 64-byte
    inter-process handles travel once through the process group, every rank maps every buffer.  After this,
    `engine.p2p_exchange(stats, xu, index_offset)` replaces pack_best + all_gather + reduce_best by ONE kernel that
    stores the packed record straight into the peers' memory over NVLink.

    Returns True when EVERY rank is connected; False (on every rank alike) when any rank could not create or map a
    buffer (no peer access between the devices, IPC handles not permitted, ...) — the caller then keeps the NCCL
    path.  The collectives are issued by every rank whatever happened locally, so a failure cannot strand the others."""
    import torch
    import torch.distributed as dist
    ok, mine, why = 1, b"\0" * 64, ""
    try:
        mine = engine.p2p_create(world, rank)
    except Exception as e:          # noqa: BLE001 - reported below, the decision is collective
        ok, why = 0, str(e)
    handles = [None] * world
    dist.all_gather_object(handles, (ok, mine), group=group)
    if all(o for o, _ in handles):
        try:
            engine.p2p_connect([h for _, h in handles])
        except Exception as e:      # noqa: BLE001
            ok, why = 0, str(e)
    else:
        ok = 0
    oks = [None] * world
    dist.all_gather_object(oks, ok, group=group)
    if not all(oks):
        if why:
            import sys
            print(f"[indy7_mpc_b200] rank {rank}: peer-memory exchange unavailable ({why}); using the NCCL path",
                  file=sys.stderr)
        return False
    dist.barrier(group=group)
    return True
