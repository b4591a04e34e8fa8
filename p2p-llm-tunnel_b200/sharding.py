"""Multi-GPU partitioning: tunnel sessions are independent (own DTLS keys, own stream-id namespace,
own pending map -- /root/reference/tunnel/src/proxy.rs:52-53, serve.rs:65), so session s lives on GPU
s mod G and nothing crosses GPUs: no collective on the data path.  torch.distributed is used only
for the barrier and the max-over-ranks of the timing."""
import numpy as np


def gpu_of_session(session, world):
    return np.asarray(session, dtype=np.uint64) % np.uint64(world)


def local_session_index(session, world):
    """index of a global session id inside its GPU's key table"""
    return (np.asarray(session, dtype=np.uint64) // np.uint64(world)).astype(np.uint32)


def global_session_id(local, rank, world):
    return (np.asarray(local, dtype=np.uint64) * np.uint64(world) + np.uint64(rank)).astype(np.uint32)


def shard_records(session, rank, world):
    """indices (ascending, so per-session record order is kept) of the records whose session is on `rank`"""
    return np.nonzero(gpu_of_session(session, world) == rank)[0]


def shard_sessions(n_sessions, rank, world):
    """global ids of the sessions held by `rank`"""
    return np.arange(rank, n_sessions, world, dtype=np.uint32)


def max_over_ranks(value, device=None):
    """max of a python float over the ranks of the default process group (identity when not initialised)"""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value, device=None):
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
