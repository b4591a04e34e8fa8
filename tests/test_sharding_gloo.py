"""CPU, world_size 2 over gloo: the N>1 path of bench.py -- sessions sharded s mod G, no data-path
collective, each rank works on its own records, timing reduced with MAX.  The per-rank work here is
done by the CPU oracle (allowed in tests) because no GPU exists on this box; what is under test is the
partitioning and the reduction plumbing."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_binding as ob


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from p2p_llm_tunnel_b200 import sharding, workload
    S, R, P = 24, 5, 300
    keys, ivs = workload.sessions_keys(S)
    w = workload.pipeline_workload(S, R, P)                       # the global, unsharded batch
    idx = sharding.shard_records(w["session"], rank, world)
    assert np.all(sharding.gpu_of_session(w["session"][idx], world) == rank)
    local = sharding.local_session_index(w["session"][idx], world)
    mine = sharding.shard_sessions(S, rank, world)
    assert np.array_equal(sharding.global_session_id(local, rank, world), w["session"][idx])
    # this rank's key table holds only its sessions
    lkeys, livs = keys[mine], ivs[mine]
    sub = {"n": len(idx), "session": local, "epoch": w["epoch"][idx], "seq48": w["seq48"][idx],
           "nonce8": w["nonce8"].reshape(-1, 8)[idx].reshape(-1), "types": w["types"][idx], "sids": w["sids"][idx],
           "payload": w["payload"].reshape(-1, P)[idx].reshape(-1),
           "pay_off": workload.fixed_offsets(len(idx), P), "rec_off": workload.fixed_offsets(len(idx), P + 42)}
    rec = ob.pipeline_seal(sub, lkeys, livs)
    np.save(os.path.join(out_dir, f"rec{rank}.npy"), rec)
    np.save(os.path.join(out_dir, f"idx{rank}.npy"), idx)
    # the timing reduction bench.py uses
    assert sharding.max_over_ranks(10.0 + rank) == 10.0 + world - 1
    assert sharding.sum_over_ranks(len(idx)) == S * R
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_cover_the_batch_exactly_once(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    from p2p_llm_tunnel_b200 import workload
    S, R, P = 24, 5, 300
    keys, ivs = workload.sessions_keys(S)
    w = workload.pipeline_workload(S, R, P)
    full = ob.pipeline_seal(w, keys, ivs).reshape(S * R, P + 42)
    seen = np.zeros(S * R, bool)
    for r in range(world):
        idx = np.load(tmp_path / f"idx{r}.npy"); rec = np.load(tmp_path / f"rec{r}.npy").reshape(-1, P + 42)
        assert not seen[idx].any()
        seen[idx] = True
        assert np.array_equal(rec, full[idx])          # sharded result == the unsharded batch's records
    assert seen.all()


def test_sharding_helpers_single_process():
    from p2p_llm_tunnel_b200 import sharding
    s = np.arange(100, dtype=np.uint32)
    for g in (1, 2, 4, 8):
        parts = [sharding.shard_sessions(100, r, g) for r in range(g)]
        assert np.array_equal(np.sort(np.concatenate(parts)), s)
        for r in range(g):
            assert np.array_equal(sharding.global_session_id(sharding.local_session_index(parts[r], g), r, g), parts[r])
    assert sharding.max_over_ranks(3.5) == 3.5
