"""Batch size at which the GPU path overtakes the CPU port, end to end with host buffers.
For n = 1 .. 65536 records of 1200 B (16 per session): wall clock of one p2pt_pipeline_seal +
p2pt_pipeline_open_demux round trip (pinned host buffers allocated once, one ctx, one thread, bare ctypes
calls) against the AES-NI port on 1 and on all host threads.
usage: python tools/crossover_probe.py > profiles/rNN_crossover.json"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (cpu_pipeline, load_cpufast)
import p2p_llm_tunnel_b200 as p2pt  # noqa: E402
from p2p_llm_tunnel_b200 import workload  # noqa: E402

P = 1200
ctx = p2pt.Context(0)
ctx.set_option("chunk_bytes", 256 << 20)
lib, h = ctx.lib, ctx.h
cpu = bench.load_cpufast()
cores = os.cpu_count() or 1
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).pin_memory()
rows = []
for n in (1, 4, 16, 64, 256, 1024, 4096, 16384, 65536):
    S = max(1, n // 16)
    w = workload.pipeline_workload(S, min(16, n), P)
    assert w["n"] == n
    keys, ivs = workload.sessions_keys(S)
    ctx.sessions_load(keys, ivs, keys, ivs)
    hb = {k: pin(w[k]) for k in ("session", "epoch", "seq48", "nonce8", "types", "sids", "payload", "pay_off", "rec_off")}
    ptr = {k: C.c_void_p(v.data_ptr()) for k, v in hb.items()}
    rec_bytes = int(w["rec_off"][-1]); ms = S * 4
    o = {"rec": torch.empty(rec_bytes, dtype=torch.uint8).pin_memory(), "status": torch.empty(n, dtype=torch.uint8).pin_memory(),
         "desc": torch.empty(24 * n, dtype=torch.uint8).pin_memory(), "uniq": torch.empty(4 * (ms + 1), dtype=torch.uint8).pin_memory(),
         "usess": torch.empty(4 * (ms + 1), dtype=torch.uint8).pin_memory(), "csr": torch.empty(8 * (ms + 1), dtype=torch.uint8).pin_memory(),
         "goff": torch.empty(8 * (ms + 1), dtype=torch.uint8).pin_memory(), "perm": torch.empty(4 * n, dtype=torch.uint8).pin_memory(),
         "gath": torch.empty(n * P, dtype=torch.uint8).pin_memory()}
    op = {k: C.c_void_p(v.data_ptr()) for k, v in o.items()}
    kk = C.c_uint32(0)

    def gpu_once():
        rc = lib.p2pt_pipeline_seal(h, n, ptr["session"], ptr["epoch"], ptr["seq48"], ptr["nonce8"], ptr["types"], ptr["sids"],
                                    ptr["payload"], ptr["pay_off"], op["rec"], ptr["rec_off"], op["status"])
        assert rc == 0, ctx.last_error() if hasattr(ctx, "last_error") else rc
        rc = lib.p2pt_pipeline_open_demux(h, n, ptr["session"], op["rec"], ptr["rec_off"], None, op["desc"], ms, op["uniq"], op["usess"],
                                          op["csr"], op["perm"], C.byref(kk), op["gath"], n * P, op["goff"])
        assert rc == 0, rc
    gpu_once(); gpu_once()
    assert kk.value == min(ms, n) and int(o["status"].sum()) == 0
    reps = 300 if n <= 256 else 50 if n <= 4096 else 10
    t0 = time.perf_counter()
    for _ in range(reps):
        gpu_once()
    t_gpu = (time.perf_counter() - t0) / reps
    rec_cpu = np.zeros(rec_bytes, np.uint8)
    t_cpu = {}
    for nt in (1, cores):
        bench.cpu_pipeline(cpu, w, keys, ivs, rec_cpu, nt)
        t_cpu[nt] = min(sum(bench.cpu_pipeline(cpu, w, keys, ivs, rec_cpu, nt)[0]) for _ in range(max(3, reps // 10)))
    assert np.array_equal(rec_cpu, o["rec"].numpy())
    rows.append({"records": n, "gpu_us": round(t_gpu * 1e6, 1), "gpu_gbs": round(n * P / t_gpu / 1e9, 3),
                 "cpu_1_thread_us": round(t_cpu[1] * 1e6, 1), "cpu_1_thread_gbs": round(n * P / t_cpu[1] / 1e9, 3),
                 f"cpu_{cores}_threads_us": round(t_cpu[cores] * 1e6, 1), f"cpu_{cores}_threads_gbs": round(n * P / t_cpu[cores] / 1e9, 3)})
    print(rows[-1], file=sys.stderr)
print(json.dumps({"payload_bytes": P, "host_cores": cores,
                  "note": "one seal + open+demux round trip per batch through the C ABI (ctypes), pinned host buffers, one ctx, one host thread; "
                          "CPU = AES-NI/PCLMULQDQ port, its own timers (threads spawned per call)", "rows": rows}, indent=1))
