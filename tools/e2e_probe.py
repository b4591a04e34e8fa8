"""Where does the host-pointer path lose time?  Runs seal-only / open-only / mixed worker sets over
8 sub-batches and prints aggregate GB/s of payload.  usage: python tools/e2e_probe.py"""
import ctypes as C
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import p2p_llm_tunnel_b200 as p2pt  # noqa: E402
from p2p_llm_tunnel_b200 import workload  # noqa: E402

S, R, P, NB = 65536, 16, 1200, int(os.environ.get("NB", "8"))
w = workload.pipeline_workload(S, R, P)
keys, ivs = workload.sessions_keys(S)
n = w["n"]; sub_n = n // NB; sub_streams = S * 4 // NB
nctx = int(os.environ.get("NCTX", "4"))
ctxs = [p2pt.Context(0) for _ in range(nctx)]
for c in ctxs:
    c.sessions_load(keys, ivs, keys, ivs)
    if os.environ.get("CHUNK_MB"):
        c.set_option("chunk_bytes", int(float(os.environ["CHUNK_MB"]) * (1 << 20)))
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).pin_memory()
h = {k: pin(w[k]) for k in ("session", "epoch", "seq48", "nonce8", "types", "sids", "payload", "pay_off", "rec_off")}
esz = {"session": 4, "epoch": 2, "seq48": 8, "nonce8": 8, "types": 1, "sids": 4, "pay_off": 8, "rec_off": 8}
hp = {k: h[k].data_ptr() for k in h}
rec_bytes = int(w["rec_off"][-1])
h_rec = torch.empty(rec_bytes, dtype=torch.uint8).pin_memory()
h_status = torch.empty(n, dtype=torch.uint8).pin_memory()
h_desc = torch.empty(n * 24, dtype=torch.uint8).pin_memory()
h_u = torch.empty(NB * (sub_streams + 1) * 8, dtype=torch.uint8).pin_memory()
h_us = torch.empty_like(h_u); h_csr = torch.empty_like(h_u); h_goff = torch.empty_like(h_u)
h_perm = torch.empty(n * 4, dtype=torch.uint8).pin_memory()
h_g = torch.empty(n * P, dtype=torch.uint8).pin_memory()
kks = (C.c_uint32 * NB)()
at = lambda name, r0: hp[name] + r0 * esz[name]


def seal(c, b):
    r0 = b * sub_n
    c.call("p2pt_pipeline_seal", sub_n, at("session", r0), at("epoch", r0), at("seq48", r0), at("nonce8", r0), at("types", r0),
           at("sids", r0), hp["payload"], at("pay_off", r0), h_rec.data_ptr(), at("rec_off", r0), h_status.data_ptr() + r0)


def open_(c, b):
    r0 = b * sub_n; so = b * (sub_streams + 1)
    c.call("p2pt_pipeline_open_demux", sub_n, at("session", r0), h_rec.data_ptr(), at("rec_off", r0), None,
           h_desc.data_ptr() + 24 * r0, sub_streams, h_u.data_ptr() + 4 * so, h_us.data_ptr() + 4 * so, h_csr.data_ptr() + 8 * so,
           h_perm.data_ptr() + 4 * r0, C.addressof(kks) + 4 * b, h_g.data_ptr() + r0 * P, sub_n * P, h_goff.data_ptr() + 8 * so)


def run(workers, reps=3):
    """workers: list of (fn, ctx index); each worker processes sub-batches b = its rank among same-fn workers, stride count"""
    groups = {}
    for fn, ci in workers:
        groups.setdefault(fn, []).append(ci)

    def body(fn, ci, rank, cnt):
        for _ in range(reps):
            for b in range(rank, NB, cnt):
                fn(ctxs[ci], b)
    ts = []
    for fn, cis in groups.items():
        for rank, ci in enumerate(cis):
            ts.append(threading.Thread(target=body, args=(fn, ci, rank, len(cis))))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for t in ts: t.start()
    for t in ts: t.join()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


for b in range(NB):
    seal(ctxs[0], b)
for b in range(NB):
    open_(ctxs[0], b)
assert all(int(k) == sub_streams for k in kks)
gb = n * P / 1e9
NW = int(os.environ.get("NW", "2"))
for name, ws in ((f"seal x{NW} + open x{NW}", [(seal, i) for i in range(NW)] + [(open_, NW + i) for i in range(NW)]),):
    t = run(ws)
    print(f"NB={NB} CHUNK_MB={os.environ.get('CHUNK_MB')} {name:22s} {t * 1e3:8.2f} ms per {n} records (each direction) -> {gb / t:6.2f} GB/s payload per flow")
sys.exit(0)
# raw DMA over the very buffers the calls use: payload up, records down, at the same time
dev = torch.device("cuda:0")
d_pay = torch.empty(n * P, dtype=torch.uint8, device=dev); d_rec = torch.empty(rec_bytes, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def both(pieces):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    step_p = (n * P) // pieces; step_r = rec_bytes // pieces
    for i in range(pieces):
        with torch.cuda.stream(s1):
            d_pay[i * step_p:(i + 1) * step_p].copy_(h["payload"][i * step_p:(i + 1) * step_p], non_blocking=True)
        with torch.cuda.stream(s2):
            h_rec[i * step_r:(i + 1) * step_r].copy_(d_rec[i * step_r:(i + 1) * step_r], non_blocking=True)
    torch.cuda.synchronize(); return time.perf_counter() - t0
for pieces in (1, 8, 80):
    t = min(both(pieces) for _ in range(3))
    print(f"raw DMA payload up + records down, {pieces:3d} pieces: {t * 1e3:7.2f} ms -> {n * P / t / 1e9:5.1f} GB/s up, {rec_bytes / t / 1e9:5.1f} GB/s down")
def one(up):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    if up: d_pay.copy_(h["payload"], non_blocking=True)
    else: h_rec.copy_(d_rec, non_blocking=True)
    torch.cuda.synchronize(); return time.perf_counter() - t0
print(f"raw DMA alone: up {n * P / min(one(True) for _ in range(3)) / 1e9:5.1f} GB/s, down {rec_bytes / min(one(False) for _ in range(3)) / 1e9:5.1f} GB/s")
