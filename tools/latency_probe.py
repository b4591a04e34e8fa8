"""Wall-clock latency of batch-of-1 host-pointer calls (the reference's per-frame call sites), with the
mapped-buffer small-call path on (default) and off.  usage: python tools/latency_probe.py"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import p2p_llm_tunnel_b200 as p2pt  # noqa: E402
from p2p_llm_tunnel_b200 import workload  # noqa: E402

ctx = p2pt.Context(0)
keys, ivs = workload.sessions_keys(4)
ctx.sessions_load(keys, ivs, keys, ivs)
lib, h = ctx.lib, ctx.h
sse = np.frombuffer(b'data: {"token": "hello"}\n\n' + b" " * 37, np.uint8).copy()
types = np.array([21], np.uint8); sids = np.array([99], np.uint32)
pay_off = np.array([0, 64], np.uint64); f_off = np.array([0, 69], np.uint64); r_off = np.array([0, 69 + 13 + 24], np.uint64)
frame = np.zeros(69, np.uint8); desc = np.zeros(1, p2pt._lib.DESC_DTYPE)
sess = np.zeros(1, np.uint32); ep = np.ones(1, np.uint16); seq = np.zeros(1, np.uint64); nonce = np.arange(8, dtype=np.uint8)
rec = np.zeros(int(r_off[1]), np.uint8); st = np.zeros(1, np.uint8); out = np.zeros(int(r_off[1]), np.uint8)
P = lambda a: C.c_void_p(a.ctypes.data)
calls = {
    "frame_encode": lambda: lib.p2pt_frame_encode_batch(h, 1, P(types), P(sids), P(sse), P(pay_off), P(frame), P(f_off)),
    "frame_decode": lambda: lib.p2pt_frame_decode_batch(h, 1, P(frame), P(f_off), P(desc)),
    "pipeline_seal": lambda: lib.p2pt_pipeline_seal(h, 1, P(sess), P(ep), P(seq), P(nonce), P(types), P(sids), P(sse), P(pay_off), P(rec), P(r_off), P(st)),
    "pipeline_open": lambda: lib.p2pt_pipeline_open(h, 1, P(sess), P(rec), P(r_off), P(out), P(r_off), P(desc)),
}
uniq = np.zeros(2, np.uint32); usess = np.zeros(2, np.uint32); csr = np.zeros(2, np.uint64); goff = np.zeros(2, np.uint64)
perm = np.zeros(1, np.uint32); kk = C.c_uint32(0); gath = np.zeros(64, np.uint8)
calls["pipeline_open_demux"] = lambda: lib.p2pt_pipeline_open_demux(h, 1, P(sess), P(rec), P(r_off), None, P(desc), 1, P(uniq), P(usess),
                                                                     P(csr), P(perm), C.byref(kk), P(gath), 64, P(goff))
desc_d = np.zeros(1, p2pt._lib.DESC_DTYPE); desc_d["pay_off"] = 5; desc_d["pay_len"] = 64; desc_d["stream_id"] = 99; desc_d["msg_type"] = 21
calls["frame_demux"] = lambda: lib.p2pt_frame_demux_batch(h, 1, P(desc_d), None, P(frame), 69, 1, P(uniq), None, P(csr), P(perm), C.byref(kk),
                                                           P(gath), 64, P(goff))
for mode in (65536, 0):
    ctx.set_option("small_call_bytes", mode)
    res = {}
    for name, fn in calls.items():
        for _ in range(20):
            assert fn() == 0
        t0 = time.perf_counter()
        for _ in range(500):
            fn()
        res[name] = round((time.perf_counter() - t0) / 500 * 1e6, 1)
    print("small_call_bytes", mode, res)
assert desc["stream_id"][0] == 99 and st[0] == 0
