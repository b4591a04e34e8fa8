"""Profiling target for the frame codec configs: encode, decode, demux + gather on a device-resident
batch.  usage: python tools/prof_codec.py [frames] [payload] [reps]   (defaults: C4 = 10 M x 64 B)"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import p2p_llm_tunnel_b200 as p2pt  # noqa: E402
from p2p_llm_tunnel_b200 import workload  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
P = int(sys.argv[2]) if len(sys.argv) > 2 else 64
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
dev = torch.device("cuda:0")
ctx = p2pt.Context(0)
if os.environ.get("COOP_SORT") is not None:
    ctx.set_option("coop_sort", int(os.environ["COOP_SORT"]))
w = workload.frames_workload(n, P, sse=(P == 64))
t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
d = {k: t(w[k]) for k in ("types", "sids", "payload", "pay_off", "frame_off")}
frames = torch.empty(n * (P + 5), dtype=torch.uint8, device=dev)
desc = torch.empty(n * 24, dtype=torch.uint8, device=dev)
sb = int(ctx.lib.p2pt_frame_demux_scratch_bytes(n))
scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
ms = 4096
uniq = torch.empty(ms + 1, dtype=torch.int32, device=dev); csr = torch.empty(ms + 1, dtype=torch.int64, device=dev)
goff = torch.empty(ms + 1, dtype=torch.int64, device=dev); perm = torch.empty(n, dtype=torch.int32, device=dev)
k = torch.zeros(1, dtype=torch.int32, device=dev); gathered = torch.empty(n * P, dtype=torch.uint8, device=dev)
s = torch.cuda.current_stream().cuda_stream
fsb = int(ctx.lib.p2pt_frame_decode_demux_scratch_bytes(n))
fscratch = torch.empty(fsb, dtype=torch.uint8, device=dev)
for _ in range(reps):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    ev[0].record()
    ctx.call("p2pt_frame_encode_batch_device", n, d["types"], d["sids"], d["payload"], d["pay_off"], frames, d["frame_off"], s)
    ev[1].record()
    ctx.call("p2pt_frame_decode_batch_device", n, frames, d["frame_off"], desc, s)
    ev[2].record()
    ctx.call("p2pt_frame_demux_batch_device", n, desc, None, frames, ms, uniq, None, csr, perm, k, gathered, n * P, goff, scratch, sb, s)
    ev[3].record()
    torch.cuda.synchronize()
    tt = [ev[i].elapsed_time(ev[i + 1]) for i in range(3)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ctx.call("p2pt_frame_decode_demux_batch_device", n, frames, d["frame_off"], None, None, ms, uniq, None, csr, None, k,
             gathered, n * P, goff, fscratch, fsb, s)
    e1.record()
    torch.cuda.synchronize()
    print(f"encode {tt[0]:.3f} ms  decode {tt[1]:.3f} ms  demux+gather {tt[2]:.3f} ms  decode+demux+gather in one call {e0.elapsed_time(e1):.3f} ms")
assert int(k.item()) == ms
print("prof codec ok", n, "frames")
