"""Small profiling target: one device-resident pipeline pass over the C5 batch (or a slice of it).
Usage: python tools/prof_target.py [sessions] [reps]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import p2p_llm_tunnel_b200 as p2pt  # noqa: E402
from p2p_llm_tunnel_b200 import workload  # noqa: E402

sessions = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = torch.device("cuda:0")
ctx = p2pt.Context(0)
w = workload.pipeline_workload(sessions, 16, 1200)
keys, ivs = workload.sessions_keys(sessions)
ctx.sessions_load(keys, ivs, keys, ivs)
n = w["n"]
t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
d = {k: t(w[k]) for k in ("session", "epoch", "seq48", "nonce8", "types", "sids", "payload", "pay_off", "rec_off")}
rec = torch.empty(int(w["rec_off"][-1]), dtype=torch.uint8, device=dev)
out = torch.empty_like(rec)
status = torch.empty(n, dtype=torch.uint8, device=dev)
desc = torch.empty(n * 24, dtype=torch.uint8, device=dev)
sb = int(ctx.lib.p2pt_frame_demux_scratch_bytes(n))
scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
ms = sessions * 4
uniq = torch.empty(ms + 1, dtype=torch.int32, device=dev); usess = torch.empty(ms + 1, dtype=torch.int32, device=dev)
csr = torch.empty(ms + 1, dtype=torch.int64, device=dev); goff = torch.empty(ms + 1, dtype=torch.int64, device=dev)
perm = torch.empty(n, dtype=torch.int32, device=dev); k = torch.zeros(1, dtype=torch.int32, device=dev)
gathered = torch.empty(n * 1200, dtype=torch.uint8, device=dev)
s = torch.cuda.current_stream().cuda_stream
print("lut_slack", ctx.get_info("lut_slack"), "threads", ctx.get_info("gcm_threads"))
gs = [int(x) for x in os.environ.get("PROF_G", "0").split(",")]
for g in gs:
    ctx.set_option("gcm_group_width", g)
    for _ in range(reps):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record()
        ctx.call("p2pt_pipeline_seal_device", n, d["session"], d["epoch"], d["seq48"], d["nonce8"], d["types"], d["sids"],
                 d["payload"], d["pay_off"], rec, d["rec_off"], status, s)
        ev[1].record()
        ctx.call("p2pt_pipeline_open_device", n, d["session"], rec, d["rec_off"], out, d["rec_off"], desc, s)
        ev[2].record()
        ctx.call("p2pt_frame_demux_batch_device", n, desc, d["session"], out, ms, uniq, usess, csr, perm, k, gathered,
                 n * 1200, goff, scratch, sb, s)
        ev[3].record()
        torch.cuda.synchronize()
    tt = [ev[i].elapsed_time(ev[i + 1]) for i in range(3)]
    print(f"G={g}: seal {tt[0]:.3f} ms ({n * 1200 / tt[0] / 1e6:.1f} GB/s)  open {tt[1]:.3f} ms ({n * 1200 / tt[1] / 1e6:.1f} GB/s)  "
          f"demux+gather {tt[2]:.3f} ms")
assert int(status.sum()) == 0 and int(k.item()) == ms
print("prof target ok", n, "records")
sys.exit(0)
for _ in range(reps):
    ctx.call("p2pt_pipeline_seal_device", n, d["session"], d["epoch"], d["seq48"], d["nonce8"], d["types"], d["sids"],
             d["payload"], d["pay_off"], rec, d["rec_off"], status, s)
    ctx.call("p2pt_pipeline_open_device", n, d["session"], rec, d["rec_off"], out, d["rec_off"], desc, s)
    ctx.call("p2pt_frame_demux_batch_device", n, desc, d["session"], out, ms, uniq, usess, csr, perm, k, gathered,
             n * 1200, goff, scratch, sb, s)
torch.cuda.synchronize()
assert int(status.sum()) == 0 and int(k.item()) == ms
print("prof target ok", n, "records")
