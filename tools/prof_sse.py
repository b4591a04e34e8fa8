"""profiling target for the SSE event scan: 4096 streams x ~293 KB of SSE text (the bookkeeping config of tools/bench_configs.py)"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import p2p_llm_tunnel_b200 as p2pt
from p2p_llm_tunnel_b200 import workload
dev = torch.device("cuda:0"); ctx = p2pt.Context(0); s = torch.cuda.current_stream().cuda_stream
t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
k = 4096
one = (workload.SSE_TEMPLATE * 2300)[:292968]
body = t(np.frombuffer(one, np.uint8).reshape(1, -1).repeat(k, 0).reshape(-1)); boff = t(workload.fixed_offsets(k, len(one)))
nev = one.count(b"\n\n") * k
ev_csr = torch.empty(k + 1, dtype=torch.int64, device=dev); ev_end = torch.empty(nev + 1, dtype=torch.int64, device=dev)
done = torch.empty(k, dtype=torch.uint8, device=dev); tot = torch.zeros(1, dtype=torch.int64, device=dev)
two_pass = os.environ.get("TWO_PASS") is not None            # the older entry point that reads the bodies twice
bb = k * len(one)
sb = int(ctx.lib.p2pt_sse_scan_scratch_bytes(k) if two_pass else ctx.lib.p2pt_sse_scan_onepass_scratch_bytes(k, bb))
scr = torch.empty(sb, dtype=torch.uint8, device=dev)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if two_pass:
        ctx.call("p2pt_sse_scan_batch_device", k, body, boff, nev, ev_csr, ev_end, done, tot, scr, sb, s)
    else:
        ctx.call("p2pt_sse_scan_onepass_batch_device", k, body, boff, bb, nev, ev_csr, ev_end, done, tot, scr, sb, s)
    e1.record(); torch.cuda.synchronize()
    print(f"sse scan {e0.elapsed_time(e1):.3f} ms for {k * len(one) / 1e9:.2f} GB, {nev} events")
assert int(tot.item()) == nev
