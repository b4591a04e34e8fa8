"""packetise / unpack times only (no result checks): for experiments that break the checksum on purpose"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import p2p_llm_tunnel_b200 as p2pt
from p2p_llm_tunnel_b200 import workload
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
L = 1200; plen = 1228
dev = torch.device("cuda:0"); ctx = p2pt.Context(0); s = torch.cuda.current_stream().cuda_stream
t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
w = workload.frames_workload(n, L)
d_msg = t(w["payload"]); d_moff = t(workload.fixed_offsets(n, L))
z16 = torch.full((n,), 5000, dtype=torch.int16, device=dev); one16 = torch.ones(n, dtype=torch.int16, device=dev)
ssn = (torch.arange(n, device=dev) & 0x7FFF).to(torch.int16); v32 = torch.full((n,), 7, dtype=torch.int32, device=dev)
pp = torch.full((n,), 53, dtype=torch.int32, device=dev); tsn = torch.arange(n, dtype=torch.int32, device=dev)
out = torch.empty(n * plen, dtype=torch.uint8, device=dev)
omoff = t(workload.fixed_offsets(n, plen)[:-1].copy()); first = t(np.arange(n + 1, dtype=np.uint64))
pkt_off = torch.empty(n + 1, dtype=torch.int64, device=dev)
info = torch.empty(n * 16, dtype=torch.uint8, device=dev); desc = torch.empty(n * 24, dtype=torch.uint8, device=dev)
pk = lambda: ctx.call("p2pt_sctp_packetize_batch_device", n, d_msg, d_moff, z16, z16, v32, one16, ssn, pp, tsn, None, None, out, omoff, first, pkt_off, s)
up = lambda: ctx.call("p2pt_sctp_unpack_batch_device", n, out, pkt_off, 0, None, info, desc, s)
def timeit(fn, reps=7):
    best = 1e9
    for _ in range(3): fn()
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best
print(f"packetize {timeit(pk):.4f} ms  unpack {timeit(up):.4f} ms")
