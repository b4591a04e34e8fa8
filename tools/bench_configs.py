"""Per-config, per-kernel device-resident timings for BASELINE.json configs C2 / C3 / C4, the SCTP layer and the
bookkeeping kernels.  bench.py imports run_all() and puts the result into its JSON line as `configs` (GPU legs
only); run as a script it also times the CPU port beside every stage and prints one JSON object.
usage: python tools/bench_configs.py > profiles/rNN_configs.json        (ONLY=substring filters the configs)"""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402
import p2p_llm_tunnel_b200 as p2pt  # noqa: E402
from p2p_llm_tunnel_b200 import workload  # noqa: E402

dev = ctx = fast = s = None
peak = 6650.0
cores = os.cpu_count() or 1
t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
p = lambda a: C.c_void_p(a.ctypes.data)


def setup(context, device, with_cpu):
    """bind the module to a p2pt context / torch device; with_cpu loads the timed CPU port (oracle/libcpufast.so)"""
    global dev, ctx, fast, s, peak
    dev, ctx = device, context
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak = json.load(open(pk)).get("hbm_gbs", 6650.0) if os.path.exists(pk) else 6650.0
    s = torch.cuda.current_stream().cuda_stream
    fast = None
    if with_cpu:
        g.build_oracle()
        fast = C.CDLL(g.CPUFAST_LIB)


def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def codec_config(n, P, sse):
    w = workload.frames_workload(n, P, sse=sse)
    d = {k: t(w[k]) for k in ("types", "sids", "payload", "pay_off", "frame_off")}
    frames = torch.empty(n * (P + 5), dtype=torch.uint8, device=dev)
    desc = torch.empty(n * 24, dtype=torch.uint8, device=dev)
    sb = int(ctx.lib.p2pt_frame_demux_scratch_bytes(n))
    scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
    ms = 4096
    uniq = torch.empty(ms + 1, dtype=torch.int32, device=dev); csr = torch.empty(ms + 1, dtype=torch.int64, device=dev)
    goff = torch.empty(ms + 1, dtype=torch.int64, device=dev); perm = torch.empty(n, dtype=torch.int32, device=dev)
    k = torch.zeros(1, dtype=torch.int32, device=dev); gathered = torch.empty(n * P, dtype=torch.uint8, device=dev)
    enc = lambda: ctx.call("p2pt_frame_encode_batch_device", n, d["types"], d["sids"], d["payload"], d["pay_off"], frames, d["frame_off"], s)
    dec = lambda: ctx.call("p2pt_frame_decode_batch_device", n, frames, d["frame_off"], desc, s)
    dmx = lambda: ctx.call("p2pt_frame_demux_batch_device", n, desc, None, frames, ms, uniq, None, csr, perm, k, None, 0, None, scratch, sb, s)
    dmg = lambda: ctx.call("p2pt_frame_demux_batch_device", n, desc, None, frames, ms, uniq, None, csr, perm, k, gathered, n * P, goff, scratch, sb, s)
    fsb = int(ctx.lib.p2pt_frame_decode_demux_scratch_bytes(n))
    fscratch = torch.empty(fsb, dtype=torch.uint8, device=dev)
    # the reader loop in one call: frames in, per-stream bodies out (no descriptors, no perm)
    fused = lambda: ctx.call("p2pt_frame_decode_demux_batch_device", n, frames, d["frame_off"], None, None, ms, uniq, None, csr, None, k,
                             gathered, n * P, goff, fscratch, fsb, s)
    t_enc, t_dec, t_dmx, t_dmg = timeit(enc), timeit(dec), timeit(dmx), timeit(dmg)
    assert int(k.item()) == 4096
    t_both = timeit(lambda: (dec(), dmg()))
    g_ref = gathered.clone(); goff_ref = goff.clone()
    gathered.zero_()
    t_fused = timeit(fused)
    assert int(k.item()) == 4096 and torch.equal(gathered, g_ref) and torch.equal(goff, goff_ref)
    ctx.set_option("direct_demux", 0)                   # the sorting path on the same batch, for comparison
    t_both_sort = timeit(lambda: (dec(), dmg()))
    assert torch.equal(gathered, g_ref)
    ctx.set_option("direct_demux", 1)
    del g_ref
    hf = np.zeros(n * (P + 5), np.uint8) if fast else None
    res = {}
    for nt in ((1, cores) if fast else ()):
        times = (C.c_double * 2)(); stats = (C.c_uint64 * 3)()
        best = [1e9, 1e9]
        for _ in range(2):
            fast.cpufast_codec(C.c_uint64(n), p(w["types"]), p(w["sids"]), p(w["payload"]), p(w["pay_off"]), p(hf), p(w["frame_off"]),
                               C.c_int(nt), times, stats)
            best = [min(best[0], times[0]), min(best[1], times[1])]
        assert stats[0] == 0 and stats[2] == n * P
        res[f"cpu_{nt}_threads"] = {"encode_gbs_payload": n * P / best[0] / 1e9, "decode_demux_gbs_payload": n * P / best[1] / 1e9,
                                     "decode_demux_frames_per_s": n / best[1]}
    if fast:
        assert np.array_equal(hf[: 1 << 20], frames[: 1 << 20].cpu().numpy())
    else:                                   # GPU-only run: check the encoder's output against numpy instead
        f2 = frames[: 4096 * (P + 5)].cpu().numpy().reshape(4096, P + 5)
        assert np.all(f2[:, 0] == 21) and np.array_equal(f2[:, 5:].reshape(-1), w["payload"][: 4096 * P])
        assert np.array_equal(f2[:, 1:5].copy().view(">u4").reshape(-1), w["sids"][:4096])
    out = {
        "frames": n, "payload_bytes": P,
        "encode": {"ms": t_enc, "alg_bytes_per_frame": 2 * P + 26, "hbm_gbs": n * (2 * P + 26) / t_enc / 1e6,
                   "frac_of_hbm_peak": n * (2 * P + 26) / t_enc / 1e6 / peak, "payload_gbs": n * P / t_enc / 1e6},
        "decode": {"ms": t_dec, "alg_bytes_per_frame": 37, "hbm_gbs": n * 37 / t_dec / 1e6, "frames_per_s": n / t_dec * 1e3},
        "demux_no_gather": {"ms": t_dmx, "alg_bytes_per_frame": 45, "frames_per_s": n / t_dmx * 1e3},
        "demux_gather": {"ms": t_dmg},
        "decode_demux_gather": {"ms": t_both, "alg_bytes_per_frame": 2 * P + 41, "hbm_gbs": n * (2 * P + 41) / t_both / 1e6,
                                "frac_of_hbm_peak": n * (2 * P + 41) / t_both / 1e6 / peak, "frames_per_s": n / t_both * 1e3,
                                "payload_gbs": n * P / t_both / 1e6},
        "decode_demux_gather_sorting_path": {"ms": t_both_sort, "frac_of_hbm_peak": n * (2 * P + 41) / t_both_sort / 1e6 / peak},
        "decode_demux_gather_one_call": {"ms": t_fused, "call": "p2pt_frame_decode_demux_batch_device, desc = perm = NULL",
                                         "alg_bytes_per_frame": 2 * P + 41, "hbm_gbs": n * (2 * P + 41) / t_fused / 1e6,
                                         "frac_of_hbm_peak": n * (2 * P + 41) / t_fused / 1e6 / peak, "frames_per_s": n / t_fused * 1e3,
                                         "payload_gbs": n * P / t_fused / 1e6},
    }
    out.update(res)
    return out


def records_config(n, L, n_sessions):
    keys, ivs = workload.sessions_keys(n_sessions, seed_shift=n_sessions)
    ctx.sessions_load(keys, ivs, keys, ivs)
    w = workload.records_workload(n, L, n_sessions=n_sessions)
    d = {k: t(w[k]) for k in ("session", "nonce8", "raw", "raw_off", "rec_off")}
    rec = torch.empty(n * (L + 37), dtype=torch.uint8, device=dev); out = torch.empty_like(rec)
    status = torch.empty(n, dtype=torch.uint8, device=dev); olen = torch.empty(n, dtype=torch.int32, device=dev)
    seal = lambda: ctx.call("p2pt_dtls_gcm_seal_batch_device", n, d["session"], d["nonce8"], d["raw"], d["raw_off"], rec, d["rec_off"], status, s)
    opn = lambda: ctx.call("p2pt_dtls_gcm_open_batch_device", n, d["session"], rec, d["rec_off"], out, d["rec_off"], olen, status, s)
    t_seal = timeit(seal); t_open = timeit(opn)
    assert int(status.sum()) == 0
    return {"records": n, "plaintext_bytes": L, "sessions": n_sessions,
            "seal": {"ms": t_seal, "gbs_plaintext": n * L / t_seal / 1e6, "records_per_s": n / t_seal * 1e3,
                     "alg_bytes_per_record": 2 * L + 79, "frac_of_hbm_peak": n * (2 * L + 79) / t_seal / 1e6 / peak},
            "open": {"ms": t_open, "gbs_plaintext": n * L / t_open / 1e6, "records_per_s": n / t_open * 1e3,
                     "alg_bytes_per_record": 2 * L + 74, "frac_of_hbm_peak": n * (2 * L + 74) / t_open / 1e6 / peak}}


def sctp_config(n, L):
    """n messages of L <= 1200 bytes -> n packets; packetise, then unpack + reassemble"""
    w = workload.frames_workload(n, L)
    msg_off = workload.fixed_offsets(n, L)
    plen = 28 + ((L + 3) & ~3)
    d_msg = t(w["payload"]); d_moff = t(msg_off)
    z16 = torch.full((n,), 5000, dtype=torch.int16, device=dev); one16 = torch.ones(n, dtype=torch.int16, device=dev)
    ssn = (torch.arange(n, device=dev) & 0x7FFF).to(torch.int16); v32 = torch.full((n,), 7, dtype=torch.int32, device=dev)
    pp = torch.full((n,), 53, dtype=torch.int32, device=dev); tsn = torch.arange(n, dtype=torch.int32, device=dev)
    out = torch.empty(n * plen, dtype=torch.uint8, device=dev)
    omoff = t(workload.fixed_offsets(n, plen)[:-1].copy()); first = t(np.arange(n + 1, dtype=np.uint64))
    pkt_off = torch.empty(n + 1, dtype=torch.int64, device=dev)
    pk = lambda: ctx.call("p2pt_sctp_packetize_batch_device", n, d_msg, d_moff, z16, z16, v32, one16, ssn, pp, tsn, None, None,
                          out, omoff, first, pkt_off, s)
    info = torch.empty(n * 16, dtype=torch.uint8, device=dev); desc = torch.empty(n * 24, dtype=torch.uint8, device=dev)
    up = lambda: ctx.call("p2pt_sctp_unpack_batch_device", n, out, pkt_off, 0, None, info, desc, s)
    sb = int(ctx.lib.p2pt_sctp_reassemble_scratch_bytes(n)); scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
    msg = torch.empty(n * L, dtype=torch.uint8, device=dev); moff = torch.empty(n + 1, dtype=torch.int64, device=dev)
    mfirst = torch.empty(n, dtype=torch.int32, device=dev); mstat = torch.empty(n, dtype=torch.uint8, device=dev)
    nm = torch.zeros(1, dtype=torch.int32, device=dev)
    ra = lambda: ctx.call("p2pt_sctp_reassemble_batch_device", n, info, desc, None, out, n, msg, n * L, moff, mfirst, mstat, nm, scratch, sb, s)
    t_pk = timeit(pk); t_up = timeit(up); t_ra = timeit(ra)
    assert int(nm.item()) == n and int(mstat.sum()) == 0 and torch.equal(msg, d_msg)
    hp = np.zeros(n * plen, np.uint8) if fast else None
    hoff = workload.fixed_offsets(n, plen)
    res = {}
    for nt in ((1, cores) if fast else ()):
        times = (C.c_double * 2)(); stats = (C.c_uint64 * 2)()
        fast.cpufast_sctp(C.c_uint64(n), p(w["payload"]), p(msg_off), p(hp), p(hoff), C.c_int(nt), times, stats)
        assert stats[0] == 0 and stats[1] == n * L
        res[f"cpu_{nt}_threads"] = {"packetize_gbs_payload": n * L / times[0] / 1e9, "unpack_verify_gbs_payload": n * L / times[1] / 1e9}
    ab_pk = L + plen + 50; ab_up = plen + 8 + 40
    out_d = {"messages": n, "payload_bytes": L,
             "packetize": {"ms": t_pk, "alg_bytes_per_packet": ab_pk, "hbm_gbs": n * ab_pk / t_pk / 1e6, "frac_of_hbm_peak": n * ab_pk / t_pk / 1e6 / peak,
                           "payload_gbs": n * L / t_pk / 1e6, "packets_per_s": n / t_pk * 1e3},
             "unpack_crc_verify": {"ms": t_up, "alg_bytes_per_packet": ab_up, "hbm_gbs": n * ab_up / t_up / 1e6, "frac_of_hbm_peak": n * ab_up / t_up / 1e6 / peak,
                                   "payload_gbs": n * L / t_up / 1e6},
             "reassemble_sort_gather": {"ms": t_ra, "payload_gbs": n * L / t_ra / 1e6}}
    out_d.update(res)
    return out_d


def bookkeeping_config(n):
    """8(f) rows 3 + 4: datagram split (2 records per datagram), replay window (65536 sessions), SSE scan"""
    rng = np.random.default_rng(7)
    plen = 600
    rec = np.zeros((n, 13 + plen), np.uint8); rec[:, 0] = 23; rec[:, 1] = 0xFE; rec[:, 2] = 0xFD
    rec[:, 11] = plen >> 8; rec[:, 12] = plen & 0xFF
    d_buf = t(rec.reshape(-1)); nd = n // 2
    d_off = t(workload.fixed_offsets(nd, 2 * (13 + plen)))
    sb = int(ctx.lib.p2pt_dtls_datagram_split_scratch_bytes(nd)); scr = torch.empty(sb, dtype=torch.uint8, device=dev)
    rec_off = torch.empty(n + 1, dtype=torch.int64, device=dev); rec_dg = torch.empty(n, dtype=torch.int32, device=dev)
    dst = torch.empty(nd, dtype=torch.uint8, device=dev); nr = torch.zeros(1, dtype=torch.int32, device=dev)
    split = lambda: ctx.call("p2pt_dtls_datagram_split_batch_device", nd, d_buf, d_off, n, rec_off, rec_dg, dst, nr, scr, sb, s)
    t_split = timeit(split)
    assert int(nr.item()) == n
    S = 65536
    session = t((np.arange(n, dtype=np.uint64) % S).astype(np.uint32)); seq = t((np.arange(n, dtype=np.uint64) // S))
    lat = torch.zeros(S, dtype=torch.int64, device=dev); msk = torch.zeros(S, dtype=torch.int64, device=dev)
    started = torch.zeros(S, dtype=torch.uint8, device=dev); acc = torch.empty(n, dtype=torch.uint8, device=dev)
    sb2 = int(ctx.lib.p2pt_dtls_replay_filter_scratch_bytes(n)); scr2 = torch.empty(sb2, dtype=torch.uint8, device=dev)
    def replay():
        started.zero_()
        ctx.call("p2pt_dtls_replay_filter_batch_device", n, session, seq, None, S, lat, msk, started, acc, scr2, sb2, s)
    t_replay = timeit(replay)
    assert int(acc.sum().item()) == n
    # SSE scan over 4096 streams x ~293 KB of SSE text
    k = 4096
    one = (workload.SSE_TEMPLATE * 2300)[:292968]
    body = t(np.frombuffer(one, np.uint8).repeat(1).reshape(1, -1).repeat(k, 0).reshape(-1)); boff = t(workload.fixed_offsets(k, len(one)))
    nev = one.count(b"\n\n") * k
    ev_csr = torch.empty(k + 1, dtype=torch.int64, device=dev); ev_end = torch.empty(nev + 1, dtype=torch.int64, device=dev)
    done = torch.empty(k, dtype=torch.uint8, device=dev); tot = torch.zeros(1, dtype=torch.int64, device=dev)
    sb3 = int(ctx.lib.p2pt_sse_scan_scratch_bytes(k)); scr3 = torch.empty(sb3, dtype=torch.uint8, device=dev)
    sse2 = lambda: ctx.call("p2pt_sse_scan_batch_device", k, body, boff, nev, ev_csr, ev_end, done, tot, scr3, sb3, s)
    t_sse2 = timeit(sse2)
    assert int(tot.item()) == nev
    ref_end = ev_end[:nev].clone(); ref_csr = ev_csr.clone(); tot.zero_(); ev_end.zero_(); ev_csr.zero_()
    bb = k * len(one)
    sb4 = int(ctx.lib.p2pt_sse_scan_onepass_scratch_bytes(k, bb)); scr4 = torch.empty(sb4, dtype=torch.uint8, device=dev)
    sse = lambda: ctx.call("p2pt_sse_scan_onepass_batch_device", k, body, boff, bb, nev, ev_csr, ev_end, done, tot, scr4, sb4, s)
    t_sse = timeit(sse)
    assert int(tot.item()) == nev and torch.equal(ev_end[:nev], ref_end) and torch.equal(ev_csr, ref_csr)
    return {"datagram_split": {"ms": t_split, "records": n, "records_per_s": n / t_split * 1e3},
            "replay_filter_65536_sessions": {"ms": t_replay, "records": n, "records_per_s": n / t_replay * 1e3},
            "sse_scan": {"ms": t_sse, "body_bytes": bb, "events": nev, "hbm_gbs": (bb + 8 * nev) / t_sse / 1e6,
                         "frac_of_hbm_peak": (bb + 8 * nev) / t_sse / 1e6 / peak, "body_gbs": bb / t_sse / 1e6,
                         "note": "one pass (p2pt_sse_scan_onepass_batch_device): algorithmic bytes = body read once + 8 per event written"},
            "sse_scan_two_pass": {"ms": t_sse2, "hbm_gbs": 2 * bb / t_sse2 / 1e6, "frac_of_hbm_peak": 2 * bb / t_sse2 / 1e6 / peak,
                                  "note": "p2pt_sse_scan_batch_device: count, scan, emit - the bodies are read twice"}}


CONFIGS = {"bookkeeping_1M_records": lambda: bookkeeping_config(1_000_000),
           "SCTP_1M_messages_1200B": lambda: sctp_config(1_000_000, 1200),
           "C2_1M_frames_1200B_4096_streams": lambda: codec_config(1_000_000, 1200, False),
           "C4_10M_frames_64B_sse": lambda: codec_config(10_000_000, 64, True),
           "C3_1M_records_1200B_1_session": lambda: records_config(1_000_000, 1200, 1),
           "C3_1M_records_1200B_4096_sessions": lambda: records_config(1_000_000, 1200, 4096)}


def run_all(context, device, with_cpu=False, only=None):
    """every config, device-resident, best of 5 with CUDA events on the launching stream; the caller must re-load its
    own sessions afterwards (the C3 configs load theirs)"""
    setup(context, device, with_cpu)
    res = {"hbm_peak_gbs": peak, "host_cores": cores, "timing": "best of 5, CUDA events on the launching stream, inputs resident in HBM"}
    for name, fn in CONFIGS.items():
        if not only or only in name:
            res[name] = fn()
            torch.cuda.empty_cache()
    return res


if __name__ == "__main__":
    print(json.dumps(run_all(p2pt.Context(0), torch.device("cuda:0"), with_cpu=True, only=os.environ.get("ONLY")), indent=1))
