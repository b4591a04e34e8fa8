"""GPU parity of the sort-free demux (csrc/demux_direct_kernels.cuh) and of the fused decode + demux entry point.

The direct path is taken when the caller's max_streams is at most 8192; it must give bit for bit what the sorting
path gives (option direct_demux = 0), what the CPU oracle gives (oracle/p2pt_oracle.c orc_frame_demux, which restates
proxy.rs:118-153 / serve.rs:112-139), and what a numpy statement of "stable group-by (session, stream_id)" gives -
including when the batch does not fit the direct path and the device falls back to the sort on its own."""
import ctypes as C

import numpy as np
import pytest

import oracle_binding as ob
from test_gpu_parity import make_frames, ragged

pytestmark = pytest.mark.gpu


def numpy_demux(desc, buf, session=None):
    """stable sort of the decodable frames by (session, stream id) + per-stream body concatenation"""
    ok = np.flatnonzero(desc["status"] == 0)
    key = desc["stream_id"][ok].astype(np.uint64)
    if session is not None:
        key |= session[ok].astype(np.uint64) << np.uint64(32)
    order = ok[np.argsort(key, kind="stable")]
    skey = np.sort(key, kind="stable")
    heads = np.flatnonzero(np.concatenate([[True], skey[1:] != skey[:-1]])) if len(ok) else np.zeros(0, np.int64)
    body = np.where(np.isin(desc["msg_type"][order], (11, 21)), desc["pay_len"][order], 0).astype(np.int64)
    excl = np.cumsum(body) - body
    src = np.repeat(desc["pay_off"][order].astype(np.int64) - excl, body) + np.arange(int(body.sum()), dtype=np.int64)
    return {"k": len(heads), "perm": order.astype(np.uint32),
            "csr": np.concatenate([heads, [len(ok)]]).astype(np.uint64),
            "uniq_sid": (skey[heads] & np.uint64(0xFFFFFFFF)).astype(np.uint32),
            "uniq_session": (skey[heads] >> np.uint64(32)).astype(np.uint32),
            "gathered": buf[src] if len(src) else np.zeros(0, np.uint8),
            "gathered_off": np.concatenate([excl[heads] if len(heads) else [], [int(body.sum())]]).astype(np.uint64)}


def same(a, b, keys, tag):
    assert a["k"] == b["k"], tag
    for k in keys:
        assert np.array_equal(a[k], b[k]), (tag, k)


KEYS = ("uniq_sid", "csr", "perm", "gathered_off", "gathered")


def random_desc(rng, n, mode, with_session):
    from p2p_llm_tunnel_b200._lib import DESC_DTYPE
    if mode == 0:   sid = rng.integers(0, 2**32, 4096, dtype=np.uint64).astype(np.uint32)[rng.integers(0, 4096, n)]   # 4096 wide ids
    elif mode == 1: sid = rng.integers(1, 5, n).astype(np.uint32)                                   # four streams: every warp is one big group
    elif mode == 2: sid = np.where(rng.random(n) < 0.9, 7, rng.integers(0, 5000, n)).astype(np.uint32)   # 90 % one stream
    elif mode == 3: sid = (rng.integers(0, 8000, n) << 8).astype(np.uint32)                         # close to the id capacity
    else:           sid = np.full(n, 0xDEADBEEF, np.uint32)                                         # a single stream
    session = None
    if with_session:
        session = (rng.integers(0, 2, n).astype(np.uint32) << 31) | rng.integers(0, 2, n).astype(np.uint32)
        if mode in (0, 3):
            session[:] = session[0]                       # keep the number of distinct keys within the direct path's reach
    lens = rng.integers(0, 90, n).astype(np.uint32)
    lens[rng.random(n) < 0.02] = 0
    big = rng.random(n) < 0.002
    lens[big] = rng.integers(257, 2000, int(big.sum()))
    gap = rng.integers(0, 7, n)
    pay_off = np.cumsum(np.concatenate([[3], (lens + gap)[:-1]])).astype(np.uint64)
    buf = rng.integers(0, 256, int(pay_off[-1] + lens[-1] + 16), dtype=np.uint8)
    desc = np.zeros(n, DESC_DTYPE)
    desc["pay_off"] = pay_off; desc["pay_len"] = lens; desc["stream_id"] = sid
    desc["msg_type"] = rng.choice(np.array([11, 21, 21, 21, 20, 22, 99], np.uint8), n)
    desc["status"] = np.where(rng.random(n) < 0.03, rng.integers(1, 3, n), 0).astype(np.uint8)
    return desc, buf, session


def test_direct_demux_matches_the_sort_and_numpy(ctx):
    """descriptor batches straight into p2pt_frame_demux_batch with max_streams = 8192: sizes around the tile (1024),
    the warp share (128) and the block share, every key shape, sessions, undecodable frames, empty and long bodies"""
    rng = np.random.default_rng(0x5EED0D1C)
    sizes = [1025, 1151, 1152, 2047, 2048, 2049, 4095, 4097, 5000, 9999, 33000, 70001, 151553, 300000, 2_000_003]
    ctx.set_option("small_demux", 0)                      # let the smallest sizes reach the direct path too
    try:
        for case, n in enumerate([1, 2, 31, 33, 129, 1023, 1024] + sizes):
            desc, buf, session = random_desc(rng, n, case % 5, case % 3 == 1)
            ctx.set_option("direct_demux", 1)
            r = ctx.frame_demux(desc, buf, session=session, max_streams=8192)
            ctx.set_option("direct_demux", 0)
            r2 = ctx.frame_demux(desc, buf, session=session, max_streams=8192)
            e = numpy_demux(desc, buf, session)
            keys = KEYS + (("uniq_session",) if session is not None else ())
            same(r, r2, keys, ("sort", case, n))
            same(r, e, keys, ("numpy", case, n))
            rn = ctx.frame_demux(desc, None, session=session, max_streams=8192, gather=False)      # perm without a gather
            same(rn, e, ("uniq_sid", "csr", "perm"), ("no gather", case, n))
    finally:
        ctx.set_option("small_demux", 1); ctx.set_option("direct_demux", 1)


def _demux_raw(ctx, desc, buf, max_streams, session=None):
    """p2pt_frame_demux_batch without the wrapper's exception: (rc, k)"""
    n = len(desc)
    uniq = np.zeros(max_streams + 1, np.uint32); usess = np.zeros(max_streams + 1, np.uint32)
    csr = np.zeros(max_streams + 1, np.uint64); goff = np.zeros(max_streams + 1, np.uint64)
    perm = np.zeros(n, np.uint32); g = np.zeros(max(len(buf), 1), np.uint8); k = C.c_uint32(0)
    rc = ctx.lib.p2pt_frame_demux_batch(ctx.h, n, desc.ctypes.data, session.ctypes.data if session is not None else None,
                                        buf.ctypes.data, len(buf), max_streams, uniq.ctypes.data,
                                        usess.ctypes.data if session is not None else None, csr.ctypes.data, perm.ctypes.data,
                                        C.addressof(k), g.ctypes.data, len(g), goff.ctypes.data)
    return rc, k.value


def test_more_streams_than_the_direct_path_holds_falls_back_on_the_device(ctx):
    """8192 distinct streams fit exactly; 8193 and 60000 do not: the sorting path takes over inside the same call and
    reports the exact stream count with the capacity error, as before"""
    from p2p_llm_tunnel_b200._lib import DESC_DTYPE
    rng = np.random.default_rng(0x5EED0D1D)
    n = 200_000
    for streams in (8192, 8193, 60000):
        desc = np.zeros(n, DESC_DTYPE)
        sid = rng.integers(0, streams, n).astype(np.uint32)
        sid[:streams] = np.arange(streams)                 # every stream at least once
        desc["stream_id"] = sid * np.uint32(65537)
        desc["pay_len"] = 8; desc["pay_off"] = np.arange(n, dtype=np.uint64) * 8; desc["msg_type"] = 21
        buf = rng.integers(0, 256, 8 * n, dtype=np.uint8)
        rc, k = _demux_raw(ctx, desc, buf, 8192)
        assert k == streams
        assert rc == (0 if streams <= 8192 else -3)
        if streams <= 8192:
            r = ctx.frame_demux(desc, buf, max_streams=8192)
            same(r, numpy_demux(desc, buf), KEYS, streams)


def test_the_key_the_table_cannot_hold(ctx):
    """(session, stream) = (0xFFFFFFFF, 0xFFFFFFFF) is the hash table's empty marker: such a batch must still come out
    right (through the fallback)"""
    rng = np.random.default_rng(0x5EED0D1E)
    n = 5000
    desc, buf, _ = random_desc(rng, n, 1, False)
    session = np.where(rng.random(n) < 0.5, 0xFFFFFFFF, 3).astype(np.uint32)
    desc["stream_id"] = np.where(rng.random(n) < 0.5, 0xFFFFFFFF, 9).astype(np.uint32)
    r = ctx.frame_demux(desc, buf, session=session, max_streams=100)
    same(r, numpy_demux(desc, buf, session), KEYS + ("uniq_session",), "all-ones key")


@pytest.mark.parametrize("n,streams,max_streams", [(1, 1, 4), (700, 5, 8192), (5000, 4096, 4096), (60000, 17, 100),
                                                   (250000, 4096, 8192), (50000, 4096, 50000)])
def test_fused_decode_demux_matches_oracle(ctx, n, streams, max_streams):
    """p2pt_frame_decode_demux_batch against oracle decode followed by oracle demux: corrupt tags, too-short frames,
    every message type; with and without descriptors / perm; small batches (single-block kernel), the direct path and
    the sorting path (max_streams above 8192)"""
    rng = np.random.default_rng(n + streams)
    buf, off = make_frames(rng, n, streams, corrupt=0.03)
    if n > 100:                                            # too-short frames: re-cut some offsets
        cut = np.unique(np.concatenate([off, off[1:n:97] - np.uint64(2)])).astype(np.uint64)
        off = cut
    d = ob.frame_decode(buf, off)
    exp = ob.frame_demux(d, buf)
    assert exp["k"] <= max_streams
    body = int(exp["gathered_off"][-1])
    r = ctx.frame_decode_demux(buf, off, max_streams, gathered_cap=body, want_desc=True)
    assert r["desc"].tobytes() == d.tobytes()
    same(r, exp, KEYS, "with desc")
    r = ctx.frame_decode_demux(buf, off, max_streams, gathered_cap=body, want_desc=False, want_perm=False)
    same(r, exp, ("uniq_sid", "csr", "gathered_off", "gathered"), "bodies only")
    r = ctx.frame_decode_demux(buf, off, max_streams)      # no gather
    same(r, exp, ("uniq_sid", "csr", "perm"), "no gather")
    r = ctx.frame_decode_demux(buf, off, max_streams, want_perm=False)      # the stream table alone
    same(r, exp, ("uniq_sid", "csr"), "table only")


def test_fused_decode_demux_with_sessions_and_fallback(ctx):
    rng = np.random.default_rng(77)
    n = 120_000
    buf, off = make_frames(rng, n, 3000, corrupt=0.01)
    session = rng.integers(0, 5, n).astype(np.uint32)       # up to 15000 (session, stream) keys: beyond the direct path
    d = ob.frame_decode(buf, off)
    exp = ob.frame_demux(d, buf, session)
    assert exp["k"] > 8192
    with pytest.raises(Exception) as e:                     # honest capacity error, exact through the fallback
        ctx.frame_decode_demux(buf, off, 8192, gathered_cap=len(buf), session=session)
    assert e.value.code == -3
    session = rng.integers(0, 2, n).astype(np.uint32)
    exp = ob.frame_demux(d, buf, session)
    assert exp["k"] <= 8192
    r = ctx.frame_decode_demux(buf, off, 8192, gathered_cap=len(buf), session=session, want_desc=True)
    same(r, exp, KEYS + ("uniq_session",), "sessions")


def test_fused_gather_into_a_buffer_that_is_too_small(ctx):
    """gathered_cap below the bodies: capacity error, nothing written past the cap (device form, canary behind)"""
    import torch
    rng = np.random.default_rng(5)
    n = 40_000
    buf, off = make_frames(rng, n, 50, corrupt=0)
    exp = ob.frame_demux(ob.frame_decode(buf, off), buf)
    total = int(exp["gathered_off"][-1])
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
    d_buf, d_off = t(buf), t(off)
    cap = total // 2
    g = torch.full((total + 64,), 0xCD, dtype=torch.uint8, device=dev)
    ms = 64
    uniq = torch.zeros(ms + 1, dtype=torch.int32, device=dev); csr = torch.zeros(ms + 1, dtype=torch.int64, device=dev)
    goff = torch.zeros(ms + 1, dtype=torch.int64, device=dev); k = torch.zeros(1, dtype=torch.int32, device=dev)
    sb = int(ctx.lib.p2pt_frame_decode_demux_scratch_bytes(n)); scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    ctx.call("p2pt_frame_decode_demux_batch_device", n, d_buf, d_off, None, None, ms, uniq, None, csr, None, k, g, cap, goff, scratch, sb, s)
    torch.cuda.synchronize()
    assert int(k.item()) == exp["k"] and int(goff[exp["k"]].item()) == total       # the caller sees what it would have needed
    gh = g.cpu().numpy()
    assert np.all(gh[cap:] == 0xCD)
    # frames that fit lie where they belong
    o = exp["gathered_off"].astype(np.int64)
    first = exp["gathered"][: int(o[1])]
    assert np.array_equal(gh[: len(first)], first)


def test_pipeline_open_demux_takes_the_direct_path(ctx):
    """records -> open -> decode -> demux -> gather in one host call, with a stream bound the direct path accepts
    (proxy.rs:109-170 from the DTLS records): same result as the oracle's open followed by its demux, one record tampered"""
    from p2p_llm_tunnel_b200 import workload
    keys, ivs = workload.sessions_keys(400, seed_shift=9)
    ctx.sessions_load(keys, ivs, keys, ivs)
    w = workload.pipeline_workload(400, 12, 200, seed_shift=9)
    rec = ob.pipeline_seal(w, keys, ivs).copy()
    rec[int(w["rec_off"][77]) + 40] ^= 4
    r = ctx.pipeline_open_demux(w["session"], rec, w["rec_off"], max_streams=8192, gathered_cap=w["n"] * 200, want_out=True)
    assert w["n"] > 1024 and list(np.nonzero(r["desc"]["status"])[0]) == [77]
    exp = ob.frame_demux(r["desc"], r["out"], w["session"])
    same(r, exp, KEYS + ("uniq_session",), "open+demux")
    ctx.set_option("direct_demux", 0)
    try:
        r2 = ctx.pipeline_open_demux(w["session"], rec, w["rec_off"], max_streams=8192, gathered_cap=w["n"] * 200, want_out=True)
    finally:
        ctx.set_option("direct_demux", 1)
    same(r, r2, KEYS + ("uniq_session",), "sorting path")
