"""GPU: SSE event scan over per-stream bodies against the oracle."""
import json

import numpy as np
import pytest

import oracle_binding as ob

pytestmark = pytest.mark.gpu


def _sse_stream(rng, n_tokens, done=True):
    # the mock upstream's stream: /root/reference/tmp/mock_llm.py:43-70
    out = b""
    for i in range(n_tokens):
        chunk = {"id": "chatcmpl-test", "object": "chat.completion.chunk",
                 "choices": [{"index": 0, "delta": {"content": "tok%d" % int(rng.integers(0, 1000))}, "finish_reason": None}]}
        out += f"data: {json.dumps(chunk)}\n\n".encode()
    if done:
        out += b"data: [DONE]\n\n"
    return out


def test_sse_scan_matches_oracle(ctx):
    rng = np.random.default_rng(51)
    streams = [_sse_stream(rng, int(rng.integers(0, 60)), done=bool(i % 3)) for i in range(500)]
    streams[3] = b""                                   # empty stream
    streams[4] = b"\n"                                 # lone LF: no event; must not pair with the LF ending stream 2 or 3
    streams[5] = b"\n\n\n\ndata: [DONE]\n\n"           # empty events, then DONE
    streams[6] = b"data: [DONE]\n"                     # truncated: neither an event nor done
    streams[7] = b"x" * 5000 + b"\n\n" + b"y" * 700    # long event, trailing partial event
    streams[8] = b"data: [DONE]\n\n"                   # DONE right at the stream start
    streams[9] = b"prefix data: [DONE]\n\n"            # not at an event start: an event, but not done
    streams[10] = rng.integers(0, 256, 100000, dtype=np.uint8).tobytes()    # binary body
    streams[11] = b"\n" * 5000                          # every byte but the first closes a pair: 512 events per row
    streams[12] = b"a" * 511 + b"\n" + b"\n" + b"b" * 1535 + b"\n" + b"\n"     # pairs that straddle row / row-group ends
    body = np.frombuffer(b"".join(streams), np.uint8)
    off = np.concatenate([[0], np.cumsum([len(s) for s in streams])]).astype(np.uint64)
    csr, ends, done = ctx.sse_scan(body, off)
    ecsr, eends, edone = ob.sse_scan(body, off)
    assert np.array_equal(csr, ecsr) and np.array_equal(ends, eends) and np.array_equal(done, edone)
    assert done[8] == 1 and done[5] == 1 and done[6] == 0 and done[9] == 0 and csr[5] - csr[4] == 0
    assert csr[12] - csr[11] == 4999 and csr[13] - csr[12] == 2
    # events reconstruct: first event of stream 0 parses as JSON
    if csr[1] > csr[0]:
        first = body[int(off[0]):int(ends[int(csr[0])])].tobytes()
        assert first.startswith(b"data: {") and first.endswith(b"\n\n")


def test_sse_scan_after_demux_gather(ctx):
    """frames of interleaved SSE streams -> decode -> demux/gather -> scan: the per-stream event boundaries
    do not depend on how the upstream chunks were cut into frames."""
    from p2p_llm_tunnel_b200.protocol import TunnelMessage, encode_batch
    rng = np.random.default_rng(52)
    bodies = {sid: _sse_stream(rng, int(rng.integers(1, 30))) for sid in range(1, 41)}
    msgs = []
    cursors = {sid: 0 for sid in bodies}
    live = list(bodies)
    while live:
        sid = live[int(rng.integers(0, len(live)))]
        n = int(rng.integers(1, 200))
        chunk = bodies[sid][cursors[sid]:cursors[sid] + n]
        cursors[sid] += n
        msgs.append(TunnelMessage.res_body(sid, chunk))
        if cursors[sid] >= len(bodies[sid]):
            live.remove(sid)
    buf, off = encode_batch(msgs, ctx)
    desc = ctx.frame_decode(buf, off)
    r = ctx.frame_demux(desc, buf)
    csr, ends, done = ctx.sse_scan(r["gathered"], r["gathered_off"])
    assert done.all()
    for s, sid in enumerate(r["uniq_sid"]):
        b = bodies[int(sid)]
        exp = [i + 2 for i in range(len(b) - 1) if b[i:i + 2] == b"\n\n"]
        got = (ends[int(csr[s]):int(csr[s + 1])] - r["gathered_off"][s]).tolist()
        assert got == exp, sid


def test_sse_scan_fuzz_newline_runs(ctx):
    """bodies of nothing but 'a' and '\\n' in random proportions, 700 streams of 0..3000 bytes back to back at every
    alignment: pairs that straddle words, vectors, lanes, rows and row groups, runs of LFs, LFs that end one stream
    and start the next (which must not pair)."""
    rng = np.random.default_rng(53)
    streams = []
    for i in range(700):
        n = int(rng.integers(0, 3000))
        p = float(rng.choice([0.02, 0.2, 0.5, 0.9]))
        streams.append(np.where(rng.random(n) < p, 10, 97).astype(np.uint8).tobytes())
    body = np.frombuffer(b"".join(streams), np.uint8)
    off = np.concatenate([[0], np.cumsum([len(s) for s in streams])]).astype(np.uint64)
    csr, ends, done = ctx.sse_scan(body, off)
    ecsr, eends, edone = ob.sse_scan(body, off)
    assert np.array_equal(csr, ecsr) and np.array_equal(ends, eends) and np.array_equal(done, edone)


def _device_scan(ctx, body, off, onepass, body_bytes=None, cap=None, lead=0):
    """the device-pointer entry points on torch buffers; `lead` shifts the body inside its allocation (byte phase)"""
    import torch
    dev = torch.device("cuda:0")
    k = len(off) - 1
    span = int(off[-1] - off[0])
    d_store = torch.zeros(lead + len(body) + 64, dtype=torch.uint8, device=dev)
    d_store[lead:lead + len(body)] = torch.from_numpy(np.array(body, dtype=np.uint8)).to(dev)
    d_body = d_store[lead:]
    d_off = torch.from_numpy(np.ascontiguousarray(off).view(np.int64)).to(dev)
    cap = (len(body) // 2 + 1) if cap is None else cap
    ev_csr = torch.full((k + 1,), -1, dtype=torch.int64, device=dev); ev_end = torch.full((cap + 1,), -1, dtype=torch.int64, device=dev)
    done = torch.full((max(k, 1),), 7, dtype=torch.uint8, device=dev); tot = torch.full((1,), -1, dtype=torch.int64, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    if onepass:
        bb = span if body_bytes is None else body_bytes
        sb = int(ctx.lib.p2pt_sse_scan_onepass_scratch_bytes(k, bb))
        scr = torch.full((sb,), 0x5A, dtype=torch.uint8, device=dev)          # the call has to clear what it needs itself
        ctx.call("p2pt_sse_scan_onepass_batch_device", k, d_body, d_off, bb, cap, ev_csr, ev_end, done, tot, scr, sb, s)
    else:
        sb = int(ctx.lib.p2pt_sse_scan_scratch_bytes(k)); scr = torch.empty(sb, dtype=torch.uint8, device=dev)
        ctx.call("p2pt_sse_scan_batch_device", k, d_body, d_off, cap, ev_csr, ev_end, done, tot, scr, sb, s)
    torch.cuda.synchronize()
    n = int(tot.item())
    return ev_csr.cpu().numpy().view(np.uint64), ev_end.cpu().numpy().view(np.uint64)[:min(n, cap)], done.cpu().numpy()[:k], n, ev_end.cpu().numpy()


@pytest.mark.parametrize("lead", [0, 1, 7, 15])
def test_one_pass_scan_over_many_tiles(ctx, lead):
    """p2pt_sse_scan_onepass_batch_device against the oracle and against the two-pass entry point: 3 MB of bodies (about 190
    tiles of 16 KiB), streams from empty to several tiles long, runs of line feeds across tile edges, streams that start
    exactly at a tile edge, a first stream that does not start at offset 0, every byte phase of the allocation, a byte
    bound larger than the span, a scratch buffer full of garbage."""
    rng = np.random.default_rng(900 + lead)
    streams = []
    total = 0
    while total < 3_000_000:
        kind = rng.random()
        if kind < 0.15:
            n = 0
        elif kind < 0.5:
            n = int(rng.integers(1, 40))
        elif kind < 0.9:
            n = int(rng.integers(40, 6000))
        else:
            n = int(rng.integers(6000, 70000))
        p = float(rng.choice([0.0, 0.03, 0.3, 0.97]))
        b = np.where(rng.random(n) < p, 10, 97).astype(np.uint8)
        if n > 40 and rng.random() < 0.3:
            at = int(rng.integers(0, n - 20)); b[at:at + 14] = np.frombuffer(b"data: [DONE]\n\n", np.uint8)
            if at >= 2 and rng.random() < 0.7:
                b[at - 2:at] = 10                                      # ... at an event start
        if rng.random() < 0.1 and n >= 14:
            b[:14] = np.frombuffer(b"data: [DONE]\n\n", np.uint8)      # DONE right at the stream start
        streams.append(b.tobytes()); total += n
    pre = 37                                                            # bytes in front of the first stream that belong to nobody
    body = np.frombuffer(b"\n" * pre + b"".join(streams), np.uint8)
    off = (np.concatenate([[0], np.cumsum([len(x) for x in streams])]) + pre).astype(np.uint64)
    # some streams start exactly where a tile starts (tile 0 starts at the 16-byte line of the first stream's first byte)
    ecsr, eends, edone = ob.sse_scan(body, off)
    for bound in (None, int(off[-1] - off[0]) + 100_000):
        csr, ends, done, n, _ = _device_scan(ctx, body, off, True, body_bytes=bound, lead=lead)
        assert n == len(eends) and np.array_equal(csr, ecsr) and np.array_equal(ends, eends) and np.array_equal(done, edone)
    csr2, ends2, done2, n2, _ = _device_scan(ctx, body, off, False, lead=lead)
    assert n2 == n and np.array_equal(csr2, ecsr) and np.array_equal(ends2, eends) and np.array_equal(done2, edone)
    assert edone.sum() > 5 and len(eends) > 10000


def test_one_pass_scan_edges(ctx):
    """no bytes at all; one stream of one byte; only empty streams; thousands of streams inside one tile; an event buffer that is
    too small (the total is still reported, nothing is written behind the buffer)."""
    for streams in ([b""], [b"\n"], [b""] * 50, [b"\n\n"] * 5000, [b"a\n\n", b"", b"\n", b"\n\n\n"]):
        body = np.frombuffer(b"".join(streams) + b"\n\n", np.uint8)     # two bytes behind the last stream: nobody's
        off = np.concatenate([[0], np.cumsum([len(x) for x in streams])]).astype(np.uint64)
        ecsr, eends, edone = ob.sse_scan(body, off)
        csr, ends, done, n, _ = _device_scan(ctx, body, off, True)
        assert n == len(eends) and np.array_equal(csr, ecsr) and np.array_equal(ends, eends) and np.array_equal(done, edone), streams[:3]
    streams = [b"x\n\n" * 3000]
    body = np.frombuffer(streams[0], np.uint8); off = np.array([0, len(body)], np.uint64)
    csr, ends, done, n, raw = _device_scan(ctx, body, off, True, cap=100)
    assert n == 3000 and len(ends) == 100 and ends.tolist() == [3 * (i + 1) for i in range(100)]
    assert raw[100] == -1                                               # the slot behind the buffer is untouched
