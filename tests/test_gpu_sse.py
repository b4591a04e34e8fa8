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
