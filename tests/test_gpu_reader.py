"""GPU: the batched reader / dispatcher against a per-frame restatement of the reference's loops
(proxy.rs:109-170, serve.rs:82-156) on random frame mixes."""
import json

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

VALID = {1, 2, 3, 4, 10, 11, 12, 20, 21, 22, 99}


def _decode(f):
    if len(f) < 5 or f[0] not in VALID:
        return None
    return f[0], int.from_bytes(f[1:5], "big"), f[5:]


def proxy_loop(frames, registered):
    """proxy.rs:109-170, one frame at a time"""
    pending = {s: [] for s in registered}; closed = {}; pongs = 0; fails = 0
    for f in frames:
        m = _decode(f)
        if m is None:
            fails += 1; continue
        t, sid, p = m
        if t == 20:
            try:
                h = json.loads(p)
            except Exception:
                continue
            if h["stream_id"] in pending:
                pending[h["stream_id"]].append(("headers", h))
        elif t == 21:
            if sid in pending:
                pending[sid].append(("body", p))
        elif t == 22:
            if sid in pending:
                q = pending.pop(sid); q.append(("end", None)); closed[sid] = q
        elif t == 99:
            if sid in pending:
                q = pending.pop(sid); q.append(("error", p.decode("utf-8", "replace"))); closed[sid] = q
        elif t == 3:
            pongs += 1
    return pending, closed, pongs, fails


def serve_loop(frames):
    """serve.rs:103-155, one frame at a time"""
    pending = {}; done = []; pongs = 0
    for f in frames:
        m = _decode(f)
        if m is None:
            continue
        t, sid, p = m
        if t == 10:
            h = json.loads(p); pending[h["stream_id"]] = (h, bytearray())
        elif t == 11:
            if sid in pending:
                pending[sid][1].extend(p)
        elif t == 12:
            if sid in pending:
                h, b = pending.pop(sid); done.append((h, bytes(b)))
        elif t == 3:
            pongs += 1
    return pending, done, pongs


def _frame(t, sid, p=b""):
    return bytes([t]) + int(sid).to_bytes(4, "big") + p


def test_proxy_reader_matches_per_frame_loop(ctx):
    from p2p_llm_tunnel_b200.reader import ProxyReader
    rng = np.random.default_rng(5)
    frames = []
    for sid in range(1, 41):
        seq = [_frame(20, sid, json.dumps({"stream_id": sid, "status": 200, "headers": {}}).encode())]
        seq += [_frame(21, sid, rng.integers(0, 256, int(rng.integers(0, 300)), dtype=np.uint8).tobytes()) for _ in range(int(rng.integers(0, 9)))]
        seq.append(_frame(22, sid) if sid % 5 else _frame(99, sid, b"upstream \xff timeout"))
        if sid % 7 == 0:
            seq.append(_frame(21, sid, b"after the end"))            # dropped: stream already removed
        frames.append(seq)
    # interleave the streams, keeping each stream's own order
    inter = []
    idx = [0] * len(frames)
    while any(i < len(s) for i, s in zip(idx, frames)):
        j = int(rng.integers(0, len(frames)))
        if idx[j] < len(frames[j]):
            inter.append(frames[j][idx[j]]); idx[j] += 1
    inter[10:10] = [_frame(3, 0), bytes([250, 0, 0, 0, 1]), b"\x15\x00", _frame(4, 0), _frame(21, 9999, b"unknown stream"),
                    _frame(20, 3, b"{not json"), _frame(20, 77, json.dumps({"stream_id": 2, "status": 404, "headers": {}}).encode())]
    registered = [s for s in range(1, 41) if s != 13]                # stream 13 was never registered
    rd = ProxyReader(ctx)
    for s in registered:
        rd.register(s)
    cut = len(inter) // 2
    rd.feed(inter[:cut]); rd.feed(inter[cut:])                       # two batches: state carries over
    pending, closed, pongs, fails = proxy_loop(inter, registered)
    assert rd.pongs_to_send == pongs == 1 and rd.decode_failures == fails == 2
    assert sorted(rd.pending) == sorted(pending) and sorted(rd.closed) == sorted(closed)
    for s, q in list(closed.items()) + list(pending.items()):
        got = (rd.closed.get(s) or rd.pending.get(s))
        assert [(e.kind, e.value) for e in got] == q, s


def test_serve_reassembler_matches_per_frame_loop(ctx):
    from p2p_llm_tunnel_b200.protocol import MAX_BODY_CHUNK
    from p2p_llm_tunnel_b200.reader import ServeReassembler
    rng = np.random.default_rng(6)
    streams = []
    for sid in range(1, 25):
        body = rng.integers(0, 256, int(rng.integers(0, 3 * MAX_BODY_CHUNK if sid == 3 else 2000)), dtype=np.uint8).tobytes()
        seq = [_frame(10, sid, json.dumps({"stream_id": sid, "method": "POST", "path": f"/v1/x{sid}", "headers": {}}).encode())]
        seq += [_frame(11, sid, body[o:o + MAX_BODY_CHUNK]) for o in range(0, len(body), MAX_BODY_CHUNK)] or []
        if sid != 9:
            seq.append(_frame(12, sid))                              # stream 9 stays pending
        streams.append(seq)
    inter = []
    idx = [0] * len(streams)
    while any(i < len(s) for i, s in zip(idx, streams)):
        j = int(rng.integers(0, len(streams)))
        if idx[j] < len(streams[j]):
            inter.append(streams[j][idx[j]]); idx[j] += 1
    inter[5:5] = [_frame(11, 555, b"body for an unknown request"), _frame(3, 0), _frame(12, 556), bytes([0, 1, 2])]
    rs = ServeReassembler(ctx)
    cut = len(inter) // 3
    done = rs.feed(inter[:cut]) + rs.feed(inter[cut:2 * cut]) + rs.feed(inter[2 * cut:])
    pending, edone, pongs = serve_loop(inter)
    assert rs.pongs_to_send == pongs == 1
    assert [(h, b) for h, b in done] == edone
    assert sorted(rs.pending_requests) == sorted(pending) == [9]
