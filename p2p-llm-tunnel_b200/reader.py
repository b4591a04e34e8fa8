"""Batched reader / dispatcher: the reference's per-frame reader loops, consuming the GPU demux output.

  ProxyReader      /root/reference/tunnel/src/proxy.rs:105-172  (PendingResponses :23-24, StreamEvent :28-37)
  ServeReassembler /root/reference/tunnel/src/serve.rs:82-156   (pending_requests :65)

The reference decodes one frame, takes a mutex, probes a HashMap and pushes one event, per frame.
Here a whole batch is decoded and demultiplexed on the GPU (p2pt_frame_decode_demux_batch) and the host walks each stream's frames once, in arrival order -- the same
per-stream event sequences come out.  Frames whose JSON stream id differs from their frame header
(headers frames are routed by the JSON field, proxy.rs:129 / serve.rs:113-118) are merged back by
arrival index so that even that corner behaves like the sequential loop.
"""
import json
from dataclasses import dataclass

import numpy as np

from . import _lib
from .protocol import MessageType, pack


@dataclass
class StreamEvent:                      # proxy.rs:28-37
    kind: str                           # "headers" | "body" | "end" | "error"
    value: object = None


def _decode_demux(ctx, frames):
    buf, off = pack([bytes(f) for f in frames])
    r = ctx.frame_decode_demux(buf, off, max_streams=max(len(frames), 1), gathered_cap=max(len(buf), 1), want_desc=True)
    return buf, r["desc"], r


def _per_target(desc, r, raw, header_type):
    """-> {target stream id: [(arrival index, msg_type, payload bytes)]} in arrival order.
    Body payloads are slices of the per-stream gathered buffer (no per-frame copies)."""
    out = {}
    gathered = r["gathered"]
    for s in range(r["k"]):
        sid = int(r["uniq_sid"][s])
        cursor = int(r["gathered_off"][s])
        for i in r["perm"][int(r["csr"][s]):int(r["csr"][s + 1])]:
            d = desc[i]
            t, n = int(d["msg_type"]), int(d["pay_len"])
            if t in (11, 21):
                payload = gathered[cursor:cursor + n]
                cursor += n
            else:
                o = int(d["pay_off"])
                payload = raw[o:o + n]
            target = sid
            if t == header_type:
                try:
                    target = int(json.loads(bytes(payload))["stream_id"])
                except Exception:
                    target = None                      # unparsable headers: error! + continue / `?` in the reference
            out.setdefault(target, []).append((int(i), t, payload))
    for k in out:
        out[k].sort(key=lambda e: e[0])
    return out


class ProxyReader:
    """Consumer side: routes RES_* frames to per-stream event queues (proxy.rs:105-172)."""

    def __init__(self, ctx=None):
        self.ctx = ctx or _lib.Context(0)
        self.pending = {}               # stream_id -> list[StreamEvent]   (the mpsc::UnboundedSender)
        self.closed = {}                # finished streams' queues, for the caller to drain
        self.pongs_to_send = 0
        self.decode_failures = 0

    def register(self, stream_id):      # proxy.rs:266-270: pending.insert(stream_id, event_tx)
        self.pending[stream_id] = []
        return self.pending[stream_id]

    def feed(self, frames):
        if not frames:
            return
        buf, desc, r = _decode_demux(self.ctx, frames)
        self.decode_failures += int((desc["status"] != 0).sum())        # warn! + continue (proxy.rs:110-116)
        raw = buf
        per = _per_target(desc, r, raw, int(MessageType.ResHeaders))
        for sid, events in per.items():
            for _, t, payload in events:
                if t == MessageType.Ping:
                    self.pongs_to_send += 1                                  # proxy.rs:154-162
                    continue
                if t == MessageType.ResHeaders:
                    if sid is None:
                        continue                                             # failed to parse response headers
                    q = self.pending.get(sid)
                    if q is not None:
                        q.append(StreamEvent("headers", json.loads(bytes(payload))))
                elif t == MessageType.ResBody:
                    q = self.pending.get(sid)
                    if q is not None:                                        # unknown stream: silently dropped (:135)
                        q.append(StreamEvent("body", bytes(payload)))
                elif t == MessageType.ResEnd:
                    q = self.pending.pop(sid, None)                          # remove + send End (:139-145)
                    if q is not None:
                        q.append(StreamEvent("end"))
                        self.closed[sid] = q
                elif t == MessageType.Error:
                    q = self.pending.pop(sid, None)                          # (:146-153), lossy UTF-8
                    if q is not None:
                        q.append(StreamEvent("error", bytes(payload).decode("utf-8", "replace")))
                        self.closed[sid] = q
                # Pong and everything else: ignored (:163-169)


class ServeReassembler:
    """Provider side: REQ_HEADERS / REQ_BODY / REQ_END -> completed requests (serve.rs:82-156)."""

    def __init__(self, ctx=None):
        self.ctx = ctx or _lib.Context(0)
        self.pending_requests = {}      # stream_id -> (headers dict, [body chunks])
        self.pongs_to_send = 0
        self.decode_failures = 0

    def feed(self, frames):
        """returns the requests completed by this batch: [(headers, body bytes)], in REQ_END arrival order"""
        if not frames:
            return []
        buf, desc, r = _decode_demux(self.ctx, frames)
        self.decode_failures += int((desc["status"] != 0).sum())        # serve.rs:103-109
        per = _per_target(desc, r, buf, int(MessageType.ReqHeaders))
        done = []
        # a stream's frames interact only with that stream's map entry, so streams can be walked one
        # at a time; completed requests are put back into arrival order of their REQ_END at the end
        for sid, events in per.items():
            for idx, t, payload in events:
                if t == MessageType.Ping:
                    self.pongs_to_send += 1
                elif t == MessageType.ReqHeaders:
                    if sid is None:
                        raise ValueError("invalid request headers JSON")     # serde `?` ends run_serve (serve.rs:113)
                    self.pending_requests[sid] = (json.loads(bytes(payload)), [])      # :112-119 insert
                elif t == MessageType.ReqBody:
                    e = self.pending_requests.get(sid)
                    if e is not None:                                        # :120-124 extend_from_slice
                        e[1].append(payload)
                elif t == MessageType.ReqEnd:
                    e = self.pending_requests.pop(sid, None)                 # :125-139 remove + spawn
                    if e is not None:
                        body = np.concatenate(e[1]).tobytes() if e[1] else b""
                        done.append((idx, e[0], body))
        done.sort(key=lambda x: x[0])
        return [(h, b) for _, h, b in done]
