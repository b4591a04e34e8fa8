"""Host-side mirror of /root/reference/tunnel/src/protocol.rs for the frame path.

Same names, argument meaning and error behaviour as the reference (`TunnelMessage::encode`,
`TunnelMessage::decode`, `MessageType::from_u8`, the convenience constructors, `MAX_BODY_CHUNK`),
but every encode / decode runs on the GPU through the C ABI (batch of 1 for the single-message
methods, like the FFI shim in INTEGRATION.md would).  JSON header payloads stay opaque bytes.
"""
import enum
import json
from dataclasses import dataclass, field

import numpy as np

from . import _lib

PROTOCOL_VERSION = 1                      # protocol.rs:6
PROTOCOL_NAME = "httptunnel"              # protocol.rs:7
MAX_FRAME_SIZE = 64 * 1024                # protocol.rs:10
MAX_BODY_CHUNK = MAX_FRAME_SIZE - 128     # protocol.rs:12


class MessageType(enum.IntEnum):          # protocol.rs:86-100
    Hello = 1
    Agree = 2
    Ping = 3
    Pong = 4
    ReqHeaders = 10
    ReqBody = 11
    ReqEnd = 12
    ResHeaders = 20
    ResBody = 21
    ResEnd = 22
    Error = 99

    @classmethod
    def from_u8(cls, v):                  # protocol.rs:103-118 -> Option<Self>
        try:
            return cls(v)
        except ValueError:
            return None


_default_ctx = None


def default_context():
    global _default_ctx
    if _default_ctx is None:
        _default_ctx = _lib.Context(0)
    return _default_ctx


@dataclass
class Hello:                              # protocol.rs:16-39 (handshake payload; host-only JSON)
    proto: str = PROTOCOL_NAME
    min_version: int = 1
    max_version: int = PROTOCOL_VERSION
    features: list = field(default_factory=lambda: ["sse"])


@dataclass
class Agree:                              # protocol.rs:24-28, :41-81
    version: int = 1
    features: list = field(default_factory=lambda: ["sse"])

    @staticmethod
    def from_hello(hello):
        if hello.proto != PROTOCOL_NAME:
            raise ValueError(f"unknown protocol: {hello.proto}")
        lo, hi = max(hello.min_version, 1), min(hello.max_version, PROTOCOL_VERSION)
        if lo > hi:
            raise ValueError(f"no compatible version: peer=[{hello.min_version},{hello.max_version}], "
                             f"ours=[1,{PROTOCOL_VERSION}]")
        return Agree(version=hi, features=[f for f in hello.features if f in ("sse",)])


@dataclass
class TunnelMessage:                      # protocol.rs:139-144
    msg_type: MessageType
    stream_id: int
    payload: bytes = b""

    # -- codec ------------------------------------------------------------------------------------
    def encode(self, ctx=None):           # protocol.rs:148-154
        out, _ = encode_batch([self], ctx)
        return out.tobytes()

    @staticmethod
    def decode(data, ctx=None):           # protocol.rs:157-172
        msgs = decode_batch([bytes(data)], ctx)
        m = msgs[0]
        if isinstance(m, Exception):
            raise m
        return m

    # -- constructors (protocol.rs:176-262) ---------------------------------------------------------
    @staticmethod
    def hello(h):
        return TunnelMessage(MessageType.Hello, 0, json.dumps(h.__dict__).encode())

    @staticmethod
    def agree(a):
        return TunnelMessage(MessageType.Agree, 0, json.dumps(a.__dict__).encode())

    @staticmethod
    def req_headers(headers):             # dict with stream_id, method, path, headers
        return TunnelMessage(MessageType.ReqHeaders, headers["stream_id"], json.dumps(headers).encode())

    @staticmethod
    def req_body(stream_id, data):
        return TunnelMessage(MessageType.ReqBody, stream_id, bytes(data))

    @staticmethod
    def req_end(stream_id):
        return TunnelMessage(MessageType.ReqEnd, stream_id, b"")

    @staticmethod
    def res_headers(headers):             # dict with stream_id, status, headers
        return TunnelMessage(MessageType.ResHeaders, headers["stream_id"], json.dumps(headers).encode())

    @staticmethod
    def res_body(stream_id, data):
        return TunnelMessage(MessageType.ResBody, stream_id, bytes(data))

    @staticmethod
    def res_end(stream_id):
        return TunnelMessage(MessageType.ResEnd, stream_id, b"")

    @staticmethod
    def error(stream_id, msg):
        return TunnelMessage(MessageType.Error, stream_id, msg.encode())

    @staticmethod
    def ping():
        return TunnelMessage(MessageType.Ping, 0, b"")

    @staticmethod
    def pong():
        return TunnelMessage(MessageType.Pong, 0, b"")


def pack(items):
    """list of bytes -> (uint8 buffer, uint64 offsets[n+1])"""
    off = np.zeros(len(items) + 1, np.uint64)
    if items:
        np.cumsum([len(b) for b in items], out=off[1:])
    buf = np.frombuffer(b"".join(items), np.uint8) if items else np.zeros(0, np.uint8)
    return buf, off


def encode_batch(msgs, ctx=None):
    """Batched TunnelMessage::encode: returns (frames buffer, offsets)."""
    ctx = ctx or default_context()
    types = np.array([int(m.msg_type) for m in msgs], np.uint8)
    sids = np.array([m.stream_id for m in msgs], np.uint32)
    payload, pay_off = pack([bytes(m.payload) for m in msgs])
    return ctx.frame_encode(types, sids, payload, pay_off)


def decode_batch(frames, ctx=None):
    """Batched TunnelMessage::decode over a list of byte strings.  Undecodable frames come back as
    ValueError instances carrying the reference's message text (the caller skips them, as
    serve.rs:103-109 / proxy.rs:110-116 do)."""
    ctx = ctx or default_context()
    buf, off = pack([bytes(f) for f in frames])
    desc = ctx.frame_decode(buf, off)
    out = []
    raw = buf.tobytes()
    for i, d in enumerate(desc):
        if d["status"] == _lib.ST_TOO_SHORT:
            out.append(ValueError(f"message too short: {len(frames[i])} bytes"))
        elif d["status"] == _lib.ST_UNKNOWN_TYPE:
            out.append(ValueError(f"unknown message type: {int(d['msg_type'])}"))
        else:
            o, n = int(d["pay_off"]), int(d["pay_len"])
            out.append(TunnelMessage(MessageType(int(d["msg_type"])), int(d["stream_id"]), raw[o:o + n]))
    return out


def demux_batch(frames, ctx=None):
    """decode + demux: {stream_id: (list of TunnelMessage in arrival order, concatenated body)} --
    the batched form of the reader loops proxy.rs:109-170 / serve.rs:82-156."""
    ctx = ctx or default_context()
    buf, off = pack([bytes(f) for f in frames])
    desc = ctx.frame_decode(buf, off)
    r = ctx.frame_demux(desc, buf)
    raw = buf.tobytes()
    res = {}
    for s in range(r["k"]):
        idx = r["perm"][int(r["csr"][s]):int(r["csr"][s + 1])]
        msgs = []
        for i in idx:
            d = desc[i]
            o, n = int(d["pay_off"]), int(d["pay_len"])
            msgs.append(TunnelMessage(MessageType(int(d["msg_type"])), int(d["stream_id"]), raw[o:o + n]))
        body = r["gathered"][int(r["gathered_off"][s]):int(r["gathered_off"][s + 1])].tobytes()
        res[int(r["uniq_sid"][s])] = (msgs, body)
    return res


def chunk_body(stream_id, data, response=True):
    """Sub-chunk a body at MAX_BODY_CHUNK into RES_BODY / REQ_BODY messages
    (serve.rs:269-276, proxy.rs:318-330)."""
    mk = TunnelMessage.res_body if response else TunnelMessage.req_body
    return [mk(stream_id, data[o:o + MAX_BODY_CHUNK]) for o in range(0, len(data), MAX_BODY_CHUNK)]
