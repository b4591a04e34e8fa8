"""p2p-llm-tunnel_b200 -- B200-native batched frame codec + DTLS 1.2 AES-128-GCM record path.

Host-side mirror of the reference's interface for this path (``protocol``: TunnelMessage /
MessageType, /root/reference/tunnel/src/protocol.rs:83-263; ``dtls``: CryptoGcm, webrtc-dtls 0.10.0)
over the C ABI in include/p2pt.h.  There is no CPU fallback: importing works anywhere (so that the
CPU test-suite can check symbols), but every compute call needs the CUDA library and a B200.
"""
from . import _lib  # noqa: F401
from ._lib import Context, P2ptError, lib_path, load_library  # noqa: F401
from .protocol import (  # noqa: F401
    MAX_BODY_CHUNK, MAX_FRAME_SIZE, MessageType, TunnelMessage, decode_batch, encode_batch,
    demux_batch, chunk_body,
)
from .dtls import CryptoGcm  # noqa: F401
from .reader import ProxyReader, ServeReassembler, StreamEvent  # noqa: F401

__all__ = [
    "Context", "P2ptError", "lib_path", "load_library", "MessageType", "TunnelMessage",
    "MAX_BODY_CHUNK", "MAX_FRAME_SIZE", "encode_batch", "decode_batch", "demux_batch", "chunk_body",
    "CryptoGcm", "ProxyReader", "ServeReassembler", "StreamEvent",
]
