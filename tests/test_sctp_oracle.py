"""CPU: the SCTP / CRC32c part of the oracle against RFC 3720 B.4 vectors and its own structure."""
import json
import os

import numpy as np

import oracle_binding as ob

G = os.path.join(os.path.dirname(__file__), "golden")


def test_crc32c_rfc3720_vectors():
    for v in json.load(open(os.path.join(G, "crc32c_kat.json")))["vectors"]:
        assert ob.crc32c(bytes.fromhex(v["data"])) == int(v["crc"], 16), v["name"]


def test_packet_layout_and_roundtrip():
    msgs = [b"", b"x", bytes(range(200)) * 6, bytes(1199), bytes([7]) * 1200, bytes([9]) * 1201, os.urandom(5000), os.urandom(65413)]
    pk, off = ob.sctp_packetize(msgs, 5000, 5001, 0xDEADBEEF, 1, np.arange(len(msgs)), 53, 1000 + 100 * np.arange(len(msgs)))
    info, po, pl = ob.sctp_unpack(pk, off)
    assert not info["status"].any()
    i = 0
    for mi, m in enumerate(msgs):
        k = max(1, -(-len(m) // 1200))
        got = b""
        for f in range(k):
            a, b = int(off[i]), int(off[i + 1])
            p = pk[a:b].tobytes()
            assert p[:2] == (5000).to_bytes(2, "big") and p[2:4] == (5001).to_bytes(2, "big") and p[4:8] == bytes.fromhex("deadbeef")
            assert p[12] == 0 and (b - a) % 4 == 0
            assert info["flags"][i] == (2 if f == 0 else 0) | (1 if f == k - 1 else 0)
            assert info["tsn"][i] == 1000 + 100 * mi + f and info["ssn"][i] == mi and info["stream_id"][i] == 1 and info["ppid"][i] == 53
            assert int.from_bytes(p[14:16], "big") == 16 + pl[i]
            assert p[28 + pl[i]:] == bytes(len(p) - 28 - pl[i])                       # zero padding
            zeroed = p[:8] + bytes(4) + p[12:]
            assert int.from_bytes(p[8:12], "little") == ob.crc32c(zeroed)            # checksum stored little-endian
            got += pk[int(po[i]):int(po[i]) + int(pl[i])].tobytes()
            i += 1
        assert got == m
    assert i == len(off) - 1


def test_unpack_error_statuses_and_dtls_headroom():
    pk, off = ob.sctp_packetize([b"hello sctp"], 1, 2, 3, 4, 5, 53, 77, epoch=1, seq48_first=9)
    assert pk[0] == 23 and pk[1] == 0xFE and pk[2] == 0xFD and int.from_bytes(pk[5:11].tobytes(), "big") == 9
    assert int.from_bytes(pk[11:13].tobytes(), "big") == len(pk) - 13
    info, po, pl = ob.sctp_unpack(pk, off, headroom=13)
    assert info["status"][0] == 0 and pk[int(po[0]):int(po[0]) + int(pl[0])].tobytes() == b"hello sctp"
    bad = pk.copy(); bad[13 + 30] ^= 1
    assert ob.sctp_unpack(bad, off, headroom=13)[0]["status"][0] == 7
    assert ob.sctp_unpack(pk[:20], np.array([0, 20], np.uint64), headroom=13)[0]["status"][0] == 8


def test_replay_window_oracle_semantics():
    """SlidingWindowDetector (window 64): duplicates and records 64 or more behind the newest are dropped;
    only authenticated records move the window."""
    st = {"latest": np.zeros(1, np.uint64), "mask": np.zeros(1, np.uint64), "started": np.zeros(1, np.uint8)}
    seqs = [0, 1, 1, 3, 2, 2, 100, 37, 36, 100, 99, 163, 100, 99, 164]
    auth = [1, 1, 1, 1, 1, 1, 1,   1,  1,  1,   1,  0,   1,   1,  1]
    got = ob.replay_filter(np.zeros(len(seqs), np.uint32), np.array(seqs, np.uint64), st, np.array(auth, np.uint8))
    #        0  1  dup 3  2  dup 100 37(63 behind: ok) 36(64 behind: old) dup 99 163(not authentic) dup dup 164
    assert list(got) == [1, 1, 0, 1, 1, 0, 1, 1, 0, 0, 1, 0, 0, 0, 1]
    assert st["latest"][0] == 164


def test_datagram_split_oracle():
    r1 = bytes([23, 0xFE, 0xFD]) + bytes(8) + (5).to_bytes(2, "big") + b"hello"
    r0 = bytes([22, 0xFE, 0xFD]) + bytes(8) + (0).to_bytes(2, "big")          # header only, zero-length record
    # [r1 r0 r1] is fine; [r1 r1-truncated] and [r1 r0] (a bare header at the very end: unpack_datagram's
    # `remaining <= 13` test) are dropped whole
    dgrams = [r1 + r0 + r1, r1 + r1[:-1], r1 + r0]
    buf = np.frombuffer(b"".join(dgrams), np.uint8)
    off = np.concatenate([[0], np.cumsum([len(d) for d in dgrams])]).astype(np.uint64)
    ro, rd, st = ob.dtls_datagram_split(buf, off)
    assert list(st) == [0, 6, 6] and list(rd) == [0, 0, 0]
    assert list(ro) == [0, len(r1), len(r1) + len(r0), len(buf)]


def test_sse_scan_oracle():
    streams = [b"data: a\n\ndata: b\n\n", b"", b"\n", b"data: [DONE]\n\n", b"x\n\n\ndata: [DONE]\n\nrest"]
    body = np.frombuffer(b"".join(streams), np.uint8)
    off = np.concatenate([[0], np.cumsum([len(s) for s in streams])]).astype(np.uint64)
    csr, ends, done = ob.sse_scan(body, off)
    assert list(csr) == [0, 2, 2, 2, 3, 6] and list(done) == [0, 0, 0, 1, 1]
    assert list(ends[:2]) == [9, 18]
