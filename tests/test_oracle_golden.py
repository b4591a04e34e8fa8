"""CPU: the oracle against every golden vector (reference literals, FIPS-197, GCM spec, DTLS records)."""
import ctypes as C
import json
import os

import numpy as np

import oracle_binding as ob

G = os.path.join(os.path.dirname(__file__), "golden")
h = bytes.fromhex


def load(name):
    with open(os.path.join(G, name)) as f:
        return json.load(f)


def _payload(v):
    if v["payload_hex"] is not None:
        return h(v["payload_hex"])
    b, n = v["payload_fill"]
    return bytes([b]) * n


def test_frame_encode_decode_reference_vectors():
    k = load("frames_kat.json")
    for v in k["ok"]:
        p = _payload(v)
        out, off = ob.frame_encode([v["type"]], [v["stream_id"]], np.frombuffer(p, np.uint8), [0, len(p)])
        assert len(out) == v["frame_len"] == 5 + len(p), v["name"]
        assert out[:32].tobytes().hex() == v["frame_head_hex"], v["name"]
        # field order and big-endian stream id, byte by byte
        assert out[0] == v["type"] and int.from_bytes(out[1:5].tobytes(), "big") == v["stream_id"]
        d = ob.frame_decode(out, off)[0]
        assert d["status"] == 0 and d["msg_type"] == v["type"] and d["stream_id"] == v["stream_id"]
        assert d["pay_off"] == 5 and d["pay_len"] == len(p)
        assert out[5:].tobytes() == p


def test_frame_decode_negative_vectors():
    k = load("frames_kat.json")
    for v in k["bad"]:
        f = h(v["frame_hex"])
        d = ob.frame_decode(np.frombuffer(f, np.uint8), [0, len(f)])[0]
        assert d["status"] == v["status"], v["name"]


def test_valid_tag_set():
    k = load("frames_kat.json")
    valid = set(k["valid_tags"])
    for b in range(256):
        assert bool(ob.lib().orc_msg_type_valid(C.c_uint8(b))) == (b in valid)


def test_aes128_fips197():
    for v in load("gcm_nist_kat.json")["aes128"]:
        rk = C.create_string_buffer(176); out = C.create_string_buffer(16)
        ob.lib().orc_aes128_expand(h(v["key"]), rk)
        ob.lib().orc_aes128_encrypt_block(rk, h(v["pt"]), out)
        assert out.raw.hex() == v["ct"], v["name"]


def test_gcm_spec_cases():
    for v in load("gcm_nist_kat.json")["gcm"]:
        pt, aad = h(v["pt"]), h(v["aad"])
        ct = C.create_string_buffer(max(len(pt), 1)); tag = C.create_string_buffer(16)
        ob.lib().orc_gcm_encrypt(h(v["key"]), h(v["iv"]), aad, C.c_size_t(len(aad)), pt, C.c_size_t(len(pt)), ct, tag)
        assert ct.raw[:len(pt)].hex() == v["ct"] and tag.raw.hex() == v["tag"], v["name"]
        back = C.create_string_buffer(max(len(pt), 1))
        rc = ob.lib().orc_gcm_decrypt(h(v["key"]), h(v["iv"]), aad, C.c_size_t(len(aad)), ct.raw[:len(pt)],
                                      C.c_size_t(len(pt)), tag.raw, back)
        assert rc == 0 and back.raw[:len(pt)] == pt
        bad = bytearray(tag.raw); bad[3] ^= 1
        assert ob.lib().orc_gcm_decrypt(h(v["key"]), h(v["iv"]), aad, C.c_size_t(len(aad)), ct.raw[:len(pt)],
                                        C.c_size_t(len(pt)), bytes(bad), back) != 0


def test_dtls_records_against_cryptography_fixtures():
    for v in load("dtls_records_kat.json")["records"]:
        raw, rec = h(v["raw"]), h(v["rec"])
        got, _, st = ob.dtls_seal([0], np.frombuffer(h(v["key"]), np.uint8), np.frombuffer(h(v["iv"]), np.uint8),
                                  np.frombuffer(h(v["nonce"]), np.uint8), np.frombuffer(raw, np.uint8), [0, len(raw)])
        assert st[0] == 0 and got.tobytes() == rec, v["name"]
        out, _, olen, st = ob.dtls_open([0], np.frombuffer(h(v["key"]), np.uint8), np.frombuffer(h(v["iv"]), np.uint8),
                                        np.frombuffer(rec, np.uint8), [0, len(rec)])
        assert st[0] == 0 and olen[0] == len(raw), v["name"]
        # header is NOT rewritten on open: it still carries the sealed length
        assert out[:int(olen[0])].tobytes() == rec[:13] + raw[13:], v["name"]


def test_dtls_open_error_paths():
    v = load("dtls_records_kat.json")["records"][0]
    key, iv = np.frombuffer(h(v["key"]), np.uint8), np.frombuffer(h(v["iv"]), np.uint8)
    rec = bytearray(h(v["rec"]))
    def st_of(r):
        return int(ob.dtls_open([0], key, iv, np.frombuffer(bytes(r), np.uint8), [0, len(r)])[3][0])
    assert st_of(rec[:12]) == 6                      # truncated header
    assert st_of(rec[:21]) == 4                      # no room for nonce
    assert st_of(rec[:30]) == 3                      # shorter than a tag
    t = bytearray(rec); t[-1] ^= 1
    assert st_of(t) == 3                             # bad tag
    t = bytearray(rec); t[25] ^= 0x80
    assert st_of(t) == 3                             # bad ciphertext
    t = bytearray(rec); t[5] ^= 1
    assert st_of(t) == 3                             # sequence number is authenticated
    t = bytearray(rec); t[2] = 0xFC
    assert st_of(t) == 6                             # unsupported version
    ccs = bytes([20, 0xFE, 0xFD, 0, 0, 0, 0, 0, 0, 0, 5, 0, 1, 1])
    out, _, olen, st = ob.dtls_open([0], key, iv, np.frombuffer(ccs, np.uint8), [0, len(ccs)])
    assert st[0] == 5 and out[:int(olen[0])].tobytes() == ccs   # ChangeCipherSpec passes through


def test_demux_semantics_small():
    # streams 7 and 3 interleaved, one corrupt frame, one ping: stable order inside a stream
    frames = [bytes([21, 0, 0, 0, 7]) + b"aa", bytes([21, 0, 0, 0, 3]) + b"b", bytes([255, 0, 0, 0, 0]),
              bytes([20, 0, 0, 0, 7]) + b"{}", bytes([21, 0, 0, 0, 7]) + b"cc", bytes([3, 0, 0, 0, 0]),
              bytes([22, 0, 0, 0, 3]), bytes([1, 2])]
    buf = np.frombuffer(b"".join(frames), np.uint8)
    off = np.concatenate([[0], np.cumsum([len(f) for f in frames])]).astype(np.uint64)
    d = ob.frame_decode(buf, off)
    r = ob.frame_demux(d, buf)
    assert list(r["uniq_sid"]) == [0, 3, 7]
    assert list(r["csr"]) == [0, 1, 3, 6]
    assert list(r["perm"]) == [5, 1, 6, 0, 3, 4]
    assert list(r["gathered_off"]) == [0, 0, 1, 5]
    assert r["gathered"].tobytes() == b"baacc"


def test_demux_session_namespaces():
    frames = [bytes([21, 0, 0, 0, 1]) + b"a", bytes([21, 0, 0, 0, 1]) + b"b", bytes([21, 0, 0, 0, 1]) + b"c"]
    buf = np.frombuffer(b"".join(frames), np.uint8)
    off = np.array([0, 6, 12, 18], np.uint64)
    d = ob.frame_decode(buf, off)
    r = ob.frame_demux(d, buf, np.array([1, 0, 1], np.uint32))
    assert list(r["uniq_session"]) == [0, 1] and list(r["uniq_sid"]) == [1, 1]
    assert list(r["perm"]) == [1, 0, 2] and r["gathered"].tobytes() == b"bac"


def test_oracle_demux_against_numpy_randomised():
    """The demux oracle itself, checked against an independent numpy statement of the same semantics (stable
    order by (session, stream id) of the decodable frames, per-stream concatenation of body payloads,
    serve.rs:120-124): the GPU parity tests lean on it."""
    from p2p_llm_tunnel_b200._lib import DESC_DTYPE
    rng = np.random.default_rng(0x5EED000B)
    for case in range(12):
        n = int(rng.integers(1, 3000))
        sid = rng.integers(0, [5, 300, 2**32][case % 3], n, dtype=np.uint64).astype(np.uint32)
        session = rng.integers(0, 4, n).astype(np.uint32) if case % 2 else None
        lens = rng.integers(0, 60, n).astype(np.uint32)
        pay_off = np.cumsum(np.concatenate([[2], (lens + rng.integers(0, 5, n))[:-1]])).astype(np.uint64)
        buf = rng.integers(0, 256, int(pay_off[-1] + lens[-1] + 8), dtype=np.uint8)
        desc = np.zeros(n, DESC_DTYPE)
        desc["pay_off"] = pay_off; desc["pay_len"] = lens; desc["stream_id"] = sid
        desc["msg_type"] = rng.choice(np.array([11, 21, 21, 20, 22, 99], np.uint8), n)
        desc["status"] = np.where(rng.random(n) < 0.05, 1, 0).astype(np.uint8)
        r = ob.frame_demux(desc, buf, session=session)
        ok = np.flatnonzero(desc["status"] == 0)
        key = sid[ok].astype(np.uint64)
        if session is not None:
            key |= session[ok].astype(np.uint64) << np.uint64(32)
        order = ok[np.argsort(key, kind="stable")]
        assert np.array_equal(np.asarray(r["perm"], np.int64)[: len(order)], order), case
        body = np.where(np.isin(desc["msg_type"][order], (11, 21)), lens[order], 0).astype(np.int64)
        excl = np.cumsum(body) - body
        src = np.repeat(pay_off[order].astype(np.int64) - excl, body) + np.arange(int(body.sum()), dtype=np.int64)
        assert np.array_equal(np.asarray(r["gathered"]), buf[src]), case
