"""CPU: the oracle's DTLS record layout against wire records captured from OpenSSL's DTLS 1.2 stack
(tests/golden/dtls_openssl_kat.json, made by tests/golden/gen_dtls_openssl.py).

This is the pin for rows a15-a17 and (f)3: webrtc-dtls 0.10.0 (/root/reference/tunnel/Cargo.lock:2822-2825,
reached at /root/reference/tunnel/src/serve.rs:274 and rtc.rs:88-92) is un-vendored, but it has to interoperate
with exactly these records (TLS_ECDHE_ECDSA_WITH_AES_128_GCM_SHA256).  The keys in the fixture come from the
key log + the TLS 1.2 PRF, not from this repository's code."""
import json
import os

import numpy as np

import oracle_binding as ob

G = os.path.join(os.path.dirname(__file__), "golden")
h = bytes.fromhex


def fixture():
    with open(os.path.join(G, "dtls_openssl_kat.json")) as f:
        return json.load(f)


def write_keys(fx, direction):
    side = "client" if direction == "c2s" else "server"
    return (np.frombuffer(h(fx["keys"][side + "_write_key"]), np.uint8),
            np.frombuffer(h(fx["keys"][side + "_write_iv"]), np.uint8))


def raw_of(rec, plain):
    """what CryptoGcm::encrypt is handed: the 13-byte header with the PLAINTEXT length, then the plaintext"""
    return rec[:11] + len(plain).to_bytes(2, "big") + plain


def test_fixture_shape():
    fx = fixture()
    assert fx["suite"] == "TLS_ECDHE_ECDSA_WITH_AES_128_GCM_SHA256"
    recs = fx["records"]
    assert {r["dir"] for r in recs} == {"c2s", "s2c"}
    lens = {len(h(r["plain"])) for r in recs if r["type"] == 23}
    assert {1, 15, 16, 17, 1200, 1205, 1228} <= lens              # block-edge sizes and the full-size record
    assert {r["type"] for r in recs} == {22, 23}                   # Finished (handshake under GCM) + application data
    assert any(d["n_records"] > 1 for d in fx["datagrams"])


def test_oracle_opens_every_captured_record():
    fx = fixture()
    for r in fx["records"]:
        key, iv = write_keys(fx, r["dir"])
        rec, plain = h(r["rec"]), h(r["plain"])
        out, _, olen, st = ob.dtls_open([0], key, iv, np.frombuffer(rec, np.uint8), [0, len(rec)])
        assert st[0] == 0, (r["dir"], r["seq"])
        assert int(olen[0]) == 13 + len(plain)
        assert out[:13].tobytes() == rec[:13]                      # header left as it came (length not rewritten)
        assert out[13:int(olen[0])].tobytes() == plain, (r["dir"], r["seq"])


def test_oracle_reseals_to_the_captured_bytes():
    fx = fixture()
    for r in fx["records"]:
        key, iv = write_keys(fx, r["dir"])
        rec, plain = h(r["rec"]), h(r["plain"])
        raw = raw_of(rec, plain)
        got, _, st = ob.dtls_seal([0], key, iv, np.frombuffer(rec[13:21], np.uint8), np.frombuffer(raw, np.uint8), [0, len(raw)])
        assert st[0] == 0 and got.tobytes() == rec, (r["dir"], r["seq"])


def test_wrong_direction_key_or_touched_header_fails():
    fx = fixture()
    r = next(x for x in fx["records"] if x["type"] == 23 and len(h(x["plain"])) == 1205)
    rec = h(r["rec"])
    other = write_keys(fx, "s2c" if r["dir"] == "c2s" else "c2s")
    _, _, _, st = ob.dtls_open([0], other[0], other[1], np.frombuffer(rec, np.uint8), [0, len(rec)])
    assert st[0] == 3
    key, iv = write_keys(fx, r["dir"])
    # type, version, epoch, seq, nonce, ciphertext, tag.  Not the length field: CryptoGcm::decrypt takes the AAD
    # length from the slice it was handed, and so does the oracle (the batch offsets, not bytes 11..13).
    for pos in (0, 2, 4, 10, 13, 20, 21, len(rec) - 1):
        t = bytearray(rec); t[pos] ^= 1
        _, _, _, st = ob.dtls_open([0], key, iv, np.frombuffer(bytes(t), np.uint8), [0, len(t)])
        assert st[0] in (3, 6), pos


def test_datagram_split_on_captured_flights():
    fx = fixture()
    for d in fx["datagrams"]:
        g = h(d["bytes"])
        off, dg, st = ob.dtls_datagram_split(np.frombuffer(g, np.uint8), [0, len(g)])
        assert st[0] == 0 and len(dg) == d["n_records"]
        # every piece is a well-formed record: version fe fd / fe ff, length field = piece length - 13
        for j in range(len(dg)):
            p = g[int(off[j]):int(off[j + 1])]
            assert p[1] == 0xFE and p[2] in (0xFD, 0xFF) and int.from_bytes(p[11:13], "big") == len(p) - 13


def test_captured_flight_through_split_then_open():
    """the server's last handshake flight: NewSessionTicket?, ChangeCipherSpec (epoch 0, passes through), Finished
    (epoch 1, sealed): split, then open what is protected"""
    fx = fixture()
    seen_ccs = seen_fin = False
    for d in fx["datagrams"]:
        g = h(d["bytes"])
        off, dg, _ = ob.dtls_datagram_split(np.frombuffer(g, np.uint8), [0, len(g)])
        key, iv = write_keys(fx, d["dir"])
        for j in range(len(dg)):
            p = g[int(off[j]):int(off[j + 1])]
            epoch = int.from_bytes(p[3:5], "big")
            if p[0] == 20:
                out, _, olen, st = ob.dtls_open([0], key, iv, np.frombuffer(p, np.uint8), [0, len(p)])
                assert st[0] == 5 and out[:int(olen[0])].tobytes() == p
                seen_ccs = True
            elif epoch >= 1:
                out, _, olen, st = ob.dtls_open([0], key, iv, np.frombuffer(p, np.uint8), [0, len(p)])
                assert st[0] == 0
                if p[0] == 22:
                    assert out[13] == 20 and int(olen[0]) == 13 + 12 + 12   # Finished: handshake header + verify_data
                    seen_fin = True
    assert seen_ccs and seen_fin


def test_application_records_carry_the_tunnel_frames():
    """records -> open -> TunnelMessage::decode (protocol.rs:157-172): what proxy.rs:109-138 sees"""
    fx = fixture()
    key, iv = write_keys(fx, "s2c")
    recs = [h(r["rec"]) for r in fx["records"] if r["dir"] == "s2c" and r["type"] == 23]
    buf = np.frombuffer(b"".join(recs), np.uint8)
    off = np.concatenate([[0], np.cumsum([len(r) for r in recs])]).astype(np.uint64)
    out, _, olen, st = ob.dtls_open(np.zeros(len(recs), np.uint32), key, iv, buf, off)
    assert (st == 0).all()
    frames = [out[int(off[i]) + 13:int(off[i]) + int(olen[i])].tobytes() for i in range(len(recs))]
    fbuf = np.frombuffer(b"".join(frames), np.uint8)
    foff = np.concatenate([[0], np.cumsum([len(f) for f in frames])]).astype(np.uint64)
    d = ob.frame_decode(fbuf, foff)
    got = [(int(x["msg_type"]), int(x["stream_id"]), int(x["status"])) for x in d]
    assert got[0] == (20, 42, 0) and got[1] == (21, 99, 0) and got[2] == (21, 42, 0)
    assert got[5] == (22, 42, 0) and got[6] == (3, 0, 0)
    assert frames[1][5:] == b'data: {"token": "hello"}\n\n'


# ---- the anti-replay window, against what a live OpenSSL DTLS 1.2 server accepted (tests/golden/replay_openssl_kat.json,
#      made by tests/golden/gen_replay_openssl.py: held records delivered with duplicates, reordering, jumps of more than
#      64, records 63 and 64 behind the newest, forged copies) ----

def replay_fixture():
    with open(os.path.join(G, "replay_openssl_kat.json")) as f:
        fx = json.load(f)
    a = fx["arrivals"]
    return (fx, np.array([x["seq"] for x in a], np.uint64), np.array([x["authentic"] for x in a], np.uint8),
            np.array([x["accepted"] for x in a], np.uint8))


def fresh_window(sessions=1):
    return {"latest": np.zeros(sessions, np.uint64), "mask": np.zeros(sessions, np.uint64), "started": np.zeros(sessions, np.uint8)}


def test_replay_fixture_shape():
    fx, seq, auth, acc = replay_fixture()
    assert fx["window"] == 64 and len(seq) > 100
    assert 0 < acc.sum() < len(acc) and (auth == 0).sum() >= 2
    assert not acc[auth == 0].any()                                  # nothing forged was delivered
    newest = np.maximum.accumulate(np.where(acc == 1, seq, 0))
    behind = newest.astype(np.int64) - seq.astype(np.int64)
    assert (behind[acc == 1] <= 63).all()                            # nothing 64 or more behind the newest got through ...
    assert 63 in set(behind[acc == 1].tolist())                      # ... the edge itself did ...
    assert 64 in set(behind[(acc == 0) & (auth == 1)].tolist())      # ... and one step beyond it did not


def test_oracle_window_decides_like_openssl():
    fx, seq, auth, acc = replay_fixture()
    st = fresh_window()
    assert list(ob.replay_filter(np.zeros(1, np.uint32), np.zeros(1, np.uint64), st)) == [1]     # seq 0: the client's Finished
    got = ob.replay_filter(np.zeros(len(seq), np.uint32), seq, st, auth)
    assert np.array_equal(got, acc), np.nonzero(got != acc)[0]
    # the same arrivals one record per call: the state carried between calls is all there is
    st = fresh_window()
    ob.replay_filter(np.zeros(1, np.uint32), np.zeros(1, np.uint64), st)
    one = [int(ob.replay_filter(np.zeros(1, np.uint32), seq[i:i + 1], st, auth[i:i + 1])[0]) for i in range(len(seq))]
    assert one == acc.tolist()


def test_sequence_numbers_of_captured_records_are_what_the_oracle_assigns():
    """Every protected record OpenSSL wrote carries the next number of its (direction, epoch), from 0, in write order:
    the rule orc_dtls_seq_assign restates (webrtc-dtls conn.rs fetch_add per record).  Both directions as two sessions."""
    fx = fixture()
    recs = [r for r in fx["records"] if r["epoch"] == 1]
    sess = np.array([0 if r["dir"] == "c2s" else 1 for r in recs], np.uint32)
    nxt = np.zeros(2, np.uint64)
    seq, ok = ob.seq_assign(sess, nxt)
    assert ok.all() and seq.tolist() == [r["seq"] for r in recs]
    assert nxt.tolist() == [int((sess == 0).sum()), int((sess == 1).sum())]
    # and the replay capture: the client's application records are numbered 1..N behind its Finished (0)
    rfx, rseq, _, _ = replay_fixture()
    nxt = np.zeros(1, np.uint64)
    seq, ok = ob.seq_assign(np.zeros(int(rseq.max()) + 1, np.uint32), nxt)
    assert set(rseq.tolist()) <= set(seq.tolist()) and seq[0] == 0


def test_seq_assign_oracle_semantics():
    rng = np.random.default_rng(11)
    S, n = 9, 5000
    sess = rng.integers(0, S + 2, n).astype(np.uint32)               # S, S + 1: not a session of the table
    nxt = rng.integers(0, 1000, S).astype(np.uint64)
    nxt[3] = (1 << 48) - 3                                           # three numbers left
    start = nxt.copy()
    seq, ok = ob.seq_assign(sess, nxt)
    for s in range(S + 2):
        idx = np.nonzero(sess == s)[0]
        if s >= S:
            assert not ok[idx].any() and not seq[idx].any()
            continue
        want = int(start[s]) + np.arange(len(idx))
        good = want <= (1 << 48) - 1
        assert np.array_equal(ok[idx], good.astype(np.uint8))
        assert np.array_equal(seq[idx][good], want[good].astype(np.uint64)) and not seq[idx][~good].any()
        assert int(nxt[s]) == min(int(start[s]) + len(idx), 1 << 48)
    assert (~ok[sess == 3].astype(bool)).sum() == (sess == 3).sum() - 3
