"""GPU: the CUDA record path against wire records captured from OpenSSL's DTLS 1.2 stack
(tests/golden/dtls_openssl_kat.json).  Same fixture as tests/test_dtls_openssl_pin.py, through the C ABI:
p2pt_dtls_gcm_open_batch / _seal_batch, p2pt_pipeline_seal / _open_demux, p2pt_dtls_datagram_split_batch.

Session 0 plays the client (local = client write key, remote = server write key), session 1 the server."""
import numpy as np
import pytest

from test_dtls_openssl_pin import fixture, h, raw_of

pytestmark = pytest.mark.gpu


def load_both_sides(ctx, fx):
    k = fx["keys"]
    ck, ci = h(k["client_write_key"]), h(k["client_write_iv"])
    sk, si = h(k["server_write_key"]), h(k["server_write_iv"])
    u8 = lambda *b: np.frombuffer(b"".join(b), np.uint8)
    ctx.sessions_load(u8(ck, sk), u8(ci, si), u8(sk, ck), u8(si, ci))


def batch(items):
    buf = np.frombuffer(b"".join(items), np.uint8)
    off = np.concatenate([[0], np.cumsum([len(x) for x in items])]).astype(np.uint64)
    return buf, off


@pytest.mark.parametrize("g", [0, 4, 8, 16, 32])
def test_open_and_reseal_every_captured_record(ctx, g):
    fx = fixture()
    load_both_sides(ctx, fx)
    recs = fx["records"]
    wire = [h(r["rec"]) for r in recs]
    plain = [h(r["plain"]) for r in recs]
    # the receiver of a c2s record is the server (session 1), of an s2c record the client (session 0)
    rx = np.array([1 if r["dir"] == "c2s" else 0 for r in recs], np.uint32)
    tx = 1 - rx
    ctx.set_option("gcm_group_width", g)
    try:
        buf, off = batch(wire)
        out, _, olen, st = ctx.dtls_open(rx, buf, off)
        assert (st == 0).all()
        for i, r in enumerate(recs):
            a = int(off[i])
            assert int(olen[i]) == 13 + len(plain[i])
            assert out[a:a + 13].tobytes() == wire[i][:13]
            assert out[a + 13:a + int(olen[i])].tobytes() == plain[i], (r["dir"], r["seq"])
        raws = [raw_of(w, p) for w, p in zip(wire, plain)]
        rbuf, roff = batch(raws)
        nonce = np.frombuffer(b"".join(w[13:21] for w in wire), np.uint8)
        got, goff, st = ctx.dtls_seal(tx, nonce, rbuf, roff)
        assert (st == 0).all() and np.array_equal(goff, off)
        assert got.tobytes() == buf.tobytes()
        # the other side's key must not open them
        _, _, _, st = ctx.dtls_open(tx, buf, off)
        assert (st == 3).all()
    finally:
        ctx.set_option("gcm_group_width", 0)


def test_fused_encode_seal_reproduces_the_captured_frames(ctx):
    """p2pt_pipeline_seal (serve.rs:269-276 + dc.send) on (type, sid, payload) + the captured epoch / seq / nonce
    gives the captured record, for every application_data record whose plaintext is a tunnel frame"""
    fx = fixture()
    load_both_sides(ctx, fx)
    valid = {1, 2, 3, 4, 10, 11, 12, 20, 21, 22, 99}
    recs = [r for r in fx["records"] if r["type"] == 23 and len(h(r["plain"])) >= 5 and h(r["plain"])[0] in valid]
    assert len(recs) >= 8
    plain = [h(r["plain"]) for r in recs]
    wire = [h(r["rec"]) for r in recs]
    pay, pay_off = batch([p[5:] for p in plain])
    if pay.size == 0:
        pay = np.zeros(1, np.uint8)
    tx = np.array([0 if r["dir"] == "c2s" else 1 for r in recs], np.uint32)
    rec, rec_off, st = ctx.pipeline_seal(tx, [r["epoch"] for r in recs], [r["seq"] for r in recs],
                                         np.frombuffer(b"".join(w[13:21] for w in wire), np.uint8),
                                         [p[0] for p in plain], [int.from_bytes(p[1:5], "big") for p in plain], pay, pay_off)
    assert (st == 0).all()
    assert rec.tobytes() == b"".join(wire)


def test_reader_loop_on_the_captured_server_records(ctx):
    """p2pt_pipeline_open_demux = proxy.rs:109-170 batched, on what the server really sent"""
    fx = fixture()
    load_both_sides(ctx, fx)
    recs = [r for r in fx["records"] if r["dir"] == "s2c" and r["type"] == 23]
    wire = [h(r["rec"]) for r in recs]
    buf, off = batch(wire)
    r = ctx.pipeline_open_demux(np.zeros(len(recs), np.uint32), buf, off, max_streams=16, gathered_cap=int(buf.size))
    d = r["desc"]
    got = [(int(x["msg_type"]), int(x["stream_id"]), int(x["status"])) for x in d]
    assert got[0] == (20, 42, 0) and got[1] == (21, 99, 0) and got[2] == (21, 42, 0) and got[5] == (22, 42, 0)
    # two raw (non-frame) 1200 / 1228-byte writes decode as whatever their first byte says; streams 42 and 99 are exact
    sid = list(r["uniq_sid"])
    s42, s99 = sid.index(42), sid.index(99)
    body = lambda s: r["gathered"][int(r["gathered_off"][s]):int(r["gathered_off"][s + 1])].tobytes()
    assert body(s99) == b'data: {"token": "hello"}\n\n'
    assert body(s42) == h(recs[2]["plain"])[5:]
    assert [int(x) for x in r["perm"][int(r["csr"][s42]):int(r["csr"][s42 + 1])]] == [0, 2, 5]


def test_datagram_split_then_open_on_captured_flights(ctx):
    """every captured datagram (handshake flights with 3-4 records each, then one record per datagram) through
    p2pt_dtls_datagram_split_batch; the protected pieces then open, ChangeCipherSpec passes through"""
    fx = fixture()
    load_both_sides(ctx, fx)
    dg = [h(d["bytes"]) for d in fx["datagrams"]]
    buf, off = batch(dg)
    rec_off, rec_dgram, st = ctx.dtls_datagram_split(buf, off)
    assert (st == 0).all()
    assert len(rec_dgram) == sum(d["n_records"] for d in fx["datagrams"])
    assert np.array_equal(np.bincount(rec_dgram, minlength=len(dg)), [d["n_records"] for d in fx["datagrams"]])
    keep = [j for j in range(len(rec_dgram)) if buf[int(rec_off[j])] == 20 or (int(buf[int(rec_off[j]) + 3]) << 8 | int(buf[int(rec_off[j]) + 4])) >= 1]
    pieces = [buf[int(rec_off[j]):int(rec_off[j + 1])].tobytes() for j in keep]
    rx = np.array([1 if fx["datagrams"][int(rec_dgram[j])]["dir"] == "c2s" else 0 for j in keep], np.uint32)
    pbuf, poff = batch(pieces)
    out, _, olen, st = ctx.dtls_open(rx, pbuf, poff)
    ccs = [i for i, p in enumerate(pieces) if p[0] == 20]
    assert len(ccs) == 2 and all(st[i] == 5 for i in ccs)
    prot = [i for i in range(len(pieces)) if i not in ccs]
    assert all(st[i] == 0 for i in prot)
    want = [h(r["plain"]) for r in fx["records"]]
    got = [out[int(poff[i]) + 13:int(poff[i]) + int(olen[i])].tobytes() for i in prot]
    assert got == want


def test_replay_window_decides_like_openssl(ctx):
    """p2pt_dtls_replay_filter_batch on the arrivals a live OpenSSL DTLS 1.2 server judged (replay_openssl_kat.json): in one
    batch, cut into batches of every size from 1 to 9, and with the same arrivals interleaved on 37 sessions."""
    from test_dtls_openssl_pin import fresh_window, replay_fixture
    fx, seq, auth, acc = replay_fixture()
    n = len(seq)
    zero = np.zeros(1, np.uint64)
    st = fresh_window()
    assert list(ctx.dtls_replay_filter(np.zeros(1, np.uint32), zero, st)) == [1]
    assert np.array_equal(ctx.dtls_replay_filter(np.zeros(n, np.uint32), seq, st, auth), acc)
    for step in range(1, 10):
        st = fresh_window()
        ctx.dtls_replay_filter(np.zeros(1, np.uint32), zero, st)
        got = np.concatenate([ctx.dtls_replay_filter(np.zeros(len(seq[i:i + step]), np.uint32), seq[i:i + step], st, auth[i:i + step])
                              for i in range(0, n, step)])
        assert np.array_equal(got, acc), step
    S = 37
    st = fresh_window(S)
    ctx.dtls_replay_filter(np.arange(S, dtype=np.uint32), np.zeros(S, np.uint64), st)
    sess = np.tile(np.arange(S, dtype=np.uint32), n)                 # arrival i of every session, then arrival i + 1 ...
    got = ctx.dtls_replay_filter(sess, np.repeat(seq, S), st, np.repeat(auth, S))
    assert np.array_equal(got.reshape(n, S), np.repeat(acc, S).reshape(n, S))
