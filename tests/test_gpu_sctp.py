"""GPU: SCTP DATA-chunk packetise / unpack / reassemble + CRC32c through the C ABI against the oracle,
and the wire-faithful chain frame -> SCTP packets -> DTLS records and back."""
import numpy as np
import pytest

import oracle_binding as ob

pytestmark = pytest.mark.gpu


def _msgs(rng, n, lo, hi):
    lens = rng.integers(lo, hi + 1, n)
    lens[:8] = [0, 1, 3, 4, 1199, 1200, 1201, 2400][:min(8, n)]
    return [rng.integers(0, 256, int(l), dtype=np.uint8).tobytes() for l in lens]


def _pack(msgs):
    off = np.zeros(len(msgs) + 1, np.uint64)
    np.cumsum([len(m) for m in msgs], out=off[1:])
    return np.frombuffer(b"".join(msgs), np.uint8), off


@pytest.mark.parametrize("dtls", [False, True])
def test_packetize_matches_oracle(ctx, dtls):
    rng = np.random.default_rng(31)
    msgs = _msgs(rng, 400, 0, 5000) + [rng.integers(0, 256, 65413, dtype=np.uint8).tobytes()]
    n = len(msgs)
    sp, dp = rng.integers(0, 65536, n), rng.integers(0, 65536, n)
    vt, sid, ssn = rng.integers(0, 2**32, n), rng.integers(0, 65536, n), rng.integers(0, 65536, n)
    tsn = rng.integers(0, 2**31, n); ppid = rng.choice([51, 53], n)
    kw = {"epoch": rng.integers(0, 65536, n), "seq48_first": rng.integers(0, 2**40, n)} if dtls else {}
    buf, off = _pack(msgs)
    got, goff = ctx.sctp_packetize(buf, off, sp, dp, vt, sid, ssn, ppid, tsn, **kw)
    exp, eoff = ob.sctp_packetize(msgs, sp, dp, vt, sid, ssn, ppid, tsn, **kw)
    assert np.array_equal(goff, eoff) and np.array_equal(got, exp)


def test_unpack_reassemble_roundtrip_shuffled_multi_session(ctx):
    rng = np.random.default_rng(32)
    S = 7
    per_sess = [_msgs(rng, 40, 0, 4000) for _ in range(S)]
    pkts, sess = [], []
    for s, msgs in enumerate(per_sess):
        lens = np.array([len(m) for m in msgs])
        frags = np.maximum(1, -(-lens // 1200))
        tsn0 = 5000 * (s + 1) + np.concatenate([[0], np.cumsum(frags)[:-1]])        # consecutive TSNs per session
        pk, off = ob.sctp_packetize(msgs, 10 + s, 20 + s, 99 + s, 1, np.arange(len(msgs)), 53, tsn0)
        for i in range(len(off) - 1):
            pkts.append(pk[int(off[i]):int(off[i + 1])].tobytes()); sess.append(s)
    order = rng.permutation(len(pkts))                                               # packets arrive in any order
    pkts = [pkts[i] for i in order]; sess = np.array(sess, np.uint32)[order]
    buf, off = _pack(pkts)
    r = ctx.sctp_unpack_reassemble(buf, off, session=sess)
    assert not r["info"]["status"].any() and not r["msg_status"].any()
    assert r["n_msgs"] == sum(len(m) for m in per_sess)
    flat = [m for msgs in per_sess for m in msgs]                                     # (session, TSN) order
    for i, m in enumerate(flat):
        assert r["msg"][int(r["msg_off"][i]):int(r["msg_off"][i + 1])].tobytes() == m, i
    einfo, _, _ = ob.sctp_unpack(buf, off)
    assert r["info"].tobytes() == einfo.tobytes()
    # msg_first points at the B fragment of each message
    assert np.all(r["info"]["flags"][r["msg_first"]] & 2)
    assert np.array_equal(r["info"]["ssn"][r["msg_first"]], np.concatenate([np.arange(40)] * S).astype(np.uint16))


def test_retransmitted_packets_are_dropped(ctx):
    """A second packet with the same (session, TSN) is a retransmission: webrtc-sctp's reassembly queue ignores
    a chunk whose TSN it already holds.  Duplicates of first, middle and last fragments - anywhere in the
    arrival order, in several sessions - must leave every message byte-identical and complete."""
    rng = np.random.default_rng(34)
    S = 5
    per_sess = [_msgs(rng, 30, 0, 5000) for _ in range(S)]
    pkts, sess = [], []
    for s, msgs in enumerate(per_sess):
        lens = np.array([len(m) for m in msgs])
        frags = np.maximum(1, -(-lens // 1200))
        tsn0 = 70000 * (s + 1) + np.concatenate([[0], np.cumsum(frags)[:-1]])
        pk, off = ob.sctp_packetize(msgs, 10 + s, 20 + s, 99 + s, 1, np.arange(len(msgs)), 53, tsn0)
        for i in range(len(off) - 1):
            pkts.append(pk[int(off[i]):int(off[i + 1])].tobytes()); sess.append(s)
    n0 = len(pkts)
    dup = rng.choice(n0, n0 // 3, replace=False)                  # a third of the packets arrive twice ...
    dup = np.concatenate([dup, dup[:20]])                          # ... and twenty of them three times
    pkts += [pkts[i] for i in dup]; sess += [sess[i] for i in dup]
    order = rng.permutation(len(pkts))
    pkts = [pkts[i] for i in order]; sess = np.array(sess, np.uint32)[order]
    buf, off = _pack(pkts)
    r = ctx.sctp_unpack_reassemble(buf, off, session=sess)
    assert not r["info"]["status"].any() and not r["msg_status"].any()
    assert r["n_msgs"] == sum(len(m) for m in per_sess)
    flat = [m for msgs in per_sess for m in msgs]
    for i, m in enumerate(flat):
        assert r["msg"][int(r["msg_off"][i]):int(r["msg_off"][i + 1])].tobytes() == m, i
    assert int(r["msg_off"][r["n_msgs"]]) == sum(len(m) for m in flat)            # no duplicate bytes anywhere
    # msg_first is the FIRST copy to arrive of each message's B fragment
    first_b = {}
    for pos in range(len(pkts)):
        inf = r["info"][pos]
        if inf["flags"] & 2:
            first_b.setdefault((int(sess[pos]), int(inf["tsn"])), pos)
    assert sorted(first_b.values()) == sorted(int(x) for x in r["msg_first"])


def test_tsn_wrap_around_inside_a_batch(ctx):
    """TSNs are serial numbers modulo 2^32 and the PEER picks where they start: messages whose fragments straddle
    0xFFFFFFFF -> 0 must reassemble like any other when the caller gives each packet's cumulative-ack base
    (webrtc-sctp orders with sna32).  Two sessions, one wrapping, shuffled arrival, a few retransmissions."""
    rng = np.random.default_rng(35)
    starts = [0xFFFFFFFF - 6, 123456]                             # session 0 wraps after seven fragments: inside its 7th message
    per_sess = [_msgs(rng, 12, 100, 6000) for _ in starts]
    pkts, sess, base = [], [], []
    for s, msgs in enumerate(per_sess):
        lens = np.array([len(m) for m in msgs])
        frags = np.maximum(1, -(-lens // 1200))
        tsn0 = (starts[s] + np.concatenate([[0], np.cumsum(frags)[:-1]])) & 0xFFFFFFFF
        pk, off = ob.sctp_packetize(msgs, 10 + s, 20 + s, 99 + s, 1, np.arange(len(msgs)), 53, tsn0)
        for i in range(len(off) - 1):
            pkts.append(pk[int(off[i]):int(off[i + 1])].tobytes()); sess.append(s); base.append((starts[s] - 1) & 0xFFFFFFFF)
    assert sum(1 for p in pkts[:len(pkts) // 2] if int.from_bytes(p[16:20], "big") < 1000) > 0      # the wrap really happens
    dup = rng.choice(len(pkts), 6, replace=False)
    pkts += [pkts[i] for i in dup]; sess += [sess[i] for i in dup]; base += [base[i] for i in dup]
    order = rng.permutation(len(pkts))
    pkts = [pkts[i] for i in order]; sess = np.array(sess, np.uint32)[order]; base = np.array(base, np.uint32)[order]
    buf, off = _pack(pkts)
    r = ctx.sctp_unpack_reassemble(buf, off, session=sess, tsn_base=base)
    assert not r["info"]["status"].any() and not r["msg_status"].any()
    flat = [m for msgs in per_sess for m in msgs]
    assert r["n_msgs"] == len(flat)
    for i, m in enumerate(flat):
        assert r["msg"][int(r["msg_off"][i]):int(r["msg_off"][i + 1])].tobytes() == m, i
    # without the base the wrapping session's first message is cut in two: the documented limit of raw ordering
    r2 = ctx.sctp_unpack_reassemble(buf, off, session=sess)
    assert (r2["msg_status"] == 9).any()


def test_unpack_flags_bad_packets_and_incomplete_messages(ctx):
    rng = np.random.default_rng(33)
    msgs = [rng.integers(0, 256, 3000, dtype=np.uint8).tobytes() for _ in range(6)]     # 3 fragments each
    pk, off = ob.sctp_packetize(msgs, 1, 2, 3, 0, np.arange(6), 53, 100 + 3 * np.arange(6))
    pkts = [pk[int(off[i]):int(off[i + 1])].tobytes() for i in range(18)]
    bad_crc = bytearray(pkts[4]); bad_crc[40] ^= 0x20; pkts[4] = bytes(bad_crc)       # message 1 loses its middle fragment
    pkts[9] = pkts[9][:-4]                                                             # message 3: first fragment truncated
    pkts.append(b"\x00" * 8)                                                           # shorter than a common header
    buf, off2 = _pack(pkts)
    r = ctx.sctp_unpack_reassemble(buf, off2)
    einfo, _, _ = ob.sctp_unpack(buf, off2)
    assert r["info"].tobytes() == einfo.tobytes()
    assert list(r["info"]["status"][[4, 9, 18]]) == [7, 7, 8]      # a truncated packet fails its checksum first
    # messages 0, 2, 4, 5 complete; message 1 incomplete (gap); message 3 has lost its B fragment: its two
    # remaining fragments are orphans that make the PREVIOUS chain inconsistent -> flagged on message 2's tail
    ok = [i for i in range(r["n_msgs"]) if r["msg_status"][i] == 0]
    got = [r["msg"][int(r["msg_off"][i]):int(r["msg_off"][i + 1])].tobytes() for i in ok]
    for m in (msgs[0], msgs[4], msgs[5]):
        assert m in got
    assert msgs[1] not in got and msgs[3] not in got
    assert 9 in set(int(x) for x in r["msg_status"])


def test_wire_chain_frames_sctp_dtls_and_back(ctx):
    """frame encode -> SCTP packetise (with DTLS headers) -> seal -> open -> unpack -> reassemble -> decode -> demux"""
    from p2p_llm_tunnel_b200 import workload
    rng = np.random.default_rng(34)
    S = 6
    keys, ivs = workload.sessions_keys(S, seed_shift=77)
    ctx.sessions_load(keys, ivs, keys, ivs)
    n = 300
    plens = rng.integers(0, 4000, n).astype(np.uint64); plens[:4] = [0, 1195, 1196, 3595]
    pay_off = np.zeros(n + 1, np.uint64); np.cumsum(plens, out=pay_off[1:])
    payload = rng.integers(0, 256, int(pay_off[-1]), dtype=np.uint8)
    sess = np.sort(rng.integers(0, S, n)).astype(np.uint32)                            # frames grouped by session
    sids = rng.integers(1, 5, n).astype(np.uint32)
    frames, f_off = ctx.frame_encode(np.full(n, 21, np.uint8), sids, payload, pay_off)
    flen = (f_off[1:] - f_off[:-1]).astype(np.int64)
    frags = np.maximum(1, -(-flen // 1200))
    first = np.concatenate([[0], np.cumsum(frags)[:-1]])
    # per-session running TSN / DTLS sequence numbers
    tsn0 = np.zeros(n, np.uint32); seq0 = np.zeros(n, np.uint64)
    run = {}
    for i in range(n):
        s = int(sess[i]); tsn0[i] = 1000 + run.get(s, 0); seq0[i] = run.get(s, 0); run[s] = run.get(s, 0) + int(frags[i])
    raw, raw_off = ctx.sctp_packetize(frames, f_off, 5000, 5000, 7 + sess, 1, np.arange(n) & 0xFFFF, 53, tsn0, epoch=1, seq48_first=seq0)
    npk = len(raw_off) - 1
    pk_sess = np.repeat(sess, frags)
    nonce = rng.integers(0, 256, 8 * npk, dtype=np.uint8)
    rec, rec_off, st = ctx.dtls_seal(pk_sess, nonce, raw, raw_off)
    assert not st.any()
    erec, _, _ = ob.dtls_seal(pk_sess, keys, ivs, nonce, raw, raw_off)
    assert np.array_equal(rec, erec)
    # network reorders packets
    order = rng.permutation(npk)
    recs = [rec[int(rec_off[i]):int(rec_off[i + 1])].tobytes() for i in order]
    rbuf, roff = _pack(recs)
    out, out_off, olen, st = ctx.dtls_open(pk_sess[order], rbuf, roff)
    assert not st.any()
    plain_off = np.concatenate([out_off[:-1], out_off[-1:]])
    # opened items are hdr13 || packet, each olen long inside a slot of the sealed size
    tight = np.concatenate([out[int(out_off[i]):int(out_off[i]) + int(olen[i])] for i in range(npk)])
    toff = np.zeros(npk + 1, np.uint64); np.cumsum(olen, out=toff[1:])
    r = ctx.sctp_unpack_reassemble(tight, toff, headroom=13, session=pk_sess[order])
    assert not r["info"]["status"].any() and not r["msg_status"].any() and r["n_msgs"] == n
    # messages come back ordered by (session, TSN) = original frame order (sess was sorted, TSNs ascending)
    assert np.array_equal(r["msg"], frames) and np.array_equal(r["msg_off"], f_off)
    desc = ctx.frame_decode(r["msg"], r["msg_off"])
    assert not desc["status"].any() and np.array_equal(desc["stream_id"], sids)
    d = ctx.frame_demux(desc, r["msg"], session=sess)
    e = ob.frame_demux(desc, r["msg"], sess)
    for key in ("uniq_sid", "uniq_session", "csr", "perm", "gathered_off", "gathered"):
        assert np.array_equal(d[key], e[key]), key
    assert np.array_equal(np.sort(d["gathered"]), np.sort(payload))
