"""GPU: DTLS record-layer bookkeeping (datagram split, anti-replay window) against the oracle."""
import numpy as np
import pytest

import oracle_binding as ob

pytestmark = pytest.mark.gpu


def _rec(rng, plen):
    return bytes([23, 0xFE, 0xFD, 0, 1]) + int(rng.integers(0, 2**40)).to_bytes(6, "big") + plen.to_bytes(2, "big") + \
        rng.integers(0, 256, plen, dtype=np.uint8).tobytes()


def test_datagram_split_matches_oracle(ctx):
    rng = np.random.default_rng(41)
    dgrams = []
    for i in range(3000):
        k = int(rng.integers(1, 5))
        d = b"".join(_rec(rng, int(rng.integers(0, 300))) for _ in range(k))
        if i % 97 == 0:
            d = d[:-1]                       # truncated last record: whole datagram dropped
        if i % 101 == 0:
            d = d + b"\x17\xfe\xfd"          # trailing garbage shorter than a header
        if i % 211 == 0:
            d = b""
        dgrams.append(d)
    buf = np.frombuffer(b"".join(dgrams), np.uint8)
    off = np.concatenate([[0], np.cumsum([len(d) for d in dgrams])]).astype(np.uint64)
    ro, rd, st = ctx.dtls_datagram_split(buf, off)
    ero, erd, est = ob.dtls_datagram_split(buf, off)
    assert np.array_equal(st, est) and st.sum() > 0
    assert np.array_equal(rd, erd) and np.array_equal(ro, ero)


def _fresh(S):
    return {"latest": np.zeros(S, np.uint64), "mask": np.zeros(S, np.uint64), "started": np.zeros(S, np.uint8)}


def test_replay_filter_matches_oracle_across_batches(ctx):
    rng = np.random.default_rng(42)
    S = 300
    st_gpu, st_cpu = _fresh(S), _fresh(S)
    nxt = np.zeros(S, np.int64)
    for batch in range(4):
        n = 20000
        session = rng.integers(0, S, n).astype(np.uint32)
        seq = np.zeros(n, np.uint64)
        for i in range(n):                                  # mostly increasing, with reordering, duplicates, old ones
            s = int(session[i]); r = rng.random()
            if r < 0.70:
                seq[i] = nxt[s]; nxt[s] += 1
            elif r < 0.85:
                seq[i] = max(0, nxt[s] - int(rng.integers(1, 70)))      # replay inside / at the edge of the window
            elif r < 0.92:
                seq[i] = max(0, nxt[s] - int(rng.integers(64, 400)))    # too old
            else:
                nxt[s] += int(rng.integers(1, 100)); seq[i] = nxt[s]; nxt[s] += 1   # a jump ahead
        auth = (rng.random(n) > 0.03).astype(np.uint8)
        if batch == 3:
            seq[:5] = [(1 << 48) - 1, 1 << 48, (1 << 48) - 1, 0, 5]     # the 48-bit limit
        got = ctx.dtls_replay_filter(session, seq, st_gpu, auth)
        exp = ob.replay_filter(session, seq, st_cpu, auth)
        assert np.array_equal(got, exp), batch
        for k in ("latest", "mask", "started"):
            assert np.array_equal(st_gpu[k], st_cpu[k]), (batch, k)
        assert 0.3 < got.mean() < 0.95
    # all-authentic shortcut and a session index outside the state table
    st_a, st_b = _fresh(4), _fresh(4)
    sess = np.array([0, 0, 9, 1, 0], np.uint32); seq = np.array([5, 5, 1, 0, 4], np.uint64)
    got = ctx.dtls_replay_filter(sess, seq, st_a)
    assert list(got) == [1, 0, 0, 1, 1]


def test_seq_assign_matches_oracle_across_batches(ctx):
    """p2pt_dtls_seq_assign_batch: consecutive numbers per session in arrival order, counters carried over batches, the
    48-bit limit, session indices outside the table, one session holding most of a batch, batches of one record."""
    rng = np.random.default_rng(77)
    S = 500
    nxt_gpu = rng.integers(0, 1 << 20, S).astype(np.uint64)
    nxt_gpu[7] = (1 << 48) - 40; nxt_gpu[8] = (1 << 48) - 1; nxt_gpu[9] = 1 << 48
    nxt_cpu = nxt_gpu.copy()
    for batch, n in enumerate([30000, 1, 2, 1024, 1025, 70000]):
        sess = rng.integers(0, S + 3, n).astype(np.uint32)
        if batch == 5:
            sess[rng.random(n) < 0.7] = 42                            # a hot session
            sess[:200] = 7
        got = ctx.dtls_seq_assign(sess, nxt_gpu)
        exp = ob.seq_assign(sess, nxt_cpu)
        assert np.array_equal(got[0], exp[0]) and np.array_equal(got[1], exp[1]), batch
        assert np.array_equal(nxt_gpu, nxt_cpu), batch
    assert int(nxt_gpu[7]) == 1 << 48 and int(nxt_gpu[9]) == 1 << 48
    assert ctx.dtls_seq_assign(np.zeros(0, np.uint32), nxt_gpu)[0].size == 0


def test_assigned_numbers_seal_open_and_pass_the_replay_window(ctx):
    """seq_assign -> fused encode+seal -> open -> replay filter, all through the C ABI: the sealed bytes equal the oracle's,
    every record comes back authentic with the number it was given and passes the window; replaying the batch is refused
    whole."""
    rng = np.random.default_rng(5)
    S, n = 64, 6000
    keys = rng.integers(0, 256, (S, 16), dtype=np.uint8); ivs = rng.integers(0, 256, (S, 4), dtype=np.uint8)
    ctx.sessions_load(keys, ivs, keys, ivs)
    sess = rng.integers(0, S, n).astype(np.uint32)
    nxt = np.zeros(S, np.uint64)
    seq, ok = ctx.dtls_seq_assign(sess, nxt)
    assert ok.all() and np.array_equal(nxt, np.bincount(sess, minlength=S).astype(np.uint64))
    lens = rng.integers(0, 300, n)
    w = {"n": n, "session": sess, "types": np.full(n, 21, np.uint8), "sids": rng.integers(1, 9, n).astype(np.uint32),
         "payload": rng.integers(0, 256, int(lens.sum()), dtype=np.uint8),
         "pay_off": np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64),
         "epoch": np.ones(n, np.uint16), "seq48": seq, "nonce8": rng.integers(0, 256, (n, 8), dtype=np.uint8)}
    rec, rec_off, status = ctx.pipeline_seal(sess, w["epoch"], seq, w["nonce8"], w["types"], w["sids"], w["payload"], w["pay_off"])
    assert not status.any()
    assert np.array_equal(rec, ob.pipeline_seal(w, keys, ivs))
    out, out_off, desc = ctx.pipeline_open(sess, rec, rec_off)
    assert not desc["status"].any() and np.array_equal(desc["stream_id"], w["sids"])
    wire_seq = np.array([int.from_bytes(bytes(rec[int(o) + 5:int(o) + 11]), "big") for o in rec_off[:-1]], np.uint64)
    assert np.array_equal(wire_seq, seq)
    st = _fresh(S)
    assert ctx.dtls_replay_filter(sess, wire_seq, st).all()
    assert not ctx.dtls_replay_filter(sess, wire_seq, st).any()
