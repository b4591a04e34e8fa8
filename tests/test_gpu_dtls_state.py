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
